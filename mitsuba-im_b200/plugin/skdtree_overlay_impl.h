// skdtree_overlay_impl.h — device scene from ShapeKDTree::getShapes() and the three interposed members (see skdtree_overlay.cpp).
#pragma once

MTS_NAMESPACE_BEGIN

namespace {

/// Geometry of the kd-tree's shapes -> libmtsb200 scene, in addShape order (= the order of the global primitive index).
Overlay *buildOverlay(const std::vector<const Shape *> &shapes) {
	if (!g_ctx) {
		if (mtsb_device_count() == 0) SLog(EError, "mtsb overlay: no CUDA device — this library has no CPU fallback");
		if (mtsb_ctx_create(0, &g_ctx) < 0) SLog(EError, "mtsb overlay: %s", mtsb_last_error());
	}
	Overlay *ov = new Overlay();
	if (mtsb_scene_create(g_ctx, &ov->scene) < 0) SLog(EError, "mtsb overlay: %s", mtsb_last_error());
	mtsb_bsdf_desc bd; memset(&bd, 0, sizeof(bd));
	bd.type = MTSB_BSDF_DIFFUSE;                 // intersection queries never look at it
	const int bid = mtsb_scene_add_bsdf(ov->scene, &bd);
	size_t prims = 0;
	for (size_t i = 0; i < shapes.size(); ++i) {
		const Shape *shape = shapes[i];
		int rc;
		if (shape->getClass()->derivesFrom(MTS_CLASS(TriMesh))) {
			const TriMesh *mesh = static_cast<const TriMesh *>(shape);
			rc = mtsb_scene_add_mesh(ov->scene, (const float *) mesh->getVertexPositions(), (const float *) mesh->getVertexNormals(),
				(const float *) mesh->getVertexTexcoords(), (const uint32_t *) mesh->getTriangles(),
				mesh->getVertexCount(), mesh->getTriangleCount(), bid, -1);
			prims += mesh->getTriangleCount();
		} else if (shape->getClass()->getName() == "Rectangle") {   // src/shapes/rectangle.cpp:79-84
			const Properties &props = shape->getProperties();
			const Matrix4x4 &mtx = props.getTransform("toWorld", Transform()).getMatrix();
			float m[16];
			for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) m[r * 4 + c] = (float) mtx(r, c);
			rc = mtsb_scene_add_rectangle(ov->scene, m, props.getBoolean("flipNormals", false) ? 1 : 0, bid, -1);
			prims += 1;
		} else if (shape->getClass()->getName() == "Disk") {        // src/shapes/disk.cpp:83-88
			const Properties &props = shape->getProperties();
			const Matrix4x4 &mtx = props.getTransform("toWorld", Transform()).getMatrix();
			float m[16];
			for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) m[r * 4 + c] = (float) mtx(r, c);
			rc = mtsb_scene_add_disk(ov->scene, m, props.getBoolean("flipNormals", false) ? 1 : 0, bid, -1);
			prims += 1;
		} else if (shape->getClass()->getName() == "Sphere") {      // src/shapes/sphere.cpp:108-131
			const Properties &props = shape->getProperties();
			Transform objectToWorld = Transform::translate(Vector(props.getPoint("center", Point(0.0f))));
			Float radius = props.getFloat("radius", 1.0f);
			if (props.hasProperty("toWorld")) {
				const Transform toWorld = props.getTransform("toWorld");
				const Float scale = toWorld(Vector(1, 0, 0)).length();
				objectToWorld = toWorld * Transform::scale(Vector(1 / scale)) * objectToWorld;
				radius *= scale;
			}
			const Matrix4x4 &mtx = objectToWorld.getMatrix();
			float m[16];
			for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) m[r * 4 + c] = (float) mtx(r, c);
			rc = mtsb_scene_add_sphere(ov->scene, m, (float) radius, props.getBoolean("flipNormals", false) ? 1 : 0, bid, -1);
			prims += 1;
		} else {
			SLog(EError, "mtsb overlay: shape \"%s\" is outside the GPU subset (triangle meshes, rectangle, disk, sphere)", shape->getClass()->getName().c_str());
			rc = -1;
		}
		if (rc < 0) SLog(EError, "mtsb overlay: %s", mtsb_last_error());
	}
	if (mtsb_scene_commit(ov->scene, MTSB_BUILDER_PLOC) < 0) SLog(EError, "mtsb overlay: %s", mtsb_last_error());
	double st[8]; mtsb_scene_build_stats(ov->scene, st);
	SLog(EInfo, "mtsb overlay: ShapeKDTree::rayIntersect x3 now run on the GPU — BVH8 over %zu primitives of %zu shapes built in %.1f ms",
		prims, shapes.size(), st[0]);
	ov->service = std::thread(serviceLoop, ov);
	return ov;
}

inline Overlay *overlayFor(const ShapeKDTree *kd) {
	static thread_local const ShapeKDTree *lastKd = NULL;
	static thread_local Overlay *lastOv = NULL;
	if (lastKd == kd) return lastOv;
	std::lock_guard<std::mutex> lock(g_mutex);
	Overlay *&ov = g_overlays[kd];
	if (!ov) ov = buildOverlay(kd->getShapes());
	lastKd = kd; lastOv = ov;
	return ov;
}

struct OverlayReport {      // batch statistics at process exit
	~OverlayReport() {
		for (std::map<const ShapeKDTree *, Overlay *>::iterator it = g_overlays.begin(); it != g_overlays.end(); ++it) {
			Overlay *ov = it->second;
			ov->stop.store(true); ov->pending.fetch_add(1); futexWake(&ov->pending, 1);
			if (ov->service.joinable()) ov->service.join();
			static const char *names[Q_COUNT] = { "closest", "light", "shadow" };
			for (int k = 0; k < Q_COUNT; ++k)
				if (ov->q[k].batches)
					fprintf(stderr, "mtsb overlay: %s queue: %llu rays in %llu batches (%.1f rays per batch)\n", names[k], ov->q[k].requests,
						ov->q[k].batches, (double) ov->q[k].requests / (double) ov->q[k].batches);
		}
	}
} g_report;

} // namespace

// ---- the interposed members ----------------------------------------------------------------------------------------------

/// src/librender/skdtree.cpp:68-110.  With the three query overloads served by the device, nothing reads the SAH kd-tree or
/// the TriAccel array any more, so the CPU build — seconds at 1 M triangles, minutes at 10 M (gkdtree.h:958-1263) — is
/// left out.  What the rest of the renderer still needs from build() is kept: the prefix sums of m_shapeMap (global
/// primitive index -> shape) and the bounding boxes (Scene::initialize takes getAABB(), scene.cpp:353-355; the sensor's
/// clip planes and the environment emitter's sphere derive from it): the tight box of all primitives and its enlargement
/// exactly as buildInternal computes them (gkdtree.h:997-1002, 1213-1220).  The node array becomes the single empty leaf
/// buildInternal makes for a scene without geometry (:972-977), so isBuilt() holds and the destructor frees it; a CPU-side
/// packet query (rayIntersectPacket), which is not interposed, finds nothing in it.  MTSB_OVERLAY_KDBUILD=1 keeps the
/// reference's build (it then runs from libmitsuba-render.so's own copy of this symbol via dlsym(RTLD_NEXT)).
void ShapeKDTree::build() {
	const char *keep = getenv("MTSB_OVERLAY_KDBUILD");
	if (keep && atoi(keep) != 0) {
		typedef void (ShapeKDTree::*BuildFn)();
		void *sym = dlsym(RTLD_NEXT, "_ZN7mitsuba11ShapeKDTree5buildEv");
		if (!sym) SLog(EError, "mtsb overlay: the reference's ShapeKDTree::build is not reachable");
		union { void *p; BuildFn f; } u; memset(&u, 0, sizeof(u)); u.p = sym;
		(this->*u.f)();
		return;
	}
	ref<Timer> timer = new Timer();
	for (size_t i = 1; i < m_shapeMap.size(); ++i)
		m_shapeMap[i] += m_shapeMap[i - 1];
	AABB aabb;
	aabb.reset();
	for (size_t i = 0; i < m_shapes.size(); ++i) {
		const Shape *shape = m_shapes[i];
		if (m_triangleFlag[i]) {
			const TriMesh *mesh = static_cast<const TriMesh *>(shape);
			const Triangle *tris = mesh->getTriangles();
			const Point *pos = mesh->getVertexPositions();
			for (size_t j = 0; j < mesh->getTriangleCount(); ++j)        // per-triangle boxes: unreferenced vertices do not count
				for (int k = 0; k < 3; ++k) aabb.expandBy(pos[tris[j].idx[k]]);
		} else {
			aabb.expandBy(shape->getAABB());
		}
	}
	m_tightAABB = aabb;
	const Float eps = MTS_KD_AABB_EPSILON;
	aabb.min -= (aabb.max - aabb.min) * eps + Vector(eps);
	aabb.max += (aabb.max - aabb.min) * eps + Vector(eps);
	m_aabb = aabb;
	m_nodes = static_cast<KDNode *>(allocAligned(sizeof(KDNode) * 2)) + 1;
	m_nodes[0].initLeafNode(0, 0);
	SLog(EInfo, "mtsb overlay: ShapeKDTree::build — bounds of %u primitives in %i ms, the CPU kd-tree is not built (queries run on the GPU)",
		(unsigned) getPrimitiveCount(), timer->getMilliseconds());
}


/// src/librender/skdtree.cpp:112-142.  Scene-box clip, floored adaptive epsilon and traversal happen on the device
/// (csrc/trace.cuh: clip_ray + trace_persistent); the record is filled by the reference's own inline code.
bool ShapeKDTree::rayIntersect(const Ray &ray, Intersection &its) const {
	Overlay *ov = overlayFor(this);
	++ovRays;
	its.t = std::numeric_limits<Float>::infinity();
	Slot &s = submit(ov, Q_CLOSEST, ray);
	const mtsb_hit h = s.out.hit;
	release(s);
	if (h.prim == MTSB_NO_HIT) return false;
	// global primitive index -> (shape, primitive in shape) through the prefix sums build() made of m_shapeMap
	const SizeType shapeIndex = (SizeType) (std::upper_bound(m_shapeMap.begin(), m_shapeMap.end(), (SizeType) h.prim) - m_shapeMap.begin()) - 1;
	uint8_t temp[MTS_KD_INTERSECTION_TEMP];
	IntersectionCache *cache = reinterpret_cast<IntersectionCache *>(temp);
	cache->shapeIndex = shapeIndex;
	cache->primIndex = (SizeType) h.prim - m_shapeMap[shapeIndex];
	// triangles: barycentrics; rectangle: the object-space x, y its own rayIntersect leaves at the same offset (rectangle.cpp:139-143)
	cache->u = h.u; cache->v = h.v;
	its.t = h.t;
	fillIntersectionRecord<true>(ray, temp, its);
	return true;
}

/// src/librender/skdtree.cpp:144-204 (the device computes the record: csrc/trace.cu MODE_LIGHT).
bool ShapeKDTree::rayIntersect(const Ray &ray, Float &t, ConstShapePtr &shape, Normal &n, Point2 &uv) const {
	Overlay *ov = overlayFor(this);
	++ovLight;
	t = std::numeric_limits<Float>::infinity();
	Slot &s = submit(ov, Q_LIGHT, ray);
	const mtsb_light_hit h = s.out.light;
	release(s);
	if (h.shape == MTSB_NO_HIT) return false;
	t = h.t;
	shape = m_shapes[h.shape];
	n = Normal(h.n[0], h.n[1], h.n[2]);
	uv = Point2(h.uv[0], h.uv[1]);
	return true;
}

/// src/librender/skdtree.cpp:207-226
bool ShapeKDTree::rayIntersect(const Ray &ray) const {
	Overlay *ov = overlayFor(this);
	++ovShadow;
	Slot &s = submit(ov, Q_ANY, ray);
	const bool hit = s.out.any != 0;
	release(s);
	return hit;
}

MTS_NAMESPACE_END
