// extract_scene.h — geometry, emitters, sensor and film of a loaded Scene -> C ABI calls (included by extract.h).
#pragma once

MTS_NAMESPACE_BEGIN

/// The scene's participating media in a fixed order (index = position): Scene::getMedia(), then whatever only the sensor
/// or a shape refers to.  uploadScene and the camera-medium lookup of the render both go through this list.
static std::vector<const Medium *> collectMedia(const Scene *scene) {
	std::vector<const Medium *> media;
	struct Add { static void one(std::vector<const Medium *> &v, const Medium *m) {
		if (m && std::find(v.begin(), v.end(), m) == v.end()) v.push_back(m); } };
	const ref_vector<Medium> &declared = scene->getMedia();
	for (size_t i = 0; i < declared.size(); ++i) Add::one(media, declared[i].get());
	if (scene->getSensor()) Add::one(media, scene->getSensor()->getMedium());
	const ref_vector<Shape> &shapes = scene->getShapes();
	for (size_t i = 0; i < shapes.size(); ++i) { Add::one(media, shapes[i]->getInteriorMedium()); Add::one(media, shapes[i]->getExteriorMedium()); }
	return media;
}
static int mediumIndex(const std::vector<const Medium *> &media, const Medium *m) {
	if (!m) return -1;
	return (int) (std::find(media.begin(), media.end(), m) - media.begin());
}

/// HomogeneousMedium -> mtsb_medium_desc.  The sampling parameters are private to the plugin class, so they are re-derived
/// from its Properties exactly as its constructor does (src/medium/homogeneous.cpp:156-226).
static void describeMedium(const Medium *medium, mtsb_medium_desc &d) {
	memset(&d, 0, sizeof(d));
	const std::string cls = medium->getClass()->getName();
	if (cls != "HomogeneousMedium")
		SLog(EError, "gpupath: medium \"%s\" is outside the GPU subset (homogeneous)", cls.c_str());
	const Properties &props = medium->getProperties();
	Spectrum sigmaA = medium->getSigmaA(), sigmaS = medium->getSigmaS(), sigmaT = medium->getSigmaT();
	const std::string strategy = props.getString("strategy", "balance");
	Float weight = props.getFloat("mediumSamplingWeight", -1);
	if (weight == -1) {
		for (int i = 0; i < SPECTRUM_SAMPLES; ++i) {
			const Float albedo = sigmaS[i] / sigmaT[i];
			if (albedo > weight && sigmaT[i] != 0) weight = albedo;
		}
		if (weight > 0) weight = std::max(weight, (Float) 0.5f);
	}
	d.medium_sampling_weight = (float) weight;
	if (strategy == "balance") d.strategy = 0;
	else if (strategy == "single") {
		int channel = 0;
		Float smallest = std::numeric_limits<Float>::infinity();
		for (int i = 0; i < SPECTRUM_SAMPLES; ++i) if (sigmaT[i] < smallest) { smallest = sigmaT[i]; channel = i; }
		channel = props.getInteger("channel", channel);
		d.strategy = 1; d.sampling_density = (float) sigmaT[channel];
		if (props.getBoolean("monochromatic", false)) SLog(EError, "gpupath: monochromatic media are outside the GPU subset");
	} else if (strategy == "manual") { d.strategy = 1; d.sampling_density = (float) props.getFloat("samplingDensity"); }
	else SLog(EError, "gpupath: medium sampling strategy \"%s\" is outside the GPU subset (balance, single, manual)", strategy.c_str());
	toRGB(sigmaA, d.sigma_a); toRGB(sigmaS, d.sigma_s);
	const PhaseFunction *phase = medium->getPhaseFunction();
	const std::string pcls = phase->getClass()->getName();
	if (pcls == "IsotropicPhaseFunction") d.phase = 0;
	else if (pcls == "HGPhaseFunction") { d.phase = 1; d.g = (float) phase->getProperties().getFloat("g", 0.8f); }
	else SLog(EError, "gpupath: phase function \"%s\" is outside the GPU subset (isotropic, hg)", pcls.c_str());
}

/// The `envmap` emitter keeps its MIP map private, so the plugin loads the image again and builds the pyramid exactly as the
/// emitter's constructor does (src/emitters/envmap.cpp:152-178), with the reference's own Bitmap / MIPMap classes: RGB,
/// half-precision storage, 2-lobed Lanczos downsampling, repeat / clamp boundaries.
static void uploadEnvmap(mtsb_scene *out, int id, const Emitter *e) {
	typedef TSpectrum<half, SPECTRUM_SAMPLES> SpectrumHalf;
	typedef TMIPMap<Spectrum, SpectrumHalf> MIPMap;
	const Properties &props = e->getProperties();
	if (!props.hasProperty("filename")) SLog(EError, "gpupath: envmap without a file name (raw bitmap data) is outside the GPU subset");
	fs::pathstr filename = Thread::getThread()->getFileResolver()->resolve(fs::pathstr(props.getString("filename")));
	ref<FileStream> stream = new FileStream(filename, FileStream::EReadOnly);
	ref<Bitmap> bitmap = new Bitmap(Bitmap::EAuto, stream);
	const Float gamma = props.getFloat("gamma", 0);
	if (gamma != 0) bitmap->setGamma(gamma);
	Properties rfilterProps("lanczos");
	rfilterProps.setInteger("lobes", 2);
	ref<ReconstructionFilter> rfilter = static_cast<ReconstructionFilter *>(PluginManager::getInstance()->createObject(
		MTS_CLASS(ReconstructionFilter), rfilterProps));
	rfilter->configure();
	MIPMap *mip = new MIPMap(bitmap, Bitmap::ERGB, Bitmap::EFloat, rfilter, ReconstructionFilter::ERepeat, ReconstructionFilter::EClamp,
		EEWA, 10.0f, fs::pathstr(), 0, std::numeric_limits<Float>::infinity(), Spectrum::EIlluminant);
	const int levels = mip->getLevels();
	std::vector<int32_t> sizes(2 * levels);
	std::vector<std::vector<float> > data(levels);
	std::vector<const float *> ptrs(levels);
	for (int l = 0; l < levels; ++l) {
		const MIPMap::Array2DType &a = mip->getArray(l);
		sizes[2 * l] = a.getSize().x; sizes[2 * l + 1] = a.getSize().y;
		data[l].resize((size_t) a.getSize().x * a.getSize().y * 3);
		float *dst = data[l].data();
		for (int y = 0; y < a.getSize().y; ++y)
			for (int x = 0; x < a.getSize().x; ++x, dst += 3) toRGB(Spectrum(a(x, y)), dst);
		ptrs[l] = data[l].data();
	}
	mtsb_envmap_desc ed; memset(&ed, 0, sizeof(ed));
	ed.levels = levels; ed.sizes = sizes.data(); ed.data = ptrs.data();
	ed.scale = (float) props.getFloat("scale", 1.0f);
	matrixToArray(e->getWorldTransform()->eval(0).getMatrix(), ed.to_world);
	MTSB_CALL(mtsb_scene_set_envmap(out, id, &ed));
	delete mip;
}

/// Uploads shapes / BSDFs / emitters in Scene order and commits (BVH build on the device).
/// A mesh that came out of the `ply` or `serialized` plugin is decoded again, on the device, from its file (mtsb_scene_add_ply /
/// mtsb_scene_add_serialized) instead of being copied out of the TriMesh's host arrays — when the file is one the device
/// decoder takes and its header agrees with the mesh the reference's loader produced.  Returns the shape index or -1 (caller
/// falls back to mtsb_scene_add_mesh).
static int addMeshFromFile(mtsb_scene *out, const TriMesh *mesh, int bid) {
	const std::string cls = mesh->getClass()->getName();
	const bool ply = cls == "PLYLoader", ser = cls == "SerializedMesh";
	if (!ply && !ser) return -1;
	const Properties &props = mesh->getProperties();
	if (props.hasProperty("maxSmoothAngle") || !props.hasProperty("filename")) return -1;     // rebuildTopology changes the arrays
	if (const char *e = getenv("MTSB_PLY_ON_DEVICE")) if (atoi(e) == 0) return -1;
	try {
		const fs::pathstr resolved = Thread::getThread()->getFileResolver()->resolve(fs::pathstr(props.getString("filename")));     // ply.cpp:76-77
		const fs::path path = fs::decode_pathstr(resolved);
		if (!fs::exists(path)) return -1;
		ref<MemoryMappedFile> file = new MemoryMappedFile(resolved, true);
		const int shapeIndex = ser ? props.getInteger("shapeIndex", 0) : 0;
		uint32_t nv = 0, nt = 0;
		const int probe = ply ? mtsb_ply_probe(file->getData(), file->getSize(), &nv, &nt)
		                      : mtsb_serialized_probe(file->getData(), file->getSize(), shapeIndex, &nv, &nt);
		if (probe != 0) return -1;
		if (nv != mesh->getVertexCount() || nt != mesh->getTriangleCount()) return -1;
		const Transform toWorld = props.getTransform("toWorld", Transform());
		float m[16], inv[16];
		matrixToArray(toWorld.getMatrix(), m); matrixToArray(toWorld.getInverseMatrix(), inv);
		const uint32_t flags = (props.getBoolean("faceNormals", false) ? MTSB_PLY_FACE_NORMALS : 0u) | (props.getBoolean("flipNormals", false) ? MTSB_PLY_FLIP_NORMALS : 0u);
		const int rc = ply ? mtsb_scene_add_ply(out, file->getData(), file->getSize(), m, inv, flags, bid, NULL, NULL)
		                   : mtsb_scene_add_serialized(out, file->getData(), file->getSize(), shapeIndex, m, inv, flags, bid, NULL, NULL);
		if (rc >= 0) SLog(EInfo, "gpupath: \"%s\" decoded on the device (%u triangles)", path.filename().string().c_str(), nt);
		return rc;
	} catch (const std::exception &) {
		return -1;
	}
}

static mtsb_scene *uploadScene(mtsb_ctx *ctx, const Scene *scene, int builder) {
	mtsb_scene *out = NULL;
	MTSB_CALL(mtsb_scene_create(ctx, &out));
	// emitters first, in Scene::getEmitters() order (that order defines the selection CDF, scene.cpp:381-386)
	const ref_vector<Emitter> &emitters = scene->getEmitters();
	std::map<const Emitter *, int> emitterIds;
	for (size_t i = 0; i < emitters.size(); ++i) {
		const Emitter *e = emitters[i].get();
		const std::string cls = e->getClass()->getName();
		const Properties &props = e->getProperties();
		mtsb_emitter_desc d; memset(&d, 0, sizeof(d));
		d.sampling_weight = (float) e->getSamplingWeight();
		if (cls == "AreaLight") {                                   // src/emitters/area.cpp:64-73
			d.type = MTSB_EMITTER_AREA;
			toRGB(props.getSpectrum("radiance", Spectrum::getD65()), d.radiance);
		} else if (cls == "PointEmitter") {                         // src/emitters/point.cpp:52-61
			d.type = MTSB_EMITTER_POINT;
			toRGB(props.getSpectrum("intensity", Spectrum::getD65()), d.radiance);
			const Point p = e->getWorldTransform()->eval(0).transformAffine(Point(0.0f));
			d.pos[0] = (float) p.x; d.pos[1] = (float) p.y; d.pos[2] = (float) p.z;
		} else if (cls == "ConstantBackgroundEmitter") {            // src/emitters/constant.cpp:44-52
			d.type = MTSB_EMITTER_CONSTANT;
			toRGB(props.getSpectrum("radiance", Spectrum::getD65()), d.radiance);
		} else if (cls == "EnvironmentMap") {                       // src/emitters/envmap.cpp:98-185
			d.type = MTSB_EMITTER_ENVMAP;
		} else {
			SLog(EError, "gpupath: emitter \"%s\" is outside the GPU subset (area, point, constant, envmap)", cls.c_str());
		}
		int id = mtsb_scene_add_emitter(out, &d);
		if (id < 0) SLog(EError, "libmtsb200: %s", mtsb_last_error());
		emitterIds[e] = id;
		if (d.type == MTSB_EMITTER_ENVMAP) uploadEnvmap(out, id, e);
	}
	const std::vector<const Medium *> media = collectMedia(scene);
	for (size_t i = 0; i < media.size(); ++i) {
		mtsb_medium_desc md; describeMedium(media[i], md);
		if (mtsb_scene_add_medium(out, &md) < 0) SLog(EError, "libmtsb200: %s", mtsb_last_error());
	}
	const ref_vector<Shape> &shapes = scene->getShapes();
	std::map<const BSDF *, int> bsdfIds;
	for (size_t i = 0; i < shapes.size(); ++i) {
		const Shape *shape = shapes[i].get();
		const BSDF *bsdf = shape->getBSDF();
		if (!bsdf) SLog(EError, "gpupath: shape %d has no BSDF", (int) i);
		if (shape->hasSubsurface())
			SLog(EError, "gpupath: subsurface scattering is outside the GPU subset");
		int bid;
		if (bsdfIds.count(bsdf)) bid = bsdfIds[bsdf];
		else {
			mtsb_bsdf_desc bd;
			std::vector<float> table;
			TextureData tex;
			if (!describeBSDF(bsdf, bd, &table, &tex))
				SLog(EError, "gpupath: BSDF \"%s\" is outside the GPU subset (diffuse, dielectric, plastic, roughplastic, conductor, roughconductor, roughdielectric, null, twosided)",
					bsdf->getClass()->getName().c_str());
			bid = mtsb_scene_add_bsdf(out, &bd);
			if (bid < 0) SLog(EError, "libmtsb200: %s", mtsb_last_error());
			if (!table.empty()) MTSB_CALL(mtsb_scene_set_bsdf_table(out, bid, table.data(), table.size()));
			if (tex.present) {
				std::vector<const float *> ptrs(tex.levels.size());
				for (size_t l = 0; l < tex.levels.size(); ++l) ptrs[l] = tex.levels[l].data();
				mtsb_texture_desc td; memset(&td, 0, sizeof(td));
				td.levels = (int32_t) tex.levels.size(); td.sizes = tex.sizes.data(); td.data = ptrs.data();
				td.wrap_u = tex.wrapU; td.wrap_v = tex.wrapV; td.filter_type = tex.filterType; td.max_anisotropy = tex.maxAnisotropy;
				td.uv_scale[0] = tex.uvScale[0]; td.uv_scale[1] = tex.uvScale[1]; td.uv_offset[0] = tex.uvOffset[0]; td.uv_offset[1] = tex.uvOffset[1];
				const int tid = mtsb_scene_add_texture(out, &td);
				if (tid < 0) SLog(EError, "libmtsb200: %s", mtsb_last_error());
				MTSB_CALL(mtsb_scene_set_bsdf_texture(out, bid, tid));
			}
			bsdfIds[bsdf] = bid;
		}
		int eid = -1;
		if (shape->isEmitter()) {
			std::map<const Emitter *, int>::const_iterator it = emitterIds.find(shape->getEmitter());
			if (it == emitterIds.end()) SLog(EError, "gpupath: shape %d carries an emitter the scene does not list", (int) i);
			eid = it->second;
		}
		int rc;
		if (shape->getClass()->derivesFrom(MTS_CLASS(TriMesh))) {
			const TriMesh *mesh = static_cast<const TriMesh *>(shape);
			rc = eid < 0 ? addMeshFromFile(out, mesh, bid) : -1;
			// Point / Normal are 3 packed floats, Point2 2 floats, Triangle 3 uint32 (include/mitsuba/core/triangle.h)
			if (rc < 0)
				rc = mtsb_scene_add_mesh(out, (const float *) mesh->getVertexPositions(), (const float *) mesh->getVertexNormals(),
					(const float *) mesh->getVertexTexcoords(), (const uint32_t *) mesh->getTriangles(),
					mesh->getVertexCount(), mesh->getTriangleCount(), bid, eid);
		} else if (shape->getClass()->getName() == "Rectangle") {   // src/shapes/rectangle.cpp:79-84
			const Properties &props = shape->getProperties();
			float m[16]; matrixToArray(props.getTransform("toWorld", Transform()).getMatrix(), m);
			rc = mtsb_scene_add_rectangle(out, m, props.getBoolean("flipNormals", false) ? 1 : 0, bid, eid);
		} else if (shape->getClass()->getName() == "Disk") {        // src/shapes/disk.cpp:83-88
			const Properties &props = shape->getProperties();
			float m[16]; matrixToArray(props.getTransform("toWorld", Transform()).getMatrix(), m);
			rc = mtsb_scene_add_disk(out, m, props.getBoolean("flipNormals", false) ? 1 : 0, bid, eid);
		} else if (shape->getClass()->getName() == "Sphere") {      // src/shapes/sphere.cpp:108-131
			const Properties &props = shape->getProperties();
			Transform objectToWorld = Transform::translate(Vector(props.getPoint("center", Point(0.0f))));
			Float radius = props.getFloat("radius", 1.0f);
			if (props.hasProperty("toWorld")) {
				const Transform toWorld = props.getTransform("toWorld");
				const Float scale = toWorld(Vector(1, 0, 0)).length();
				objectToWorld = toWorld * Transform::scale(Vector(1 / scale)) * objectToWorld;     // the scale moves into the radius
				radius *= scale;
			}
			float m[16]; matrixToArray(objectToWorld.getMatrix(), m);
			rc = mtsb_scene_add_sphere(out, m, (float) radius, props.getBoolean("flipNormals", false) ? 1 : 0, bid, eid);
		} else {
			SLog(EError, "gpupath: shape \"%s\" is outside the GPU subset (triangle meshes, rectangle, disk, sphere)", shape->getClass()->getName().c_str());
			rc = -1;
		}
		if (rc < 0) SLog(EError, "libmtsb200: %s", mtsb_last_error());
		if (shape->isMediumTransition())
			MTSB_CALL(mtsb_scene_set_shape_media(out, (int) i, mediumIndex(media, shape->getInteriorMedium()), mediumIndex(media, shape->getExteriorMedium())));
	}
	if (scene->getSensor()) {       // the part of Scene::getAABB() the environment emitter saw beside the geometry (scene.cpp:394-404)
		const AABB sa = scene->getSensor()->getAABB();
		if (sa.isValid()) {
			const float lo[3] = { (float) sa.min.x, (float) sa.min.y, (float) sa.min.z }, hi[3] = { (float) sa.max.x, (float) sa.max.y, (float) sa.max.z };
			MTSB_CALL(mtsb_scene_expand_bounds(out, lo, hi));
		}
	}
	MTSB_CALL(mtsb_scene_commit(out, builder));
	for (std::map<const BSDF *, int>::const_iterator it = bsdfIds.begin(); it != bsdfIds.end(); ++it)
		selfCheckBSDF(out, it->second, it->first);
	return out;
}

/// PerspectiveCameraImpl::configure (src/sensors/perspective.cpp:126-180) restated with the reference's own Transform class.
static void describeCamera(const Sensor *sensor, mtsb_camera &cam) {
	const std::string sensorCls = sensor->getClass()->getName();
	const bool thinlens = sensorCls == "ThinLens";
	if (!sensor->getClass()->derivesFrom(MTS_CLASS(PerspectiveCamera)) || (sensor->needsApertureSample() && !thinlens) || sensor->needsTimeSample())
		SLog(EError, "gpupath: sensor \"%s\" is outside the GPU subset (perspective, thinlens)", sensorCls.c_str());
	const PerspectiveCamera *pc = static_cast<const PerspectiveCamera *>(sensor);
	const Film *film = sensor->getFilm();
	const Vector2i filmSize = film->getSize(), cropSize = film->getCropSize();
	const Point2i cropOffset = film->getCropOffset();
	const Float aspect = pc->getAspect();
	Vector2 relSize((Float) cropSize.x / (Float) filmSize.x, (Float) cropSize.y / (Float) filmSize.y);
	Point2 relOffset((Float) cropOffset.x / (Float) filmSize.x, (Float) cropOffset.y / (Float) filmSize.y);
	Transform cameraToSample =
		  Transform::scale(Vector(1.0f / relSize.x, 1.0f / relSize.y, 1.0f))
		* Transform::translate(Vector(-relOffset.x, -relOffset.y, 0.0f))
		* Transform::scale(Vector(-0.5f, -0.5f * aspect, 1.0f))
		* Transform::translate(Vector(-1.0f, -1.0f / aspect, 0.0f))
		* Transform::perspective(pc->getXFov(), pc->getNearClip(), pc->getFarClip());
	matrixToArray(cameraToSample.inverse().getMatrix(), cam.sample_to_camera);
	matrixToArray(sensor->getWorldTransform()->eval(0).getMatrix(), cam.camera_to_world);
	cam.near_clip = (float) pc->getNearClip(); cam.far_clip = (float) pc->getFarClip();
	cam.inv_resolution[0] = 1.0f / cropSize.x; cam.inv_resolution[1] = 1.0f / cropSize.y;
	cam.aperture_radius = 0.0f; cam.focus_distance = (float) pc->getFocusDistance();
	if (thinlens) {                                                 // src/sensors/thinlens.cpp:124-138
		Float r = sensor->getProperties().getFloat("apertureRadius");
		cam.aperture_radius = (float) (r == 0 ? Epsilon : r);
	}
}

/// The float layout of an ImageBlock + its filter table (include/mitsuba/core/rfilter.h:76-77).
static void describeFilm(const ImageBlock &block, const ReconstructionFilter *rf, const Film *film, mtsb_film_desc &fd) {
	memset(&fd, 0, sizeof(fd));
	fd.width = block.getSize().x; fd.height = block.getSize().y;
	fd.offset[0] = block.getOffset().x; fd.offset[1] = block.getOffset().y;
	fd.border = block.getBorderSize();
	fd.channels = block.getChannelCount();
	fd.filter_radius = (float) rf->getRadius();
	for (int i = 0; i < 32; ++i)     // evalDiscretized(x) = m_values[min(int(|x| * 31 / r), 31)]
		fd.filter_table[i] = (float) rf->evalDiscretized((i + 0.5f) * rf->getRadius() / 31.0f);
	fd.crop_offset[0] = film->getCropOffset().x; fd.crop_offset[1] = film->getCropOffset().y;
	fd.crop_size[0] = film->getCropSize().x; fd.crop_size[1] = film->getCropSize().y;
}

MTS_NAMESPACE_END
