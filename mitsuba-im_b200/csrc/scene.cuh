// scene.cuh — host-side scene object behind the mtsb_scene handle.
#pragma once
#include "common.cuh"
#include "bvh.cuh"
#include <mutex>
#include <algorithm>

namespace mtsb { struct TracePipe; struct PlyMesh; }

struct mtsb_scene {
	mtsb_ctx *ctx = nullptr;
	bool committed = false;
	bool lean = false;                    // every BSDF is diffuse / null / dielectric / GGX roughconductor with visible-normal sampling (bsdf.cuh: LEAN)
	bool volLean = false;                 // LEAN for the volumetric kernels: the same set plus the index-matched `null` boundary (render_vol.cuh)
	bool plain = true;                    // no environment map, no textures, no area light on an analytic sphere / disk: kernels compiled without those branches
	bool curved = false;                  // holds analytic disks / spheres: traversal kernels with their leaf test (trace.cuh)

	// host staging (filled by mtsb_scene_add_*)
	std::vector<mtsb::ShapeRec> shapes;
	std::vector<mtsb::RectRec> rects;
	std::vector<mtsb::BsdfRec> bsdfs;
	std::vector<float> bsdfTables;                 // per-BSDF lookup tables, each stored as {count, values...} (roughplastic)
	std::vector<long> bsdfTableOffset;             // per BSDF: offset into bsdfTables or -1
	mtsb::DevBuf<float> dBsdfTables;
	// Scene::m_specialShapes of the subset (scene.cpp:394-422): caller-added aperture disks, then the environment spheres
	std::vector<mtsb::RectRec> specialRecs; std::vector<uint32_t> specialTypes;
	mtsb::DevBuf<mtsb::RectRec> dSpecialRecs; mtsb::DevBuf<uint32_t> dSpecialTypes;
	size_t specialUploaded = (size_t) -1;          // number of caller-added specials the device copy was made for
	// `envmap` emitter (at most one): pyramid + sampling tables, and the device record that refers to them (envmap.cuh)
	int envmapEmitter = -1;
	std::vector<float> envData, envCdfRows, envCdfCols, envRowWeights;
	mtsb::DevBuf<float> dEnvData, dEnvCdfRows, dEnvCdfCols, dEnvRowWeights;
	mtsb::DevBuf<unsigned char> dEnvmap;           // one EnvmapDev
	std::vector<unsigned char> envRecord;          // host copy of the EnvmapDev (pointers patched at commit)
	// `bitmap` textures: one EnvmapDev-shaped record each (envmap.cuh), all texels in one buffer
	std::vector<std::vector<unsigned char> > texRecords; std::vector<size_t> texDataOffset;
	std::vector<float> texData;
	std::vector<int> bsdfTexture;                  // per BSDF: texture bound to its reflectance or -1
	mtsb::DevBuf<float> dTexData; mtsb::DevBuf<unsigned char> dTexRecords;
	std::vector<mtsb::EmitterRec> emitters;
	std::vector<mtsb::MediumRec> media;
	std::vector<float> emitterWeights;
	std::vector<float> positions, normals, texcoords;
	std::vector<uint32_t> indices;
	std::vector<uint32_t> shapeOffsets;   // n+1 prefix sums of primitive counts
	uint32_t primCount = 0;
	// vertex / triangle totals over host-staged AND device-decoded meshes (a shape's vtxBase / triBase index the final device
	// arrays); per shape, where its host-staged arrays start in positions / indices ((size_t) -1: not host-staged)
	size_t vtxTotal = 0, triTotal = 0;
	std::vector<std::pair<size_t, size_t> > meshHost;
	float plyDecodeMs = 0;                         // device time of the decode kernels at commit (CUDA events), mtsb_scene_build_stats out[7]
	std::vector<mtsb::PlyMesh *> plyMeshes;        // binary PLY images waiting on the device for commit (ply.cu)

	// device copies
	mtsb::DevBuf<mtsb::ShapeRec> dShapes;
	mtsb::DevBuf<mtsb::RectRec> dRects;
	mtsb::DevBuf<mtsb::BsdfRec> dBsdfs;
	mtsb::DevBuf<mtsb::EmitterRec> dEmitters;
	mtsb::DevBuf<mtsb::MediumRec> dMedia;
	mtsb::DevBuf<float> dPositions, dNormals, dTexcoords, dEmitterCdf, dTriAreaCdf;
	mtsb::DevBuf<uint32_t> dIndices, dTriAreaCdfOffset;
	mtsb::DevBuf<unsigned long long> dCounters;
	mtsb::BvhResult bvh;

	mtsb::SceneDev dev;                   // kernel parameter block
	float aabbMin[3], aabbMax[3];         // enlarged (ShapeKDTree::getAABB)
	// what Scene::getAABB() holds beside the geometry when the environment emitter sizes its sphere: the sensor's AABB
	// (scene.cpp:394-404, constant.cpp:69-80) — mtsb_scene_expand_bounds
	float extraMin[3] = { 0, 0, 0 }, extraMax[3] = { 0, 0, 0 }; bool hasExtra = false;

	// staging ring + streams of the host-pointer ray queries (trace.cu), created on first use
	std::mutex pipeMutex;
	mtsb::TracePipe *pipe = nullptr;
	size_t l2Bytes = 0;                   // persisting-L2 claim of this scene (l2_claim below), 0 = none
};

// ply.cu
namespace mtsb {
int ply_stage(mtsb_scene *scene, const void *file, size_t bytes, const float *toWorld, const float *toWorldInv, uint32_t flags, int bsdf,
	uint32_t *vertexCountOut, uint32_t *triangleCountOut, bool *hasUVOut, bool *hasNormalsOut, PlyMesh **out);
void ply_release(PlyMesh *M);
int ply_probe(const void *file, size_t bytes, uint32_t *nv, uint32_t *nt);
int ser_probe(const void *file, size_t bytes, int index, uint32_t *nv, uint32_t *nt);
int ser_stage(mtsb_scene *scene, const void *file, size_t bytes, int index, const float *toWorld, const float *toWorldInv, uint32_t flags,
	uint32_t *vertexCountOut, uint32_t *triangleCountOut, bool *hasUVOut, bool *hasNormalsOut, PlyMesh **out);
void ply_place(PlyMesh *M, int shape, uint32_t vtxBase, uint32_t triBase);
int ply_decode(mtsb_scene *scene, PlyMesh *M, float *positions, float *normals, float *texcoords, uint32_t *indices, uint32_t *scratch);
}

namespace mtsb {

// L2 residency of the acceleration structure.  A scene whose structure (nodes + TriAccel records) fits the device's
// persisting-L2 ceiling claims that many bytes at commit.  The device-wide set-aside (cudaLimitPersistingL2CacheSize) is
// raised to the largest live claim only while BATCHED RAY QUERIES run (l2_phase(scene, true), trace.cu) and dropped to
// zero for path-tracing passes (l2_phase(scene, false), render_host.cuh): every byte of it is L2 taken away from the
// streaming kernels — with a 78 MB set-aside the shade kernel of config 1 ran 2.4x slower (11.2 -> 18.0 ms per frame),
// while the traversal-only batches of config 2 gain 10 % from it.  Larger scenes claim nothing: a partial window measured
// no gain on the 10.5 M-triangle scene.
inline void l2_apply(mtsb_ctx *ctx, size_t want) {       // caller holds ctx->l2Mutex
	if (want == ctx->l2Persist) return;
	if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want) == cudaSuccess) {
		size_t got = 0; cudaDeviceGetLimit(&got, cudaLimitPersistingL2CacheSize);
		ctx->l2Persist = got;
		if (want == 0) cudaCtxResetPersistingL2Cache();
	} else cudaGetLastError();
}

inline void l2_claim(mtsb_scene *scene, size_t bytes) {
	mtsb_ctx *ctx = scene->ctx;
	std::lock_guard<std::mutex> lock(ctx->l2Mutex);
	auto &v = ctx->l2Claims;
	for (size_t i = 0; i < v.size(); ++i) if (v[i].first == scene) { v.erase(v.begin() + i); break; }
	scene->l2Bytes = 0;
	if (bytes > 0 && bytes <= ctx->l2PersistMax && bytes <= ctx->l2WindowMax) { v.push_back(std::make_pair((const void *) scene, bytes)); scene->l2Bytes = bytes; }
	if (v.empty()) l2_apply(ctx, 0);
}

// queries = true: a batched query on `scene` is about to be launched; false: a path-tracing pass is.
inline void l2_phase(mtsb_scene *scene, bool queries) {
	mtsb_ctx *ctx = scene->ctx;
	if (ctx->l2PersistMax == 0) return;
	if (!queries && ctx->l2Persist == 0) return;                       // the common case of a render loop: nothing to do
	if (queries && (scene->l2Bytes == 0 || ctx->l2Persist >= scene->l2Bytes)) return;
	std::lock_guard<std::mutex> lock(ctx->l2Mutex);
	size_t want = 0;
	if (queries) for (auto &c : ctx->l2Claims) want = std::max(want, c.second);
	l2_apply(ctx, want);
}

// Launch attribute of the traversal kernels: hits inside the scene's structure are marked persisting; everything outside
// the window (rays, hit records, path state) takes the normal, evictable path.  0 attributes when the scene made no claim.
inline int l2_window(const mtsb_scene *scene, cudaLaunchAttribute *attr) {
	const mtsb_ctx *ctx = scene->ctx;
	const size_t bytes = scene->l2Bytes;
	if (bytes == 0 || ctx->l2Persist < bytes) return 0;
	attr->id = cudaLaunchAttributeAccessPolicyWindow;
	cudaAccessPolicyWindow &w = attr->val.accessPolicyWindow;
	w.base_ptr = (void *) scene->bvh.accel.ptr;
	w.num_bytes = bytes;
	w.hitRatio = 1.0f;
	w.hitProp = cudaAccessPropertyPersisting;
	w.missProp = cudaAccessPropertyStreaming;
	return 1;
}

} // namespace mtsb
