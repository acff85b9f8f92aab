// scene.cu — context + scene assembly half of the C ABI (include/mtsb200.h).
//
// Stands in for ShapeKDTree::addShape/build (src/librender/skdtree.cpp:49-110) and for the emitter
// distribution Scene::initialize builds (src/librender/scene.cpp:370-389).  Geometry is staged on the host
// in the order shapes are added (that order defines shape indices and the global primitive index, as
// m_shapeMap does), uploaded once, and every derived structure is built on the device.
#include "scene.cuh"
#include "envmap.cuh"
#include <stdlib.h>
#include <stdarg.h>
#include <string.h>
#include <math.h>
#include <algorithm>

namespace mtsb {

static thread_local char g_error[1024] = "";
void set_error(const char *fmt, ...) {
	va_list ap; va_start(ap, fmt);
	vsnprintf(g_error, sizeof(g_error), fmt, ap);
	va_end(ap);
}

static void invert_affine(const float m[16], float inv[12]) {
	// inverse of the upper 3x4 affine part (general 3x3 inverse)
	double a = m[0], b = m[1], c = m[2], d = m[4], e = m[5], f = m[6], g = m[8], h = m[9], i = m[10];
	double A = e*i - f*h, B = -(d*i - f*g), C = d*h - e*g;
	double det = a*A + b*B + c*C;
	double id = 1.0 / det;
	double r[9] = { A*id, -(b*i - c*h)*id, (b*f - c*e)*id,
	                B*id, (a*i - c*g)*id, -(a*f - c*d)*id,
	                C*id, -(a*h - b*g)*id, (a*e - b*d)*id };
	double tx = m[3], ty = m[7], tz = m[11];
	for (int k = 0; k < 3; ++k) {
		inv[k*4+0] = (float) r[k*3+0]; inv[k*4+1] = (float) r[k*3+1]; inv[k*4+2] = (float) r[k*3+2];
		inv[k*4+3] = (float) -(r[k*3+0]*tx + r[k*3+1]*ty + r[k*3+2]*tz);
	}
}

} // namespace mtsb

using namespace mtsb;
void mtsb_release_pipe(mtsb_scene *scene);   // trace.cu

extern "C" {

int mtsb_abi_version(void) { return MTSB_ABI_VERSION; }
const char *mtsb_last_error(void) { return mtsb::g_error; }

int mtsb_device_count(void) {
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
	return n;
}

int mtsb_ctx_create(int device, mtsb_ctx **out) {
	if (!out) MTSB_FAIL(MTSB_EINVAL, "mtsb_ctx_create: out is NULL");
	*out = nullptr;
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
		cudaGetLastError();
		MTSB_FAIL(MTSB_ENODEVICE, "mtsb_ctx_create: no CUDA device available (this library has no CPU fallback)");
	}
	if (device < 0 || device >= n) MTSB_FAIL(MTSB_EINVAL, "mtsb_ctx_create: device %d out of range [0,%d)", device, n);
	MTSB_CUDA(cudaSetDevice(device));
	cudaDeviceProp prop;
	MTSB_CUDA(cudaGetDeviceProperties(&prop, device));
	if (prop.major < 10)
		MTSB_FAIL(MTSB_ENODEVICE, "mtsb_ctx_create: device %d is sm_%d%d; this build targets sm_100a (B200) only", device, prop.major, prop.minor);
	mtsb_ctx *ctx = new mtsb_ctx();
	ctx->device = device; ctx->smCount = prop.multiProcessorCount;
	// experiment knobs: the environment is read here, once per context, and nowhere else in the library
	if (const char *e = getenv("MTSB_HOST_CHUNK")) { if (atol(e) >= 1024) ctx->hostChunk = (size_t) atol(e); }
	if (const char *e = getenv("MTSB_HOST_TAPER")) ctx->hostTaper = (size_t) std::max(0L, atol(e));
	if (ctx->hostTaper == 0) ctx->hostTaper = ctx->hostChunk;
	if (const char *e = getenv("MTSB_FETCH_MIN_IDLE")) { const int v = atoi(e); if (v >= 1 && v <= 32) ctx->fetchMinIdle = v; }
	if (const char *e = getenv("MTSB_TRI_MIN_LANES")) { const int v = atoi(e); if (v >= 1 && v <= 33) ctx->triMinLanes = v; }
	if (const char *e = getenv("MTSB_COLLAPSE_DP")) ctx->collapseDp = atoi(e) != 0;
	if (const char *e = getenv("MTSB_DP_CNODE")) ctx->dpCNode = (float) atof(e);
	if (const char *e = getenv("MTSB_DP_CPRIM")) ctx->dpCPrim = (float) atof(e);
	ctx->tailPaths = 1 << 17;
	if (const char *e = getenv("MTSB_TAIL_PATHS")) ctx->tailPaths = std::max(0, atoi(e));
	if (const char *e = getenv("MTSB_GRAPH")) ctx->useGraph = atoi(e) != 0;
	if (const char *e = getenv("MTSB_SHADE_LEAN")) ctx->shadeLean = atoi(e) != 0;
	if (const char *e = getenv("MTSB_SUSPEND_AFTER")) ctx->suspendAfter = std::max(0, atoi(e));
	if (const char *e = getenv("MTSB_SHADE_WAVES")) ctx->shadeWaves = std::max(0, atoi(e));
	// L2 residency: set aside as much persisting L2 as the device grants (B200: 126 MB of L2); the traversal launches
	// L2 residency of acceleration structures (scene.cuh: l2_window / l2_claim): only the ceiling is fixed here; the
	// persisting set-aside itself is sized by the scenes that can use it, because every byte of it is taken from the L2
	// that the streaming kernels (shade, splat) live on — a fixed maximal set-aside cost config 1 a third of its speed
	ctx->l2PersistMax = (size_t) prop.persistingL2CacheMaxSize;
	if (const char *e = getenv("MTSB_L2_PERSIST")) ctx->l2PersistMax = std::min<size_t>(ctx->l2PersistMax, (size_t) std::max(0L, atol(e)) << 20);
	ctx->l2WindowMax = (size_t) prop.accessPolicyMaxWindowSize;
	cudaError_t e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
	if (e != cudaSuccess) { delete ctx; MTSB_FAIL(MTSB_ECUDA, "cudaStreamCreate: %s", cudaGetErrorString(e)); }
	*out = ctx;
	return 0;
}

void mtsb_ctx_destroy(mtsb_ctx *ctx) {
	if (!ctx) return;
	cudaSetDevice(ctx->device);
	if (ctx->stream) cudaStreamDestroy(ctx->stream);
	delete ctx;
}

int mtsb_ctx_sync(mtsb_ctx *ctx) {
	if (!ctx) MTSB_FAIL(MTSB_EINVAL, "mtsb_ctx_sync: NULL context");
	MTSB_CUDA(cudaSetDevice(ctx->device));
	MTSB_CUDA(cudaStreamSynchronize(ctx->stream));
	return 0;
}

void *mtsb_ctx_stream(mtsb_ctx *ctx) { return ctx ? (void *) ctx->stream : nullptr; }

int mtsb_scene_create(mtsb_ctx *ctx, mtsb_scene **out) {
	if (!ctx || !out) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_create: NULL argument");
	mtsb_scene *s = new mtsb_scene();
	s->ctx = ctx;
	s->shapeOffsets.push_back(0);
	memset(&s->dev, 0, sizeof(s->dev));
	*out = s;
	return 0;
}

void mtsb_scene_destroy(mtsb_scene *scene) {
	if (!scene) return;
	cudaSetDevice(scene->ctx->device);
	mtsb_release_pipe(scene);
	mtsb::l2_claim(scene, 0);
	for (mtsb::PlyMesh *M : scene->plyMeshes) mtsb::ply_release(M);
	delete scene;
}

int mtsb_scene_add_bsdf(mtsb_scene *scene, const mtsb_bsdf_desc *desc) {
	if (!scene || !desc) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_bsdf: NULL argument");
	if (scene->committed) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_bsdf: scene already committed");
	int base = desc->type & 0xff;
	if (base < MTSB_BSDF_DIFFUSE || base > MTSB_BSDF_ROUGHPLASTIC)
		MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_scene_add_bsdf: BSDF type %d is outside the GPU subset", desc->type);
	BsdfRec r; r.type = desc->type; memcpy(r.p, desc->p, sizeof(r.p));
	scene->bsdfs.push_back(r);
	scene->bsdfTableOffset.push_back(-1);
	return (int) scene->bsdfs.size() - 1;
}

int mtsb_scene_set_bsdf_table(mtsb_scene *scene, int bsdf, const float *table, size_t n) {
	if (!scene || !table) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_set_bsdf_table: NULL argument");
	if (scene->committed) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_set_bsdf_table: scene already committed");
	if (bsdf < 0 || bsdf >= (int) scene->bsdfs.size()) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_set_bsdf_table: bsdf id %d unknown", bsdf);
	if ((scene->bsdfs[bsdf].type & 0xff) != MTSB_BSDF_ROUGHPLASTIC) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_set_bsdf_table: bsdf %d takes no table", bsdf);
	if (n < 2 || n > 4096) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_set_bsdf_table: %zu entries (2..4096 expected)", n);
	scene->bsdfTableOffset[bsdf] = (long) scene->bsdfTables.size();
	scene->bsdfTables.push_back((float) n);
	scene->bsdfTables.insert(scene->bsdfTables.end(), table, table + n);
	return 0;
}

int mtsb_scene_add_emitter(mtsb_scene *scene, const mtsb_emitter_desc *desc) {
	if (!scene || !desc) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_emitter: NULL argument");
	if (scene->committed) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_emitter: scene already committed");
	if (desc->type < MTSB_EMITTER_AREA || desc->type > MTSB_EMITTER_ENVMAP)
		MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_scene_add_emitter: emitter type %d is outside the GPU subset", desc->type);
	EmitterRec r; memset(&r, 0, sizeof(r));
	r.type = desc->type; r.shape = desc->shape;
	r.radiance = v3(desc->radiance[0], desc->radiance[1], desc->radiance[2]);
	r.pos = v3(desc->pos[0], desc->pos[1], desc->pos[2]);
	r.samplingWeight = desc->sampling_weight;
	scene->emitters.push_back(r);
	scene->emitterWeights.push_back(desc->sampling_weight);
	return (int) scene->emitters.size() - 1;
}

// levels of a MIP pyramid -> record + texel store (sizes, offsets, size ratios of mipmap.h:243, 265-267, the EWA weight table)
static int fill_pyramid(EnvmapDev &E, int levels, const int32_t *sizes, const float *const *data, std::vector<float> &store, const char *who) {
	if (levels < 1 || levels > MTSB_ENVMAP_MAX_LEVELS) MTSB_FAIL(MTSB_EINVAL, "%s: %d levels", who, levels);
	const size_t base = store.size() / 3;
	E.levels = levels;
	for (int l = 0; l < levels; ++l) {
		const int w = sizes[2 * l], h = sizes[2 * l + 1];
		if (w < 1 || h < 1 || w > 0xFFFF || h > 0xFFFF || !data[l]) MTSB_FAIL(MTSB_EINVAL, "%s: level %d is %dx%d", who, l, w, h);
		E.w[l] = w; E.h[l] = h; E.offset[l] = (unsigned) (store.size() / 3 - base);
		E.ratioX[l] = (float) w / (float) sizes[0]; E.ratioY[l] = (float) h / (float) sizes[1];
		store.insert(store.end(), data[l], data[l] + (size_t) w * h * 3);
	}
	for (int i = 0; i < MTSB_MIPMAP_LUT_SIZE; ++i) {                           // mipmap.h:297-303 (math::fastexp via double)
		const float r2 = (float) i / (float) (MTSB_MIPMAP_LUT_SIZE - 1);
		E.lut[i] = (float) exp((double) (-2.0f * r2)) - (float) exp((double) -2.0f);
	}
	E.uvScaleX = E.uvScaleY = 1.0f;
	return 0;
}

int mtsb_scene_add_texture(mtsb_scene *scene, const mtsb_texture_desc *desc) {
	if (!scene || !desc || !desc->sizes || !desc->data) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_texture: NULL argument");
	if (scene->committed) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_texture: scene already committed");
	if (desc->wrap_u < 0 || desc->wrap_u > 4 || desc->wrap_v < 0 || desc->wrap_v > 4 || desc->filter_type < 0 || desc->filter_type > 3)
		MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_texture: wrap mode / filter type out of range");
	EnvmapDev E; memset(&E, 0, sizeof(E));
	scene->texDataOffset.push_back(scene->texData.size());
	MTSB_CHECK(fill_pyramid(E, desc->levels, desc->sizes, desc->data, scene->texData, "mtsb_scene_add_texture"));
	E.bcu = desc->wrap_u; E.bcv = desc->wrap_v; E.filterType = desc->filter_type;
	E.maxAnisotropy = desc->filter_type == 3 ? desc->max_anisotropy : 1.0f;                 // bitmap.cpp:234-237
	E.uvScaleX = desc->uv_scale[0]; E.uvScaleY = desc->uv_scale[1]; E.uvOffsetX = desc->uv_offset[0]; E.uvOffsetY = desc->uv_offset[1];
	scene->texRecords.push_back(std::vector<unsigned char>((const unsigned char *) &E, (const unsigned char *) &E + sizeof(E)));
	return (int) scene->texRecords.size() - 1;
}

int mtsb_scene_set_bsdf_texture(mtsb_scene *scene, int bsdf, int texture) {
	if (!scene) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_set_bsdf_texture: NULL scene");
	if (scene->committed) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_set_bsdf_texture: scene already committed");
	if (bsdf < 0 || bsdf >= (int) scene->bsdfs.size() ||
	    ((scene->bsdfs[bsdf].type & 0xff) != MTSB_BSDF_DIFFUSE && (scene->bsdfs[bsdf].type & 0xff) != MTSB_BSDF_PLASTIC))
		MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_scene_set_bsdf_texture: only the reflectance of a diffuse BSDF and the diffuseReflectance of a plastic BSDF take a texture (bsdf %d)", bsdf);
	if (texture < 0 || texture >= (int) scene->texRecords.size()) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_set_bsdf_texture: texture %d unknown", texture);
	scene->bsdfTexture.resize(scene->bsdfs.size(), -1);
	scene->bsdfTexture[bsdf] = texture;
	return 0;
}

// EnvironmentMap::configure (src/emitters/envmap.cpp:276-333) restated: marginal / conditional CDFs over luminance x sin(theta)
int mtsb_scene_set_envmap(mtsb_scene *scene, int emitter, const mtsb_envmap_desc *desc) {
	if (!scene || !desc || !desc->sizes || !desc->data) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_set_envmap: NULL argument");
	if (scene->committed) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_set_envmap: scene already committed");
	if (emitter < 0 || emitter >= (int) scene->emitters.size() || scene->emitters[emitter].type != MTSB_EMITTER_ENVMAP)
		MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_set_envmap: emitter %d is not an MTSB_EMITTER_ENVMAP", emitter);
	if (scene->envmapEmitter >= 0 && scene->envmapEmitter != emitter) MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_scene_set_envmap: one environment map per scene");
	EnvmapDev E; memset(&E, 0, sizeof(E));
	scene->envData.clear();
	MTSB_CHECK(fill_pyramid(E, desc->levels, desc->sizes, desc->data, scene->envData, "mtsb_scene_set_envmap"));
	E.bcu = 0; E.bcv = 1; E.filterType = 3; E.maxAnisotropy = 10.0f;                       // envmap.cpp:143-144, 175-176
	const int W = E.w[0], H = E.h[0];
	const float *img = scene->envData.data();
	scene->envCdfCols.assign((size_t) (W + 1) * H, 0.0f); scene->envCdfRows.assign(H + 1, 0.0f); scene->envRowWeights.assign(H, 0.0f);
	size_t colPos = 0, rowPos = 0;
	float rowSum = 0.0f;
	scene->envCdfRows[rowPos++] = 0;
	for (int y = 0; y < H; ++y) {
		float colSum = 0;
		scene->envCdfCols[colPos++] = 0;
		for (int x = 0; x < W; ++x) {
			const float *p = img + ((size_t) y * W + x) * 3;
			colSum += 0.212671f * p[0] + 0.715160f * p[1] + 0.072169f * p[2];      // Spectrum::getLuminance (RGB build)
			scene->envCdfCols[colPos++] = colSum;
		}
		const float normalization = 1.0f / colSum;
		for (int x = 1; x < W; ++x) scene->envCdfCols[colPos - x - 1] *= normalization;
		scene->envCdfCols[colPos - 1] = 1.0f;
		const float weight = sinf((y + 0.5f) * ENVMAP_PI / H);
		scene->envRowWeights[y] = weight;
		rowSum += colSum * weight;
		scene->envCdfRows[rowPos++] = rowSum;
	}
	const float normalization = 1.0f / rowSum;
	for (int y = 1; y < H; ++y) scene->envCdfRows[rowPos - y - 1] *= normalization;
	scene->envCdfRows[rowPos - 1] = 1.0f;
	if (rowSum == 0) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_set_envmap: the environment map is completely black -- this is not allowed.");
	if (!std::isfinite(rowSum)) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_set_envmap: the environment map contains an invalid floating point value (nan/inf)");
	E.normalization = 1.0f / (rowSum * (2 * ENVMAP_PI / W) * (ENVMAP_PI / H));
	E.pixelSizeX = 2 * ENVMAP_PI / W; E.pixelSizeY = ENVMAP_PI / H;
	E.scale = desc->scale;
	float inv[12];
	invert_affine(desc->to_world, inv);
	for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) { E.toWorld[r * 3 + c] = desc->to_world[r * 4 + c]; E.toLocal[r * 3 + c] = inv[r * 4 + c]; }
	scene->envRecord.assign((const unsigned char *) &E, (const unsigned char *) &E + sizeof(E));
	scene->envmapEmitter = emitter;
	return 0;
}

static int check_refs(mtsb_scene *scene, int bsdf, int emitter, const char *who) {
	if (scene->committed) MTSB_FAIL(MTSB_EINVAL, "%s: scene already committed", who);
	if (bsdf < 0 || bsdf >= (int) scene->bsdfs.size()) MTSB_FAIL(MTSB_EINVAL, "%s: bsdf id %d unknown", who, bsdf);
	if (emitter < -1 || emitter >= (int) scene->emitters.size()) MTSB_FAIL(MTSB_EINVAL, "%s: emitter id %d unknown", who, emitter);
	return 0;
}

int mtsb_scene_add_mesh(mtsb_scene *scene, const float *positions, const float *normals, const float *texcoords,
		const uint32_t *triangles, size_t vertex_count, size_t triangle_count, int bsdf, int emitter) {
	if (!scene) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_mesh: NULL scene");
	MTSB_CHECK(check_refs(scene, bsdf, emitter, "mtsb_scene_add_mesh"));
	if (triangle_count == 0 || vertex_count == 0 || !positions || !triangles)
		MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_mesh: empty triangle mesh");   // TriMesh::prepareSamplingTable logs the same error
	if ((uint64_t) scene->primCount + triangle_count > 0x7fffffffull)
		MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_mesh: more than 2^31 primitives");
	for (size_t i = 0; i < triangle_count * 3; ++i)
		if (triangles[i] >= vertex_count)
			MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_mesh: triangle %zu references vertex %u >= %zu", i / 3, triangles[i], vertex_count);
	ShapeRec sh; memset(&sh, 0, sizeof(sh));
	sh.type = SHAPE_MESH;
	sh.primOffset = scene->primCount; sh.primCount = (uint32_t) triangle_count;
	if (scene->vtxTotal + vertex_count > 0xffffffffull) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_mesh: more than 2^32 vertices");
	sh.vtxBase = (uint32_t) scene->vtxTotal;
	sh.triBase = (uint32_t) scene->triTotal;
	sh.flags = (normals ? SHAPE_HAS_NORMALS : 0) | (texcoords ? SHAPE_HAS_UV : 0);
	sh.bsdf = bsdf; sh.emitter = emitter; sh.interior = -1; sh.exterior = -1;
	scene->meshHost.resize(scene->shapes.size() + 1, std::make_pair((size_t) -1, (size_t) -1));
	scene->meshHost.back() = std::make_pair(scene->positions.size() / 3, scene->indices.size() / 3);
	scene->vtxTotal += vertex_count; scene->triTotal += triangle_count;
	scene->positions.insert(scene->positions.end(), positions, positions + vertex_count * 3);
	if (normals) scene->normals.insert(scene->normals.end(), normals, normals + vertex_count * 3);
	else scene->normals.resize(scene->normals.size() + vertex_count * 3, 0.0f);
	if (texcoords) scene->texcoords.insert(scene->texcoords.end(), texcoords, texcoords + vertex_count * 2);
	else scene->texcoords.resize(scene->texcoords.size() + vertex_count * 2, 0.0f);
	scene->indices.insert(scene->indices.end(), triangles, triangles + triangle_count * 3);
	scene->shapes.push_back(sh);
	scene->primCount += (uint32_t) triangle_count;
	scene->shapeOffsets.push_back(scene->primCount);
	if (emitter >= 0) scene->emitters[emitter].shape = (uint32_t) scene->shapes.size() - 1;
	return (int) scene->shapes.size() - 1;
}

static int add_decoded_mesh(mtsb_scene *scene, PlyMesh *M, uint32_t nv, uint32_t nt, bool hasUV, uint32_t flags, int bsdf,
		uint32_t *vertex_count, uint32_t *triangle_count, const char *who) {
	if ((uint64_t) scene->primCount + nt > 0x7fffffffull || scene->vtxTotal + nv > 0xffffffffull) {
		ply_release(M); MTSB_FAIL(MTSB_EINVAL, "%s: more than 2^31 primitives", who);
	}
	ShapeRec sh; memset(&sh, 0, sizeof(sh));
	sh.type = SHAPE_MESH;
	sh.primOffset = scene->primCount; sh.primCount = nt;
	sh.vtxBase = (uint32_t) scene->vtxTotal; sh.triBase = (uint32_t) scene->triTotal;
	// a file without normals gets computed ones unless faceNormals is set (TriMesh::computeNormals)
	sh.flags = ((flags & MTSB_PLY_FACE_NORMALS) ? 0 : SHAPE_HAS_NORMALS) | (hasUV ? SHAPE_HAS_UV : 0);
	sh.bsdf = bsdf; sh.emitter = -1; sh.interior = -1; sh.exterior = -1;
	scene->vtxTotal += nv; scene->triTotal += nt;
	scene->shapes.push_back(sh);
	scene->meshHost.resize(scene->shapes.size(), std::make_pair((size_t) -1, (size_t) -1));
	scene->primCount += nt;
	scene->shapeOffsets.push_back(scene->primCount);
	ply_place(M, (int) scene->shapes.size() - 1, sh.vtxBase, sh.triBase);
	scene->plyMeshes.push_back(M);
	if (vertex_count) *vertex_count = nv;
	if (triangle_count) *triangle_count = nt;
	return (int) scene->shapes.size() - 1;
}

int mtsb_scene_add_ply(mtsb_scene *scene, const void *file, size_t bytes, const float to_world[16], const float to_world_inv[16],
		uint32_t flags, int bsdf, uint32_t *vertex_count, uint32_t *triangle_count) {
	if (!scene || !file) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_ply: NULL argument");
	MTSB_CHECK(check_refs(scene, bsdf, -1, "mtsb_scene_add_ply"));
	if (flags & ~(MTSB_PLY_FACE_NORMALS | MTSB_PLY_FLIP_NORMALS)) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_ply: unknown flags %#x", flags);
	uint32_t nv = 0, nt = 0; bool hasUV = false, hasNormals = false; PlyMesh *M = nullptr;
	MTSB_CHECK(ply_stage(scene, file, bytes, to_world, to_world_inv, flags, bsdf, &nv, &nt, &hasUV, &hasNormals, &M));
	return add_decoded_mesh(scene, M, nv, nt, hasUV, flags, bsdf, vertex_count, triangle_count, "mtsb_scene_add_ply");
}

int mtsb_scene_add_serialized(mtsb_scene *scene, const void *file, size_t bytes, int shape_index, const float to_world[16],
		const float to_world_inv[16], uint32_t flags, int bsdf, uint32_t *vertex_count, uint32_t *triangle_count) {
	if (!scene || !file) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_serialized: NULL argument");
	MTSB_CHECK(check_refs(scene, bsdf, -1, "mtsb_scene_add_serialized"));
	if (flags & ~(MTSB_PLY_FACE_NORMALS | MTSB_PLY_FLIP_NORMALS)) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_serialized: unknown flags %#x", flags);
	uint32_t nv = 0, nt = 0; bool hasUV = false, hasNormals = false; PlyMesh *M = nullptr;
	MTSB_CHECK(ser_stage(scene, file, bytes, shape_index, to_world, to_world_inv, flags, &nv, &nt, &hasUV, &hasNormals, &M));
	return add_decoded_mesh(scene, M, nv, nt, hasUV, flags, bsdf, vertex_count, triangle_count, "mtsb_scene_add_serialized");
}

int mtsb_serialized_probe(const void *file, size_t bytes, int shape_index, uint32_t *vertex_count, uint32_t *triangle_count) {
	if (!file) MTSB_FAIL(MTSB_EINVAL, "mtsb_serialized_probe: NULL argument");
	return ser_probe(file, bytes, shape_index, vertex_count, triangle_count);
}

int mtsb_ply_probe(const void *file, size_t bytes, uint32_t *vertex_count, uint32_t *triangle_count) {
	if (!file) MTSB_FAIL(MTSB_EINVAL, "mtsb_ply_probe: NULL argument");
	return ply_probe(file, bytes, vertex_count, triangle_count);
}

int mtsb_scene_read_mesh(const mtsb_scene *scene, int shape, float *positions, float *normals, float *texcoords, uint32_t *triangles) {
	if (!scene || !scene->committed) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_read_mesh: scene not committed");
	if (shape < 0 || shape >= (int) scene->shapes.size() || scene->shapes[shape].type != SHAPE_MESH) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_read_mesh: shape %d is not a triangle mesh", shape);
	const ShapeRec &sh = scene->shapes[shape];
	// the vertex range of a mesh ends where the next mesh's begins
	size_t vEnd = scene->vtxTotal;
	for (size_t s = (size_t) shape + 1; s < scene->shapes.size(); ++s) if (scene->shapes[s].type == SHAPE_MESH) { vEnd = scene->shapes[s].vtxBase; break; }
	const size_t V = vEnd - sh.vtxBase, T = sh.primCount;
	MTSB_CUDA(cudaSetDevice(scene->ctx->device));
	MTSB_CUDA(cudaStreamSynchronize(scene->ctx->stream));
	if (positions) MTSB_CUDA(cudaMemcpy(positions, scene->dPositions.ptr + (size_t) sh.vtxBase * 3, V * 3 * sizeof(float), cudaMemcpyDeviceToHost));
	if (normals) MTSB_CUDA(cudaMemcpy(normals, scene->dNormals.ptr + (size_t) sh.vtxBase * 3, V * 3 * sizeof(float), cudaMemcpyDeviceToHost));
	if (texcoords) MTSB_CUDA(cudaMemcpy(texcoords, scene->dTexcoords.ptr + (size_t) sh.vtxBase * 2, V * 2 * sizeof(float), cudaMemcpyDeviceToHost));
	if (triangles) MTSB_CUDA(cudaMemcpy(triangles, scene->dIndices.ptr + (size_t) sh.triBase * 3, T * 3 * sizeof(uint32_t), cudaMemcpyDeviceToHost));
	return 0;
}

int mtsb_scene_add_rectangle(mtsb_scene *scene, const float m[16], int flip_normals, int bsdf, int emitter) {
	if (!scene || !m) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_rectangle: NULL argument");
	MTSB_CHECK(check_refs(scene, bsdf, emitter, "mtsb_scene_add_rectangle"));
	float ow[16]; memcpy(ow, m, sizeof(ow));
	if (flip_normals)   // m_objectToWorld * scale(1,1,-1)  (rectangle.cpp:81-83)
		for (int r = 0; r < 4; ++r) ow[r*4+2] = -ow[r*4+2];
	RectRec rr; memset(&rr, 0, sizeof(rr));
	for (int r = 0; r < 3; ++r) for (int c = 0; c < 4; ++c) rr.toWorld[r*4+c] = ow[r*4+c];
	invert_affine(ow, rr.toObject);
	// configure(): dpdu = toWorld(Vector(2,0,0)), dpdv = toWorld(Vector(0,2,0)), n = normalize(toWorld(Normal(0,0,1)))
	rr.dpdu = v3(2*ow[0], 2*ow[4], 2*ow[8]);
	rr.dpdv = v3(2*ow[1], 2*ow[5], 2*ow[9]);
	// normals transform with the inverse transpose: n_i = sum_j inv[j][i] * (0,0,1)_j = inv[2][i]
	V3 n = v3(rr.toObject[8], rr.toObject[9], rr.toObject[10]);
	float len = sqrtf(n.x*n.x + n.y*n.y + n.z*n.z);
	rr.n = v3(n.x/len, n.y/len, n.z/len);
	float lu = sqrtf(dot(rr.dpdu, rr.dpdu)), lv = sqrtf(dot(rr.dpdv, rr.dpdv));
	if (fabsf(dot(rr.dpdu, rr.dpdv) / (lu * lv)) > MTS_EPSILON)
		MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_rectangle: 'toWorld' transformation contains shear!");   // rectangle.cpp:106-107
	rr.invSurfaceArea = 1.0f / (lu * lv);
	ShapeRec sh; memset(&sh, 0, sizeof(sh));
	sh.type = SHAPE_RECT; sh.primOffset = scene->primCount; sh.primCount = 1;
	sh.bsdf = bsdf; sh.emitter = emitter; sh.interior = -1; sh.exterior = -1; sh.rect = (uint32_t) scene->rects.size();
	sh.invArea = rr.invSurfaceArea;
	scene->rects.push_back(rr);
	scene->shapes.push_back(sh);
	scene->primCount += 1;
	scene->shapeOffsets.push_back(scene->primCount);
	if (emitter >= 0) scene->emitters[emitter].shape = (uint32_t) scene->shapes.size() - 1;
	return (int) scene->shapes.size() - 1;
}

int mtsb_scene_add_medium(mtsb_scene *scene, const mtsb_medium_desc *desc) {
	if (!scene || !desc) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_medium: NULL argument");
	if (scene->committed) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_medium: scene already committed");
	if (desc->strategy < 0 || desc->strategy > 1 || desc->phase < 0 || desc->phase > 1)
		MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_scene_add_medium: sampling strategy %d / phase function %d is outside the GPU subset", desc->strategy, desc->phase);
	MediumRec m; memset(&m, 0, sizeof(m));
	m.sigmaA = v3(desc->sigma_a[0], desc->sigma_a[1], desc->sigma_a[2]);
	m.sigmaS = v3(desc->sigma_s[0], desc->sigma_s[1], desc->sigma_s[2]);
	m.sigmaT = m.sigmaA + m.sigmaS;                       // Medium::Medium (src/librender/medium.cpp)
	m.strategy = desc->strategy; m.samplingDensity = desc->sampling_density; m.mediumSamplingWeight = desc->medium_sampling_weight;
	m.phase = desc->phase; m.g = desc->g;
	scene->media.push_back(m);
	return (int) scene->media.size() - 1;
}

int mtsb_scene_set_shape_media(mtsb_scene *scene, int shape, int interior, int exterior) {
	if (!scene) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_set_shape_media: NULL argument");
	if (scene->committed) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_set_shape_media: scene already committed");
	const int n = (int) scene->media.size();
	if (shape < 0 || shape >= (int) scene->shapes.size() || interior < -1 || interior >= n || exterior < -1 || exterior >= n)
		MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_set_shape_media: shape %d / media %d, %d unknown", shape, interior, exterior);
	scene->shapes[shape].interior = interior; scene->shapes[shape].exterior = exterior;
	return 0;
}

int mtsb_scene_expand_bounds(mtsb_scene *scene, const float min[3], const float max[3]) {
	if (!scene || !min || !max) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_expand_bounds: NULL argument");
	if (scene->committed) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_expand_bounds: scene already committed");
	for (int a = 0; a < 3; ++a) {
		scene->extraMin[a] = scene->hasExtra ? fminf(scene->extraMin[a], min[a]) : min[a];
		scene->extraMax[a] = scene->hasExtra ? fmaxf(scene->extraMax[a], max[a]) : max[a];
	}
	scene->hasExtra = true;
	return 0;
}

int mtsb_scene_add_disk(mtsb_scene *scene, const float m[16], int flip_normals, int bsdf, int emitter) {
	if (!scene || !m) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_disk: NULL argument");
	MTSB_CHECK(check_refs(scene, bsdf, emitter, "mtsb_scene_add_disk"));
	float ow[16]; memcpy(ow, m, sizeof(ow));
	if (flip_normals)   // m_objectToWorld->prependScale(Vector(1, 1, -1))  (disk.cpp:86-87)
		for (int r = 0; r < 4; ++r) ow[r*4+2] = -ow[r*4+2];
	RectRec rr; memset(&rr, 0, sizeof(rr));
	for (int r = 0; r < 3; ++r) for (int c = 0; c < 4; ++c) rr.toWorld[r*4+c] = ow[r*4+c];
	invert_affine(ow, rr.toObject);
	rr.dpdu = v3(ow[0], ow[4], ow[8]);
	rr.dpdv = v3(ow[1], ow[5], ow[9]);
	V3 n = v3(rr.toObject[8], rr.toObject[9], rr.toObject[10]);      // trafo(Normal(0,0,1)): inverse transpose
	float len = sqrtf(n.x*n.x + n.y*n.y + n.z*n.z);
	rr.n = v3(n.x/len, n.y/len, n.z/len);
	const float lu = sqrtf(dot(rr.dpdu, rr.dpdu)), lv = sqrtf(dot(rr.dpdv, rr.dpdv));
	if (fabsf(dot(rr.dpdu, rr.dpdv) / (lu * lv)) > 1e-3f) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_disk: 'toWorld' transformation contains shear!");           // disk.cpp:108-109
	if (fabsf(lu / lv - 1) > 1e-3f) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_disk: 'toWorld' transformation contains a non-uniform scale!");                  // disk.cpp:111-112
	rr.invSurfaceArea = 1.0f / (3.14159265358979323846f * lu * lu);
	ShapeRec sh; memset(&sh, 0, sizeof(sh));
	scene->curved = true;
	sh.type = SHAPE_DISK; sh.primOffset = scene->primCount; sh.primCount = 1;
	sh.bsdf = bsdf; sh.emitter = emitter; sh.interior = -1; sh.exterior = -1; sh.rect = (uint32_t) scene->rects.size();
	sh.invArea = rr.invSurfaceArea;
	scene->rects.push_back(rr);
	scene->shapes.push_back(sh);
	scene->primCount += 1;
	scene->shapeOffsets.push_back(scene->primCount);
	if (emitter >= 0) scene->emitters[emitter].shape = (uint32_t) scene->shapes.size() - 1;
	return (int) scene->shapes.size() - 1;
}

int mtsb_scene_add_special_disk(mtsb_scene *scene, const float m[16]) {
	if (!scene || !m) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_special_disk: NULL argument");
	RectRec rr; memset(&rr, 0, sizeof(rr));
	for (int r = 0; r < 3; ++r) for (int c = 0; c < 4; ++c) rr.toWorld[r*4+c] = m[r*4+c];
	invert_affine(m, rr.toObject);
	rr.dpdu = v3(m[0], m[4], m[8]);
	rr.dpdv = v3(m[1], m[5], m[9]);
	V3 n = v3(rr.toObject[8], rr.toObject[9], rr.toObject[10]);
	const float len = sqrtf(n.x*n.x + n.y*n.y + n.z*n.z);
	rr.n = v3(n.x/len, n.y/len, n.z/len);
	scene->specialRecs.push_back(rr);
	scene->specialTypes.push_back(SHAPE_DISK);
	return (int) scene->specialRecs.size() - 1;
}

int mtsb_scene_add_sphere(mtsb_scene *scene, const float object_to_world[16], float radius, int flip_normals, int bsdf, int emitter) {
	if (!scene || !object_to_world) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_sphere: NULL argument");
	MTSB_CHECK(check_refs(scene, bsdf, emitter, "mtsb_scene_add_sphere"));
	if (!(radius > 0)) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_add_sphere: cannot create spheres of radius <= 0");                                                  // sphere.cpp:130-131
	RectRec rr; memset(&rr, 0, sizeof(rr));
	for (int r = 0; r < 3; ++r) for (int c = 0; c < 4; ++c) rr.toWorld[r*4+c] = object_to_world[r*4+c];
	invert_affine(object_to_world, rr.toObject);
	rr.dpdu = v3(object_to_world[3], object_to_world[7], object_to_world[11]);     // m_center = m_objectToWorld(Point(0,0,0))
	rr.dpdv = v3(radius, 0.0f, 0.0f);
	rr.pad[0] = flip_normals ? 1u : 0u;
	rr.invSurfaceArea = 1.0f / (4 * 3.14159265358979323846f * radius * radius);
	ShapeRec sh; memset(&sh, 0, sizeof(sh));
	scene->curved = true;
	sh.type = SHAPE_SPHERE; sh.primOffset = scene->primCount; sh.primCount = 1;
	sh.bsdf = bsdf; sh.emitter = emitter; sh.interior = -1; sh.exterior = -1; sh.rect = (uint32_t) scene->rects.size();
	sh.invArea = rr.invSurfaceArea;
	scene->rects.push_back(rr);
	scene->shapes.push_back(sh);
	scene->primCount += 1;
	scene->shapeOffsets.push_back(scene->primCount);
	if (emitter >= 0) scene->emitters[emitter].shape = (uint32_t) scene->shapes.size() - 1;
	return (int) scene->shapes.size() - 1;
}

int mtsb_scene_commit(mtsb_scene *scene, int builder) {
	if (!scene) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_commit: NULL scene");
	if (scene->committed) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_commit: already committed");
	if (scene->primCount == 0) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_commit: scene has no primitives");
	if (builder != MTSB_BUILDER_LBVH && builder != MTSB_BUILDER_PLOC) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_commit: unknown builder %d", builder);
	mtsb_ctx *ctx = scene->ctx;
	MTSB_CUDA(cudaSetDevice(ctx->device));
	cudaStream_t st = ctx->stream;

	// area-light tables: TriMesh::prepareSamplingTable (src/librender/trimesh.cpp:389-404) in float, like the reference
	std::vector<float> triAreaCdf;
	std::vector<uint32_t> triAreaCdfOffset(scene->shapes.size(), 0xffffffffu);
	for (size_t s = 0; s < scene->shapes.size(); ++s) {
		ShapeRec &sh = scene->shapes[s];
		if (sh.type != SHAPE_MESH) continue;
		const bool needTable = sh.emitter >= 0;
		if (!needTable) continue;
		triAreaCdfOffset[s] = (uint32_t) triAreaCdf.size();
		size_t base = triAreaCdf.size();
		triAreaCdf.push_back(0.0f);
		for (uint32_t t = 0; t < sh.primCount; ++t) {
			const size_t hv = scene->meshHost[s].first, ht = scene->meshHost[s].second;      // emitter meshes are host-staged
			const uint32_t *ix = &scene->indices[(ht + t) * 3];
			const float *p0 = &scene->positions[(hv + ix[0]) * 3];
			const float *p1 = &scene->positions[(hv + ix[1]) * 3];
			const float *p2 = &scene->positions[(hv + ix[2]) * 3];
			V3 a = v3(p1[0]-p0[0], p1[1]-p0[1], p1[2]-p0[2]), b = v3(p2[0]-p0[0], p2[1]-p0[1], p2[2]-p0[2]);
			V3 c = cross(a, b);
			float area = 0.5f * sqrtf(dot(c, c));
			triAreaCdf.push_back(triAreaCdf.back() + area);
		}
		float sum = triAreaCdf.back();
		if (sum > 0) {
			float norm = 1.0f / sum;
			for (size_t i = base + 1; i < triAreaCdf.size(); ++i) triAreaCdf[i] *= norm;
			triAreaCdf.back() = 1.0f;
			sh.invArea = 1.0f / sum;
		}
	}

	// emitter selection distribution (DiscreteDistribution, include/mitsuba/core/pmf.h)
	std::vector<float> emitterCdf(1, 0.0f);
	for (size_t i = 0; i < scene->emitters.size(); ++i) emitterCdf.push_back(emitterCdf.back() + scene->emitterWeights[i]);
	float emitterNorm = 0;
	if (emitterCdf.back() > 0) {
		emitterNorm = 1.0f / emitterCdf.back();
		for (size_t i = 1; i < emitterCdf.size(); ++i) emitterCdf[i] *= emitterNorm;
		emitterCdf.back() = 1.0f;
	}
	int envEmitter = -1;
	for (size_t i = 0; i < scene->emitters.size(); ++i) {
		EmitterRec &e = scene->emitters[i];
		if (e.type == MTSB_EMITTER_CONSTANT || e.type == MTSB_EMITTER_ENVMAP) envEmitter = (int) i;
		if (e.type == MTSB_EMITTER_AREA) {
			if (e.shape >= scene->shapes.size() || scene->shapes[e.shape].emitter != (int) i)
				MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_commit: area emitter %zu is not attached to a shape", i);
			e.invSurfaceArea = scene->shapes[e.shape].invArea;
		}
	}

	MTSB_CHECK(scene->dShapes.upload(scene->shapes, st));
	MTSB_CHECK(scene->dRects.upload(scene->rects, st));
	// bitmap textures: texels and records to the device; a textured diffuse record carries the record's address in p[13..14], p[12] = 1
	if (!scene->texRecords.empty()) {
		MTSB_CHECK(scene->dTexData.upload(scene->texData, st));
		std::vector<unsigned char> recs;
		for (size_t t = 0; t < scene->texRecords.size(); ++t) {
			EnvmapDev *E = (EnvmapDev *) scene->texRecords[t].data();
			E->data = scene->dTexData.ptr + scene->texDataOffset[t];
			recs.insert(recs.end(), scene->texRecords[t].begin(), scene->texRecords[t].end());
		}
		MTSB_CHECK(scene->dTexRecords.upload(recs, st));
		scene->bsdfTexture.resize(scene->bsdfs.size(), -1);
		for (size_t i = 0; i < scene->bsdfs.size(); ++i) {
			if (scene->bsdfTexture[i] < 0) continue;
			const unsigned char *dev = scene->dTexRecords.ptr + (size_t) scene->bsdfTexture[i] * sizeof(EnvmapDev);
			memcpy(&scene->bsdfs[i].p[13], &dev, sizeof(dev));
			scene->bsdfs[i].p[12] = 1.0f;
		}
		for (const ShapeRec &sh : scene->shapes)
			if (scene->bsdfTexture[sh.bsdf] >= 0 && sh.type != SHAPE_MESH && sh.type != SHAPE_RECT)
				MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_scene_commit: textured BSDFs go on triangle meshes and rectangles");
	}
	// lookup tables first: a roughplastic record carries the device address of its table in p[13..14]
	if (!scene->bsdfTables.empty()) MTSB_CHECK(scene->dBsdfTables.upload(scene->bsdfTables, st));
	for (size_t i = 0; i < scene->bsdfs.size(); ++i) {
		if ((scene->bsdfs[i].type & 0xff) != MTSB_BSDF_ROUGHPLASTIC) continue;
		if (scene->bsdfTableOffset[i] < 0) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_commit: roughplastic BSDF %zu has no transmittance table (mtsb_scene_set_bsdf_table)", i);
		const float *dev = scene->dBsdfTables.ptr + scene->bsdfTableOffset[i];
		memcpy(&scene->bsdfs[i].p[13], &dev, sizeof(dev));
	}
	MTSB_CHECK(scene->dBsdfs.upload(scene->bsdfs, st));
	if (scene->plyMeshes.empty()) {
		MTSB_CHECK(scene->dPositions.upload(scene->positions, st));
		MTSB_CHECK(scene->dNormals.upload(scene->normals, st));
		MTSB_CHECK(scene->dTexcoords.upload(scene->texcoords, st));
		MTSB_CHECK(scene->dIndices.upload(scene->indices, st));
	} else {
		// host-staged meshes are copied to their place, PLY images are decoded into theirs by kernels (ply.cu)
		MTSB_CHECK(scene->dPositions.alloc(scene->vtxTotal * 3)); MTSB_CHECK(scene->dNormals.alloc(scene->vtxTotal * 3));
		MTSB_CHECK(scene->dTexcoords.alloc(scene->vtxTotal * 2)); MTSB_CHECK(scene->dIndices.alloc(scene->triTotal * 3));
		scene->meshHost.resize(scene->shapes.size(), std::make_pair((size_t) -1, (size_t) -1));
		for (size_t s = 0; s < scene->shapes.size(); ++s) {
			const ShapeRec &sh = scene->shapes[s];
			if (sh.type != SHAPE_MESH || scene->meshHost[s].first == (size_t) -1) continue;
			const size_t hv = scene->meshHost[s].first, ht = scene->meshHost[s].second;
			size_t hvEnd = scene->positions.size() / 3;
			for (size_t k = s + 1; k < scene->shapes.size(); ++k) if (scene->meshHost[k].first != (size_t) -1) { hvEnd = scene->meshHost[k].first; break; }
			const size_t V = hvEnd - hv, T = sh.primCount;
			MTSB_CUDA(cudaMemcpyAsync(scene->dPositions.ptr + (size_t) sh.vtxBase * 3, &scene->positions[hv * 3], V * 3 * sizeof(float), cudaMemcpyHostToDevice, st));
			MTSB_CUDA(cudaMemcpyAsync(scene->dNormals.ptr + (size_t) sh.vtxBase * 3, &scene->normals[hv * 3], V * 3 * sizeof(float), cudaMemcpyHostToDevice, st));
			MTSB_CUDA(cudaMemcpyAsync(scene->dTexcoords.ptr + (size_t) sh.vtxBase * 2, &scene->texcoords[hv * 2], V * 2 * sizeof(float), cudaMemcpyHostToDevice, st));
			MTSB_CUDA(cudaMemcpyAsync(scene->dIndices.ptr + (size_t) sh.triBase * 3, &scene->indices[ht * 3], T * 3 * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
		}
		DevBuf<uint32_t> plyScratch; MTSB_CHECK(plyScratch.alloc(2));
		for (PlyMesh *M : scene->plyMeshes)
			MTSB_CHECK(ply_decode(scene, M, scene->dPositions.ptr, scene->dNormals.ptr, scene->dTexcoords.ptr, scene->dIndices.ptr, plyScratch.ptr));
		for (PlyMesh *M : scene->plyMeshes) ply_release(M);
		scene->plyMeshes.clear();
	}
	MTSB_CHECK(scene->dEmitterCdf.upload(emitterCdf, st));
	MTSB_CHECK(scene->dTriAreaCdf.upload(triAreaCdf, st));
	MTSB_CHECK(scene->dTriAreaCdfOffset.upload(triAreaCdfOffset, st));
	MTSB_CHECK(scene->dCounters.alloc(4));
	MTSB_CUDA(cudaMemsetAsync(scene->dCounters.ptr, 0, 4 * sizeof(unsigned long long), st));

	SceneDev &D = scene->dev;
	memset(&D, 0, sizeof(D));
	D.shapes = scene->dShapes.ptr; D.rects = scene->dRects.ptr; D.bsdfs = scene->dBsdfs.ptr;
	D.positions = scene->dPositions.ptr; D.normals = scene->dNormals.ptr; D.texcoords = scene->dTexcoords.ptr;
	D.indices = scene->dIndices.ptr;
	D.emitterCdf = scene->dEmitterCdf.ptr; D.triAreaCdf = scene->dTriAreaCdf.ptr; D.triAreaCdfOffset = scene->dTriAreaCdfOffset.ptr;
	D.shapeCount = (uint32_t) scene->shapes.size(); D.emitterCount = (uint32_t) scene->emitters.size();
	D.envEmitter = envEmitter; D.emitterNorm = emitterNorm;
	D.counters = nullptr;
	// tuning knob of the triangle phase (trace.cuh); MTSB_TRI_MIN_LANES overrides for experiments
	D.triMinLanes = ctx->triMinLanes;
	D.oneBits = 0x3F800000u;
	D.fetchMinIdle = ctx->fetchMinIdle;

	MTSB_CHECK(build_bvh(ctx, D, scene->primCount, builder, scene->bvh));
	D.nodes = scene->bvh.nodes; D.tris = scene->bvh.tris; D.verts = scene->bvh.verts.ptr;
	D.nodeCount = scene->bvh.nodeCount; D.slotCount = scene->bvh.slotCount;

	// "Slightly enlarge the bounding box" — include/mitsuba/render/gkdtree.h:1213-1220 (max uses the already moved min)
	for (int a = 0; a < 3; ++a) {
		float mn = scene->bvh.tightMin[a], mx = scene->bvh.tightMax[a];
		const float eps = MTS_KD_AABB_EPS;
		mn -= (mx - mn) * eps + eps;
		mx += (mx - mn) * eps + eps;
		scene->aabbMin[a] = mn; scene->aabbMax[a] = mx;
	}
	D.aabbMin = v3(scene->aabbMin[0], scene->aabbMin[1], scene->aabbMin[2]);
	D.aabbMax = v3(scene->aabbMax[0], scene->aabbMax[1], scene->aabbMax[2]);

	// constant environment: bounding sphere of the (kd-tree) AABB, radius x1.5 (src/emitters/constant.cpp:69-80;
	// AABB::getBSphere = centre + distance to max corner)
	for (size_t i = 0; i < scene->emitters.size(); ++i) {
		EmitterRec &e = scene->emitters[i];
		if (e.type != MTSB_EMITTER_CONSTANT && e.type != MTSB_EMITTER_ENVMAP) continue;
		V3 lo = D.aabbMin, hi = D.aabbMax;
		if (scene->hasExtra) {
			lo = vmin3(lo, v3(scene->extraMin[0], scene->extraMin[1], scene->extraMin[2]));
			hi = vmax3(hi, v3(scene->extraMax[0], scene->extraMax[1], scene->extraMax[2]));
		}
		V3 c = 0.5f * (lo + hi);
		V3 h = hi - c;
		e.sphereCenter = c;
		e.sphereRadius = fmaxf(MTS_EPSILON, sqrtf(dot(h, h)) * 1.5f);
	}
	// Scene::getBSphere() as Scene::initializeBidirectional leaves m_aabb (scene.cpp:394-422): kd-tree box + sensor box
	// (mtsb_scene_expand_bounds) + the boxes of the emitters (a point emitter adds its position; area emitters lie inside
	// the kd-tree box; the constant environment reports an invalid box).  The opacity query of volpath_simple ends at 2 R.
	{
		V3 lo = D.aabbMin, hi = D.aabbMax;
		if (scene->hasExtra) {
			lo = vmin3(lo, v3(scene->extraMin[0], scene->extraMin[1], scene->extraMin[2]));
			hi = vmax3(hi, v3(scene->extraMax[0], scene->extraMax[1], scene->extraMax[2]));
		}
		for (const EmitterRec &e : scene->emitters)
			if (e.type == MTSB_EMITTER_POINT) { lo = vmin3(lo, e.pos); hi = vmax3(hi, e.pos); }
		const V3 c = 0.5f * (lo + hi), hd = hi - c;
		D.sceneDiameter = 2.0f * sqrtf(dot(hd, hd));
	}
	// the environment map: tables to the device, their addresses into the EnvmapDev record, its address into the emitter
	for (size_t i = 0; i < scene->emitters.size(); ++i) {
		if (scene->emitters[i].type != MTSB_EMITTER_ENVMAP) continue;
		if ((int) i != scene->envmapEmitter) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_commit: envmap emitter %zu has no image (mtsb_scene_set_envmap)", i);
		MTSB_CHECK(scene->dEnvData.upload(scene->envData, st));
		MTSB_CHECK(scene->dEnvCdfRows.upload(scene->envCdfRows, st));
		MTSB_CHECK(scene->dEnvCdfCols.upload(scene->envCdfCols, st));
		MTSB_CHECK(scene->dEnvRowWeights.upload(scene->envRowWeights, st));
		EnvmapDev *E = (EnvmapDev *) scene->envRecord.data();
		E->data = scene->dEnvData.ptr; E->cdfRows = scene->dEnvCdfRows.ptr; E->cdfCols = scene->dEnvCdfCols.ptr; E->rowWeights = scene->dEnvRowWeights.ptr;
		MTSB_CHECK(scene->dEnvmap.upload(scene->envRecord, st));
		const unsigned char *dev = scene->dEnvmap.ptr;
		memcpy(&scene->emitters[i].pos.x, &dev, sizeof(dev));
	}
	MTSB_CHECK(scene->dMedia.upload(scene->media, st));
	D.media = scene->dMedia.ptr;
	MTSB_CHECK(scene->dEmitters.upload(scene->emitters, st));
	D.emitters = scene->dEmitters.ptr;
	MTSB_CUDA(cudaStreamSynchronize(st));
	scene->plain = scene->envmapEmitter < 0 && scene->texRecords.empty();
	for (const ShapeRec &sh : scene->shapes)
		if (sh.emitter >= 0 && (sh.type == SHAPE_SPHERE || sh.type == SHAPE_DISK)) scene->plain = false;     // emitter.cuh: LEAN
	scene->lean = !(ctx->shadeLean == 0) && scene->plain;
	scene->volLean = scene->lean;
	for (const BsdfRec &b : scene->bsdfs) {
		const int t = b.type & 0xff;
		const bool ok = t == MTSB_BSDF_DIFFUSE || t == MTSB_BSDF_NULL || t == MTSB_BSDF_DIELECTRIC ||
			(t == MTSB_BSDF_ROUGHCONDUCTOR && b.p[11] == 1.0f && b.p[12] != 0.0f);
		if (!ok) scene->lean = false;
		if (!ok && t != MTSB_BSDF_PASSTHROUGH) scene->volLean = false;
	}
	mtsb::l2_claim(scene, scene->bvh.accel.bytes());
	scene->committed = true;
	return 0;
}

int mtsb_scene_aabb(const mtsb_scene *scene, float mn[3], float mx[3]) {
	if (!scene || !scene->committed) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_aabb: scene not committed");
	for (int a = 0; a < 3; ++a) { mn[a] = scene->aabbMin[a]; mx[a] = scene->aabbMax[a]; }
	return 0;
}

size_t mtsb_scene_primitive_count(const mtsb_scene *scene) { return scene ? scene->primCount : 0; }
size_t mtsb_scene_shape_count(const mtsb_scene *scene) { return scene ? scene->shapes.size() : 0; }

int mtsb_scene_shape_offsets(const mtsb_scene *scene, uint32_t *out) {
	if (!scene || !out) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_shape_offsets: NULL argument");
	memcpy(out, scene->shapeOffsets.data(), scene->shapeOffsets.size() * sizeof(uint32_t));
	return 0;
}

int mtsb_scene_build_stats(const mtsb_scene *scene, double out[8]) {
	if (!scene || !scene->committed) MTSB_FAIL(MTSB_EINVAL, "mtsb_scene_build_stats: scene not committed");
	out[0] = scene->bvh.buildMs; out[1] = scene->bvh.nodeCount; out[2] = (double) scene->bvh.nodeBytes;
	out[3] = (double) scene->bvh.triBytes; out[4] = scene->bvh.sahCost; out[5] = scene->bvh.leafCount;
	out[6] = scene->bvh.depth; out[7] = scene->plyDecodeMs;
	return 0;
}

} // extern "C"
