// common.cuh — internal types of libmtsb200 (device scene layout, small vector math, error plumbing).
// Nothing here crosses the C ABI (include/mtsb200.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <mutex>
#include <utility>
#include <vector>
#include "../../include/mtsb200.h"

namespace mtsb {

// ----------------------------------------------------------------------------- errors
void set_error(const char *fmt, ...);
#define MTSB_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { \
	mtsb::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); return MTSB_ECUDA; } } while (0)
#define MTSB_CHECK(expr) do { int r__ = (expr); if (r__ < 0) return r__; } while (0)
#define MTSB_FAIL(code, ...) do { mtsb::set_error(__VA_ARGS__); return (code); } while (0)

// ----------------------------------------------------------------------------- constants of the reference
// include/mitsuba/core/constants.h:26-29
#define MTS_EPSILON        1e-4f
#define MTS_SHADOW_EPSILON 1e-3f
#define MTS_KD_AABB_EPS    1e-3f   /* include/mitsuba/render/gkdtree.h:34 MTS_KD_AABB_EPSILON */
#define K_NO_TRIANGLE      0xFFFFFFFFu  /* KNoTriangleFlag, include/mitsuba/render/triaccel.h */

// ----------------------------------------------------------------------------- vector math
struct V3 { float x, y, z; };
__host__ __device__ inline V3 v3(float x, float y, float z) { V3 r = { x, y, z }; return r; }
__host__ __device__ inline V3 v3(float s) { V3 r = { s, s, s }; return r; }
__host__ __device__ inline V3 operator+(V3 a, V3 b) { return v3(a.x+b.x, a.y+b.y, a.z+b.z); }
__host__ __device__ inline V3 operator-(V3 a, V3 b) { return v3(a.x-b.x, a.y-b.y, a.z-b.z); }
__host__ __device__ inline V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
__host__ __device__ inline V3 operator*(V3 a, V3 b) { return v3(a.x*b.x, a.y*b.y, a.z*b.z); }
__host__ __device__ inline V3 operator*(V3 a, float s) { return v3(a.x*s, a.y*s, a.z*s); }
__host__ __device__ inline V3 operator*(float s, V3 a) { return v3(a.x*s, a.y*s, a.z*s); }
__host__ __device__ inline V3 operator/(V3 a, float s) { return v3(a.x/s, a.y/s, a.z/s); }
__host__ __device__ inline V3 operator/(V3 a, V3 b) { return v3(a.x/b.x, a.y/b.y, a.z/b.z); }
__host__ __device__ inline V3 &operator+=(V3 &a, V3 b) { a.x+=b.x; a.y+=b.y; a.z+=b.z; return a; }
__host__ __device__ inline V3 &operator*=(V3 &a, V3 b) { a.x*=b.x; a.y*=b.y; a.z*=b.z; return a; }
__host__ __device__ inline V3 &operator*=(V3 &a, float s) { a.x*=s; a.y*=s; a.z*=s; return a; }
__host__ __device__ inline V3 &operator/=(V3 &a, float s) { a.x/=s; a.y/=s; a.z/=s; return a; }
__host__ __device__ inline float dot(V3 a, V3 b) { return a.x*b.x + a.y*b.y + a.z*b.z; }
__host__ __device__ inline V3 cross(V3 a, V3 b) { return v3(a.y*b.z - a.z*b.y, a.z*b.x - a.x*b.z, a.x*b.y - a.y*b.x); }
__host__ __device__ inline float length(V3 a) { return sqrtf(dot(a, a)); }
__host__ __device__ inline V3 normalize(V3 a) { return a / length(a); }
__host__ __device__ inline float vmax(V3 a) { return fmaxf(a.x, fmaxf(a.y, a.z)); }
__host__ __device__ inline bool is_zero(V3 a) { return a.x == 0 && a.y == 0 && a.z == 0; }
__host__ __device__ inline float comp(V3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }
__host__ __device__ inline V3 vmin3(V3 a, V3 b) { return v3(fminf(a.x,b.x), fminf(a.y,b.y), fminf(a.z,b.z)); }
__host__ __device__ inline V3 vmax3(V3 a, V3 b) { return v3(fmaxf(a.x,b.x), fmaxf(a.y,b.y), fmaxf(a.z,b.z)); }

// ----------------------------------------------------------------------------- device scene layout (HBM)
// Per-shape record; shape order = Scene::getShapes() order = ShapeKDTree::addShape order.
enum { SHAPE_MESH = 0, SHAPE_RECT = 1, SHAPE_DISK = 2, SHAPE_SPHERE = 3 };
struct ShapeRec {
	uint32_t type;
	uint32_t primOffset;   // global primitive index of first primitive (m_shapeMap prefix)
	uint32_t primCount;
	uint32_t vtxBase;      // first vertex in the concatenated vertex arrays (mesh)
	uint32_t triBase;      // first triangle in the concatenated index array (mesh)
	uint32_t flags;        // bit0 normals, bit1 texcoords
	int32_t  bsdf;
	int32_t  emitter;
	uint32_t rect;         // index into rect table (SHAPE_RECT)
	float    invArea;      // 1/surface area (Shape::pdfPosition)
	int32_t  interior, exterior;   // participating media on the two sides (-1: none); Shape::isMediumTransition() = either set
};
enum { SHAPE_HAS_NORMALS = 1, SHAPE_HAS_UV = 2 };

// Analytic shapes share one record.  Rectangle (src/shapes/rectangle.cpp) and disk (src/shapes/disk.cpp): both transforms +
// the precomputed frame.  Sphere (src/shapes/sphere.cpp): toWorld / toObject = m_objectToWorld / m_worldToObject (scale
// removed, sphere.cpp:113-121), dpdu = m_center, dpdv.x = m_radius, pad[0] = flipNormals.
struct RectRec {
	float toWorld[12];     // 3x4 affine object->world
	float toObject[12];    // 3x4 affine world->object
	V3 dpdu, dpdv;         // m_dpdu, m_dpdv (rectangle.cpp:88-91)
	V3 n;                  // m_frame.n = normalize(cross(dpdu,dpdv)) (flipped if flipNormals)
	float invSurfaceArea;
	uint32_t pad[2];
};

struct EmitterRec {
	int32_t type; uint32_t shape;
	V3 radiance; V3 pos;
	V3 sphereCenter; float sphereRadius;   // constant env: scene bounding sphere (constant.cpp:69-94)
	float invSurfaceArea;                  // area
	float samplingWeight;                  // Emitter::m_samplingWeight
};

struct BsdfRec { int32_t type; float p[15]; };

// HomogeneousMedium + its phase function (src/medium/homogeneous.cpp, src/phase/{isotropic,hg}.cpp)
struct MediumRec {
	V3 sigmaA, sigmaS, sigmaT;
	int32_t strategy; float samplingDensity, mediumSamplingWeight;
	int32_t phase; float g;
};

// Everything the kernels read, passed by value as a kernel parameter.
struct SceneDev {
	// BVH8: 5 x uint4 per node (layout in bvh.cuh)
	const uint4  *nodes;
	// TriAccel slots in BVH leaf order: 3 x float4 per slot
	//   [0] = {k (bits), n_u, n_v, n_d}  [1] = {a_u, a_v, b_nu, b_nv}  [2] = {c_nu, c_nv, shapeIndex(bits), primIndex(bits)}
	const float4 *tris;
	// de-indexed shading record per slot (leaf order), 5 x float4 = 80 B:
	//   {p0.xyz p1.x} {p1.yz p2.xy} {p2.z n0.xyz} {n1.xyz n2.x} {n2.yz shape(bits) prim(bits)}
	// one 80-byte gather instead of the TriAccel tail -> shape -> indices -> 3 positions + 3 normals chain
	const float4 *verts;
	const ShapeRec *shapes;
	const RectRec  *rects;
	const float    *positions;   // 3 floats per vertex, all meshes
	const float    *normals;     // 3 floats per vertex (zeros where a mesh has none)
	const float    *texcoords;   // 2 floats per vertex
	const uint32_t *indices;     // 3 per triangle, local to the mesh (add vtxBase)
	const BsdfRec  *bsdfs;
	const EmitterRec *emitters;
	const MediumRec *media;
	const float    *emitterCdf;  // DiscreteDistribution m_cdf (n+1 entries), normalised
	const float    *triAreaCdf;  // per emitting mesh: concatenated normalised area CDFs (T+1 entries each)
	const uint32_t *triAreaCdfOffset; // per shape: offset into triAreaCdf or 0xFFFFFFFF
	uint32_t nodeCount, slotCount, shapeCount, emitterCount;
	int32_t  envEmitter;         // index of the constant environment emitter or -1
	float    emitterNorm;        // DiscreteDistribution normalisation (1/sum of weights)
	V3 aabbMin, aabbMax;         // enlarged kd-tree AABB (gkdtree.h:1217-1220)
	float sceneDiameter;         // 2 x Scene::getBSphere().radius (scene.cpp:394-422): far end of the opacity query, records.inl:128-136
	unsigned long long *counters; // optional traversal counters (nodes, tris, rays) or NULL
	uint32_t oneBits;            // 0x3F800000, read as a constant-bank operand by the box test (trace.cuh: q_to_unit)
	int fetchMinIdle;            // idle lanes a warp waits for before it pulls new rays from the work counter
	int triMinLanes;             // triangle phase: lanes needed for a second/third round in one iteration (trace.cuh)
};

// ----------------------------------------------------------------------------- RAII device buffer
template <typename T> struct DevBuf {
	T *ptr = nullptr; size_t count = 0;
	DevBuf() {}
	DevBuf(const DevBuf &) = delete; DevBuf &operator=(const DevBuf &) = delete;
	~DevBuf() { release(); }
	void release() { if (ptr) cudaFree(ptr); ptr = nullptr; count = 0; }
	int alloc(size_t n) {
		release();
		if (n == 0) n = 1;
		cudaError_t e = cudaMalloc((void **) &ptr, n * sizeof(T));
		if (e != cudaSuccess) { set_error("cudaMalloc(%zu bytes): %s", n * sizeof(T), cudaGetErrorString(e)); return MTSB_ENOMEM; }
		count = n; return 0;
	}
	int upload(const T *host, size_t n, cudaStream_t s) {
		MTSB_CHECK(alloc(n));
		if (n) MTSB_CUDA(cudaMemcpyAsync(ptr, host, n * sizeof(T), cudaMemcpyHostToDevice, s));
		return 0;
	}
	int upload(const std::vector<T> &v, cudaStream_t s) { return upload(v.data(), v.size(), s); }
	size_t bytes() const { return count * sizeof(T); }
};

} // namespace mtsb

struct mtsb_ctx {
	int device = 0;
	int smCount = 0;
	cudaStream_t stream = nullptr;
	// Experiment knobs, read ONCE from the environment in mtsb_ctx_create (never from a kernel-launch path: the plugin
	// calls the library from several mitsuba::Threads) and immutable afterwards.
	size_t hostChunk = (size_t) 1 << 21;   // MTSB_HOST_CHUNK: rays per pipelined chunk of the host-pointer queries
	size_t hostTaper = (size_t) 1 << 19;   // MTSB_HOST_TAPER: smallest piece at the end of a host batch (0 = no taper)
	int fetchMinIdle = 8;                  // MTSB_FETCH_MIN_IDLE (trace.cuh)
	int triMinLanes = 8;                   // MTSB_TRI_MIN_LANES (trace.cuh)
	int collapseDp = 1;                    // MTSB_COLLAPSE_DP: cost-driven BVH8 collapse (0 = greedy)
	float dpCNode = 1.0f, dpCPrim = 0.7f;  // MTSB_DP_CNODE / MTSB_DP_CPRIM
	int tailPaths = 0;                     // MTSB_TAIL_PATHS: live-path count below which the late-bounce kernel takes over (render)
	int shadeWaves = 0;                    // MTSB_SHADE_WAVES: shade grid = SMs x resident CTAs x this; 0 = one thread per path slot
	                                       // (measured on config 3: 277 M samples/s against 248-255 for 1-16 waves of a grid-stride loop)
	int suspendAfter = 0;                  // MTSB_SUSPEND_AFTER: warp iterations before the extend stage parks its remaining rays (0 = off,
	                                       // the default: measured +3 % on config 5 at 24, nothing elsewhere — DESIGN.md 10.2)
	int shadeLean = 1;                     // MTSB_SHADE_LEAN: scenes whose BSDFs allow it shade with the kernel compiled without the rest (bsdf.cuh)
	int useGraph = 1;                      // MTSB_GRAPH: bounce loop as a CUDA graph with a device-driven WHILE node (0 = plain launches)
	// L2 residency of the acceleration structure (cudaAccessPolicyWindow, see l2_window in scene.cuh)
	size_t l2PersistMax = 0;               // ceiling of the persisting set-aside (device maximum, or MTSB_L2_PERSIST in MB; 0 = off)
	size_t l2Persist = 0;                  // current cudaLimitPersistingL2CacheSize = the largest claim of a live scene
	size_t l2WindowMax = 0;                // cudaDevAttrMaxAccessPolicyWindowSize
	std::mutex l2Mutex;
	std::vector<std::pair<const void *, size_t> > l2Claims;   // (scene, bytes) of committed scenes whose structure fits
};
