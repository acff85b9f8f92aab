/*
 * mtsb200.h — flat C ABI of libmtsb200.so, the B200-native (sm_100a CUDA) implementation of Mitsuba's
 * ray-intersection + path-tracing hot path.
 *
 * This is the drop-in boundary: plain pointers and sizes, opaque handles, no C++/torch/Mitsuba types.
 * Every entry point cites the reference interface it stands in for (paths relative to the reference
 * tree, li779/mitsuba-im).  The Mitsuba-side adapter (an Integrator / ResponsiveIntegrator plugin) lives
 * in mitsuba-im_b200/plugin/ and is described in INTEGRATION.md.
 *
 * Conventions
 *   - every function returning int returns 0 on success and a negative MTSB_E* code on failure;
 *     mtsb_last_error() returns a thread-local description of the last failure.
 *   - the caller owns every host buffer passed in; the library copies what it needs before returning
 *     (mtsb_scene_* calls) or before mtsb_*_sync() returns (trace / film calls).
 *   - a handle may be used from one host thread at a time; different handles are independent.
 *   - there is NO CPU fallback: without a CUDA device every call fails with MTSB_ENODEVICE.
 *   - Float = float32, Spectrum = 3 x float32 linear RGB (reference build: SINGLE_PRECISION,
 *     SPECTRUM_SAMPLES=3).
 */
#ifndef MTSB200_H
#define MTSB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#  define MTSB_API
#else
#  define MTSB_API __attribute__((visibility("default")))
#endif

#define MTSB_ABI_VERSION 1

enum {
	MTSB_OK          =  0,
	MTSB_EINVAL      = -1,   /* bad argument / call order */
	MTSB_ENODEVICE   = -2,   /* no CUDA device / wrong architecture */
	MTSB_ECUDA       = -3,   /* CUDA runtime error (message in mtsb_last_error) */
	MTSB_EUNSUPPORTED= -4,   /* plugin outside the supported subset */
	MTSB_ENOMEM      = -5
};

typedef struct mtsb_ctx    mtsb_ctx;
typedef struct mtsb_scene  mtsb_scene;
typedef struct mtsb_render mtsb_render;

/* ---------------------------------------------------------------- records ------------------------ */

/* A ray, the fields of TRay that intersection queries read (include/mitsuba/core/ray.h:36-134):
 * o, mint, d, maxt.  dRcp is recomputed on the device; time is unused on this path. 32 bytes. */
typedef struct mtsb_ray {
	float o[3]; float mint;
	float d[3]; float maxt;
} mtsb_ray;

/* Closest-hit record: what ShapeKDTree's IntersectionCache holds after rayIntersectHavran
 * (include/mitsuba/render/skdtree.h:237-241) plus t.  `prim` is the kd-tree's GLOBAL primitive index
 * (shape prefix sum + index inside the shape, ShapeKDTree::build, src/librender/skdtree.cpp:68-70);
 * 0xFFFFFFFF and t = +inf on a miss.  For an analytic rectangle u,v are the object-space x,y of the
 * hit (Rectangle::rayIntersect temp record, src/shapes/rectangle.cpp:125-147).  16 bytes. */
typedef struct mtsb_hit {
	float t, u, v;
	uint32_t prim;
} mtsb_hit;
#define MTSB_NO_HIT 0xFFFFFFFFu

/* Full surface record = the fields of Intersection that MIPathTracer reads after
 * ShapeKDTree::fillIntersectionRecord<true> (include/mitsuba/render/skdtree.h:343-428). */
typedef struct mtsb_surface {
	float t, u, v; uint32_t prim;      /* as mtsb_hit                                        */
	uint32_t shape, prim_local;         /* shape index (addShape order) and index in the shape */
	float p[3];                         /* barycentric hit point                               */
	float ng[3];                        /* geoFrame.n                                          */
	float ns[3], s[3], tt[3];           /* shFrame.n / .s / .t                                 */
	float uv[2];                        /* its.uv                                              */
} mtsb_surface;

/* Light record: what ShapeKDTree::rayIntersect(ray, t, shape, n, uv) returns (src/librender/skdtree.cpp:144-204; used by
 * Scene::rayIntersect(ray, t, shape, n, uv), include/mitsuba/render/scene.h:216-222, for bidirectional vertex queries):
 * the UN-FLOORED adaptive epsilon of the shadow-ray overload, closest hit, the face normal normalize((p1-p0) x (p2-p0))
 * (rectangle: its frame normal) and the interpolated texture coordinates — (0,0) for a mesh without any.  32 bytes. */
typedef struct mtsb_light_hit {
	float t;                            /* +inf on a miss                                      */
	uint32_t shape;                     /* shape index (addShape order), MTSB_NO_HIT on a miss */
	float n[3];
	float uv[2];
	uint32_t pad;
} mtsb_light_hit;

/* BSDF subset (src/bsdfs/{diffuse,dielectric,plastic,roughplastic,roughconductor,conductor,roughdielectric,null,twosided}.cpp), constant textures only. */
enum {
	MTSB_BSDF_DIFFUSE        = 0,  /* p[0..2] reflectance                                                    */
	MTSB_BSDF_DIELECTRIC     = 1,  /* p[0] eta=intIOR/extIOR, p[1..3] specularReflectance, p[4..6] specularTransmittance */
	MTSB_BSDF_PLASTIC        = 2,  /* p[0] eta, p[1..3] diffuseReflectance, p[4..6] specularReflectance,
	                                  p[7] fdrInt, p[8] fdrExt, p[9] specularSamplingWeight, p[10] nonlinear(0/1) */
	MTSB_BSDF_ROUGHCONDUCTOR = 3,  /* p[0] alphaU, p[1] alphaV, p[2..4] eta, p[5..7] k, p[8..10] specularReflectance,
	                                  p[11] distribution (0 beckmann, 1 ggx), p[12] sampleVisible(0/1)           */
	MTSB_BSDF_NULL           = 4,  /* black diffuse given to BSDF-less emitter shapes (src/librender/shape.cpp:48-58) */
	MTSB_BSDF_CONDUCTOR      = 5,  /* p[0..2] eta, p[3..5] k (both already divided by extEta), p[6..8] specularReflectance
	                                  (src/bsdfs/conductor.cpp:167-199)                                       */
	MTSB_BSDF_ROUGHDIELECTRIC = 6, /* p[0] alphaU, p[1] alphaV, p[2] eta=intIOR/extIOR, p[3..5] specularReflectance,
	                                  p[6..8] specularTransmittance, p[11] distribution, p[12] sampleVisible
	                                  (src/bsdfs/roughdielectric.cpp:186-214)                                 */
	MTSB_BSDF_PASSTHROUGH    = 7,  /* `null` (src/bsdfs/null.cpp): index-matched boundary of a participating medium — the only
	                                  lobe is wo = -wi with weight 1 (ENull); emitter sampling sees through it              */
	MTSB_BSDF_ROUGHPLASTIC   = 8,  /* `roughplastic` (src/bsdfs/roughplastic.cpp:199-520), constant alpha: p[0] alpha, p[1] eta=intIOR/extIOR,
	                                  p[2..4] diffuseReflectance, p[5..7] specularReflectance, p[8] specularSamplingWeight,
	                                  p[9] Fdr = 1 - internal diffuse rough transmittance (RoughTransmittance::evalDiffuse of the
	                                  1/eta slice), p[10] nonlinear(0/1), p[11] distribution (0 beckmann, 1 ggx), p[12] sampleVisible;
	                                  p[13..14] are filled in by the library.  Needs mtsb_scene_set_bsdf_table.            */
	MTSB_BSDF_TWOSIDED_FLAG  = 0x100 /* OR-ed in: wrapped by `twosided` (src/bsdfs/twosided.cpp), same BSDF both sides */
};
typedef struct mtsb_bsdf_desc {
	int32_t type;
	float   p[15];
} mtsb_bsdf_desc;

/* Emitter subset (src/emitters/{area,point,constant}.cpp). */
enum {
	MTSB_EMITTER_AREA     = 0,  /* radiance; attached to shape `shape`            */
	MTSB_EMITTER_POINT    = 1,  /* intensity; position pos                        */
	MTSB_EMITTER_CONSTANT = 2,  /* radiance; bounding sphere filled in at commit  */
	MTSB_EMITTER_ENVMAP   = 3   /* `envmap` (src/emitters/envmap.cpp): latitude-longitude radiance map, see mtsb_scene_set_envmap */
};
typedef struct mtsb_emitter_desc {
	int32_t  type;
	uint32_t shape;              /* AREA: shape index                                   */
	float    radiance[3];        /* radiance / intensity, linear RGB                    */
	float    pos[3];             /* POINT: world position                               */
	float    sampling_weight;    /* Emitter::m_samplingWeight (default 1)               */
} mtsb_emitter_desc;

/* Perspective / thin-lens sensor: the matrices PerspectiveCameraImpl::configure builds
 * (src/sensors/perspective.cpp:126-180), row-major 4x4. */
typedef struct mtsb_camera {
	float sample_to_camera[16];  /* m_sampleToCamera (maps [0,1]^2 x near plane to camera space) */
	float camera_to_world[16];   /* m_worldTransform->eval(t) — assumed static                    */
	float near_clip, far_clip;
	float inv_resolution[2];     /* 1/film size (m_invResolution)                                 */
	float aperture_radius;       /* `thinlens` (src/sensors/thinlens.cpp:293-322): world-space lens radius;
	                                0 = pinhole `perspective`, no aperture sample is drawn         */
	float focus_distance;        /* ProjectiveCamera::getFocusDistance(), used when aperture_radius > 0 */
} mtsb_camera;

/* Film: the float bitmap of an ImageBlock (src/librender/imageblock.cpp:22-37) + its filter table. */
typedef struct mtsb_film_desc {
	int32_t width, height;       /* ImageBlock size (without border)                           */
	int32_t offset[2];           /* ImageBlock::getOffset()                                    */
	int32_t border;              /* ImageBlock::getBorderSize() = ceil(radius - 0.5)           */
	int32_t channels;            /* 4 = ESpectrumAlpha/ERGBA, 5 = ESpectrumAlphaWeight         */
	float   filter_radius;       /* ReconstructionFilter::getRadius()                          */
	float   filter_table[32];    /* m_values[0..31] (include/mitsuba/core/rfilter.h:76-77)     */
	int32_t crop_offset[2];      /* film crop window in sensor pixels                          */
	int32_t crop_size[2];
} mtsb_film_desc;

/* MTSB_INTEGRATOR_VOLPATH_SIMPLE: SimpleVolumetricPathTracer (src/integrators/path/volpath_simple.cpp:92-300) — path tracing
 * with homogeneous participating media, emitter sampling attenuated through index-matched boundaries, no MIS. */
/* MTSB_INTEGRATOR_VOLPATH: VolumetricPathTracer (src/integrators/path/volpath.cpp:84-425) — the same with multiple importance
 * sampling of emitter and phase-function / BSDF samples, emitters looked up through index-matched boundaries. */
enum { MTSB_INTEGRATOR_PATH = 0, MTSB_INTEGRATOR_DIRECT = 1, MTSB_INTEGRATOR_VOLPATH_SIMPLE = 2, MTSB_INTEGRATOR_VOLPATH = 3 };
typedef struct mtsb_integrator_desc {
	int32_t type;
	int32_t max_depth;           /* MonteCarloIntegrator maxDepth (-1 = unbounded)  (src/librender/integrator.cpp:191-226) */
	int32_t rr_depth;            /* rrDepth (default 5)                              */
	int32_t strict_normals;
	int32_t hide_emitters;
	int32_t emitter_samples;     /* direct only (src/integrators/direct/direct.cpp)  */
	int32_t bsdf_samples;        /* direct only                                      */
} mtsb_integrator_desc;

enum { MTSB_BUILDER_LBVH = 0, MTSB_BUILDER_PLOC = 1 };

/* ---------------------------------------------------------------- context ------------------------ */

MTSB_API int         mtsb_abi_version(void);
MTSB_API const char *mtsb_last_error(void);
MTSB_API int         mtsb_device_count(void);
/* One context per GPU (one CUDA stream + scratch allocator). */
MTSB_API int  mtsb_ctx_create(int device, mtsb_ctx **out);
MTSB_API void mtsb_ctx_destroy(mtsb_ctx *ctx);
MTSB_API int  mtsb_ctx_sync(mtsb_ctx *ctx);
/* cudaStream_t of the context, for callers that time with their own events. */
MTSB_API void *mtsb_ctx_stream(mtsb_ctx *ctx);

/* ---------------------------------------------------------------- scene -------------------------- */
/* Replaces ShapeKDTree::addShape/build (src/librender/skdtree.cpp:49-110) and the emitter/BSDF lookups
 * Scene::initialize performs (src/librender/scene.cpp:330-392).  Shapes must be added in the order of
 * Scene::getShapes(); that order defines shape indices and the global primitive index. */
MTSB_API int  mtsb_scene_create(mtsb_ctx *ctx, mtsb_scene **out);
MTSB_API void mtsb_scene_destroy(mtsb_scene *scene);
MTSB_API int  mtsb_scene_add_bsdf(mtsb_scene *scene, const mtsb_bsdf_desc *desc);          /* returns id >= 0 */
/* The external rough-transmittance table of a `roughplastic` BSDF: RoughTransmittance after setEta(eta) and setAlpha(alpha)
 * (src/bsdfs/rtrans.h:277-377) — n values over the warped angle cosTheta^(1/4) in [0, 1], evaluated on the device with
 * evalCubicInterp1D (src/libcore/spline.cpp:23-60).  The plugin builds it from the reference's own data/microfacet/{beckmann,ggx}.dat. */
MTSB_API int  mtsb_scene_set_bsdf_table(mtsb_scene *scene, int bsdf, const float *table, size_t n);
/* The image of an MTSB_EMITTER_ENVMAP emitter as the MIP pyramid the reference builds from it (envmap.cpp:152-178: RGB,
 * 2-lobed Lanczos downsampling, repeat / clamp boundaries, half-precision texels): `levels` images, level l of sizes[2l] x
 * sizes[2l+1] pixels, row-major RGB float triples.  The library derives the sampling tables of EnvironmentMap::configure
 * (:276-333) from level 0.  scale = m_scale; to_world = the emitter's world transform (a rotation). */
typedef struct mtsb_envmap_desc {
	int32_t levels;
	const int32_t *sizes;
	const float *const *data;
	float scale;
	float to_world[16];
} mtsb_envmap_desc;
/* `bitmap` texture (src/textures/bitmap.cpp over Texture2D, src/librender/texture.cpp:112-121) as the MIP pyramid the reference
 * builds (bitmap.cpp:283-305: RGB float texels, 2-lobed Lanczos downsampling under the texture's wrap modes).  wrap_u / wrap_v:
 * 0 repeat, 1 clamp, 2 mirror, 3 zero, 4 one; filter_type: 0 nearest, 1 bilinear, 2 trilinear, 3 ewa (max_anisotropy, 20 by
 * default; 1 for the other filters).  Sensor rays carry differentials: the first hit of a path computes uv partials
 * (Intersection::computePartials, src/librender/intersection.cpp:5-76) and takes the filtered lookup, later hits the
 * full-resolution one — as `path` / `direct` / `volpath*` do in the reference.  mtsb_scene_set_bsdf_texture binds a texture to
 * the `reflectance` of an MTSB_BSDF_DIFFUSE or to the `diffuseReflectance` of an MTSB_BSDF_PLASTIC (whose p[1..3] then only
 * matter for p[9]: the reference weighs the lobes by the texture's average, plastic.cpp:202-205); textured BSDFs go on triangle
 * meshes and rectangles. */
typedef struct mtsb_texture_desc {
	int32_t levels;
	const int32_t *sizes;
	const float *const *data;
	int32_t wrap_u, wrap_v, filter_type;
	float max_anisotropy;
	float uv_scale[2], uv_offset[2];
} mtsb_texture_desc;
MTSB_API int  mtsb_scene_add_texture(mtsb_scene *scene, const mtsb_texture_desc *desc);
MTSB_API int  mtsb_scene_set_bsdf_texture(mtsb_scene *scene, int bsdf, int texture);
MTSB_API int  mtsb_scene_set_envmap(mtsb_scene *scene, int emitter, const mtsb_envmap_desc *desc);
MTSB_API int  mtsb_scene_add_emitter(mtsb_scene *scene, const mtsb_emitter_desc *desc);    /* returns id >= 0 */
/* TriMesh (include/mitsuba/render/trimesh.h): positions 3V, normals 3V or NULL, texcoords 2V or NULL,
 * triangles 3T uint32.  emitter = -1 if none.  Returns the shape index. */
MTSB_API int  mtsb_scene_add_mesh(mtsb_scene *scene, const float *positions, const float *normals,
                                  const float *texcoords, const uint32_t *triangles,
                                  size_t vertex_count, size_t triangle_count, int bsdf, int emitter);
/* A binary PLY file decoded ON THE DEVICE (replaces, for the on-disk geometry path, the host-side parse of
 * src/shapes/ply.cpp:146-290 + TriMesh::configure/computeNormals, src/librender/trimesh.cpp:363-387,608-681): `file` is the
 * whole file image (header and body) in host memory, `bytes` its size; the call parses the text header, copies the image to
 * the device and returns; the vertex / face records are decoded by kernels at mtsb_scene_commit straight into the scene's
 * arrays.  to_world / to_world_inv: row-major 4x4 `toWorld` and its inverse (Transform holds both), NULL = identity.
 * flags: MTSB_PLY_FACE_NORMALS (`faceNormals`: drop / do not compute vertex normals), MTSB_PLY_FLIP_NORMALS (`flipNormals`).
 * Reads what the reference binds: float32 x y z, nx ny nz, u|s|texture_u, v|t|texture_v of `vertex`, the
 * vertex_indices / vertex_index list of `face` (uchar|uint count, int|uint index); other scalar properties are skipped;
 * a file without normals gets the angle-weighted vertex normals of computeNormals.  MTSB_EUNSUPPORTED (use the host
 * parser + mtsb_scene_add_mesh): ascii files, other list properties (binary_big_endian files are byte-swapped on the
 * device; for a file that mixes triangles and quads the host makes one pass over the count fields to index the records).
 * Vertex colours are not kept (the device subset has no `vertexcolors` texture).  The shape cannot carry an area emitter.
 * vertex_count / triangle_count (may be NULL) receive the sizes.  Returns the shape index. */
#define MTSB_PLY_FACE_NORMALS 1u
#define MTSB_PLY_FLIP_NORMALS 2u
MTSB_API int  mtsb_scene_add_ply(mtsb_scene *scene, const void *file, size_t bytes, const float to_world[16],
                                 const float to_world_inv[16], uint32_t flags, int bsdf,
                                 uint32_t *vertex_count, uint32_t *triangle_count);
/* A `serialized` mesh file (TriMesh::loadCompressed, src/librender/trimesh.cpp:176-296, and the `serialized` shape plugin,
 * src/shapes/serialized.cpp:148-210): `file` is the whole file image, shape_index the mesh inside it (`shapeIndex`, offset
 * dictionary of format versions 3 and 4).  The deflate stream is inflated on the host (zlib, sequential by nature); the
 * payload — one array per attribute, float or double — is converted, sent through to_world (a mirroring transform swaps the
 * winding, stored normals are re-normalised only behind a transform, as the plugin does), triangulated and given computed
 * normals by the same kernels as mtsb_scene_add_ply at commit.  flags as for mtsb_scene_add_ply.  Vertex colours are skipped. */
MTSB_API int  mtsb_scene_add_serialized(mtsb_scene *scene, const void *file, size_t bytes, int shape_index,
                                        const float to_world[16], const float to_world_inv[16], uint32_t flags, int bsdf,
                                        uint32_t *vertex_count, uint32_t *triangle_count);
MTSB_API int  mtsb_serialized_probe(const void *file, size_t bytes, int shape_index, uint32_t *vertex_count, uint32_t *triangle_count);
/* Header of a binary PLY file image only: the vertex and triangle counts mtsb_scene_add_ply would produce, or the error it
 * would return (no device work, no scene) — lets a caller that already holds the host-parsed mesh check the two agree. */
MTSB_API int  mtsb_ply_probe(const void *file, size_t bytes, uint32_t *vertex_count, uint32_t *triangle_count);
/* The decoded arrays of a mesh shape, copied back to the host after commit (checks, and callers that need the geometry
 * the device decoded): positions 3V, normals 3V, texcoords 2V, triangles 3T; any pointer may be NULL. */
MTSB_API int  mtsb_scene_read_mesh(const mtsb_scene *scene, int shape, float *positions, float *normals, float *texcoords,
                                   uint32_t *triangles);
/* Analytic rectangle (src/shapes/rectangle.cpp): object-to-world 4x4 row-major, flipNormals. */
MTSB_API int  mtsb_scene_add_rectangle(mtsb_scene *scene, const float object_to_world[16],
                                       int flip_normals, int bsdf, int emitter);
/* Analytic `disk` (src/shapes/disk.cpp: unit disk in the object-space xy plane, to_world without shear / non-uniform scale)
 * and `sphere` (src/shapes/sphere.cpp: object_to_world = m_objectToWorld with the scale removed as the constructor does,
 * :113-121; the quadratic is solved in double precision like the reference's).  Leaf primitives of the BVH like the
 * rectangle; as area emitters they are outside the GPU subset (MTSB_EUNSUPPORTED). */
MTSB_API int  mtsb_scene_add_disk(mtsb_scene *scene, const float to_world[16], int flip_normals, int bsdf, int emitter);
MTSB_API int  mtsb_scene_add_sphere(mtsb_scene *scene, const float object_to_world[16], float radius, int flip_normals, int bsdf, int emitter);
/* Before commit: an AABB that Scene::getAABB() holds beside the geometry when the constant environment emitter sizes its
 * bounding sphere — the sensor's AABB (Scene::initializeBidirectional, src/librender/scene.cpp:394-404;
 * ConstantBackgroundEmitter::createShape, src/emitters/constant.cpp:69-80).  Only the far end of environment shadow rays
 * (DirectSamplingRecord::dist) depends on it. */
MTSB_API int  mtsb_scene_expand_bounds(mtsb_scene *scene, const float min[3], const float max[3]);
/* Homogeneous participating medium (src/medium/homogeneous.cpp:156-400) with an isotropic or Henyey-Greenstein phase function
 * (src/phase/isotropic.cpp, hg.cpp).  sigma_a / sigma_s already carry the `scale` (densityMultiplier) of the Medium base
 * class.  strategy: 0 = balance (a random channel per distance sample), 1 = single / manual (sampling_density given);
 * medium_sampling_weight as the constructor derives it (:171-186). */
typedef struct mtsb_medium_desc {
	float   sigma_a[3], sigma_s[3];
	int32_t strategy;
	float   sampling_density;
	float   medium_sampling_weight;
	int32_t phase;               /* 0 isotropic, 1 Henyey-Greenstein */
	float   g;
} mtsb_medium_desc;
MTSB_API int  mtsb_scene_add_medium(mtsb_scene *scene, const mtsb_medium_desc *desc);
/* Shape::getInteriorMedium / getExteriorMedium (include/mitsuba/render/shape.h:427-431); -1 = none (vacuum) */
MTSB_API int  mtsb_scene_set_shape_media(mtsb_scene *scene, int shape, int interior, int exterior);
/* Builds TriAccel records, the BVH and the emitter tables on the device. */
MTSB_API int  mtsb_scene_commit(mtsb_scene *scene, int builder);
/* Scene AABB as ShapeKDTree::getAABB() returns it (enlarged, include/mitsuba/render/gkdtree.h:1217-1220). */
MTSB_API int  mtsb_scene_aabb(const mtsb_scene *scene, float min_out[3], float max_out[3]);
MTSB_API size_t mtsb_scene_primitive_count(const mtsb_scene *scene);
MTSB_API size_t mtsb_scene_shape_count(const mtsb_scene *scene);
/* shape_offsets[i] = global primitive index of shape i's first primitive; n+1 entries. */
MTSB_API int  mtsb_scene_shape_offsets(const mtsb_scene *scene, uint32_t *offsets_out);
/* Build statistics: out[0]=build ms (device), out[1]=wide nodes, out[2]=bytes of nodes, out[3]=bytes of
 * triangle records, out[4]=SAH cost of the wide tree, out[5]=leaf count, out[6]=max depth, out[7]=device ms of the PLY
 * decode kernels (mtsb_scene_add_ply), 0 without any. */
MTSB_API int  mtsb_scene_build_stats(const mtsb_scene *scene, double out[8]);

/* ---------------------------------------------------------------- ray queries -------------------- */
/* Batched ShapeKDTree::rayIntersect(ray, its) (src/librender/skdtree.cpp:112-142) without the record
 * fill: scene-AABB clip, adaptive epsilon (mint == Epsilon is scaled by max(|o|, Epsilon)), closest
 * primitive, inclusive t range.  rays/hits are HOST pointers unless on_device != 0.  Asynchronous on
 * the context stream when on_device != 0; synchronous (returns with hits filled) otherwise. */
MTSB_API int  mtsb_trace_closest(mtsb_scene *scene, const mtsb_ray *rays, size_t n, mtsb_hit *hits, int on_device);
/* Batched ShapeKDTree::rayIntersect(ray) (src/librender/skdtree.cpp:207-226): any hit; out[i] = 0/1.
 * The epsilon scale has no Epsilon floor in this overload. */
MTSB_API int  mtsb_trace_any(mtsb_scene *scene, const mtsb_ray *rays, size_t n, uint8_t *out, int on_device);
/* Batched closest hit + ShapeKDTree::fillIntersectionRecord<true> (include/mitsuba/render/skdtree.h:343-428). */
MTSB_API int  mtsb_trace_surface(mtsb_scene *scene, const mtsb_ray *rays, size_t n, mtsb_surface *out, int on_device);
/* replaces ShapeKDTree::rayIntersect(ray, t, shape, n, uv), src/librender/skdtree.cpp:144-204 */
MTSB_API int  mtsb_trace_light(mtsb_scene *scene, const mtsb_ray *rays, size_t n, mtsb_light_hit *out, int on_device);
/* mtsb_trace_closest without any device allocation.  NO COUNTERPART IN THE REFERENCE.
 * Same arguments, same results (bit for bit) as mtsb_trace_closest, but the call makes no cudaMalloc /
 * cudaFree / allocating library call: the persistent kernel's work counter and — for host pointers — the device staging
 * of rays and hit records live in one fixed static __device__ array of the library (3 MiB + 256 bytes per device).
 * Limit: n <= MTSB_NOALLOC_MAX_RAYS rays a call; a larger n returns MTSB_EINVAL with a message naming the limit.
 * The array is shared by every call on a device (whatever the scene or context) and keeps the last call's contents;
 * the call clears what must start at zero.  Calls that overlap on the GPU (two contexts on one device, two host threads)
 * RACE on it — serialise them; calls on one context are ordered by its stream.  on_device != 0: rays/hits are device
 * pointers, asynchronous on the context stream; otherwise host pointers, synchronous. */
#define MTSB_NOALLOC_MAX_RAYS 65536u
MTSB_API int  mtsb_trace_closest_noalloc(mtsb_scene *scene, const mtsb_ray *rays, size_t n, mtsb_hit *hits, int on_device);
/* Scene::rayIntersectAll x2 (src/librender/scene.cpp:716-730, 769-792): the kd-tree query, then the scene's "special
 * shapes" — the shapes Scene::initializeBidirectional creates for the sensor and the emitters (scene.cpp:394-422), which are
 * not in the kd-tree — tested linearly with the floored adaptive epsilon and no scene-box clip.  In the GPU subset these
 * are the aperture disk of a `thinlens` sensor (thinlens.cpp:522-534: mtsb_scene_add_special_disk with toWorld x
 * scale(apertureRadius), added by the caller, any time) and the bounding sphere of the constant environment emitter
 * (constant.cpp:69-94; known to the library after commit, appended after the caller's shapes as Scene does).  A hit on
 * special shape i reports prim = MTSB_SPECIAL_PRIM + i (u, v of the disk as for mtsb_scene_add_disk; 0, 0 on the sphere).
 * Used by bidirectional integrators only; `path` / `direct` / `volpath*` never call it. */
#define MTSB_SPECIAL_PRIM 0xF0000000u
MTSB_API int  mtsb_scene_add_special_disk(mtsb_scene *scene, const float to_world[16]);
MTSB_API int  mtsb_trace_closest_all(mtsb_scene *scene, const mtsb_ray *rays, size_t n, mtsb_hit *hits, int on_device);
MTSB_API int  mtsb_trace_any_all(mtsb_scene *scene, const mtsb_ray *rays, size_t n, uint8_t *out, int on_device);
/* Traversal counters of the last trace call when built with counters enabled, else zeros:
 * out[0]=nodes visited, out[1]=primitive tests, out[2]=rays. */
MTSB_API int  mtsb_trace_counters(mtsb_scene *scene, uint64_t out[4]);
/* Re-traces a batch with the counters enabled (closest or any hit) and returns them; the analogue of the
 * "Expected traversals/query" / "Expected prim. visits/query" figures the kd-tree build logs
 * (include/mitsuba/render/gkdtree.h:1252-1255), measured rather than estimated. */
MTSB_API int  mtsb_trace_count(mtsb_scene *scene, const mtsb_ray *rays, size_t n, int any_hit, int on_device, uint64_t out[4]);
/* Page-locked host memory so that host-pointer calls copy at PCIe speed (plain malloc memory also works). */
MTSB_API void *mtsb_host_alloc(size_t bytes);
MTSB_API void  mtsb_host_free(void *p);

/* Pointwise plugin math, for parity tests against BSDF::eval/pdf/sample
 * (include/mitsuba/render/bsdf.h:215-445) and Scene::sampleEmitterDirect / pdfEmitterDirect
 * (src/librender/scene.cpp:860-884,981-984) *without* the visibility test.  Host pointers.
 *   bsdf_in  : n x 8  {wi.xyz, wo.xyz, s1, s2}   local frame
 *   bsdf_out : n x 16 {eval.rgb, pdf, weight.rgb, wo'.xyz, pdf', eta, sampledType, 0,0,0}
 *   em_in    : n x 8  {p.xyz, refN.xyz, s1, s2}
 *   em_out   : n x 12 {value.rgb, pdf, d.xyz, dist, measure, pdfEmitterDirect, emitter id, 0} */
MTSB_API int  mtsb_eval_bsdf(mtsb_scene *scene, int bsdf, const float *bsdf_in, size_t n, float *bsdf_out);
/* Same, with the extra draw that EUsesSampler plugins take from bRec.sampler inside sample() (roughdielectric's
 * lobe selection, src/bsdfs/roughdielectric.cpp:562): s3 = n floats, NULL = 0.5 for every row. */
/* Scene::evalTransmittance (src/librender/scene.cpp:650-713) for a batch of segments: the transmittance between p1 and p2
 * through the scene's participating media, walking through index-matched (`null` BSDF) boundaries with the light-record
 * overload and switching media at each of them; 0 behind an occluder or on a "medium inconsistency".  flags bit 0 / bit 1 =
 * p1OnSurface / p2OnSurface (Epsilon at the start, the shadow-ray shortening at the end), medium = the medium at p1 (-1 =
 * none), max_interactions = the caller's `interactions` bound (-1 = none).  out_rgb: 3 floats per segment.  The same kernel
 * (connect_vol_kernel, csrc/render_vol.cuh) serves the volumetric integrators' connection queue. */
typedef struct mtsb_segment { float p1[3]; uint32_t flags; float p2[3]; int32_t medium; } mtsb_segment;
MTSB_API int  mtsb_eval_transmittance(mtsb_scene *scene, const mtsb_segment *segments, size_t n, int max_interactions, float *out_rgb);
/* Test support: the draws the render's sampler (independent or ldsampler) hands to samples first .. first+n-1 of pixel (px, py):
 * per sample `draws` next2D calls, then `draws` next1D calls; out[n][draws][3] = (2D.x, 2D.y, 1D). */
MTSB_API int  mtsb_render_sampler_dump(mtsb_render *r, uint32_t px, uint32_t py, uint32_t first, uint32_t n, uint32_t draws, float *out);
MTSB_API int  mtsb_eval_bsdf_s3(mtsb_scene *scene, int bsdf, const float *bsdf_in, const float *s3, size_t n, float *bsdf_out);
MTSB_API int  mtsb_eval_emitter(mtsb_scene *scene, const float *em_in, size_t n, float *em_out);

/* ---------------------------------------------------------------- rendering ---------------------- */
/* Replaces the per-sample loop SamplingIntegrator::renderBlock / ImageOrderIntegrator::render
 * (src/librender/integrator.cpp:141-189,336-402) + MIPathTracer::Li (src/integrators/path/path.cpp:119-294)
 * / MIDirectIntegrator::Li (src/integrators/direct/direct.cpp:146-309) + ImageBlock::put
 * (include/mitsuba/render/imageblock.h:161-221) with a wavefront pipeline.  The film lives on the
 * device as un-normalised sums in ImageBlock layout: (H+2b) x (W+2b) x channels float32. */
MTSB_API int  mtsb_render_create(mtsb_scene *scene, const mtsb_camera *camera, const mtsb_film_desc *film,
                                 const mtsb_integrator_desc *integrator, uint64_t seed, mtsb_render **out);
MTSB_API void mtsb_render_destroy(mtsb_render *r);
MTSB_API int  mtsb_render_set_camera(mtsb_render *r, const mtsb_camera *camera);
MTSB_API int  mtsb_render_clear(mtsb_render *r);
/* Switches the render from the independent sampler to the `sobol` sampler (src/samplers/sobol.cpp:84-251, scramble = 0): sample
 * index s of pixel (x, y) is the s-th point of the Sobol' sequence that falls into that pixel (sobol::look_up,
 * src/samplers/sobolseq.h:97-131, over the film size rounded up to a power of two), its first two dimensions place the sample
 * inside the pixel, every later draw of the path takes the next dimension.  tables = the reference's direction numbers as
 * 32-bit words: Matrices::matrices32 [1024 x 52], then rows m = 1..16 of vdc_sobol_matrices and of vdc_sobol_matrices_inv
 * ([52] uint64 each, low word first) — 56576 words; the plugin reads them out of the reference's sobol plugin.  Integrators
 * that request sample arrays (`direct` with emitterSamples / bsdfSamples other than 1) are refused with this sampler. */
#define MTSB_SOBOL_TABLE_WORDS (1024 * 52 + 2 * 16 * 52 * 2)
MTSB_API int  mtsb_render_set_sobol(mtsb_render *r, const uint32_t *tables, size_t words);
/* Switches the render to the `ldsampler` (src/samplers/ldsampler.cpp:78-239): for the first `dimension` 1D and 2D draws of a
 * sample the pixel's sample_count (rounded up to a power of two) points are a randomly scrambled van der Corput sequence / (0,2)
 * sequence (radicalInverse2Single, sample02Single, include/mitsuba/core/qmc.h) in shuffled order; later draws are independent.
 * The reference scrambles and shuffles from a sequential Random per pixel in scheduler order and documents that two of its own
 * runs differ; here both are keyed by a counter-based generator on (pixel, dimension, seed): the same point-set construction,
 * another realisation (per-sample parity through the reference's integrators with the oracle's twin sampler,
 * statistical parity against the stock plugin).  sample_count = the render's total samples per pixel. */
MTSB_API int  mtsb_render_set_ldsampler(mtsb_render *r, int dimension, size_t sample_count);
/* Sensor::getMedium(): the medium the camera sits in (-1 = none, the default); volpath_simple starts its paths there. */
MTSB_API int  mtsb_render_set_camera_medium(mtsb_render *r, int medium);
/* Gives the render a CUDA stream of its own (default: the context stream).  Renders are independent — own path state,
 * own film — so passes of several renders on one GPU then overlap; this is what lets the `threadCount` host threads of
 * ResponsiveIntegrator::render (include/mitsuba/render/integrator2.h:49-100), one render each, fill the device together. */
MTSB_API int  mtsb_render_private_stream(mtsb_render *r);
/* Optional, before the first pass: allocates the path state for passes of `spp_per_pass` samples per pixel over the whole
 * crop window and runs one 8x4-pixel pass (kernel loading, graph instantiation), then clears the film — the GPU
 * counterpart of the work the reference does in preprocess(), outside its render timer.  Blocks until done. */
MTSB_API int  mtsb_render_reserve(mtsb_render *r, uint32_t spp_per_pass);
/* Live-path count below which one persistent kernel carries the surviving paths through all their remaining bounces
 * instead of one wavefront round per bounce (csrc/render_tail.cuh).  0 = wavefront stages only, < 0 = the context's
 * default (131072).  Results do not depend on it (same estimator, same random numbers); it exists so that tests can
 * drive every path through either route. */
MTSB_API int  mtsb_render_set_tail(mtsb_render *r, int paths);
/* Render sample indices [spp_begin, spp_begin + spp_count) of every pixel in `rect` = {x0,y0,x1,y1}
 * (NULL = whole crop window); spp_total = the sampler's sampleCount (sets diffScaleFactor).
 * sample_stride/sample_phase select indices s with s % stride == phase (multi-GPU sample-pass split;
 * pass 1,0 for all).  Asynchronous on the context stream: the call queues its kernels and returns without waiting for
 * the device (no host polling of queue counts, whatever maxDepth is); mtsb_render_sync / mtsb_film_to_host /
 * mtsb_render_stats wait. */
MTSB_API int  mtsb_render_pass(mtsb_render *r, uint32_t spp_begin, uint32_t spp_count, uint32_t spp_total,
                               const int32_t rect[4], uint32_t sample_stride, uint32_t sample_phase);
MTSB_API int  mtsb_render_sync(mtsb_render *r);
/* Headless stand-in for the frame the im-mts preview puts on screen (src/im-mts/preview_opengl2.cpp:188-270; the GL path
 * itself is unavailable on a compute-only box): the un-normalised films of `n` renders (the worker framebuffers) are
 * stacked as the preview's additive blend does — rgb = sum_k film_k.rgb / max(sum_k spp[k], 1), alpha = 1 — on the device
 * and land as float RGBA in host memory (`pixel_count` = pixels of the film INCLUDING its border, as the GUI uploads it). */
MTSB_API int  mtsb_preview_stack(mtsb_render *const *renders, const double *spp, int n, float *rgba_out, size_t pixel_count);
/* Device pointer + float count of the film accumulation buffer (for NCCL reduce by the caller). */
MTSB_API int  mtsb_film_device_ptr(mtsb_render *r, void **ptr_out, size_t *float_count_out);
/* dst[i] += film[i] (accumulate != 0) or dst[i] = film[i]; dst is ImageBlock::getBitmap()->getFloatData(). */
MTSB_API int  mtsb_film_to_host(mtsb_render *r, float *dst, size_t float_count, int accumulate);
/* Copies the film into a page-locked staging buffer owned by the render (allocated on first use), waits for it and
 * returns the pointer: for callers that merge the film into memory they share with other threads under their own lock
 * (the ResponsiveIntegrator adapter; the reference's counterpart is ImageBlock::putAtomic, imageblock.h:223-318).
 * The pointer stays valid until the next mtsb_film_stage / mtsb_film_to_host(accumulate) / mtsb_render_destroy. */
MTSB_API int  mtsb_film_stage(mtsb_render *r, const float **stage_out, size_t *float_count_out);
/* The same hand-over split in two, so that the host's share of it overlaps the next pass: _begin queues, in stream order,
 * the copy of the film into the render's page-locked staging buffer and the clearing of the film — passes queued afterwards
 * accumulate into the cleared film; _wait blocks until the COPY has landed (not until the stream is idle) and returns the
 * staging buffer.  One flush at a time per render.  (ResponsiveIntegrator: un-normalised sums are added to the caller's
 * ImageBlock while the device renders on, include/mitsuba/render/integrator2.h:49-100.) */
MTSB_API int  mtsb_film_flush_begin(mtsb_render *r);
MTSB_API int  mtsb_film_flush_wait(mtsb_render *r, const float **stage_out, size_t *float_count_out);
/* Statistics of the passes since the last clear: out[0]=camera paths, out[1]=closest-hit rays,
 * out[2]=shadow rays, out[3]=sum of path depths (avgPathLength numerator, path.cpp:290-291),
 * out[4]=kernel launches. */
MTSB_API int  mtsb_render_stats(mtsb_render *r, uint64_t out[8]);
/* Radiance of single camera samples, for parity tests against MIPathTracer::Li with the `ctr`
 * sampler: samples n x 4 uint32 {px, py, sampleIndex, 0}; out n x 4 float {L.rgb, alpha}; host ptrs. */
MTSB_API int  mtsb_render_samples(mtsb_render *r, const uint32_t *samples, size_t n, uint32_t spp_total, float *out);
/* Same, and pos_out (n x 2 float, may be NULL) receives each sample's film position px + u (integrator.cpp:172), the
 * argument ImageBlock::put sees — so that a test can restate the splat on the host. */
MTSB_API int  mtsb_render_samples_ex(mtsb_render *r, const uint32_t *samples, size_t n, uint32_t spp_total, float *out, float *pos_out);

/* ---------------------------------------------------------------- multi-GPU film reduction ------- */
/* The path shards by SAMPLE INDEX: every GPU holds the scene + BVH and renders the indices s = phase (mod stride) of
 * every pixel into its own full-size film (mtsb_render_pass); nothing is exchanged while tracing.  The one exchange
 * step is the sum of the per-GPU films, the device-side form of Film::put(block) -> ImageBlock::put(const ImageBlock*)
 * -> Bitmap::accumulate (src/films/hdrfilm.cpp:391-393, include/mitsuba/render/imageblock.h:140-143) and of the
 * responsive caller summing its per-thread targets (src/mitsuba/im_render.cpp:225-247).  It runs as ONE NCCL
 * reduce over NVLink on the context streams.  NCCL is bound at run time (libnccl.so.2); without it these calls fail
 * with MTSB_EUNSUPPORTED — there is no host-side substitute. */
typedef struct mtsb_comm mtsb_comm;
#define MTSB_COMM_ID_BYTES 128
/* NCCL version code (e.g. 22809) or a negative error. */
MTSB_API int  mtsb_comm_version(void);
/* One process per GPU (torchrun): rank 0 makes an id, ships it to the others by any side channel, everyone joins. */
MTSB_API int  mtsb_comm_unique_id(uint8_t id[MTSB_COMM_ID_BYTES]);
MTSB_API int  mtsb_comm_create(mtsb_ctx *ctx, int nranks, int rank, const uint8_t id[MTSB_COMM_ID_BYTES], mtsb_comm **out);
/* One process driving n GPUs (the Mitsuba plugin with gpus=n): out receives n communicators, rank i on ctxs[i]. */
MTSB_API int  mtsb_comm_create_all(mtsb_ctx *const *ctxs, int n, mtsb_comm **out);
MTSB_API void mtsb_comm_destroy(mtsb_comm *comm);
/* film(root) = sum over ranks of film(rank); root < 0 = every rank receives the sum.  Asynchronous on the context
 * stream, ordered after the passes already queued there; mtsb_render_sync / mtsb_film_to_host wait for it. */
MTSB_API int  mtsb_film_reduce(mtsb_render *r, mtsb_comm *comm, int root);
/* Single-process form: the n per-GPU reduces issued as one NCCL group. */
MTSB_API int  mtsb_film_reduce_all(mtsb_render *const *renders, mtsb_comm *const *comms, int n, int root);

#ifdef __cplusplus
}
#endif
#endif /* MTSB200_H */
