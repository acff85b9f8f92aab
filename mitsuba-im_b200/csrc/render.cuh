// render.cuh — state of the wavefront path tracer behind the mtsb_render handle.
//
// Data layout in HBM (structure of arrays, one slot per path in flight, 16-byte elements):
//   rayO   {o.xyz, mint}          rayD  {d.xyz, maxt}            written by generate / shade, read by extend
//   hit    {t, u, v, slot}                                        written by extend, read by shade
//   thr    {throughput.rgb, eta}  L     {Li.rgb, alpha}
//   prev   {refN.xyz, bsdfPdf}    misc  {samplePos.xy, -, -}  (written by generate, read by splat)
//   rng    {px, py, sampleIndex, dim | flags << 16 | depth << 20}
//   shO/shD shadow ray            shL   {contribution.rgb, path slot}   shadow queue written by shade, read by connect
// Queues of path slots (uint32) are compacted with warp-aggregated atomics.
#pragma once
#include "scene.cuh"

namespace mtsb {

struct PathState {
	float4 *rayO, *rayD, *hit, *thr, *L, *prev, *misc;
	uint4  *rng;
	float4 *shO, *shD, *shL;
	float4 *shX;                   // fourth word of the connection queue (volpath only, render_vol.cuh), else NULL
	uint32_t *queueA, *queueB;     // live path slots (ping-pong); entries may carry QUEUE_RESUME (render.cu)
	uint2  *susp;                  // suspension records of the extend stage, MTSB_SUSP_WORDS per path slot (trace.cuh: Trav::save)
	uint32_t *counters;            // 16 words, see the CTR_* enum below
	unsigned long long *work;      // work counters of the persistent kernels [0] extend, [1] connect, [2] tail
	unsigned long long *stats;     // [0] camera paths, [1] closest rays, [2] shadow rays, [3] depth sum, [5] kernels run inside the bounce graph
};

// counters[]: the bounce loop is driven from the device (render_tail.cuh: begin_kernel / advance_kernel), so which queue
// is the live one and how deep the wavefront is live here, not in kernel arguments
enum { CTR_COUNT_A = 0, CTR_COUNT_B = 1, CTR_SHADOW = 2, CTR_PARITY = 3,
       CTR_TAIL = 4, CTR_TAIL_COUNT = 5, CTR_TAIL_PARITY = 6, CTR_TAIL_BOUNCE = 7, CTR_BOUNCE = 8, CTR_WORDS = 16 };

enum { PF_FROM_BSDF = 1, PF_SCATTERED = 2, PF_DELTA = 4 };
// rng.w of a path slot: sampler dimension (16 bits) | PF_* flags (4 bits) | depth (12 bits, saturating)
#define PATH_DIM(w)    ((w) & 0xffffu)
#define PATH_FLAGS(w)  (((w) >> 16) & 0xfu)
#define PATH_DEPTH(w)  ((w) >> 20)
#define PATH_PACK(dim, flags, depth) (((dim) & 0xffffu) | (((flags) & 0xfu) << 16) | (min((uint32_t) (depth), 4095u) << 20))

struct CameraDev {
	float s2c[16], c2w[16];
	float nearClip, farClip, invResX, invResY;
	float apertureRadius, focusDistance;   // thinlens; apertureRadius == 0: pinhole
};

struct FilmDev {
	float *data;
	int width, height, offX, offY, border, channels;
	float radius, scale;           // filter radius, m_scaleFactor = 31 / radius
	float table[32];
	int cropW, cropH;
};

struct IntegratorDev {
	int type, maxDepth, rrDepth, strictNormals, hideEmitters, emitterSamples, bsdfSamples, cameraMedium;
	const uint32_t *sobol;         // Sobol' direction numbers when the render uses that sampler (mtsb_render_set_sobol), else NULL
};

} // namespace mtsb

struct mtsb_render {
	mtsb_scene *scene = nullptr;
	mtsb::CameraDev cam;
	mtsb::FilmDev film;
	mtsb::IntegratorDev integ;
	uint32_t seed = 0;
	mtsb::DevBuf<uint32_t> dSobol; // matrices32 [1024 x 52], then vdc_sobol_matrices / _inv rows for m = 1..16 as (lo, hi) words
	int sobolLogRes = 0;
	int cameraMedium = -1;         // Sensor::getMedium() (volpath_simple starts its paths in it)
	cudaStream_t stream = nullptr; // the context stream, or a stream of its own (mtsb_render_private_stream)
	bool ownStream = false;
	size_t capacity = 0;           // paths in flight
	mtsb::PathState ps;
	mtsb::DevBuf<float> dFilm;
	mtsb::DevBuf<float4> bufF4;    // one allocation for all float4 arrays
	mtsb::DevBuf<uint4> bufRng;
	mtsb::DevBuf<uint32_t> bufQueue, bufCounters;
	mtsb::DevBuf<uint2> bufSusp;
	mtsb::DevBuf<unsigned long long> bufWork;
	unsigned long long launches = 0;
	int tailPaths = 0;                  // live-path count below which the tail kernel takes over (0 = never; mtsb_render_set_tail)
	// the bounce loop as a CUDA graph with a device-driven WHILE node (render_host.cuh: bounce_graph); rebuilt when the
	// path-state arrays move or the tail threshold changes
	cudaGraph_t bounceGraph = nullptr;
	cudaGraphExec_t bounceExec = nullptr;
	const void *graphKey = nullptr; int graphTail = -1;
	bool graphFailed = false;
	float4 *previewDev = nullptr, *previewHost = nullptr; size_t previewCount = 0;   // mtsb_preview_stack
	cudaEvent_t flushDone = nullptr;   // mtsb_film_flush_begin / _wait
	float *hostStage = nullptr;    // page-locked staging for mtsb_film_to_host(accumulate = 1), allocated on first use
};
