// render.cuh — state of the wavefront path tracer behind the mtsb_render handle.
//
// Data layout in HBM (structure of arrays, one slot per path in flight, 16-byte elements):
//   rayO   {o.xyz, mint}          rayD  {d.xyz, maxt}            written by generate / shade, read by extend
//   hit    {t, u, v, slot}                                        written by extend, read by shade
//   thr    {throughput.rgb, eta}  L     {Li.rgb, alpha}
//   prev   {refN.xyz, bsdfPdf}    misc  {samplePos.xy, flags, depth}
//   rng    {px, py, sampleIndex, dim}
//   shO/shD shadow ray            shL   {contribution.rgb, path slot}   shadow queue written by shade, read by connect
// Queues of path slots (uint32) are compacted with warp-aggregated atomics.
#pragma once
#include "scene.cuh"

namespace mtsb {

struct PathState {
	float4 *rayO, *rayD, *hit, *thr, *L, *prev, *misc;
	uint4  *rng;
	float4 *shO, *shD, *shL;
	uint32_t *queueA, *queueB;     // live path slots (ping-pong)
	uint32_t *counters;            // [0] queue A count, [1] queue B count, [2] shadow count, [4..7] hand-over to the tail kernel (render_tail.cuh)
	unsigned long long *work;      // work counters of the persistent kernels [0] extend, [1] connect, [2] tail
	volatile uint32_t *feedback;   // pinned host word (mapped): bounce at which the last chunk went to the tail kernel
	unsigned long long *stats;     // [0] camera paths, [1] closest rays, [2] shadow rays, [3] depth sum
};

enum { PF_FROM_BSDF = 1, PF_SCATTERED = 2, PF_DELTA = 4 };

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

struct IntegratorDev { int type, maxDepth, rrDepth, strictNormals, hideEmitters, emitterSamples, bsdfSamples; };

} // namespace mtsb

struct mtsb_render {
	mtsb_scene *scene = nullptr;
	mtsb::CameraDev cam;
	mtsb::FilmDev film;
	mtsb::IntegratorDev integ;
	uint32_t seed = 0;
	size_t capacity = 0;           // paths in flight
	mtsb::PathState ps;
	mtsb::DevBuf<float> dFilm;
	mtsb::DevBuf<float4> bufF4;    // one allocation for all float4 arrays
	mtsb::DevBuf<uint4> bufRng;
	mtsb::DevBuf<uint32_t> bufQueue, bufCounters;
	mtsb::DevBuf<unsigned long long> bufWork;
	unsigned long long launches = 0;
	uint32_t *hostFeedback = nullptr;   // pinned + mapped, see PathState::feedback
	int tailPaths = 0;                  // live-path count below which the tail kernel takes over (0 = never; mtsb_render_set_tail)
	int waveBounces = 0;                // wavefront bounces enqueued per chunk before the tail kernel (adapted from hostFeedback)
	float *hostStage = nullptr;    // page-locked staging for mtsb_film_to_host(accumulate = 1), allocated on first use
};
