// render.cu — wavefront path tracer: generate -> (extend -> shade -> connect)* -> splat.
//
// Replaces, for the GPU plugin subset, the per-sample loop of SamplingIntegrator::renderBlock
// (src/librender/integrator.cpp:141-189), PerspectiveCameraImpl::sampleRayDifferential
// (src/sensors/perspective.cpp:271-298), MIPathTracer::Li (src/integrators/path/path.cpp:119-294),
// MIDirectIntegrator::Li (src/integrators/direct/direct.cpp:146-309, via the same kernels with maxDepth = 2) and
// ImageBlock::put (include/mitsuba/render/imageblock.h:161-221).
//
// One loop iteration of Li is split at its ray cast: `shade` runs the tail of iteration k-1 (emitter hit by the BSDF
// sample + MIS, Russian roulette) followed by the head of iteration k (emitted radiance, emitter sampling, BSDF
// sampling).  Random numbers are consumed in the reference's order from a counter-based stream, so a path sees the
// same numbers as MIPathTracer does under oracle/ctrsampler.cpp.
#include "render.cuh"
#include "shading.cuh"
#include <algorithm>
#include <string.h>

namespace mtsb {

#define RENDER_BLOCK 128
#define INVALID_SLOT 0xffffffffu

// warp-aggregated append: one atomic per warp, ranks from the active mask
__device__ __forceinline__ uint32_t warp_append(uint32_t *counter) {
	const unsigned m = __activemask();
	const unsigned lane = threadIdx.x & 31u;
	const int leader = __ffs(m) - 1;
	uint32_t base = 0;
	if ((int) lane == leader) base = atomicAdd(counter, (uint32_t) __popc(m));
	base = __shfl_sync(m, base, leader);
	return base + __popc(m & ((1u << lane) - 1u));
}

__device__ __forceinline__ V3 xform_affine_point(const float *m, const V3 &p) {
	return v3(m[0]*p.x + m[1]*p.y + m[2]*p.z + m[3], m[4]*p.x + m[5]*p.y + m[6]*p.z + m[7], m[8]*p.x + m[9]*p.y + m[10]*p.z + m[11]);
}
__device__ __forceinline__ V3 xform_vector(const float *m, const V3 &v) {
	return v3(m[0]*v.x + m[1]*v.y + m[2]*v.z, m[4]*v.x + m[5]*v.y + m[6]*v.z, m[8]*v.x + m[9]*v.y + m[10]*v.z);
}
__device__ __forceinline__ V3 xform_projective_point(const float *m, const V3 &p) {   // Transform::operator()(Point)
	const float x = m[0]*p.x + m[1]*p.y + m[2]*p.z + m[3], y = m[4]*p.x + m[5]*p.y + m[6]*p.z + m[7];
	const float z = m[8]*p.x + m[9]*p.y + m[10]*p.z + m[11], w = m[12]*p.x + m[13]*p.y + m[14]*p.z + m[15];
	if (w == 1.0f) return v3(x, y, z);
	return v3(x, y, z) / w;
}

// ---------------------------------------------------------------------------------------------------
// generate: sample position + sensor ray for every path of the chunk
// ---------------------------------------------------------------------------------------------------
struct GenParams {
	uint32_t count;            // path slots in this chunk (padded to whole 8x4 pixel tiles)
	unsigned long long first;  // index of the chunk's first path within the pass
	int x0, y0, w, h;          // pixel rectangle, relative to the crop window
	uint32_t tilesX, pixPadded;
	uint32_t sppBegin, sppStride, sppPhase, seed;
	const uint32_t *samples;   // explicit {px,py,sampleIndex,0} list (mtsb_render_samples) or NULL
	const uint32_t *sobol;     // Sobol' sampler: direction numbers + enumeration tables (mtsb_render_set_sobol) or NULL
	uint32_t sobolLogRes;      // m_logResolution: log2 of the film size rounded up to a power of two (sobol.cpp:147-158)
	float diffScale;           // > 0: the scene looks ray differentials up (an `envmap` seen by sensor rays): 1 / sqrt(spp), integrator.cpp:145-146
};

__global__ void __launch_bounds__(RENDER_BLOCK)
generate_kernel(GenParams g, CameraDev cam, PathState ps, uint32_t *queue) {
	const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
	uint32_t px = 0, py = 0, sidx = 0;
	bool valid = j < g.count;
	if (valid) {
		if (g.samples) {
			px = g.samples[4 * j]; py = g.samples[4 * j + 1]; sidx = g.samples[4 * j + 2];
		} else {
			const unsigned long long gi = g.first + j;
			const uint32_t k = (uint32_t) (gi / g.pixPadded), q = (uint32_t) (gi % g.pixPadded);
			const uint32_t tile = q >> 5, within = q & 31u;
			const uint32_t lx = (tile % g.tilesX) * 8 + (within & 7u), ly = (tile / g.tilesX) * 4 + (within >> 3);
			valid = lx < (uint32_t) g.w && ly < (uint32_t) g.h;
			px = g.x0 + lx; py = g.y0 + ly;
			sidx = g.sppBegin + g.sppPhase + k * g.sppStride;
			if (!valid) ps.rng[j] = make_uint4(INVALID_SLOT, 0, 0, 0);
		}
	}
	// queue slot: ONE atomic per block (a per-warp append to the single counter serialises in L2 — a short kernel that
	// does nothing but stores waited on it: long-scoreboard stall 30-45 cycles per issue, profiles/r02a_ncu_full_render.json)
	__shared__ uint32_t warpCount[RENDER_BLOCK / 32], blockBase;
	const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	const unsigned vm = __ballot_sync(0xffffffffu, valid);
	if (lane == 0) warpCount[warp] = __popc(vm);
	__syncthreads();
	if (threadIdx.x == 0) {
		uint32_t total = 0;
		#pragma unroll
		for (int w = 0; w < RENDER_BLOCK / 32; ++w) { const uint32_t c = warpCount[w]; warpCount[w] = total; total += c; }
		blockBase = total ? atomicAdd(ps.counters + 0, total) : 0u;
	}
	__syncthreads();
	if (!valid) return;
	const uint32_t pos = blockBase + warpCount[warp] + __popc(vm & ((1u << lane) - 1u));
	SamplerState rs; rs.px = px; rs.py = py; rs.sample = sidx; rs.dim = 0; rs.seed = g.seed; rs.sobol = g.sobol;
	float u1, u2;
	uint32_t rngY = py, rngZ = sidx;
	if (g.sobol && !ld_tag(g.sobol)) {
		// SobolSampler::generate / setSampleIndex (sobol.cpp:172-215): the sidx-th point of the sequence that falls into this pixel
		// (sobol::look_up, sobolseq.h:97-131), whose first two dimensions give the position inside the pixel (:240-243)
		const uint32_t m = g.sobolLogRes;
		unsigned long long index = sidx;
		if (m > 1) {
			const uint32_t *vdc = g.sobol + 1024 * 52 + (m - 1) * 104, *inv = g.sobol + 1024 * 52 + 16 * 104 + (m - 1) * 104;
			index = (unsigned long long) sidx << (2 * m);
			unsigned long long delta = 0;
			uint32_t frame = sidx;
			for (uint32_t c = 0; frame; frame >>= 1, ++c)
				if (frame & 1) delta ^= (unsigned long long) vdc[2 * c] | ((unsigned long long) vdc[2 * c + 1] << 32);
			unsigned long long b = (((unsigned long long) px << m) | py) ^ delta;
			for (uint32_t c = 0; b; b >>= 1, ++c)
				if (b & 1) index ^= (unsigned long long) inv[2 * c] | ((unsigned long long) inv[2 * c + 1] << 32);
		}
		rngY = (uint32_t) (index >> 32); rngZ = (uint32_t) index;
		rs.py = rngY; rs.sample = rngZ;
		u1 = next1D(rs); u2 = next1D(rs);
		if (index != (unsigned long long) sidx) {
			const float res = (float) (1u << m);
			u1 = u1 * res - (float) px; u2 = u2 * res - (float) py;
		}
	} else next2D(rs, u1, u2);
	const float sx = (float) px + u1, sy = (float) py + u2;            // samplePos (integrator.cpp:172)
	const V3 nearP = xform_projective_point(cam.s2c, v3(sx * cam.invResX, sy * cam.invResY, 0.0f));
	V3 d, o;
	if (cam.apertureRadius > 0) {                                      // ThinLens::sampleRay (thinlens.cpp:293-322)
		float a1, a2; next2D(rs, a1, a2);                              // apertureSample (integrator.cpp:174-175)
		float ax, ay; square_to_disk_concentric(a1, a2, ax, ay);
		const V3 apertureP = v3(ax * cam.apertureRadius, ay * cam.apertureRadius, 0.0f);
		const V3 focusP = nearP * (cam.focusDistance / nearP.z);
		d = normalize(focusP - apertureP);
		o = v3(cam.c2w[0] * apertureP.x + cam.c2w[1] * apertureP.y + cam.c2w[3],
		       cam.c2w[4] * apertureP.x + cam.c2w[5] * apertureP.y + cam.c2w[7],
		       cam.c2w[8] * apertureP.x + cam.c2w[9] * apertureP.y + cam.c2w[11]);
	} else {                                                           // PerspectiveCameraImpl::sampleRay (perspective.cpp:208-231)
		d = normalize(nearP);
		o = v3(cam.c2w[3], cam.c2w[7], cam.c2w[11]);
	}
	const float invZ = 1.0f / d.z;
	const V3 dw = xform_vector(cam.c2w, d);
	float4 prev0 = make_float4(0, 0, 0, 0);
	if (g.diffScale > 0) {
		// sampleRayDifferential (perspective.cpp:271-298, thinlens.cpp:325-360) + RayDifferential::scaleDifferential (ray.h:163-168):
		// the directions through the neighbouring pixels, kept in `prev` / `L` until the first shade of the path reads them
		const V3 p0 = xform_projective_point(cam.s2c, v3(0.0f, 0.0f, 0.0f));
		const V3 dxv = xform_projective_point(cam.s2c, v3(cam.invResX, 0.0f, 0.0f)) - p0, dyv = xform_projective_point(cam.s2c, v3(0.0f, cam.invResY, 0.0f)) - p0;
		V3 rx, ry;
		if (cam.apertureRadius > 0) {
			float a1, a2; SamplerState r2 = rs; r2.dim -= 2; next2D(r2, a1, a2);       // the aperture sample drawn above
			float ax, ay; square_to_disk_concentric(a1, a2, ax, ay);
			const V3 apertureP = v3(ax * cam.apertureRadius, ay * cam.apertureRadius, 0.0f);
			const float fDist = cam.focusDistance / nearP.z;
			rx = normalize((nearP + dxv) * fDist - apertureP); ry = normalize((nearP + dyv) * fDist - apertureP);
		} else { rx = normalize(nearP + dxv); ry = normalize(nearP + dyv); }
		rx = xform_vector(cam.c2w, rx); ry = xform_vector(cam.c2w, ry);
		rx = dw + (rx - dw) * g.diffScale; ry = dw + (ry - dw) * g.diffScale;
		prev0 = make_float4(rx.x, rx.y, rx.z, 0.0f);
		ps.L[j] = make_float4(ry.x, ry.y, ry.z, 0.0f);
	}
	ps.rayO[j] = make_float4(o.x, o.y, o.z, cam.nearClip * invZ);
	ps.rayD[j] = make_float4(dw.x, dw.y, dw.z, cam.farClip * invZ);
	ps.thr[j] = make_float4(1, 1, 1, 1);
	ps.prev[j] = prev0;
	ps.misc[j] = make_float4(sx, sy, 0.0f, 0.0f);
	ps.rng[j] = make_uint4(px, rngY, rngZ, PATH_PACK(rs.dim, 0u, 1u));              // flags = 0, depth = 1; L is started by the first shade
	queue[pos] = j;
}

// ---------------------------------------------------------------------------------------------------
// extend / connect: persistent traversal over the queues
// ---------------------------------------------------------------------------------------------------
// Queue entries of the extend stage: a path slot, with QUEUE_RESUME set when the slot's ray was suspended by the
// previous launch (its traversal state waits in ps.susp) — the shade kernel passes such slots on instead of shading them.
#define QUEUE_RESUME   0x80000000u
#define QUEUE_SLOT(q)  ((q) & 0x7fffffffu)
#define SUSPENDED_SLOT 0xfffffffeu      // hit.w of a slot whose ray is parked

struct ExtendIO {
	static constexpr bool CAN_SUSPEND = true;
	const SceneDev &S;                 // the kernel parameter itself (constant bank), not a copy
	const PathState &ps;               // likewise: pointers stay in the constant bank
	uint32_t qOff;                     // 0: queue A, capacity: queue B (the two queues are one allocation, render_host.cuh)
	mutable unsigned fresh;            // rays started by this thread (resumed ones were counted by the launch that started them)
	__device__ __forceinline__ int load(unsigned long long i, V3 &o, V3 &d, float &tmin, float &tmax, V3 &rcp) const {
		const uint32_t q = ps.queueA[qOff + i], slot = QUEUE_SLOT(q);
		const float4 r0 = ps.rayO[slot], r1 = ps.rayD[slot];
		o = v3(r0.x, r0.y, r0.z); d = v3(r1.x, r1.y, r1.z);
		if (q & QUEUE_RESUME) return 2;
		++fresh;
		if (clip_ray(S, o, d, r0.w, r1.w, true, tmin, tmax, rcp)) return 1;
		ps.hit[slot] = make_float4(INFINITY, 0, 0, __uint_as_float(INVALID_SLOT));
		return 0;
	}
	__device__ __forceinline__ void resume(unsigned long long i, Trav &tv) const {
		tv.restore(ps.susp + (size_t) QUEUE_SLOT(ps.queueA[qOff + i]) * MTSB_SUSP_WORDS);
	}
	__device__ __forceinline__ void suspend(unsigned long long i, const Trav &tv) const {
		const uint32_t slot = QUEUE_SLOT(ps.queueA[qOff + i]);
		tv.save(ps.susp + (size_t) slot * MTSB_SUSP_WORDS);
		ps.hit[slot] = make_float4(0, 0, 0, __uint_as_float(SUSPENDED_SLOT));
	}
	__device__ __forceinline__ void store(unsigned long long i, bool found, const HitRec &h, const V3 &, const V3 &) const {
		const uint32_t slot = QUEUE_SLOT(ps.queueA[qOff + i]);
		ps.hit[slot] = found ? make_float4(h.t, h.u, h.v, __uint_as_float(h.slot))
		                     : make_float4(INFINITY, 0, 0, __uint_as_float(INVALID_SLOT));
	}
};

struct ConnectIO {
	static constexpr bool CAN_SUSPEND = false;
	SceneDev S; PathState ps;
	__device__ __forceinline__ void add(unsigned long long i) const {
		const float4 c = ps.shL[i];
		const uint32_t slot = __float_as_uint(c.w);
		float4 L = ps.L[slot];
		L.x += c.x; L.y += c.y; L.z += c.z;
		ps.L[slot] = L;
	}
	__device__ __forceinline__ bool load(unsigned long long i, V3 &o, V3 &d, float &tmin, float &tmax, V3 &rcp) const {
		const float4 r0 = ps.shO[i], r1 = ps.shD[i];
		o = v3(r0.x, r0.y, r0.z); d = v3(r1.x, r1.y, r1.z);
		if (clip_ray(S, o, d, r0.w, r1.w, false, tmin, tmax, rcp)) return true;
		add(i);                 // nothing can occlude a segment outside the scene box
		return false;
	}
	__device__ __forceinline__ void store(unsigned long long i, bool found, const HitRec &, const V3 &, const V3 &) const {
		if (!found) add(i);
	}
};

// Both kernels find the live queue through counters[CTR_PARITY] (flipped by advance_kernel between bounces), so the same
// launch — the same graph node — serves every bounce.
// suspendAfter: warp iterations a warp may spend on its remaining rays once the queue is empty before it parks them
// (0 = never: every ray is finished by the launch that started it)
template <bool CURVED>
__global__ void __launch_bounds__(MTSB_TRACE_BLOCK, MTSB_TRACE_BLOCKS_PER_SM)
extend_kernel(SceneDev S, PathState ps, uint32_t suspendAfter) {
	__shared__ uint2 shStack[MTSB_STACK_SHARED * MTSB_TRACE_BLOCK];
	const uint32_t parity = ps.counters[CTR_PARITY] & 1u;
	const unsigned long long n = ps.counters[parity];
	if (n == 0) return;
	ExtendIO io = { S, ps, parity ? (uint32_t) (ps.queueB - ps.queueA) : 0u, 0u };
	trace_persistent<false, false, CURVED>(S, io, n, ps.work + 0, shStack + threadIdx.x, suspendAfter);
	// "Normal rays traced" (skdtree.cpp:46,118): rays STARTED here
	unsigned fresh = io.fresh;
	#pragma unroll
	for (int o = 16; o > 0; o >>= 1) fresh += __shfl_xor_sync(0xffffffffu, fresh, o);
	if ((threadIdx.x & 31u) == 0 && fresh) atomicAdd(ps.stats + 1, (unsigned long long) fresh);
}

template <bool CURVED>
__global__ void __launch_bounds__(MTSB_TRACE_BLOCK, MTSB_TRACE_BLOCKS_PER_SM)
connect_kernel(SceneDev S, PathState ps) {
	__shared__ uint2 shStack[MTSB_STACK_SHARED * MTSB_TRACE_BLOCK];
	const unsigned long long n = ps.counters[CTR_SHADOW];
	if (n == 0) return;
	ConnectIO io; io.S = S; io.ps = ps;
	trace_persistent<true, false, CURVED>(S, io, n, ps.work + 1, shStack + threadIdx.x);
	if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(ps.stats + 2, n);
}

} // namespace mtsb
#include "render_shade.cuh"
#include "render_tail.cuh"
#include "render_direct.cuh"
#include "render_vol.cuh"
#include "render_host.cuh"
