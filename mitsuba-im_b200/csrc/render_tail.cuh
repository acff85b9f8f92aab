// render_tail.cuh — the late-bounce kernel of the wavefront path tracer (included by render.cu).
//
// The wavefront stages pay a fixed price per bounce however few paths are left: a persistent traversal launch lasts as
// long as its LONGEST ray (≈ 0.2 ms of dependent node fetches on the 10 M-triangle scene), twice per
// bounce, plus the launch gaps.  Once the live queue is short, one kernel therefore takes every surviving path through
// ALL its remaining bounces: a warp owns 32 paths and alternates, warp-synchronously,
//     extend (closest hit, trace_warp<false>) -> shade (shade_path) -> connect (any hit, trace_warp<true>)
// refilling finished lanes from the queue, until the queue is drained.  The lanes of a warp may be at different
// depths; the floors of the remaining bounces overlap instead of adding up.  Same device functions as the wavefront
// stages, same random-number stream, so a path's result does not depend on which route it took (tests: the per-sample
// Li goldens run with the tail kernel forced on for every path and forced off).
#pragma once

namespace mtsb {

// counters[] slots shared with render_host.cuh
enum { CTR_COUNT_A = 0, CTR_COUNT_B = 1, CTR_SHADOW = 2, CTR_TAIL = 4, CTR_TAIL_COUNT = 5, CTR_TAIL_PARITY = 6, CTR_TAIL_BOUNCE = 7 };

// Between bounces (one thread): resets the work counters and the output counts; when the live count has dropped below
// `tailBelow` the queue is handed to the tail kernel — its count moves to counters[CTR_TAIL_COUNT] and the wavefront
// stages of this and all later bounces find an empty queue.  The bounce at which that happened goes to pinned host
// memory so that the host enqueues about that many wavefront bounces next time (a hint only: the tail kernel finishes
// whatever is left, however many paths that is).
__global__ void prep_kernel(PathState ps, uint32_t *countIn, uint32_t *countOut, uint32_t *countShadow, uint32_t parity,
		uint32_t bounce, uint32_t tailBelow, uint32_t force) {
	ps.work[0] = 0; ps.work[1] = 0;
	if (countOut) *countOut = 0;
	if (countShadow) *countShadow = 0;
	uint32_t *c = ps.counters;
	if (countIn && c[CTR_TAIL] == 0) {
		const uint32_t live = *countIn;
		if (force || live < tailBelow) {
			c[CTR_TAIL] = 1; c[CTR_TAIL_COUNT] = live; c[CTR_TAIL_PARITY] = parity; c[CTR_TAIL_BOUNCE] = bounce;
			*countIn = 0;
			ps.work[2] = 0;
			if (ps.feedback) { ps.feedback[1] = live; ps.feedback[0] = bounce | (force ? 0x80000000u : 0u); }
		}
	}
}

#ifndef MTSB_TAIL_MIN_BLOCKS
#define MTSB_TAIL_MIN_BLOCKS 4
#endif
__global__ void __launch_bounds__(MTSB_TRACE_BLOCK, MTSB_TAIL_MIN_BLOCKS)
tail_kernel(SceneDev S, PathState ps, IntegratorDev integ, uint32_t seed) {
	__shared__ uint2 shStack[MTSB_STACK_SHARED * MTSB_TRACE_BLOCK];
	const uint32_t *c = ps.counters;
	if (c[CTR_TAIL] == 0) return;
	const uint32_t n = c[CTR_TAIL_COUNT];
	if (n == 0) return;
	const uint32_t *queue = c[CTR_TAIL_PARITY] ? ps.queueB : ps.queueA;
	const unsigned lane = threadIdx.x & 31u, ltMask = (1u << lane) - 1u, FULL = 0xffffffffu;
	uint2 locStack[MTSB_STACK_LOCAL];
	Trav tv; tv.init(shStack + threadIdx.x, locStack);
	bool have = false, exhausted = false;
	uint32_t slot = 0;
	unsigned nClosest = 0, nShadow = 0;

	for (;;) {
		// ---- refill: lanes whose path has ended take the next queue entries (one atomic per warp)
		const unsigned idleMask = __ballot_sync(FULL, !have);
		if (!exhausted && idleMask) {
			const unsigned nIdle = __popc(idleMask);
			unsigned long long base = 0;
			if (lane == 0) base = atomicAdd(ps.work + 2, (unsigned long long) nIdle);
			base = __shfl_sync(FULL, base, 0);
			if (!have) {
				const unsigned long long idx = base + __popc(idleMask & ltMask);
				if (idx < n) { slot = queue[idx]; have = true; }
			}
			if (base + nIdle >= n) exhausted = true;
		}
		if (!__any_sync(FULL, have)) break;

		// ---- extend: ShapeKDTree::rayIntersect(ray, its) for the path's current ray
		if (have) {
			const float4 r0 = ps.rayO[slot], r1 = ps.rayD[slot];
			tv.o = v3(r0.x, r0.y, r0.z); tv.d = v3(r1.x, r1.y, r1.z);
			float tmax; V3 rcp;
			++nClosest;
			if (clip_ray(S, tv.o, tv.d, r0.w, r1.w, true, tv.tmin, tmax, rcp)) tv.begin(tmax, rcp);
			else tv.found = false;
		}
		trace_warp<false>(S, tv);
		ShadeOut so; so.alive = false; so.shadow = false;
		if (have) {
			ps.hit[slot] = tv.found ? make_float4(tv.hit.t, tv.hit.u, tv.hit.v, __uint_as_float(tv.hit.slot))
			                        : make_float4(INFINITY, 0, 0, __uint_as_float(INVALID_SLOT));
			// ---- shade
			shade_path(S, ps, integ, seed, slot, so);
		}
		// ---- connect: the emitter sample's visibility test
		bool unoccluded = false;
		if (so.shadow) {
			tv.o = v3(so.shO.x, so.shO.y, so.shO.z); tv.d = v3(so.shD.x, so.shD.y, so.shD.z);
			float tmax; V3 rcp;
			++nShadow;
			tv.found = false;
			if (clip_ray(S, tv.o, tv.d, so.shO.w, so.shD.w, false, tv.tmin, tmax, rcp)) tv.begin(tmax, rcp);
		}
		if (__any_sync(FULL, so.shadow)) trace_warp<true>(S, tv);
		if (so.shadow) unoccluded = !tv.found;
		if (unoccluded) {
			float4 L = ps.L[slot];
			L.x += so.shL.x; L.y += so.shL.y; L.z += so.shL.z;
			ps.L[slot] = L;
		}
		if (have && !so.alive) have = false;
	}
	// statistics (Normal rays traced / Shadow rays traced, skdtree.cpp:46-47)
	#pragma unroll
	for (int o = 16; o > 0; o >>= 1) { nClosest += __shfl_xor_sync(FULL, nClosest, o); nShadow += __shfl_xor_sync(FULL, nShadow, o); }
	if (lane == 0) { atomicAdd(ps.stats + 1, (unsigned long long) nClosest); atomicAdd(ps.stats + 2, (unsigned long long) nShadow); }
}

} // namespace mtsb
