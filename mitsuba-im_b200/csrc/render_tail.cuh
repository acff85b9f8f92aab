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

// ---- the device-driven bounce loop -----------------------------------------------------------------------------------
// One wavefront bounce is  extend -> shade -> connect -> advance.  Whether there is another one is decided ON THE DEVICE:
// `advance` (one thread) looks at the live count the shade kernel left behind and either flips the queues for the next
// bounce or hands the survivors to the tail kernel, and writes that decision into the condition of the CUDA-graph WHILE
// node that holds the four kernels (render_host.cuh: bounce_graph).  The host launches ONE graph per chunk and never
// reads a count back.  The same kernels also run as plain launches (MTSB_GRAPH=0): every bounded bounce is enqueued and
// the ones after the hand-over find empty queues.
struct LoopCtl {
	uint32_t tailBelow;                  // live-path count below which the tail kernel takes over (0: never)
	uint32_t maxBounces;                 // wavefront bounces at most (the integrator's depth bound, or "unbounded")
	cudaGraphConditionalHandle cond;     // WHILE-node condition
	uint32_t inGraph;                    // 0: plain launches, `cond` is not valid
};

__device__ __forceinline__ void loop_decide(const PathState &ps, const LoopCtl &ctl, uint32_t live, uint32_t parity, uint32_t bounce) {
	uint32_t *c = ps.counters;
	const bool go = live > 0 && live >= ctl.tailBelow && bounce < ctl.maxBounces;
	if (!go && c[CTR_TAIL] == 0) {
		// hand-over: the wavefront stages of any later (plain-launch) bounce find an empty queue
		c[CTR_TAIL] = 1; c[CTR_TAIL_COUNT] = live; c[CTR_TAIL_PARITY] = parity; c[CTR_TAIL_BOUNCE] = bounce;
		c[parity] = 0;
		ps.work[2] = 0;
	}
	if (ctl.inGraph) cudaGraphSetConditional(ctl.cond, go ? 1u : 0u);
}

// before the first bounce (counters were zeroed by a memset, generate_kernel filled queue A)
__global__ void begin_kernel(PathState ps, LoopCtl ctl) {
	ps.work[0] = 0; ps.work[1] = 0; ps.work[2] = 0;
	loop_decide(ps, ctl, ps.counters[CTR_COUNT_A], 0u, 0u);
}

// after the connect kernel of every bounce
__global__ void advance_kernel(PathState ps, LoopCtl ctl) {
	uint32_t *c = ps.counters;
	const uint32_t parity = c[CTR_PARITY] & 1u;
	const uint32_t live = c[parity ^ 1u];          // what shade appended to the other queue
	c[parity] = 0; c[CTR_SHADOW] = 0;              // this bounce's input queue becomes the next bounce's output queue
	ps.work[0] = 0; ps.work[1] = 0;
	c[CTR_PARITY] = parity ^ 1u;
	const uint32_t bounce = ++c[CTR_BOUNCE];
	ps.stats[5] += 4;
	loop_decide(ps, ctl, live, parity ^ 1u, bounce);
}

#ifndef MTSB_TAIL_MIN_BLOCKS
#define MTSB_TAIL_MIN_BLOCKS 4
#endif
template <bool CURVED>
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
				if (idx < n) { slot = queue[idx]; have = true; }      // may carry QUEUE_RESUME: first extend below picks the parked ray up
			}
			if (base + nIdle >= n) exhausted = true;
		}
		if (!__any_sync(FULL, have)) break;

		// ---- extend: ShapeKDTree::rayIntersect(ray, its) for the path's current ray
		if (have) {
			const bool parked = (slot & QUEUE_RESUME) != 0;
			slot = QUEUE_SLOT(slot);
			const float4 r0 = ps.rayO[slot], r1 = ps.rayD[slot];
			tv.o = v3(r0.x, r0.y, r0.z); tv.d = v3(r1.x, r1.y, r1.z);
			float tmax; V3 rcp;
			if (parked) tv.restore(ps.susp + (size_t) slot * MTSB_SUSP_WORDS);
			else {
				++nClosest;
				if (clip_ray(S, tv.o, tv.d, r0.w, r1.w, true, tv.tmin, tmax, rcp)) tv.begin(tmax, rcp);
				else tv.found = false;
			}
		}
		trace_warp<false, CURVED>(S, tv);
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
		if (__any_sync(FULL, so.shadow)) trace_warp<true, CURVED>(S, tv);
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
