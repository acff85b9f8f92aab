// render_direct.cuh — MIDirectIntegrator::Li with any emitterSamples / bsdfSamples (included by render.cu).
//
// src/integrators/direct/direct.cpp:146-309.  With emitterSamples = bsdfSamples = 1 the estimator is the first bounce of
// the path tracer and runs through the same kernels (render_shade.cuh); for every other combination one camera path
// fans out into N shadow rays and M BSDF rays, which the wavefront handles as
//
//     generate -> extend (camera rays, queue A) -> direct_head -> connect (N shadow entries per path, atomic adds)
//              -> extend (BSDF rays: "secondary" slots behind the camera slots, queue B) -> direct_tail -> splat
//
// Weights: emitter sample  miWeight(pdf_em * fracLum, pdf_bsdf * fracBSDF) * weightLum,
//          BSDF sample     miWeight(pdf_bsdf * fracBSDF, pdf_em * fracLum) * weightBSDF          (direct.cpp:128-137,236-239,299-303)
// Random numbers: a technique with ONE sample draws it from the stream (rRec.nextSample2D()), with several from the
// sampler's pre-generated 2D ARRAY (Sampler::next2DArray; requested in the order emitter, BSDF: direct.cpp:140-145).  The
// counter sampler defines array entry j of array a as Philox(ctr = {px, py, sample, j}, key = {seed, 0x32446100 + a})
// (oracle/ctrsampler.cpp), so the reference's `direct` plugin sees the same numbers.
#pragma once

namespace mtsb {

__device__ inline void array2D(const SamplerState &rs, uint32_t arrayIdx, uint32_t j, float &a, float &b) {
	uint32_t w[4];
	philox4x32_10(rs.px, rs.py, rs.sample, j, rs.seed, 0x32446100u + arrayIdx, w);
	a = __uint_as_float((w[0] >> 9) | 0x3f800000u) - 1.0f;
	b = __uint_as_float((w[1] >> 9) | 0x3f800000u) - 1.0f;
}

struct DirectFan {
	uint32_t capacity;        // camera slots; secondary slot of (path, k) = capacity + path * M + k
	float fracLum, fracBSDF, weightLum, weightBSDF;
};

__device__ __forceinline__ float mi_weight_frac(float pdfA, float pdfB) {
	pdfA *= pdfA; pdfB *= pdfB;
	return pdfA / (pdfA + pdfB);
}

// after the camera rays have been traced: emitted radiance, N emitter samples (shadow queue), M BSDF samples (queue B)
__global__ void __launch_bounds__(RENDER_BLOCK)
direct_head_kernel(SceneDev S, PathState ps, IntegratorDev integ, uint32_t seed, DirectFan fan) {
	const uint32_t n = ps.counters[CTR_COUNT_A];
	const uint32_t N = (uint32_t) integ.emitterSamples, M = (uint32_t) integ.bsdfSamples;
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
		const uint32_t slot = QUEUE_SLOT(ps.queueA[i]);
		const float4 r1 = ps.rayD[slot], r0 = ps.rayO[slot], h = ps.hit[slot];
		const uint4 rg = ps.rng[slot];
		SamplerState rs; rs.px = rg.x; rs.py = rg.y; rs.sample = rg.z; rs.dim = PATH_DIM(rg.w); rs.seed = seed; rs.sobol = integ.sobol;
		const V3 o = v3(r0.x, r0.y, r0.z), d = v3(r1.x, r1.y, r1.z);
		const uint32_t hitSlot = __float_as_uint(h.w);
		V3 Li = v3(0.0f);
		float alpha = 0.0f;
		if (hitSlot == INVALID_SLOT) {                       // :159-166
			if (!integ.hideEmitters && S.envEmitter >= 0) {
				V3 rx = v3(0.0f), ry = v3(0.0f);             // the sensor ray's differentials (generate_kernel: prev / L)
				if (S.emitters[S.envEmitter].type == MTSB_EMITTER_ENVMAP) { const float4 a = ps.prev[slot], l = ps.L[slot]; rx = v3(a.x, a.y, a.z); ry = v3(l.x, l.y, l.z); }
				Li = env_radiance(S, d, true, rx, ry);
			}
		} else {
			alpha = 1.0f;
			Surf sf; fill_surface(S, hitSlot, h.x, h.y, h.z, o, d, sf);
			const int emitterId = S.shapes[sf.shape].emitter, bsdfId = S.shapes[sf.shape].bsdf;
			if (emitterId >= 0 && !integ.hideEmitters) Li += area_emitter_eval(S, emitterId, sf.ns, -d);      // :169-170
			const V3 wi = to_local(sf, -d);
			if (!(integ.strictNormals && dot(d, sf.ng) * wi.z >= 0)) {                                          // :178-191
				const BsdfRec bsdf = load_bsdf(S, bsdfId, sf, hitSlot, o, true, ps.prev[slot], ps.L + slot);
				const uint32_t btype = bsdf_type(bsdf);
				const V3 refN = (btype & (B_TRANSMISSION | B_BACK)) == 0 ? sf.ns : v3(0.0f);
				// ---- emitter sampling (:211-244)
				float sx = 0, sy = 0;
				if (N <= 1) next2D(rs, sx, sy);
				if ((btype & B_SMOOTH) && S.emitterCount > 0) {
					for (uint32_t k = 0; k < N; ++k) {
						if (N > 1) array2D(rs, 0u, k, sx, sy);
						DirectSample ds;
						const V3 value = scene_sample_emitter_direct(S, sf.p, refN, sx, sy, ds);
						if (is_zero(value)) continue;
						const V3 wo = to_local(sf, ds.d);
						const V3 bsdfVal = bsdf_eval(bsdf, wi, wo);
						if (is_zero(bsdfVal) || (integ.strictNormals && !(dot(sf.ng, ds.d) * wo.z > 0))) continue;
						const bool onSurface = S.emitters[ds.emitter].type != MTSB_EMITTER_POINT;
						const float bp = onSurface ? bsdf_pdf(bsdf, wi, wo) : 0;
						const V3 c = value * bsdfVal * (mi_weight_frac(ds.pdf * fan.fracLum, bp * fan.fracBSDF) * fan.weightLum);
						const uint32_t q = atomicAdd(ps.counters + CTR_SHADOW, 1u);
						ps.shO[q] = make_float4(sf.p.x, sf.p.y, sf.p.z, MTS_EPSILON);
						ps.shD[q] = make_float4(ds.d.x, ds.d.y, ds.d.z, ds.dist * (1 - MTS_SHADOW_EPSILON));
						ps.shL[q] = make_float4(c.x, c.y, c.z, __uint_as_float(slot));
					}
				}
				// ---- BSDF sampling (:250-306), first half: the rays
				if (M <= 1) next2D(rs, sx, sy);
				for (uint32_t k = 0; k < M; ++k) {
					if (M > 1) array2D(rs, N > 1 ? 1u : 0u, k, sx, sy);
					const float sz = bsdf_uses_sampler(bsdf) ? next1D(rs) : 0.0f;
					BsdfSample bs;
					const V3 bsdfVal = bsdf_sample(bsdf, wi, sx, sy, sz, bs);
					if (is_zero(bsdfVal)) continue;
					const V3 wo = to_world(sf, bs.wo);
					if (integ.strictNormals && dot(sf.ng, wo) * bs.wo.z <= 0) continue;
					const uint32_t s2 = fan.capacity + slot * M + k;
					ps.rayO[s2] = make_float4(sf.p.x, sf.p.y, sf.p.z, MTS_EPSILON);
					ps.rayD[s2] = make_float4(wo.x, wo.y, wo.z, INFINITY);
					ps.thr[s2] = make_float4(bsdfVal.x, bsdfVal.y, bsdfVal.z, 0.0f);
					ps.prev[s2] = make_float4(refN.x, refN.y, refN.z, bs.pdf);
					ps.misc[s2] = make_float4(0.0f, 0.0f, __uint_as_float((bs.type & B_DELTA) ? (uint32_t) PF_DELTA : 0u), __uint_as_float(slot));
					ps.queueB[atomicAdd(ps.counters + CTR_COUNT_B, 1u)] = s2;
				}
			}
		}
		ps.L[slot] = make_float4(Li.x, Li.y, Li.z, alpha);
	}
}

// after the BSDF rays have been traced (:274-303): emitter hit or environment, MIS against the emitter technique
__global__ void __launch_bounds__(RENDER_BLOCK)
direct_tail_kernel(SceneDev S, PathState ps, IntegratorDev integ, DirectFan fan) {
	const uint32_t n = ps.counters[CTR_COUNT_B];
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
		const uint32_t s2 = ps.queueB[i];
		const float4 r0 = ps.rayO[s2], r1 = ps.rayD[s2], h = ps.hit[s2], w = ps.thr[s2], prev = ps.prev[s2], misc = ps.misc[s2];
		const V3 o = v3(r0.x, r0.y, r0.z), d = v3(r1.x, r1.y, r1.z), refN = v3(prev.x, prev.y, prev.z);
		const uint32_t flags = __float_as_uint(misc.z), owner = __float_as_uint(misc.w), hitSlot = __float_as_uint(h.w);
		V3 value; float lumPdf = 0;
		if (hitSlot != INVALID_SLOT) {
			Surf sf; fill_surface(S, hitSlot, h.x, h.y, h.z, o, d, sf);
			const int emitterId = S.shapes[sf.shape].emitter;
			if (emitterId < 0) continue;
			value = area_emitter_eval(S, emitterId, sf.ns, -d);
			if (!(flags & PF_DELTA))
				lumPdf = emitter_pdf_direct(S, emitterId, refN, d, h.x, sf.ns, M_SOLID_ANGLE, o) * emitter_pdf_discrete(S, emitterId);
		} else {
			if (S.envEmitter < 0) continue;
			const EmitterRec &env = S.emitters[S.envEmitter];
			value = env_radiance(S, d);
			float nearT, farT;                            // fillDirectSamplingRecord (constant.cpp:238-256)
			if (!sphere_intersect(env.sphereCenter, env.sphereRadius, o, d, nearT, farT) || nearT > 0 || farT < 0) continue;
			if (!(flags & PF_DELTA)) {
				const V3 nrm = normalize(env.sphereCenter - (o + d * farT));
				lumPdf = emitter_pdf_direct(S, S.envEmitter, refN, d, farT, nrm, M_SOLID_ANGLE) * emitter_pdf_discrete(S, S.envEmitter);
			}
		}
		const float weight = mi_weight_frac(prev.w * fan.fracBSDF, lumPdf * fan.fracLum) * fan.weightBSDF;
		const V3 c = value * v3(w.x, w.y, w.z) * weight;
		float *L = (float *) (ps.L + owner);
		atomicAdd(L + 0, c.x); atomicAdd(L + 1, c.y); atomicAdd(L + 2, c.z);
	}
}

// between the two halves: queue B (the BSDF rays) becomes the live queue of the second extend launch
__global__ void fan_flip_kernel(PathState ps) {
	ps.counters[CTR_PARITY] = 1u; ps.counters[CTR_COUNT_A] = 0;
	ps.work[0] = 0; ps.work[1] = 0;
}

// connect with several shadow entries per path slot: the unoccluded contributions are added atomically
struct ConnectFanIO {
	static constexpr bool CAN_SUSPEND = false;
	SceneDev S; PathState ps;
	__device__ __forceinline__ void add(unsigned long long i) const {
		const float4 c = ps.shL[i];
		float *L = (float *) (ps.L + __float_as_uint(c.w));
		atomicAdd(L + 0, c.x); atomicAdd(L + 1, c.y); atomicAdd(L + 2, c.z);
	}
	__device__ __forceinline__ bool load(unsigned long long i, V3 &o, V3 &d, float &tmin, float &tmax, V3 &rcp) const {
		const float4 r0 = ps.shO[i], r1 = ps.shD[i];
		o = v3(r0.x, r0.y, r0.z); d = v3(r1.x, r1.y, r1.z);
		if (clip_ray(S, o, d, r0.w, r1.w, false, tmin, tmax, rcp)) return true;
		add(i);
		return false;
	}
	__device__ __forceinline__ void store(unsigned long long i, bool found, const HitRec &, const V3 &, const V3 &) const {
		if (!found) add(i);
	}
};

template <bool CURVED>
__global__ void __launch_bounds__(MTSB_TRACE_BLOCK, MTSB_TRACE_BLOCKS_PER_SM)
connect_fan_kernel(SceneDev S, PathState ps) {
	__shared__ uint2 shStack[MTSB_STACK_SHARED * MTSB_TRACE_BLOCK];
	const unsigned long long n = ps.counters[CTR_SHADOW];
	if (n == 0) return;
	ConnectFanIO io; io.S = S; io.ps = ps;
	trace_persistent<true, false, CURVED>(S, io, n, ps.work + 1, shStack + threadIdx.x);
	if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(ps.stats + 2, n);
}

} // namespace mtsb
