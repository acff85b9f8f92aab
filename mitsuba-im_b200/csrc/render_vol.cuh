// render_vol.cuh — participating media on the wavefront (included by render.cu): SimpleVolumetricPathTracer::Li
// (src/integrators/path/volpath_simple.cpp:88-300) over homogeneous media (src/medium/homogeneous.cpp:238-330) with the
// isotropic / Henyey-Greenstein phase functions (src/phase/isotropic.cpp:62-68, hg.cpp:73-104), and the attenuated
// emitter connection Scene::sampleAttenuatedEmitterDirect -> Scene::evalTransmittance (src/librender/scene.cpp:650-713,
// 886-928), which walks through index-matched (`null` BSDF) medium boundaries.
//
// One loop iteration of Li = one `volshade` per `extend` result: this integrator has no multiple importance sampling, so
// nothing of iteration k waits for the ray of iteration k+1 (unlike MIPathTracer, render_shade.cuh).  The stages are
//     extend (closest hit, the same kernel as `path`) -> volshade -> connect_vol (transmittance march) -> advance
// Path state: the slot arrays of render.cuh; `prev` (the MIS record of `path`) carries the current medium and the
// RadianceQueryRecord flags instead.  Random numbers are drawn in the reference's order from the same counter stream.
#pragma once

namespace mtsb {

// prev.y of a path slot (bit mask)
enum { VF_EMITTED = 1,        // rRec.type & EEmittedRadiance
       VF_REST = 2,           // rRec.type & ERadianceNoEmission (direct/indirect surface and medium radiance travel together)
       VF_NULLCHAIN = 4, VF_SCATTERED = 8,
       // `volpath` only: what the previous iteration left for this one (volmis_path)
       VF_LOOK = 16,          // its phase / BSDF sample looks for an emitter (MIS-weighted): prev.z = sample pdf, {prev.w, misc.zw} = refN
       VF_DELTA = 32,         // ... and was drawn from a delta lobe (emitter pdf counts as 0)
       VF_CONTINUE = 64 };    // it ended with `continue` through an index-matched boundary: no Russian roulette, `scattered` unchanged

// entries of the connection queue (shO.w bits): {p1.xyz, bits} {p2.xyz, -} {contribution.rgb, path slot}
#define VQ_P1_SURFACE 1u
#define VQ_P2_SURFACE 2u
#define VQ_ALPHA      4u       // opacity query of RadianceQueryRecord::rayIntersect (records.inl:117-145): L.w = 1 - average(T)
#define VQ_LOOK       0x8000u  // rayIntersectAndLookForEmitter of `volpath` (volpath.cpp:369-425) beyond its first hit: {o, bits} {d, sample pdf}
                               // {throughput x transmittance so far, slot} {refN, -}; bit 31 of the pdf's sign = the sample was a delta lobe
#define VQ_MEDIUM(b)  ((int) (((b) >> 3) & 0xfffu) - 1)
#define VQ_MAXI(b)    ((b) >> 16)                  // 0xffff: no bound
#define VQ_PACK(flags, medium, maxi) ((uint32_t) (flags) | ((uint32_t) ((medium) + 1) << 3) | ((uint32_t) (maxi) << 16))

// math::fastexp / fastlog on Linux x86_64 go through double precision (include/mitsuba/core/math.h:177-199)
#ifdef MTSB_VOL_FP32_EXP
__device__ inline float mts_exp(float x) { return expf(x); }
__device__ inline float mts_log(float x) { return logf(x); }
#else
__device__ inline float mts_exp(float x) { return (float) exp((double) x); }
__device__ inline float mts_log(float x) { return (float) log((double) x); }
#endif

// HomogeneousMedium::evalTransmittance over a length (homogeneous.cpp:238-245)
__device__ inline V3 medium_transmittance(const MediumRec &m, float length) {
	const float neg = -length;
	return v3(m.sigmaT.x != 0 ? mts_exp(m.sigmaT.x * neg) : 1.0f, m.sigmaT.y != 0 ? mts_exp(m.sigmaT.y * neg) : 1.0f,
	          m.sigmaT.z != 0 ? mts_exp(m.sigmaT.z * neg) : 1.0f);
}

struct MediumSample { V3 p, transmittance; float pdfSuccess, pdfFailure; };

// HomogeneousMedium::sampleDistance for Ray(ray, 0, maxt) (homogeneous.cpp:247-330); strategy 0 = balance, 1 = single / manual
__device__ inline bool medium_sample_distance(const MediumRec &m, const V3 &o, const V3 &d, float maxt, SamplerState &rs, MediumSample &out) {
	float rand = next1D(rs), sampledDistance;
	float samplingDensity = m.samplingDensity;
	if (rand < m.mediumSamplingWeight) {
		rand /= m.mediumSamplingWeight;
		if (m.strategy == 0) {
			const int channel = min((int) (next1D(rs) * 3), 2);
			samplingDensity = comp(m.sigmaT, channel);
		}
		sampledDistance = -mts_log(1 - rand) / samplingDensity;
	} else sampledDistance = INFINITY;
	const float distSurf = maxt;
	bool success = true;
	out.p = o;
	if (sampledDistance < distSurf) {
		out.p = o + d * sampledDistance;
		if (out.p.x == o.x && out.p.y == o.y && out.p.z == o.z) success = false;      // no forward progress
	} else { sampledDistance = distSurf; success = false; }
	float pdfFailure, pdfSuccess;
	if (m.strategy == 0) {
		pdfFailure = 0; pdfSuccess = 0;
		#pragma unroll
		for (int i = 0; i < 3; ++i) {
			const float tmp = mts_exp(-comp(m.sigmaT, i) * sampledDistance);
			pdfFailure += tmp; pdfSuccess += comp(m.sigmaT, i) * tmp;
		}
		pdfFailure /= 3; pdfSuccess /= 3;
	} else {
		pdfFailure = mts_exp(-samplingDensity * sampledDistance);
		pdfSuccess = samplingDensity * pdfFailure;
	}
	out.transmittance = v3(mts_exp(m.sigmaT.x * -sampledDistance), mts_exp(m.sigmaT.y * -sampledDistance), mts_exp(m.sigmaT.z * -sampledDistance));
	out.pdfSuccess = pdfSuccess * m.mediumSamplingWeight;
	out.pdfFailure = m.mediumSamplingWeight * pdfFailure + (1 - m.mediumSamplingWeight);
	if (vmax(out.transmittance) < 1e-20f) out.transmittance = v3(0.0f);
	return success;
}

__device__ inline float phase_eval(const MediumRec &m, const V3 &wi, const V3 &wo) {
	const float inv4pi = 1.0f / (4.0f * MTS_PI);
	if (m.phase == 0) return inv4pi;                                      // isotropic.cpp:76-78
	const float g = m.g, temp = 1.0f + g * g + 2.0f * g * dot(wi, wo);     // hg.cpp:106-109
	return inv4pi * (1 - g * g) / (temp * sqrtf(temp));
}

// PhaseFunction::sample(pRec, sampler): direction only, the weight is 1 for both phase functions
__device__ inline V3 phase_sample(const MediumRec &m, const V3 &wi, float sx, float sy) {
	if (m.phase == 0) return square_to_uniform_sphere(sx, sy);
	const float g = m.g;
	float cosTheta;
	if (fabsf(g) < MTS_EPSILON) cosTheta = 1 - 2 * sx;
	else {
		const float sqrTerm = (1 - g * g) / (1 - g + 2 * g * sx);
		cosTheta = (1 + g * g - sqrTerm * sqrTerm) / (2 * g);
	}
	const float sinTheta = safe_sqrt(1.0f - cosTheta * cosTheta);
	float sinPhi, cosPhi; sincosf(2 * MTS_PI * sy, &sinPhi, &cosPhi);
	const V3 n = v3(-wi.x, -wi.y, -wi.z);                                  // Frame(-pRec.wi)
	V3 s, t; coordinate_system(n, s, t);
	return s * (sinTheta * cosPhi) + t * (sinTheta * sinPhi) + n * cosTheta;
}

// Intersection::getTargetMedium(d) (include/mitsuba/render/shape.h, records.inl): exterior where d leaves along the normal
__device__ __forceinline__ int target_medium(const ShapeRec &sh, const V3 &ng, const V3 &d) {
	return dot(d, ng) > 0 ? sh.exterior : sh.interior;
}

struct VolOut { bool alive; int nReq; float4 qO[2], qD[2], qL[2], qX[2]; };

__device__ __forceinline__ void vol_request(VolOut &out, const V3 &p1, const V3 &p2, uint32_t bits, const V3 &c, uint32_t slot) {
	const int k = out.nReq++;
	out.qO[k] = make_float4(p1.x, p1.y, p1.z, __uint_as_float(bits));
	out.qD[k] = make_float4(p2.x, p2.y, p2.z, 0.0f);
	out.qL[k] = make_float4(c.x, c.y, c.z, __uint_as_float(slot));
	out.qX[k] = make_float4(0, 0, 0, 0);
}
__device__ __forceinline__ void vol_look_request(VolOut &out, const V3 &o, const V3 &d, uint32_t bits, const V3 &c, uint32_t slot,
		float samplePdf, bool delta, const V3 &refN) {
	const int k = out.nReq++;
	out.qO[k] = make_float4(o.x, o.y, o.z, __uint_as_float(bits | VQ_LOOK));
	out.qD[k] = make_float4(d.x, d.y, d.z, delta ? -samplePdf : samplePdf);
	out.qL[k] = make_float4(c.x, c.y, c.z, __uint_as_float(slot));
	out.qX[k] = make_float4(refN.x, refN.y, refN.z, 0.0f);
}

__device__ __forceinline__ uint32_t vol_max_interactions(const IntegratorDev &integ, uint32_t depth) {
	const int m = integ.maxDepth - (int) depth - 1;        // volpath_simple.cpp:133,206; negative: never reached
	return m < 0 ? 0xffffu : (uint32_t) min(m, 0xfffe);
}

// One iteration of the while loop of SimpleVolumetricPathTracer::Li for the path in `slot`.
template <bool PLAIN = false, bool LEAN = false>
__device__ __forceinline__ void volshade_path(const SceneDev &S, const PathState &ps, const IntegratorDev &integ, uint32_t seed,
		uint32_t slot, VolOut &out) {
	const float4 r0 = ps.rayO[slot], r1 = ps.rayD[slot], h = ps.hit[slot];
	float4 thr4 = ps.thr[slot];
	const float4 prev = ps.prev[slot];
	const uint4 rg = ps.rng[slot];
	SamplerState rs; rs.px = rg.x; rs.py = rg.y; rs.sample = rg.z; rs.dim = PATH_DIM(rg.w); rs.seed = seed; rs.sobol = LEAN ? nullptr : integ.sobol;
	const V3 o = v3(r0.x, r0.y, r0.z), d = v3(r1.x, r1.y, r1.z);
	V3 thr = v3(thr4.x, thr4.y, thr4.z), Li = v3(0.0f);
	float eta = thr4.w;
	uint32_t depth = PATH_DEPTH(rg.w);
	const bool first = !(PATH_FLAGS(rg.w) & PF_FROM_BSDF);
	int medium = first ? integ.cameraMedium : __float_as_int(prev.x);
	uint32_t vf = first ? (uint32_t) (VF_EMITTED | VF_REST | VF_NULLCHAIN) : __float_as_uint(prev.y);
	const uint32_t hitSlot = __float_as_uint(h.w);
	const bool valid = hitSlot != INVALID_SLOT;
	out.alive = true; out.nReq = 0;

	Surf sf;
	ShapeRec sh;
	if (valid) { fill_surface(S, hitSlot, h.x, h.y, h.z, o, d, sf); sh = S.shapes[sf.shape]; }

	float alpha = 0.0f;
	if (first) {
		// RadianceQueryRecord::rayIntersect (records.inl:117-145): opacity through index-matched boundaries / the camera's medium
		const V3 far = o + d * S.sceneDiameter;
		if (valid) {
			if (sh.interior < 0 && sh.exterior < 0) alpha = 1.0f;
			else vol_request(out, sf.p, far, VQ_PACK(VQ_P1_SURFACE | VQ_ALPHA, target_medium(sh, sf.ng, d), 0xffffu), v3(0.0f), slot);
		} else if (medium >= 0) vol_request(out, o, far, VQ_PACK(VQ_ALPHA, medium, 0xffffu), v3(0.0f), slot);
		if (integ.maxDepth == 1) vf &= VF_EMITTED;                  // :106-107
	}

	MediumSample ms;
	bool inMedium = false;
	if (medium >= 0) inMedium = medium_sample_distance(S.media[medium], o, d, h.x, rs, ms);
	if (inMedium) {
		// ---- medium interaction (:119-165)
		const MediumRec &m = S.media[medium];
		thr *= m.sigmaS * ms.transmittance / ms.pdfSuccess;
		if (vf & VF_REST) {
			float sx, sy; next2D(rs, sx, sy);
			if (S.emitterCount > 0) {
				DirectSample ds;
				const V3 value = scene_sample_emitter_direct<PLAIN>(S, ms.p, v3(0.0f), sx, sy, ds);
				if (ds.pdf != 0 && !is_zero(value)) {
					const V3 c = thr * value * phase_eval(m, -d, ds.d);
					const bool onSurface = S.emitters[ds.emitter].type != MTSB_EMITTER_POINT;
					vol_request(out, ms.p, ds.p, VQ_PACK(onSurface ? VQ_P2_SURFACE : 0u, medium, vol_max_interactions(integ, depth)), c, slot);
				}
			}
		}
		if (((int) depth + 1 >= integ.maxDepth && integ.maxDepth > 0) || !(vf & VF_REST)) out.alive = false;
		else {
			float sx, sy; next2D(rs, sx, sy);
			const V3 wo = phase_sample(m, -d, sx, sy);                  // weight 1
			ps.rayO[slot] = make_float4(ms.p.x, ms.p.y, ms.p.z, 0.0f);  // ray.mint = 0 (:160)
			ps.rayD[slot] = make_float4(wo.x, wo.y, wo.z, INFINITY);
			vf &= ~VF_NULLCHAIN; vf |= VF_SCATTERED;
		}
	} else {
		// ---- surface interaction, reached with probability pdfFailure (:166-270)
		if (medium >= 0) thr *= ms.transmittance / ms.pdfFailure;
		if (!valid) {
			if ((vf & VF_EMITTED) && (!integ.hideEmitters || (vf & VF_SCATTERED)) && S.envEmitter >= 0) {
				V3 rx = v3(0.0f), ry = v3(0.0f);           // only the sensor ray carries differentials (generate_kernel: prev / L)
				if (!PLAIN && first && S.emitters[S.envEmitter].type == MTSB_EMITTER_ENVMAP) { const float4 l = ps.L[slot]; rx = v3(prev.x, prev.y, prev.z); ry = v3(l.x, l.y, l.z); }
				V3 value = thr * env_radiance<PLAIN>(S, d, first, rx, ry);
				// rRec.medium->evalTransmittance(ray): the ray's own [mint, maxt] = an infinite segment
				if (medium >= 0) value *= medium_transmittance(S.media[medium], INFINITY);
				Li += value;
			}
			out.alive = false;
		} else {
			if (sh.emitter >= 0 && (vf & VF_EMITTED) && (!integ.hideEmitters || (vf & VF_SCATTERED)))
				Li += thr * area_emitter_eval(S, sh.emitter, sf.ns, -d);
			const V3 wi = to_local(sf, -d);
			if (integ.strictNormals && -dot(sf.ng, d) * wi.z < 0) out.alive = false;          // :192-195 (no depth test here: recursiveType below ends the path)
			if (out.alive) {
				const BsdfRec bsdf = load_bsdf<PLAIN>(S, sh.bsdf, sf, hitSlot, o, first, prev, ps.L + slot);
				const uint32_t btype = bsdf_type(bsdf);
				// ---- direct illumination (:201-227): no MIS, one attenuated emitter sample
				if ((vf & VF_REST) && (btype & B_SMOOTH)) {
					const V3 refN = (btype & (B_TRANSMISSION | B_BACK)) == 0 ? sf.ns : v3(0.0f);
					float sx, sy; next2D(rs, sx, sy);
					if (S.emitterCount > 0) {
						DirectSample ds;
						const V3 value = scene_sample_emitter_direct<PLAIN>(S, sf.p, refN, sx, sy, ds);
						if (ds.pdf != 0 && !is_zero(value)) {
							const V3 wo = to_local(sf, ds.d);
							const V3 bsdfVal = bsdf_eval<LEAN>(bsdf, wi, wo);
							if (!is_zero(bsdfVal) && (!integ.strictNormals || dot(sf.ng, ds.d) * wo.z > 0)) {
								const int med = (sh.interior >= 0 || sh.exterior >= 0) ? target_medium(sh, sf.ng, ds.d) : medium;   // scene.cpp:919-920
								const bool onSurface = S.emitters[ds.emitter].type != MTSB_EMITTER_POINT;
								vol_request(out, sf.p, ds.p, VQ_PACK(VQ_P1_SURFACE | (onSurface ? VQ_P2_SURFACE : 0u), med, vol_max_interactions(integ, depth)),
									thr * value * bsdfVal, slot);
							}
						}
					}
				}
				// ---- BSDF sampling (:233-270)
				float sx, sy; next2D(rs, sx, sy);
				const float sz = bsdf_uses_sampler(bsdf) ? next1D(rs) : 0.0f;
				BsdfSample bs;
				const V3 bsdfVal = bsdf_sample<LEAN>(bsdf, wi, sx, sy, sz, bs);
				if (is_zero(bsdfVal)) out.alive = false;
				else {
					uint32_t recursive = 0;
					if (((int) depth + 1 < integ.maxDepth || integ.maxDepth < 0) && (vf & VF_REST)) recursive |= VF_REST;
					if (((int) depth < integ.maxDepth || integ.maxDepth < 0) && (vf & VF_REST) && (bs.type & B_DELTA) &&
							(!(bs.type & B_NULL) || (vf & VF_NULLCHAIN))) {
						recursive |= VF_EMITTED; vf |= VF_NULLCHAIN;
					} else if (bs.type != B_NULL) vf &= ~VF_NULLCHAIN;
					if (recursive == 0) out.alive = false;
					else {
						vf = (vf & ~(uint32_t) (VF_EMITTED | VF_REST)) | recursive;
						const V3 wo = to_world(sf, bs.wo);
						if (dot(sf.ng, wo) * bs.wo.z <= 0 && integ.strictNormals) out.alive = false;
						else {
							thr *= bsdfVal; eta *= bs.eta;
							if (sh.interior >= 0 || sh.exterior >= 0) medium = target_medium(sh, sf.ng, wo);
							ps.rayO[slot] = make_float4(sf.p.x, sf.p.y, sf.p.z, MTS_EPSILON);
							ps.rayD[slot] = make_float4(wo.x, wo.y, wo.z, INFINITY);
							if (bs.type != B_NULL) vf |= VF_SCATTERED;
						}
					}
				}
			}
		}
	}
	if (out.alive) {
		if (depth++ >= (uint32_t) integ.rrDepth) {                       // :272-283
			const float q = fminf(vmax(thr) * eta * eta, 0.95f);
			if (next1D(rs) >= q) out.alive = false;
			else thr /= q;
		}
	}
	if (out.alive && !((int) depth <= integ.maxDepth || integ.maxDepth < 0)) out.alive = false;     // while condition (:113)

	if (first) ps.L[slot] = make_float4(Li.x, Li.y, Li.z, alpha);
	else if (!is_zero(Li)) {
		// component-wise: a connection of the previous bounce never overlaps this kernel, but keep L.w untouched
		float4 L4 = ps.L[slot];
		L4.x += Li.x; L4.y += Li.y; L4.z += Li.z;
		ps.L[slot] = L4;
	}
	ps.rng[slot] = make_uint4(rg.x, rg.y, rg.z, PATH_PACK(rs.dim, PF_FROM_BSDF, depth));
	if (out.alive) {
		ps.thr[slot] = make_float4(thr.x, thr.y, thr.z, eta);
		ps.prev[slot] = make_float4(__int_as_float(medium), __uint_as_float(vf), 0.0f, 0.0f);
	}
}

// One iteration of VolumetricPathTracer::Li (src/integrators/path/volpath.cpp:84-343) — the variant WITH multiple importance
// sampling.  Its phase / BSDF sample is followed by rayIntersectAndLookForEmitter (:369-425), which (1) finds the first
// intersection, the one the next iteration starts from, and (2) keeps walking through index-matched boundaries for an
// emitter whose radiance is weighted against the emitter-sampling density.  Step (1) is the wavefront's `extend`; this
// function therefore begins with the TAIL of the previous iteration — the emitter found at (or behind) the first hit, then
// the Russian roulette that closes that iteration — as shade_path does for `path`.  Behind an index-matched first hit the walk
// goes on in connect_vol_kernel (a VQ_LOOK queue entry).
template <bool PLAIN = false, bool LEAN = false>
__device__ __forceinline__ void volmis_path(const SceneDev &S, const PathState &ps, const IntegratorDev &integ, uint32_t seed,
		uint32_t slot, VolOut &out) {
	const float4 r0 = ps.rayO[slot], r1 = ps.rayD[slot], h = ps.hit[slot];
	float4 thr4 = ps.thr[slot];
	const float4 prev = ps.prev[slot];
	const uint4 rg = ps.rng[slot];
	SamplerState rs; rs.px = rg.x; rs.py = rg.y; rs.sample = rg.z; rs.dim = PATH_DIM(rg.w); rs.seed = seed; rs.sobol = LEAN ? nullptr : integ.sobol;
	const V3 o = v3(r0.x, r0.y, r0.z), d = v3(r1.x, r1.y, r1.z);
	V3 thr = v3(thr4.x, thr4.y, thr4.z), Li = v3(0.0f);
	float eta = thr4.w;
	uint32_t depth = PATH_DEPTH(rg.w);
	const bool first = !(PATH_FLAGS(rg.w) & PF_FROM_BSDF);
	int medium = first ? integ.cameraMedium : __float_as_int(prev.x);
	uint32_t vf = first ? (uint32_t) (VF_EMITTED | VF_REST) : __float_as_uint(prev.y);
	const uint32_t hitSlot = __float_as_uint(h.w);
	const bool valid = hitSlot != INVALID_SLOT;
	out.alive = true; out.nReq = 0;

	Surf sf;
	ShapeRec sh;
	bool transition = false;
	if (valid) { fill_surface(S, hitSlot, h.x, h.y, h.z, o, d, sf); sh = S.shapes[sf.shape]; transition = sh.interior >= 0 || sh.exterior >= 0; }

	float alpha = 0.0f;
	if (first) {
		const V3 far = o + d * S.sceneDiameter;                    // opacity, records.inl:117-145 (as in volshade_path)
		if (valid) {
			if (!transition) alpha = 1.0f;
			else vol_request(out, sf.p, far, VQ_PACK(VQ_P1_SURFACE | VQ_ALPHA, target_medium(sh, sf.ng, d), 0xffffu), v3(0.0f), slot);
		} else if (medium >= 0) vol_request(out, o, far, VQ_PACK(VQ_ALPHA, medium, 0xffffu), v3(0.0f), slot);
	} else {
		// ---- tail of the previous iteration
		if (vf & VF_LOOK) {
			// rayIntersectAndLookForEmitter with maxInteractions = m_maxDepth - rRec.depth - 1 (depth not yet incremented)
			const float samplePdf = prev.z;
			const bool delta = (vf & VF_DELTA) != 0;
			const float4 misc = ps.misc[slot];
			const V3 refN = v3(prev.w, misc.z, misc.w);
			const uint32_t maxI = vol_max_interactions(integ, depth);
			const V3 T0 = medium >= 0 ? medium_transmittance(S.media[medium], valid ? h.x : INFINITY) : v3(1.0f);
			if (valid) {
				const bool isNull = (bsdf_type(S.bsdfs[sh.bsdf]) & B_NULL) != 0, isEmitter = sh.emitter >= 0;
				if (maxI == 0 || !isNull || isEmitter) {
					if (isEmitter) {                                  // dRec.setQuery(ray, its); value = transmittance * Le
						const V3 value = T0 * area_emitter_eval(S, sh.emitter, sf.ns, -d);
						if (!is_zero(value)) {
							const float emitterPdf = delta ? 0.0f : emitter_pdf_direct<PLAIN>(S, sh.emitter, refN, d, h.x, sf.ns, M_SOLID_ANGLE, o) * emitter_pdf_discrete(S, sh.emitter);
							Li += thr * value * mi_weight(samplePdf, emitterPdf);
						}
					}
				} else if (!is_zero(T0)) {
					// an index-matched boundary: the walk goes on from ray(its.t) in the medium behind it
					const int med2 = transition ? target_medium(sh, sf.ng, d) : medium;
					vol_look_request(out, o + d * h.x, d, VQ_PACK(0u, med2, maxI), thr * T0, slot, samplePdf, delta, refN);
				}
			} else if (S.envEmitter >= 0) {
				const EmitterRec &env = S.emitters[S.envEmitter];
				float nearT, farT;                                  // fillDirectSamplingRecord (constant.cpp:238-256)
				if (sphere_intersect(env.sphereCenter, env.sphereRadius, o, d, nearT, farT) && !(nearT > 0) && !(farT < 0)) {
					const V3 value = T0 * env_radiance<PLAIN>(S, d);
					if (!is_zero(value)) {
						const V3 n = normalize(env.sphereCenter - (o + d * farT));
						const float emitterPdf = delta ? 0.0f : emitter_pdf_direct<PLAIN>(S, S.envEmitter, refN, d, farT, n, M_SOLID_ANGLE) * emitter_pdf_discrete(S, S.envEmitter);
						Li += thr * value * mi_weight(samplePdf, emitterPdf);
					}
				}
			}
		}
		if (!(vf & VF_CONTINUE)) {
			if (depth++ >= (uint32_t) integ.rrDepth) {                   // :324-336
				const float q = fminf(vmax(thr) * eta * eta, 0.95f);
				if (next1D(rs) >= q) out.alive = false;
				else thr /= q;
			}
			vf |= VF_SCATTERED;                                          // :338
		}
		vf &= ~(uint32_t) (VF_LOOK | VF_DELTA | VF_CONTINUE);
		if (out.alive && !((int) depth <= integ.maxDepth || integ.maxDepth < 0)) out.alive = false;     // while condition (:100)
	}

	float samplePdf = 0.0f;
	V3 sampleRefN = v3(0.0f);
	if (out.alive) {
		MediumSample ms;
		bool inMedium = false;
		if (medium >= 0) inMedium = medium_sample_distance(S.media[medium], o, d, h.x, rs, ms);
		if (inMedium) {
			// ---- medium interaction (:104-168)
			const MediumRec &m = S.media[medium];
			if ((int) depth >= integ.maxDepth && integ.maxDepth != -1) out.alive = false;
			else {
				thr *= m.sigmaS * ms.transmittance / ms.pdfSuccess;
				if (vf & VF_REST) {
					float sx, sy; next2D(rs, sx, sy);
					if (S.emitterCount > 0) {
						DirectSample ds;
						const V3 value = scene_sample_emitter_direct<PLAIN>(S, ms.p, v3(0.0f), sx, sy, ds);
						if (ds.pdf != 0 && !is_zero(value)) {
							const float phaseVal = phase_eval(m, -d, ds.d);
							if (phaseVal != 0) {
								const bool onSurface = S.emitters[ds.emitter].type != MTSB_EMITTER_POINT;
								const float phasePdf = (onSurface && ds.measure == M_SOLID_ANGLE) ? phaseVal : 0.0f;   // PhaseFunction::pdf = eval
								vol_request(out, ms.p, ds.p, VQ_PACK(onSurface ? VQ_P2_SURFACE : 0u, medium, vol_max_interactions(integ, depth)),
									thr * value * (phaseVal * mi_weight(ds.pdf, phasePdf)), slot);
							}
						}
					}
				}
				float sx, sy; next2D(rs, sx, sy);
				const V3 wo = phase_sample(m, -d, sx, sy);
				samplePdf = phase_eval(m, -d, wo);                          // sample(pRec, pdf, sampler): pdf = eval(pRec)
				ps.rayO[slot] = make_float4(ms.p.x, ms.p.y, ms.p.z, 0.0f);
				ps.rayD[slot] = make_float4(wo.x, wo.y, wo.z, INFINITY);
				if (!(vf & VF_REST)) out.alive = false;                    // :159-167: nothing is gathered without the medium radiance bits
				else vf = (vf & ~(uint32_t) VF_EMITTED) | VF_LOOK;         // rRec.type = ERadianceNoEmission
			}
		} else {
			// ---- surface interaction (:169-322)
			if (medium >= 0) thr *= ms.transmittance / ms.pdfFailure;
			if (!valid) {
				if ((vf & VF_EMITTED) && (!integ.hideEmitters || (vf & VF_SCATTERED)) && S.envEmitter >= 0) {
					V3 rx = v3(0.0f), ry = v3(0.0f);           // only the sensor ray carries differentials (generate_kernel: prev / L)
					if (!PLAIN && first && S.emitters[S.envEmitter].type == MTSB_EMITTER_ENVMAP) { const float4 l = ps.L[slot]; rx = v3(prev.x, prev.y, prev.z); ry = v3(l.x, l.y, l.z); }
					V3 value = thr * env_radiance<PLAIN>(S, d, first, rx, ry);
					if (medium >= 0) value *= medium_transmittance(S.media[medium], INFINITY);
					Li += value;
				}
				out.alive = false;
			} else {
				if (sh.emitter >= 0 && (vf & VF_EMITTED) && (!integ.hideEmitters || (vf & VF_SCATTERED)))
					Li += thr * area_emitter_eval(S, sh.emitter, sf.ns, -d);
				if ((int) depth >= integ.maxDepth && integ.maxDepth != -1) out.alive = false;      // :199-200
				const V3 wi = to_local(sf, -d);
				if (out.alive && integ.strictNormals && -dot(sf.ng, d) * wi.z < 0) out.alive = false;
				if (out.alive) {
					const BsdfRec bsdf = load_bsdf<PLAIN>(S, sh.bsdf, sf, hitSlot, o, first, prev, ps.L + slot);
					const uint32_t btype = bsdf_type(bsdf);
					const V3 refN = (btype & (B_TRANSMISSION | B_BACK)) == 0 ? sf.ns : v3(0.0f);
					// ---- emitter sampling with MIS (:212-247)
					if ((vf & VF_REST) && (btype & B_SMOOTH)) {
						float sx, sy; next2D(rs, sx, sy);
						if (S.emitterCount > 0) {
							DirectSample ds;
							const V3 value = scene_sample_emitter_direct<PLAIN>(S, sf.p, refN, sx, sy, ds);
							if (ds.pdf != 0 && !is_zero(value)) {
								const V3 wo = to_local(sf, ds.d);
								const V3 bsdfVal = bsdf_eval<LEAN>(bsdf, wi, wo);
								if (!is_zero(bsdfVal) && (!integ.strictNormals || dot(sf.ng, ds.d) * wo.z > 0)) {
									const bool onSurface = S.emitters[ds.emitter].type != MTSB_EMITTER_POINT;
									const float bp = (onSurface && ds.measure == M_SOLID_ANGLE) ? bsdf_pdf<LEAN>(bsdf, wi, wo) : 0.0f;
									const int med = transition ? target_medium(sh, sf.ng, ds.d) : medium;
									vol_request(out, sf.p, ds.p, VQ_PACK(VQ_P1_SURFACE | (onSurface ? VQ_P2_SURFACE : 0u), med, vol_max_interactions(integ, depth)),
										thr * value * bsdfVal * mi_weight(ds.pdf, bp), slot);
								}
							}
						}
					}
					// ---- BSDF sampling (:254-322)
					float sx, sy; next2D(rs, sx, sy);
					const float sz = bsdf_uses_sampler(bsdf) ? next1D(rs) : 0.0f;
					BsdfSample bs;
					const V3 bsdfWeight = bsdf_sample<LEAN>(bsdf, wi, sx, sy, sz, bs);
					if (is_zero(bsdfWeight)) out.alive = false;
					else {
						const V3 wo = to_world(sf, bs.wo);
						if (dot(sf.ng, wo) * bs.wo.z <= 0 && integ.strictNormals) out.alive = false;
						else {
							ps.rayO[slot] = make_float4(sf.p.x, sf.p.y, sf.p.z, MTS_EPSILON);
							ps.rayD[slot] = make_float4(wo.x, wo.y, wo.z, INFINITY);
							thr *= bsdfWeight; eta *= bs.eta;
							if (transition) medium = target_medium(sh, sf.ng, wo);
							if (bs.type == B_NULL) {
								// an index-matched transition (:283-292): next iteration at once — no emitter look-up, no roulette
								if (!(vf & VF_REST)) out.alive = false;
								else {
									vf = (vf & ~(uint32_t) (VF_EMITTED | VF_REST)) | ((vf & VF_SCATTERED) ? (uint32_t) VF_REST : (uint32_t) (VF_EMITTED | VF_REST)) | VF_CONTINUE;
									++depth;
									if (!((int) depth <= integ.maxDepth || integ.maxDepth < 0)) out.alive = false;
								}
							} else {
								if (!(vf & VF_REST)) out.alive = false;        // :299-310: neither the look-up nor the recursion is wanted
								else {
									vf = (vf & ~(uint32_t) VF_EMITTED) | VF_LOOK | ((bs.type & B_DELTA) ? (uint32_t) VF_DELTA : 0u);
									samplePdf = bs.pdf; sampleRefN = refN;
								}
							}
						}
					}
				}
			}
		}
	}

	if (first) ps.L[slot] = make_float4(Li.x, Li.y, Li.z, alpha);
	else if (!is_zero(Li)) {
		float *L = (float *) (ps.L + slot);          // component-wise: a look-up of this path may be finishing in connect_vol_kernel
		L[0] += Li.x; L[1] += Li.y; L[2] += Li.z;
	}
	ps.rng[slot] = make_uint4(rg.x, rg.y, rg.z, PATH_PACK(rs.dim, PF_FROM_BSDF, depth));
	if (out.alive) {
		ps.thr[slot] = make_float4(thr.x, thr.y, thr.z, eta);
		ps.prev[slot] = make_float4(__int_as_float(medium), __uint_as_float(vf), samplePdf, sampleRefN.x);
		if (vf & VF_LOOK) {
			float4 misc = ps.misc[slot];
			misc.z = sampleRefN.y; misc.w = sampleRefN.z;
			ps.misc[slot] = misc;
		}
	}
}

#ifndef MTSB_VOLSHADE_MIN_BLOCKS
#define MTSB_VOLSHADE_MIN_BLOCKS 1
#endif
// MIS = 0: volpath_simple (volshade_path), 1: volpath (volmis_path); PLAIN: scene->plain (no envmap / textures / analytic lights)
// LEAN (implies PLAIN): every BSDF is diffuse / null / dielectric / GGX roughconductor with visible-normal sampling / the index-matched
// `null` boundary, and the sampler is the independent one — the other BSDFs' and samplers' code is left out of the kernel, which is
// bound by instruction fetch (profiles/r02v_ncu_full_vol.json: stall_no_instruction 8.5-9.2 cycles per issue)
template <int MIS = 0, bool PLAIN = false, bool LEAN = false>
__global__ void __launch_bounds__(RENDER_BLOCK, MTSB_VOLSHADE_MIN_BLOCKS)
volshade_kernel(SceneDev S, PathState ps, IntegratorDev integ, uint32_t seed) {
	const uint32_t parity = ps.counters[CTR_PARITY] & 1u;
	const uint32_t n = ps.counters[parity];
	const uint32_t *queueIn = parity ? ps.queueB : ps.queueA;
	uint32_t *queueOut = parity ? ps.queueA : ps.queueB;
	uint32_t *countOut = ps.counters + (parity ^ 1u), *countShadow = ps.counters + CTR_SHADOW;
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
		const uint32_t slot = QUEUE_SLOT(queueIn[i]);
		if (__float_as_uint(ps.hit[slot].w) == SUSPENDED_SLOT) { queueOut[warp_append(countOut)] = slot | QUEUE_RESUME; continue; }
		VolOut vo;
		if (MIS) volmis_path<PLAIN, LEAN>(S, ps, integ, seed, slot, vo);
		else volshade_path<PLAIN, LEAN>(S, ps, integ, seed, slot, vo);
		if (vo.alive) queueOut[warp_append(countOut)] = slot;
		for (int k = 0; k < 2; ++k)
			if (k < vo.nReq) {
				const uint32_t q = warp_append(countShadow);
				ps.shO[q] = vo.qO[k]; ps.shD[q] = vo.qD[k]; ps.shL[q] = vo.qL[k];
				if (MIS) ps.shX[q] = vo.qX[k];
			}
	}
}

// Scene::evalTransmittance (scene.cpp:650-713) for every entry of the connection queue.  A warp holds 32 connections and
// alternates, warp-synchronously, one light-record query (ShapeKDTree::rayIntersect(ray, t, shape, n, uv), the un-floored
// epsilon of skdtree.cpp:144-204) with the bookkeeping of the segment it closed; lanes whose connection has ended pull the
// next queue entry (one atomic per warp).
template <bool CURVED>
__global__ void __launch_bounds__(MTSB_TRACE_BLOCK, MTSB_TAIL_MIN_BLOCKS)
connect_vol_kernel(SceneDev S, PathState ps) {
	__shared__ uint2 shStack[MTSB_STACK_SHARED * MTSB_TRACE_BLOCK];
	const uint32_t n = ps.counters[CTR_SHADOW];
	if (n == 0) return;
	const unsigned lane = threadIdx.x & 31u, ltMask = (1u << lane) - 1u, FULL = 0xffffffffu;
	uint2 locStack[MTSB_STACK_LOCAL];
	Trav tv; tv.init(shStack + threadIdx.x, locStack);
	bool have = false, exhausted = false;
	uint32_t entry = 0, bits = 0;
	V3 o = v3(0.0f), d = v3(0.0f), T = v3(1.0f);
	float remaining = 0, lengthFactor = 1, mint = 0;
	int medium = -1;
	uint32_t interactions = 0;
	float samplePdf = 0;             // VQ_LOOK entries: density of the phase / BSDF sample (sign bit: delta lobe) and refN of its vertex
	V3 refN = v3(0.0f);
	unsigned nRays = 0;
	// A path may have two entries in one bounce (the opacity query of the sensor ray and its emitter sample), handled by
	// different lanes: the radiance is added component-wise with atomics, the opacity is a plain store to L.w.
	auto finish = [&](const V3 &tr) {
		const float4 c = ps.shL[entry];
		const uint32_t slot = __float_as_uint(c.w);
		float *L = (float *) (ps.L + slot);
		// 1 - Spectrum::average(): the product is rounded before the subtraction (no FMA), so T = 1 gives exactly 0 — a
		// contracted 1 - 3 * 0.33333334 would be -3e-8, and ImageBlock::put rejects samples with a negative channel
		if (bits & VQ_ALPHA) L[3] = 1.0f - __fmul_rn(tr.x + tr.y + tr.z, 1.0f / 3.0f);
		else if (!is_zero(tr)) { atomicAdd(L, c.x * tr.x); atomicAdd(L + 1, c.y * tr.y); atomicAdd(L + 2, c.z * tr.z); }
	};

	for (;;) {
		const unsigned idleMask = __ballot_sync(FULL, !have);
		if (!exhausted && idleMask) {
			const unsigned nIdle = __popc(idleMask);
			unsigned long long base = 0;
			if (lane == 0) base = atomicAdd(ps.work + 1, (unsigned long long) nIdle);
			base = __shfl_sync(FULL, base, 0);
			if (!have) {
				const unsigned long long idx = base + __popc(idleMask & ltMask);
				if (idx < n) {
					entry = (uint32_t) idx; have = true;
					const float4 a = ps.shO[entry], b = ps.shD[entry];
					bits = __float_as_uint(a.w);
					o = v3(a.x, a.y, a.z);
					d = v3(b.x, b.y, b.z) - o;                               // :652-654
					remaining = length(d);
					d /= remaining;
					lengthFactor = (bits & VQ_P2_SURFACE) ? (1 - MTS_SHADOW_EPSILON) : 1.0f;
					mint = (bits & VQ_P1_SURFACE) ? MTS_EPSILON : 0.0f;
					medium = VQ_MEDIUM(bits);
					interactions = 0;
					T = v3(1.0f);
					if (bits & VQ_LOOK) {
						// the walk of rayIntersectAndLookForEmitter behind an index-matched first hit: a ray, not a segment
						d = v3(b.x, b.y, b.z); samplePdf = b.w;
						const float4 x = ps.shX[entry];
						refN = v3(x.x, x.y, x.z);
						remaining = INFINITY; lengthFactor = 1.0f; mint = MTS_EPSILON; interactions = 1;
					} else
					if (!(remaining > 0)) { finish(v3(1.0f)); have = false; }      // while (remaining > 0): the loop body never runs
				}
			}
			if (base + nIdle >= n) exhausted = true;
		}
		if (!__any_sync(FULL, have)) break;

		// ---- one segment: closest surface along the rest of the connection
		tv.found = false;
		if (have) {
			float tmax; V3 rcp;
			tv.o = o; tv.d = d;
			++nRays;
			// evalTransmittance asks the light-record overload (un-floored epsilon), the emitter look-up Scene::rayIntersect(ray, its)
			if (clip_ray(S, o, d, mint, remaining * lengthFactor, (bits & VQ_LOOK) != 0, tv.tmin, tmax, rcp)) tv.begin(tmax, rcp);
		}
		trace_warp<false, CURVED>(S, tv);
		if (have) {
			bool done = false;
			const bool surface = tv.found;
			const float t = surface ? tv.hit.t : INFINITY;
			Surf sf; ShapeRec sh;
			bool isNull = false;
			if (surface) {
				fill_surface(S, tv.hit.slot, tv.hit.t, tv.hit.u, tv.hit.v, o, d, sf);
				sh = S.shapes[sf.shape];
				isNull = (bsdf_type(S.bsdfs[sh.bsdf]) & B_NULL) != 0;
			}
			if (bits & VQ_LOOK) {
				// volpath.cpp:376-424
				if (medium >= 0) T *= medium_transmittance(S.media[medium], t);
				const bool delta = (__float_as_uint(samplePdf) >> 31) != 0;
				const float pdfS = fabsf(samplePdf);
				V3 value = v3(0.0f); float emitterPdf = 0.0f;
				if (surface && (interactions == VQ_MAXI(bits) || !isNull || sh.emitter >= 0)) {
					if (sh.emitter >= 0) {
						value = T * area_emitter_eval(S, sh.emitter, sf.ns, -d);
						if (!delta) emitterPdf = emitter_pdf_direct(S, sh.emitter, refN, d, t, sf.ns, M_SOLID_ANGLE, o) * emitter_pdf_discrete(S, sh.emitter);
					}
					done = true;
				} else if (!surface) {
					if (S.envEmitter >= 0) {
						const EmitterRec &env = S.emitters[S.envEmitter];
						float nearT, farT;
						if (sphere_intersect(env.sphereCenter, env.sphereRadius, o, d, nearT, farT) && !(nearT > 0) && !(farT < 0)) {
							value = T * env_radiance(S, d);
							const V3 n = normalize(env.sphereCenter - (o + d * farT));
							if (!delta) emitterPdf = emitter_pdf_direct(S, S.envEmitter, refN, d, farT, n, M_SOLID_ANGLE) * emitter_pdf_discrete(S, S.envEmitter);
						}
					}
					done = true;
				} else if (is_zero(T)) done = true;
				else {
					if (sh.interior >= 0 || sh.exterior >= 0) medium = target_medium(sh, sf.ng, d);
					o = o + d * t;
					mint = MTS_EPSILON;
					if (++interactions > 100) done = true;
				}
				if (done) {
					if (!is_zero(value)) finish(value * mi_weight(pdfS, emitterPdf));
					have = false; done = false;
				}
			} else
			if (surface && (interactions == VQ_MAXI(bits) || !isNull)) { T = v3(0.0f); done = true; }      // an occluder
			else {
				if (medium >= 0) T *= medium_transmittance(S.media[medium], fminf(t, remaining));
				if (!surface || is_zero(T)) done = true;
				else {
					// the null BSDF's eval(ENull, EDiscrete) is 1 (null.cpp:47-49)
					if (sh.interior >= 0 || sh.exterior >= 0) {
						if (medium != target_medium(sh, sf.ng, -d)) { T = v3(0.0f); done = true; }        // "medium inconsistency"
						else medium = target_medium(sh, sf.ng, d);
					}
					if (!done) {
						if (++interactions > 100) done = true;
						else {
							o = o + d * t;
							remaining -= t;
							mint = MTS_EPSILON;
							if (!(remaining > 0)) done = true;
						}
					}
				}
			}
			if (done) { finish(T); have = false; }
		}
	}
	#pragma unroll
	for (int k = 16; k > 0; k >>= 1) nRays += __shfl_xor_sync(FULL, nRays, k);
	if (lane == 0 && nRays) atomicAdd(ps.stats + 2, (unsigned long long) nRays);          // "Shadow rays traced" (skdtree.cpp:47,151)
}

} // namespace mtsb
