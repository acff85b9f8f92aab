// bsdf.cuh — the closed BSDF set, dispatched by tag.  Semantics of BSDF::eval / pdf / sample with the default
// arguments MIPathTracer uses (measure = ESolidAngle, typeMask = EAll, component = -1, mode = ERadiance;
// include/mitsuba/render/bsdf.h:215-445):
//   eval   -> f * cos(theta_o)          pdf -> solid-angle density          sample -> f * cos / pdf, sets wo, eta, type
//   diffuse        src/bsdfs/diffuse.cpp:112-153
//   dielectric     src/bsdfs/dielectric.cpp:231-343
//   plastic        src/bsdfs/plastic.cpp:248-455
//   roughconductor src/bsdfs/roughconductor.cpp:260-413
//   twosided       src/bsdfs/twosided.cpp:106-170 (same nested BRDF on both sides)
//   roughplastic   src/bsdfs/roughplastic.cpp:330-520 (constant alpha; RoughTransmittance table, src/bsdfs/rtrans.h)
//   null           the black diffuse BSDF Shape::configure gives emitters without one (src/librender/shape.cpp:48-58)
#pragma once

namespace mtsb {

struct BsdfSample { V3 wo; float pdf, eta; uint32_t type; };

__device__ inline uint32_t bsdf_type(const BsdfRec &b) {   // m_combinedType of each plugin
	uint32_t t;
	switch (b.type & 0xff) {
		case MTSB_BSDF_DIFFUSE: case MTSB_BSDF_NULL: t = B_DIFFUSE_REFL | B_FRONT; break;
		case MTSB_BSDF_DIELECTRIC: t = B_DELTA_REFL | B_DELTA_TRANS | B_FRONT | B_BACK; break;
		case MTSB_BSDF_PLASTIC: t = B_DELTA_REFL | B_DIFFUSE_REFL | B_FRONT; break;
		case MTSB_BSDF_CONDUCTOR: t = B_DELTA_REFL | B_FRONT; break;
		case MTSB_BSDF_ROUGHDIELECTRIC: t = B_GLOSSY_REFL | B_GLOSSY_TRANS | B_FRONT | B_BACK; break;
		case MTSB_BSDF_PASSTHROUGH: t = B_NULL | B_FRONT | B_BACK; break;        // null.cpp:35-38
		case MTSB_BSDF_ROUGHPLASTIC: t = B_GLOSSY_REFL | B_DIFFUSE_REFL | B_FRONT; break;   // roughplastic.cpp:264-269
		default: t = B_GLOSSY_REFL | B_FRONT; break;
	}
	if (b.type & MTSB_BSDF_TWOSIDED_FLAG) t |= B_BACK;
	return t;
}

__device__ inline V3 plastic_diffuse(const BsdfRec &b) {
	V3 diff = v3(b.p[1], b.p[2], b.p[3]);
	const float fdrInt = b.p[7];
	if (b.p[10] != 0) diff = diff / (v3(1.0f) - diff * fdrInt); else diff = diff / (1 - fdrInt);
	return diff;
}
__device__ inline float plastic_prob_specular(const BsdfRec &b, float Fi) {
	const float w = b.p[9];
	return (Fi * w) / (Fi * w + (1 - Fi) * (1 - w));
}

// --- one-sided evaluation (wi, wo in the local shading frame) ---
// --- roughdielectric (src/bsdfs/roughdielectric.cpp): p[0] alphaU, p[1] alphaV, p[2] eta, p[3..5] specularReflectance,
// p[6..8] specularTransmittance, p[11] distribution, p[12] sampleVisible.  ERadiance transport mode.
__device__ inline float signum1(float v) { return copysignf(1.0f, v); }          // math::signum (core/math.h:269-280)

__device__ inline V3 roughdielectric_eval(const BsdfRec &b, const V3 &wi, const V3 &wo) {          // :282-351
	if (wi.z == 0) return v3(0.0f);
	const float etaI = b.p[2], invEta = 1 / etaI;
	const bool reflect = wi.z * wo.z > 0;
	const float eta = wi.z > 0 ? etaI : invEta;
	V3 H = reflect ? normalize(wo + wi) : normalize(wi + wo * eta);
	H = H * signum1(H.z);
	Microfacet distr; distr.init((int) b.p[11], b.p[0], b.p[1], b.p[12] != 0);
	const float D = distr.eval(H);
	if (D == 0) return v3(0.0f);
	const float F = fresnel_dielectric_ext(dot(wi, H), etaI);
	const float G = distr.G(wi, wo, H);
	if (reflect) {
		const float value = F * D * G / (4.0f * fabsf(wi.z));
		return v3(b.p[3], b.p[4], b.p[5]) * value;
	}
	const float sqrtDenom = dot(wi, H) + eta * dot(wo, H);
	const float value = ((1 - F) * D * G * eta * eta * dot(wi, H) * dot(wo, H)) / (wi.z * sqrtDenom * sqrtDenom);
	const float factor = wi.z > 0 ? invEta : etaI;
	return v3(b.p[6], b.p[7], b.p[8]) * fabsf(value * factor * factor);
}

__device__ inline float roughdielectric_pdf(const BsdfRec &b, const V3 &wi, const V3 &wo) {         // :353-424
	const float etaI = b.p[2], invEta = 1 / etaI;
	const bool reflect = wi.z * wo.z > 0;
	V3 H; float dwh_dwo;
	if (reflect) {
		H = normalize(wo + wi);
		dwh_dwo = 1.0f / (4.0f * dot(wo, H));
	} else {
		const float eta = wi.z > 0 ? etaI : invEta;
		H = normalize(wi + wo * eta);
		const float sqrtDenom = dot(wi, H) + eta * dot(wo, H);
		dwh_dwo = (eta * eta * dot(wo, H)) / (sqrtDenom * sqrtDenom);
	}
	H = H * signum1(H.z);
	Microfacet sampleDistr; sampleDistr.init((int) b.p[11], b.p[0], b.p[1], b.p[12] != 0);
	if (!sampleDistr.sampleVisible) {
		const float sc = 1.2f - 0.2f * sqrtf(fabsf(wi.z));
		sampleDistr.alphaU *= sc; sampleDistr.alphaV *= sc;
	}
	float prob = sampleDistr.pdf(wi * signum1(wi.z), H);
	const float F = fresnel_dielectric_ext(dot(wi, H), etaI);
	prob *= reflect ? F : (1 - F);
	return fabsf(prob * dwh_dwo);
}

// `sz` is the lobe-selection draw the plugin takes from bRec.sampler->next1D() (:562)
__device__ inline V3 roughdielectric_sample(const BsdfRec &b, const V3 &wi, float sx, float sy, float sz, BsdfSample &out) {   // :529-620
	const float etaI = b.p[2], invEta = 1 / etaI;
	Microfacet distr; distr.init((int) b.p[11], b.p[0], b.p[1], b.p[12] != 0);
	Microfacet sampleDistr = distr;
	if (!distr.sampleVisible) {
		const float sc = 1.2f - 0.2f * sqrtf(fabsf(wi.z));
		sampleDistr.alphaU *= sc; sampleDistr.alphaV *= sc;
	}
	float microfacetPDF;
	const V3 m = sampleDistr.sample(wi * signum1(wi.z), sx, sy, microfacetPDF);
	if (microfacetPDF == 0) return v3(0.0f);
	float pdf = microfacetPDF;
	float cosThetaT;
	const float F = fresnel_dielectric_ext(dot(wi, m), cosThetaT, etaI);
	V3 weight = v3(1.0f);
	bool sampleReflection = true;
	if (sz > F) { sampleReflection = false; pdf *= 1 - F; } else pdf *= F;
	float dwh_dwo;
	if (sampleReflection) {
		out.wo = 2 * dot(wi, m) * m - wi;
		out.eta = 1.0f; out.type = B_GLOSSY_REFL;
		if (wi.z * out.wo.z <= 0) return v3(0.0f);
		weight = v3(b.p[3], b.p[4], b.p[5]);
		dwh_dwo = 1.0f / (4.0f * dot(out.wo, m));
	} else {
		if (cosThetaT == 0) return v3(0.0f);
		const float e = cosThetaT < 0 ? invEta : etaI;                     // refract(wi, m, eta, cosThetaT), util.cpp:769-774
		out.wo = m * (dot(wi, m) * e + cosThetaT) - wi * e;
		out.eta = cosThetaT < 0 ? etaI : invEta; out.type = B_GLOSSY_TRANS;
		if (wi.z * out.wo.z >= 0) return v3(0.0f);
		const float factor = cosThetaT < 0 ? invEta : etaI;
		weight = v3(b.p[6], b.p[7], b.p[8]) * (factor * factor);
		const float sqrtDenom = dot(wi, m) + out.eta * dot(out.wo, m);
		dwh_dwo = (out.eta * out.eta * dot(out.wo, m)) / (sqrtDenom * sqrtDenom);
	}
	if (distr.sampleVisible) weight = weight * distr.smithG1(out.wo, m);
	else weight = weight * fabsf(distr.eval(m) * distr.G(wi, out.wo, m) * dot(wi, m) / (microfacetPDF * wi.z));
	out.pdf = pdf * fabsf(dwh_dwo);
	return weight;
}


// ---- roughplastic (src/bsdfs/roughplastic.cpp) ----
// RoughTransmittance::eval with eta and alpha fixed (rtrans.h:179-191, 243): evalCubicInterp1D (spline.cpp:23-60) of the
// table over cosTheta^(1/4), clamped to [0, 1].  The record carries the device address of {count, values...} in p[13..14].
__device__ inline float roughplastic_T(const BsdfRec &b, float cosTheta) {
	if (!(cosTheta >= 0)) return 0.0f;
	const unsigned long long addr = (unsigned long long) __float_as_uint(b.p[13]) | ((unsigned long long) __float_as_uint(b.p[14]) << 32);
	const float *tab = (const float *) addr;
	const int size = (int) tab[0];
	const float *values = tab + 1;
	const float x = powf(fabsf(cosTheta), 0.25f);
	float result = 0.0f;
	if (x >= 0.0f && x <= 1.0f) {
		float t = x * (float) (size - 1);
		const int k = max(0, min((int) t, size - 2));
		const float f0 = values[k], f1 = values[k + 1];
		const float d0 = k > 0 ? 0.5f * (values[k + 1] - values[k - 1]) : values[k + 1] - values[k];
		const float d1 = k + 2 < size ? 0.5f * (values[k + 2] - values[k]) : values[k + 1] - values[k];
		t = t - (float) k;
		const float t2 = t * t, t3 = t2 * t;
		result = (2 * t3 - 3 * t2 + 1) * f0 + (-2 * t3 + 3 * t2) * f1 + (t3 - 2 * t2 + t) * d0 + (t3 - t2) * d1;
	}
	return fminf(1.0f, fmaxf(0.0f, result));
}
__device__ inline float roughplastic_prob_specular(const BsdfRec &b, float cosThetaI) {    // :395-402, :449-456
	float probSpecular = 1 - roughplastic_T(b, cosThetaI);
	const float ssw = b.p[8];
	return (probSpecular * ssw) / (probSpecular * ssw + (1 - probSpecular) * (1 - ssw));
}
__device__ inline V3 roughplastic_eval(const BsdfRec &b, const V3 &wi, const V3 &wo) {       // :330-381
	if (wi.z <= 0 || wo.z <= 0) return v3(0.0f);
	Microfacet distr; distr.init((int) b.p[11], b.p[0], b.p[0], b.p[12] != 0);
	const V3 H = normalize(wo + wi);
	const float D = distr.eval(H);
	const float F = fresnel_dielectric_ext(dot(wi, H), b.p[1]);
	const float G = distr.G(wi, wo, H);
	const float value = F * D * G / (4.0f * wi.z);
	V3 result = v3(b.p[5], b.p[6], b.p[7]) * value;
	V3 diff = v3(b.p[2], b.p[3], b.p[4]);
	const float T12 = roughplastic_T(b, wi.z), T21 = roughplastic_T(b, wo.z), Fdr = b.p[9];
	if (b.p[10] != 0) diff = v3(diff.x / (1 - diff.x * Fdr), diff.y / (1 - diff.y * Fdr), diff.z / (1 - diff.z * Fdr));
	else diff /= 1 - Fdr;
	const float invEta2 = 1.0f / (b.p[1] * b.p[1]);
	return result + diff * (MTS_INV_PI * wo.z * T12 * T21 * invEta2);
}
__device__ inline float roughplastic_pdf(const BsdfRec &b, const V3 &wi, const V3 &wo) {    // :383-432
	if (wi.z <= 0 || wo.z <= 0) return 0.0f;
	Microfacet distr; distr.init((int) b.p[11], b.p[0], b.p[0], b.p[12] != 0);
	const V3 H = normalize(wo + wi);
	const float probSpecular = roughplastic_prob_specular(b, wi.z), probDiffuse = 1 - probSpecular;
	const float dwh_dwo = 1.0f / (4.0f * dot(wo, H));
	const float prob = distr.pdf(wi, H);
	return prob * dwh_dwo * probSpecular + probDiffuse * (MTS_INV_PI * wo.z);
}
__device__ inline V3 roughplastic_sample(const BsdfRec &b, const V3 &wi, float sx, float sy, BsdfSample &out) {   // :434-495
	if (wi.z <= 0) return v3(0.0f);
	Microfacet distr; distr.init((int) b.p[11], b.p[0], b.p[0], b.p[12] != 0);
	bool choseSpecular = true;
	const float probSpecular = roughplastic_prob_specular(b, wi.z);
	if (sy < probSpecular) sy /= probSpecular;
	else { sy = (sy - probSpecular) / (1 - probSpecular); choseSpecular = false; }
	if (choseSpecular) {
		float mpdf;
		const V3 m = distr.sample(wi, sx, sy, mpdf);
		out.wo = 2 * dot(wi, m) * m - wi;
		out.type = B_GLOSSY_REFL;
		if (out.wo.z <= 0) return v3(0.0f);
	} else {
		out.type = B_DIFFUSE_REFL;
		out.wo = square_to_cosine_hemisphere(sx, sy);
	}
	out.eta = 1.0f;
	out.pdf = roughplastic_pdf(b, wi, out.wo);
	if (out.pdf == 0) return v3(0.0f);
	return roughplastic_eval(b, wi, out.wo) / out.pdf;
}

// LEAN: the scene's BSDFs are all of {diffuse, null, dielectric, roughconductor with GGX + visible-normal sampling} (the
// set of configs 1, 3, 4, 5).  The other cases are then unreachable and their code — rough dielectric, Beckmann's
// visible-normal inversion, plastic, smooth conductor — is left out of the kernel: the full shade kernel stalls 6-11
// cycles per issue on instruction fetch on config 3 (profiles/r02m_ncu_full_render.json).
template <bool LEAN = false>
__device__ inline V3 bsdf_eval_1(const BsdfRec &b, const V3 &wi, const V3 &wo) {
	switch (b.type & 0xff) {
		case MTSB_BSDF_DIFFUSE: case MTSB_BSDF_NULL: {
			if (wi.z <= 0 || wo.z <= 0) return v3(0.0f);
			return v3(b.p[0], b.p[1], b.p[2]) * (MTS_INV_PI * wo.z);
		}
		case MTSB_BSDF_DIELECTRIC: case MTSB_BSDF_CONDUCTOR: case MTSB_BSDF_PASSTHROUGH: return v3(0.0f);     // discrete measure only
		case MTSB_BSDF_ROUGHDIELECTRIC: if (LEAN) __builtin_unreachable(); return roughdielectric_eval(b, wi, wo);
		case MTSB_BSDF_ROUGHPLASTIC: if (LEAN) __builtin_unreachable(); return roughplastic_eval(b, wi, wo);
		case MTSB_BSDF_PLASTIC: {
			if (LEAN) __builtin_unreachable();
			if (wo.z <= 0 || wi.z <= 0) return v3(0.0f);
			const float eta = b.p[0];
			const float Fi = fresnel_dielectric_ext(wi.z, eta), Fo = fresnel_dielectric_ext(wo.z, eta);
			return plastic_diffuse(b) * ((MTS_INV_PI * wo.z) * (1 / (eta * eta)) * (1 - Fi) * (1 - Fo));
		}
		default: {
			if (wi.z <= 0 || wo.z <= 0) return v3(0.0f);
			const V3 H = normalize(wo + wi);
			Microfacet distr; distr.init(LEAN ? 1 : (int) b.p[11], b.p[0], b.p[1], LEAN ? true : b.p[12] != 0);
			const float D = distr.eval(H);
			if (D == 0) return v3(0.0f);
			const V3 F = fresnel_conductor_exact(dot(wi, H), v3(b.p[2], b.p[3], b.p[4]), v3(b.p[5], b.p[6], b.p[7])) * v3(b.p[8], b.p[9], b.p[10]);
			const float G = distr.G(wi, wo, H);
			return F * (D * G / (4.0f * wi.z));
		}
	}
}

template <bool LEAN = false>
__device__ inline float bsdf_pdf_1(const BsdfRec &b, const V3 &wi, const V3 &wo) {
	switch (b.type & 0xff) {
		case MTSB_BSDF_DIFFUSE: case MTSB_BSDF_NULL:
			if (wi.z <= 0 || wo.z <= 0) return 0.0f;
			return MTS_INV_PI * wo.z;
		case MTSB_BSDF_DIELECTRIC: case MTSB_BSDF_CONDUCTOR: case MTSB_BSDF_PASSTHROUGH: return 0.0f;
		case MTSB_BSDF_ROUGHDIELECTRIC: if (LEAN) __builtin_unreachable(); return roughdielectric_pdf(b, wi, wo);
		case MTSB_BSDF_ROUGHPLASTIC: if (LEAN) __builtin_unreachable(); return roughplastic_pdf(b, wi, wo);
		case MTSB_BSDF_PLASTIC: {
			if (LEAN) __builtin_unreachable();
			if (wo.z <= 0 || wi.z <= 0) return 0.0f;
			const float Fi = fresnel_dielectric_ext(wi.z, b.p[0]);
			return MTS_INV_PI * wo.z * (1 - plastic_prob_specular(b, Fi));
		}
		default: {
			if (wi.z <= 0 || wo.z <= 0) return 0.0f;
			const V3 H = normalize(wo + wi);
			Microfacet distr; distr.init((int) b.p[11], b.p[0], b.p[1], b.p[12] != 0);
			if (distr.sampleVisible) return distr.eval(H) * distr.smithG1(wi, H) / (4.0f * wi.z);
			return distr.pdf(wi, H) / (4 * fabsf(dot(wo, H)));
		}
	}
}

// returns the importance weight (f cos / pdf); zero weight = failed sample
template <bool LEAN = false>
__device__ inline V3 bsdf_sample_1(const BsdfRec &b, const V3 &wi, float sx, float sy, float sz, BsdfSample &out) {
	out.pdf = 0; out.eta = 1.0f; out.type = 0; out.wo = v3(0.0f);
	switch (b.type & 0xff) {
		case MTSB_BSDF_DIFFUSE: case MTSB_BSDF_NULL: {
			if (wi.z <= 0) return v3(0.0f);
			out.wo = square_to_cosine_hemisphere(sx, sy);
			out.type = B_DIFFUSE_REFL;
			out.pdf = MTS_INV_PI * out.wo.z;
			return v3(b.p[0], b.p[1], b.p[2]);
		}
		case MTSB_BSDF_PASSTHROUGH: {                              // null.cpp:55-78 (part of the LEAN set: the volumetric kernels meet it)
			out.wo = v3(-wi.x, -wi.y, -wi.z); out.type = B_NULL; out.pdf = 1.0f;
			return v3(1.0f);
		}
		case MTSB_BSDF_DIELECTRIC: {
			const float eta = b.p[0], invEta = 1 / eta;
			float cosThetaT;
			const float F = fresnel_dielectric_ext(wi.z, cosThetaT, eta);
			if (sx <= F) {
				out.type = B_DELTA_REFL; out.wo = v3(-wi.x, -wi.y, wi.z); out.pdf = F;
				return v3(b.p[1], b.p[2], b.p[3]);
			}
			out.type = B_DELTA_TRANS;
			const float scale = -(cosThetaT < 0 ? invEta : eta);
			out.wo = v3(scale * wi.x, scale * wi.y, cosThetaT);
			out.eta = cosThetaT < 0 ? eta : invEta;
			out.pdf = 1 - F;
			const float factor = cosThetaT < 0 ? invEta : eta;
			return v3(b.p[4], b.p[5], b.p[6]) * (factor * factor);
		}
		case MTSB_BSDF_PLASTIC: {
			if (LEAN) __builtin_unreachable();
			if (wi.z <= 0) return v3(0.0f);
			const float eta = b.p[0];
			const float Fi = fresnel_dielectric_ext(wi.z, eta);
			const float probSpecular = plastic_prob_specular(b, Fi);
			if (sx < probSpecular) {
				out.type = B_DELTA_REFL; out.wo = v3(-wi.x, -wi.y, wi.z); out.pdf = probSpecular;
				return v3(b.p[4], b.p[5], b.p[6]) * Fi / probSpecular;
			}
			out.type = B_DIFFUSE_REFL;
			out.wo = square_to_cosine_hemisphere((sx - probSpecular) / (1 - probSpecular), sy);
			const float Fo = fresnel_dielectric_ext(out.wo.z, eta);
			out.pdf = (1 - probSpecular) * (MTS_INV_PI * out.wo.z);
			return plastic_diffuse(b) * ((1 / (eta * eta)) * (1 - Fi) * (1 - Fo) / (1 - probSpecular));
		}
		case MTSB_BSDF_ROUGHDIELECTRIC: if (LEAN) __builtin_unreachable(); return roughdielectric_sample(b, wi, sx, sy, sz, out);
		case MTSB_BSDF_ROUGHPLASTIC: if (LEAN) __builtin_unreachable(); return roughplastic_sample(b, wi, sx, sy, out);
		case MTSB_BSDF_CONDUCTOR: {                                // conductor.cpp:214-235
			if (LEAN) __builtin_unreachable();
			if (wi.z <= 0) return v3(0.0f);
			out.type = B_DELTA_REFL; out.wo = v3(-wi.x, -wi.y, wi.z); out.pdf = 1.0f;
			return v3(b.p[6], b.p[7], b.p[8]) * fresnel_conductor_exact(wi.z, v3(b.p[0], b.p[1], b.p[2]), v3(b.p[3], b.p[4], b.p[5]));
		}
		default: {
			if (wi.z < 0) return v3(0.0f);
			Microfacet distr; distr.init(LEAN ? 1 : (int) b.p[11], b.p[0], b.p[1], LEAN ? true : b.p[12] != 0);
			float pdf;
			const V3 m = distr.sample(wi, sx, sy, pdf);
			if (pdf == 0) return v3(0.0f);
			out.wo = 2 * dot(wi, m) * m - wi;
			out.type = B_GLOSSY_REFL;
			if (out.wo.z <= 0) return v3(0.0f);
			const V3 F = fresnel_conductor_exact(dot(wi, m), v3(b.p[2], b.p[3], b.p[4]), v3(b.p[5], b.p[6], b.p[7])) * v3(b.p[8], b.p[9], b.p[10]);
			float weight;
			if (distr.sampleVisible) weight = distr.smithG1(out.wo, m);
			else weight = distr.eval(m) * distr.G(wi, out.wo, m) * dot(wi, m) / (pdf * wi.z);
			out.pdf = pdf / (4.0f * dot(out.wo, m));
			return F * weight;
		}
	}
}

// --- public entry points: add the twosided adapter ---
template <bool LEAN = false>
__device__ inline V3 bsdf_eval(const BsdfRec &b, V3 wi, V3 wo) {
	if ((b.type & MTSB_BSDF_TWOSIDED_FLAG) && !(wi.z > 0)) { wi.z = -wi.z; wo.z = -wo.z; }
	return bsdf_eval_1<LEAN>(b, wi, wo);
}
template <bool LEAN = false>
__device__ inline float bsdf_pdf(const BsdfRec &b, V3 wi, V3 wo) {
	if ((b.type & MTSB_BSDF_TWOSIDED_FLAG) && !(wi.z > 0)) { wi.z = -wi.z; wo.z = -wo.z; }
	return bsdf_pdf_1<LEAN>(b, wi, wo);
}
// true for the plugins whose sample() takes one more draw from bRec.sampler (EUsesSampler); the caller draws it
// right after the 2D sample, which is where the plugin's next1D() lands in the stream
__device__ inline bool bsdf_uses_sampler(const BsdfRec &b) { return (b.type & 0xff) == MTSB_BSDF_ROUGHDIELECTRIC; }

template <bool LEAN = false>
__device__ inline V3 bsdf_sample(const BsdfRec &b, V3 wi, float sx, float sy, float sz, BsdfSample &out) {
	bool flipped = false;
	if ((b.type & MTSB_BSDF_TWOSIDED_FLAG) && wi.z < 0) { wi.z = -wi.z; flipped = true; }
	const V3 w = bsdf_sample_1<LEAN>(b, wi, sx, sy, sz, out);
	if (flipped && !is_zero(w) && out.pdf != 0) out.wo.z = -out.wo.z;
	return w;
}

} // namespace mtsb
