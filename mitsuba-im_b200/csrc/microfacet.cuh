// microfacet.cuh — MicrofacetDistribution (src/bsdfs/microfacet.h) for Beckmann and GGX, including visible-normal
// sampling (Heitz & d'Eon), as used by roughconductor.
// NOTE: `sample_visible11` below is a statement-for-statement transliteration of MicrofacetDistribution::sampleVisible11
// (src/bsdfs/microfacet.h:572-696) — the same locals, the same Newton / bisection loop, the same polynomial fit
// constants — not a design of its own: pointwise parity of the sampled directions (5e-4) needs exactly those fits.
#pragma once

namespace mtsb {

struct Microfacet {
	int type;            // 0 beckmann, 1 ggx
	float alphaU, alphaV;
	bool sampleVisible;

	__device__ inline void init(int t, float au, float av, bool sv) {   // microfacet.h:75-82 (alpha clamped to >= 1e-4)
		type = t; alphaU = fmaxf(au, 1e-4f); alphaV = fmaxf(av, 1e-4f); sampleVisible = sv;
	}
	__device__ inline bool isotropic() const { return alphaU == alphaV; }

	// microfacet.h:190-239
	__device__ inline float eval(const V3 &m) const {
		if (m.z <= 0) return 0.0f;
		const float cosTheta2 = m.z * m.z;
		const float beckmannExponent = ((m.x * m.x) / (alphaU * alphaU) + (m.y * m.y) / (alphaV * alphaV)) / cosTheta2;
		float result;
		if (type == 0) {
			result = expf(-beckmannExponent) / (MTS_PI * alphaU * alphaV * cosTheta2 * cosTheta2);
		} else {
			const float root = (1.0f + beckmannExponent) * cosTheta2;
			result = 1.0f / (MTS_PI * alphaU * alphaV * root * root);
		}
		if (result * m.z < 1e-20f) result = 0;
		return result;
	}

	// microfacet.h:536-544
	__device__ inline float project_roughness(const V3 &v) const {
		const float invSinTheta2 = 1 / (1 - v.z * v.z);
		if (isotropic() || invSinTheta2 <= 0) return alphaU;
		const float cosPhi2 = v.x * v.x * invSinTheta2, sinPhi2 = v.y * v.y * invSinTheta2;
		return sqrtf(cosPhi2 * alphaU * alphaU + sinPhi2 * alphaV * alphaV);
	}

	// microfacet.h:476-513
	__device__ inline float smithG1(const V3 &v, const V3 &m) const {
		if (dot(v, m) * v.z <= 0) return 0.0f;
		const float temp = 1 - v.z * v.z;
		const float tanTheta = temp <= 0.0f ? 0.0f : fabsf(sqrtf(temp) / v.z);
		if (tanTheta == 0.0f) return 1.0f;
		const float alpha = project_roughness(v);
		if (type == 0) {
			const float a = 1.0f / (alpha * tanTheta);
			if (a >= 1.6f) return 1.0f;
			const float aSqr = a * a;
			return (3.535f * a + 2.181f * aSqr) / (1.0f + 2.276f * a + 2.577f * aSqr);
		} else {
			const float root = alpha * tanTheta;
			return 2.0f / (1.0f + hypotf(1.0f, root));
		}
	}
	__device__ inline float G(const V3 &wi, const V3 &wo, const V3 &m) const { return smithG1(wi, m) * smithG1(wo, m); }

	// microfacet.h:465-471
	__device__ inline float pdf_visible(const V3 &wi, const V3 &m) const {
		if (wi.z == 0) return 0.0f;
		return smithG1(wi, m) * fabsf(dot(wi, m)) * eval(m) / fabsf(wi.z);
	}
	__device__ inline float pdf(const V3 &wi, const V3 &m) const { return sampleVisible ? pdf_visible(wi, m) : eval(m) * m.z; }

	// microfacet.h:284-395 (Beckmann and GGX branches)
	__device__ inline V3 sample_all(float sx, float sy, float &pdf) const {
		float sinPhiM, cosPhiM, alphaSqr, cosThetaM;
		if (isotropic()) {
			sincosf((2.0f * MTS_PI) * sy, &sinPhiM, &cosPhiM);
			alphaSqr = alphaU * alphaU;
		} else {
			const float phiM = atanf(alphaV / alphaU * tanf(MTS_PI + 2 * MTS_PI * sy)) + MTS_PI * floorf(2 * sy + 0.5f);
			sincosf(phiM, &sinPhiM, &cosPhiM);
			const float cosSc = cosPhiM / alphaU, sinSc = sinPhiM / alphaV;
			alphaSqr = 1.0f / (cosSc * cosSc + sinSc * sinSc);
		}
		if (type == 0) {
			const float tanThetaMSqr = alphaSqr * -logf(1.0f - sx);
			cosThetaM = 1.0f / sqrtf(1.0f + tanThetaMSqr);
			pdf = (1.0f - sx) / (MTS_PI * alphaU * alphaV * cosThetaM * cosThetaM * cosThetaM);
		} else {
			const float tanThetaMSqr = alphaSqr * sx / (1.0f - sx);
			cosThetaM = 1.0f / sqrtf(1.0f + tanThetaMSqr);
			const float temp = 1 + tanThetaMSqr / alphaSqr;
			pdf = MTS_INV_PI / (alphaU * alphaV * cosThetaM * cosThetaM * cosThetaM * temp * temp);
		}
		if (pdf < 1e-20f) pdf = 0;
		const float sinThetaM = sqrtf(fmaxf(0.0f, 1 - cosThetaM * cosThetaM));
		return v3(sinThetaM * cosPhiM, sinThetaM * sinPhiM, cosThetaM);
	}

	// microfacet.h:572-696: slopes of the visible normals for alpha = 1
	__device__ inline void sample_visible11(float thetaI, float sx, float sy, float &slopeX, float &slopeY) const {
		const float SQRT_PI_INV = 0.5641895835477563f;
		if (type == 0) {
			if (thetaI < 1e-4f) {
				float s, c; const float r = sqrtf(-logf(1.0f - sx));
				sincosf(2 * MTS_PI * sy, &s, &c);
				slopeX = r * c; slopeY = r * s; return;
			}
			const float tanThetaI = tanf(thetaI), cotThetaI = 1 / tanThetaI;
			float a = -1, c = erff(cotThetaI);
			const float sample_x = fmaxf(sx, 1e-6f);
			const float fit = 1 + thetaI * (-0.876f + thetaI * (0.4265f - 0.0594f * thetaI));
			float b = c - (1 + c) * powf(1 - sample_x, fit);
			const float normalization = 1 / (1 + c + SQRT_PI_INV * tanThetaI * expf(-cotThetaI * cotThetaI));
			int it = 0;
			while (++it < 10) {
				if (!(b >= a && b <= c)) b = 0.5f * (a + c);
				const float invErf = erfinvf(b);
				const float value = normalization * (1 + b + SQRT_PI_INV * tanThetaI * expf(-invErf * invErf)) - sample_x;
				const float derivative = normalization * (1 - invErf * tanThetaI);
				if (fabsf(value) < 1e-5f) break;
				if (value > 0) c = b; else a = b;
				b -= value / derivative;
			}
			slopeX = erfinvf(b);
			slopeY = erfinvf(2.0f * fmaxf(sy, 1e-6f) - 1.0f);
		} else {
			if (thetaI < 1e-4f) {
				float s, c; const float r = safe_sqrt(sx / (1 - sx));
				sincosf(2 * MTS_PI * sy, &s, &c);
				slopeX = r * c; slopeY = r * s; return;
			}
			const float tanThetaI = tanf(thetaI), a = 1 / tanThetaI;
			const float G1 = 2.0f / (1.0f + safe_sqrt(1.0f + 1.0f / (a * a)));
			float A = 2.0f * sx / G1 - 1.0f;
			if (fabsf(A) == 1) A -= (A < 0 ? -1.0f : 1.0f) * MTS_EPSILON;
			const float tmp = 1.0f / (A * A - 1.0f);
			const float B = tanThetaI;
			const float D = safe_sqrt(B * B * tmp * tmp - (A * A - B * B) * tmp);
			const float slope_x_1 = B * tmp - D, slope_x_2 = B * tmp + D;
			slopeX = (A < 0.0f || slope_x_2 > 1.0f / tanThetaI) ? slope_x_1 : slope_x_2;
			float S;
			if (sy > 0.5f) { S = 1.0f; sy = 2.0f * (sy - 0.5f); } else { S = -1.0f; sy = 2.0f * (0.5f - sy); }
			const float z = (sy * (sy * (sy * (-0.365728915865723f) + 0.790235037209296f) - 0.424965825137544f) + 0.000152998850436920f) /
				(sy * (sy * (sy * (sy * 0.169507819808272f - 0.397203533833404f) - 0.232500544458471f) + 1.0f) - 0.539825872510702f);
			slopeY = S * z * sqrtf(1.0f + slopeX * slopeX);
		}
	}

	// microfacet.h:420-458
	__device__ inline V3 sample_visible(const V3 &_wi, float sx, float sy) const {
		const V3 wi = normalize(v3(alphaU * _wi.x, alphaV * _wi.y, _wi.z));
		float theta = 0, phi = 0;
		if (wi.z < 0.99999f) { theta = acosf(wi.z); phi = atan2f(wi.y, wi.x); }
		float sinPhi, cosPhi; sincosf(phi, &sinPhi, &cosPhi);
		float slx, sly; sample_visible11(theta, sx, sy, slx, sly);
		float rx = cosPhi * slx - sinPhi * sly, ry = sinPhi * slx + cosPhi * sly;
		rx *= alphaU; ry *= alphaV;
		const float normalization = 1.0f / sqrtf(rx * rx + ry * ry + 1.0f);
		return v3(-rx * normalization, -ry * normalization, normalization);
	}

	// microfacet.h:241-249
	__device__ inline V3 sample(const V3 &wi, float sx, float sy, float &pdfOut) const {
		if (sampleVisible) { const V3 m = sample_visible(wi, sx, sy); pdfOut = pdf_visible(wi, m); return m; }
		return sample_all(sx, sy, pdfOut);
	}
};

} // namespace mtsb
