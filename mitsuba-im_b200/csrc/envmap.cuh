// envmap.cuh — the `envmap` emitter (src/emitters/envmap.cpp) on the device: latitude-longitude radiance map behind the MIP
// pyramid of include/mitsuba/render/mipmap.h (repeat in u, clamp in v; bilinear level-0 lookups for secondary rays, the
// EWA-filtered lookup of MIPMap::eval for rays that carry differentials, i.e. sensor rays), importance sampling from the
// row / column CDFs weighted by sin(theta) with the tent-filtered sub-pixel offset.  The pyramid itself comes from the host:
// the reference builds it with a 2-lobed Lanczos filter and stores it in half precision (envmap.cpp:94-96, 152-178), and the
// values passed in are those texels.
#pragma once

namespace mtsb {

#define MTSB_ENVMAP_MAX_LEVELS 17          // images are smaller than 65536 pixels a side (envmap.cpp:159-161)
#define MTSB_MIPMAP_LUT_SIZE 64
#define ENVMAP_PI     3.14159265358979323846f
#define ENVMAP_INV_PI 0.31830988618379067154f
__host__ __device__ inline float env_safe_sqrt(float x) { return sqrtf(fmaxf(x, 0.0f)); }

struct EnvmapDev {
	int levels;
	int w[MTSB_ENVMAP_MAX_LEVELS], h[MTSB_ENVMAP_MAX_LEVELS];
	unsigned offset[MTSB_ENVMAP_MAX_LEVELS];       // first texel of a level in `data` (in texels)
	float ratioX[MTSB_ENVMAP_MAX_LEVELS], ratioY[MTSB_ENVMAP_MAX_LEVELS];   // m_sizeRatio (mipmap.h:243, 265-267)
	const float *data;                             // RGB triples, all levels
	const float *cdfRows, *cdfCols, *rowWeights;   // envmap.cpp:283-321
	float normalization, scale, pixelSizeX, pixelSizeY;
	float toWorld[9], toLocal[9];                  // linear parts of m_worldTransform and of its inverse
	float lut[MTSB_MIPMAP_LUT_SIZE];               // mipmap.h:297-303
	// the same record serves `bitmap` textures (src/textures/bitmap.cpp); an envmap is repeat / clamp, EWA, anisotropy 10
	int bcu, bcv;                                  // ReconstructionFilter::EBoundaryCondition: 0 repeat, 1 clamp, 2 mirror, 3 zero, 4 one
	int filterType;                                // EMIPFilterType: 0 nearest, 1 bilinear, 2 trilinear, 3 EWA
	float maxAnisotropy;
	float uvScaleX, uvScaleY, uvOffsetX, uvOffsetY;   // Texture2D (src/librender/texture.cpp:112-121)
};

// an ENVMAP emitter record carries the device address of its EnvmapDev in pos.x / pos.y (scene.cu)
__device__ __forceinline__ const EnvmapDev *envmap_of(const EmitterRec &em) {
	const unsigned long long a = (unsigned long long) __float_as_uint(em.pos.x) | ((unsigned long long) __float_as_uint(em.pos.y) << 32);
	return (const EnvmapDev *) a;
}

__device__ __forceinline__ float rgb_luminance(const V3 &c) { return 0.212671f * c.x + 0.715160f * c.y + 0.072169f * c.z; }
__device__ __forceinline__ V3 mat3_mul(const float *m, const V3 &v) {
	return v3(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[3] * v.x + m[4] * v.y + m[5] * v.z, m[6] * v.x + m[7] * v.y + m[8] * v.z);
}
__device__ __forceinline__ float safe_acos(float x) { return acosf(fminf(1.0f, fmaxf(-1.0f, x))); }

// MIPMap::evalTexel (mipmap.h:504-564)
__device__ __forceinline__ bool env_wrap(int bc, int &x, int size, float &constant) {
	if (x >= 0 && x < size) return true;
	switch (bc) {
		case 0: { const int r = x % size; x = r < 0 ? r + size : r; return true; }
		case 1: x = min(max(x, 0), size - 1); return true;
		case 2: { int r = x % (2 * size); if (r < 0) r += 2 * size; x = r >= size ? 2 * size - r - 1 : r; return true; }
		case 3: constant = 0.0f; return false;
		default: constant = 1.0f; return false;
	}
}
__device__ inline V3 env_texel(const EnvmapDev &E, int level, int x, int y) {
	const int w = E.w[level], h = E.h[level];
	float c = 0.0f;
	if (!env_wrap(E.bcu, x, w, c)) return v3(c);
	if (!env_wrap(E.bcv, y, h, c)) return v3(c);
	const float *p = E.data + ((size_t) E.offset[level] + (size_t) y * w + x) * 3;
	return v3(p[0], p[1], p[2]);
}
__device__ inline V3 env_box(const EnvmapDev &E, int level, float uvx, float uvy) {                    // :567-570
	return env_texel(E, level, (int) floorf(uvx * E.w[level]), (int) floorf(uvy * E.h[level]));
}
__device__ inline V3 env_bilinear(const EnvmapDev &E, int level, float uvx, float uvy) {               // :576-597
	if (!isfinite(uvx) || !isfinite(uvy)) return v3(0.0f);
	if (level >= E.levels) return env_box(E, E.levels - 1, uvx, uvy);
	const float u = uvx * E.w[level] - 0.5f, v = uvy * E.h[level] - 0.5f;
	const int xPos = (int) floorf(u), yPos = (int) floorf(v);
	const float dx1 = u - xPos, dx2 = 1.0f - dx1, dy1 = v - yPos, dy2 = 1.0f - dy1;
	return env_texel(E, level, xPos, yPos) * (dx2 * dy2) + env_texel(E, level, xPos, yPos + 1) * (dx2 * dy1)
	     + env_texel(E, level, xPos + 1, yPos) * (dx1 * dy2) + env_texel(E, level, xPos + 1, yPos + 1) * (dx1 * dy1);
}
__device__ inline V3 env_ewa(const EnvmapDev &E, int level, float uvx, float uvy, float A, float B, float C) {   // :765-832
	if (!isfinite(A + B + C + uvx + uvy)) return v3(0.0f);
	if (level >= E.levels) return env_box(E, E.levels - 1, uvx, uvy);
	const float u = uvx * E.w[level] - 0.5f, v = uvy * E.h[level] - 0.5f;
	const float rx = E.ratioX[level], ry = E.ratioY[level];
	A /= rx * rx; B /= rx * ry; C /= ry * ry;
	const float invDet = 1.0f / (-B * B + 4.0f * A * C), deltaU = 2.0f * sqrtf(C * invDet), deltaV = 2.0f * sqrtf(A * invDet);
	const int u0 = (int) ceilf(u - deltaU), u1 = (int) floorf(u + deltaU), v0 = (int) ceilf(v - deltaV), v1 = (int) floorf(v + deltaV);
	const float As = A * MTSB_MIPMAP_LUT_SIZE, Bs = B * MTSB_MIPMAP_LUT_SIZE, Cs = C * MTSB_MIPMAP_LUT_SIZE;
	V3 result = v3(0.0f);
	float denominator = 0.0f;
	const float ddq = 2 * As, uu0 = (float) u0 - u;
	for (int vt = v0; vt <= v1; ++vt) {
		const float vv = (float) vt - v;
		float q = As * uu0 * uu0 + (Bs * uu0 + Cs * vv) * vv;
		float dq = As * (2 * uu0 + 1) + Bs * vv;
		for (int ut = u0; ut <= u1; ++ut) {
			if (q < (float) MTSB_MIPMAP_LUT_SIZE) {
				const unsigned qi = (unsigned) q;
				if (qi < MTSB_MIPMAP_LUT_SIZE) {
					const float weight = E.lut[(int) q];
					result += env_texel(E, level, ut, vt) * weight;
					denominator += weight;
				}
			}
			q += dq; dq += ddq;
		}
	}
	if (denominator == 0) return env_bilinear(E, level, uvx, uvy);
	return result / denominator;
}
// MIPMap::eval (mipmap.h:630-716); an envmap: EWA with maxAnisotropy = 10 (envmap.cpp:143-144)
__device__ inline V3 env_filtered(const EnvmapDev &E, float uvx, float uvy, float d0x, float d0y, float d1x, float d1y) {
	if (E.filterType == 0) return env_box(E, 0, uvx, uvy);
	if (E.filterType == 1) return env_bilinear(E, 0, uvx, uvy);
	const float du0 = d0x * E.w[0], dv0 = d0y * E.h[0], du1 = d1x * E.w[0], dv1 = d1y * E.h[0];
	float A = dv0 * dv0 + dv1 * dv1, B = -2.0f * (du0 * dv0 + du1 * dv1), C = du0 * du0 + du1 * du1, F = A * C - B * B * 0.25f;
	const float root = hypotf(A - C, B), Aprime = 0.5f * (A + C - root), Cprime = 0.5f * (A + C + root);
	float majorRadius = Aprime != 0 ? sqrtf(F / Aprime) : 0, minorRadius = Cprime != 0 ? sqrtf(F / Cprime) : 0;
	if (E.filterType == 2 || !(minorRadius > 0) || !(majorRadius > 0) || F < 0) {
		const float level = log2f(fmaxf(majorRadius, MTS_EPSILON));
		const int ilevel = (int) floorf(level);
		if (ilevel < 0) return env_bilinear(E, 0, uvx, uvy);
		const float a = level - ilevel;
		return env_bilinear(E, ilevel, uvx, uvy) * (1.0f - a) + env_bilinear(E, ilevel + 1, uvx, uvy) * a;
	}
	const float maxAnisotropy = E.maxAnisotropy;
	if (minorRadius * maxAnisotropy < majorRadius) {
		minorRadius = majorRadius / maxAnisotropy;
		const float theta = 0.5f * atanf(B / (A - C));
		float sinTheta, cosTheta; sincosf(theta, &sinTheta, &cosTheta);
		const float a2 = majorRadius * majorRadius, b2 = minorRadius * minorRadius, sinTheta2 = sinTheta * sinTheta,
			cosTheta2 = cosTheta * cosTheta, sin2Theta = 2 * sinTheta * cosTheta;
		A = a2 * cosTheta2 + b2 * sinTheta2; B = (a2 - b2) * sin2Theta; C = a2 * sinTheta2 + b2 * cosTheta2; F = a2 * b2;
	}
	const float sc = 1.0f / F;
	A *= sc; B *= sc; C *= sc;
	const float level = fmaxf(0.0f, log2f(minorRadius));
	const int ilevel = (int) level;
	const float a = level - ilevel;
	if (majorRadius < 1 || !(A > 0 && C > 0)) return env_bilinear(E, ilevel, uvx, uvy);
	return env_ewa(E, ilevel, uvx, uvy, A, B, C) * (1.0f - a) + env_ewa(E, ilevel + 1, uvx, uvy, A, B, C) * a;
}

// EnvironmentMap::evalEnvironment (envmap.cpp:383-414): d, rx, ry in world space; rx / ry only for rays with differentials
__device__ inline V3 envmap_eval(const EnvmapDev &E, const V3 &d, bool hasDifferentials, const V3 &rx, const V3 &ry) {
	const V3 v = mat3_mul(E.toLocal, d);
	const float inv2pi = 1.0f / (2.0f * ENVMAP_PI);
	const float uvx = atan2f(v.x, -v.z) * inv2pi, uvy = safe_acos(v.y) * ENVMAP_INV_PI;
	V3 value;
	if (!hasDifferentials) value = env_bilinear(E, 0, uvx, uvy);
	else {
		const V3 dvdx = mat3_mul(E.toLocal, rx) - v, dvdy = mat3_mul(E.toLocal, ry) - v;
		const float t1 = inv2pi / (v.x * v.x + v.z * v.z), t2 = -ENVMAP_INV_PI / fmaxf(env_safe_sqrt(1.0f - v.y * v.y), MTS_EPSILON);
		value = env_filtered(E, uvx, uvy, t1 * (dvdx.z * v.x - dvdx.x * v.z), t2 * dvdx.y, t1 * (dvdy.z * v.x - dvdy.x * v.z), t2 * dvdy.y);
	}
	return value * E.scale;
}

// Texture2D::eval(its, filter = true) over BitmapTexture::eval (texture.cpp:112-121, bitmap.cpp:434-502): the filtered lookup
// where the intersection carries uv partials, the full-resolution one (bilinear, or box for `nearest`) otherwise
__device__ inline V3 texture_eval(const EnvmapDev &T, float u, float v, bool hasPartials, float dudx, float dvdx, float dudy, float dvdy) {
	const float uvx = u * T.uvScaleX + T.uvOffsetX, uvy = v * T.uvScaleY + T.uvOffsetY;
	if (hasPartials) return env_filtered(T, uvx, uvy, dudx * T.uvScaleX, dvdx * T.uvScaleY, dudy * T.uvScaleX, dvdy * T.uvScaleY);
	return T.filterType != 0 ? env_bilinear(T, 0, uvx, uvy) : env_box(T, 0, uvx, uvy);
}

// sampleReuse of envmap.cpp:663-668 (std::lower_bound over size + 1 entries)
__device__ inline uint32_t env_sample_reuse(const float *cdf, uint32_t size, float &sample) {
	uint32_t lo = 0, hi = size + 1;
	while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (cdf[mid] < sample) lo = mid + 1; else hi = mid; }
	const long e = (long) lo - 1;
	const uint32_t index = (uint32_t) min((long) size - 1, max(0l, e));
	sample = (sample - cdf[index]) / (cdf[index + 1] - cdf[index]);
	return index;
}
__device__ inline float interval_to_tent(float sample) {                      // warp.cpp:143-155
	float sign;
	if (sample < 0.5f) { sign = 1; sample *= 2; } else { sign = -1; sample = 2 * (sample - 0.5f); }
	return sign * (1 - sqrtf(sample));
}
// internalSampleDirection (envmap.cpp:566-606): direction in EMITTER space
__device__ inline void envmap_sample_direction(const EnvmapDev &E, float sx, float sy, V3 &d, V3 &value, float &pdf) {
	const int W = E.w[0], H = E.h[0];
	const uint32_t row = env_sample_reuse(E.cdfRows, (uint32_t) H, sy);
	const uint32_t col = env_sample_reuse(E.cdfCols + (size_t) row * (W + 1), (uint32_t) W, sx);
	const float px = (float) col + interval_to_tent(sx), py = (float) row + interval_to_tent(sy);
	const int xPos = (int) floorf(px), yPos = (int) floorf(py);
	const float dx1 = px - xPos, dx2 = 1.0f - dx1, dy1 = py - yPos, dy2 = 1.0f - dy1;
	const V3 value1 = env_texel(E, 0, xPos, yPos) * (dx2 * dy2) + env_texel(E, 0, xPos + 1, yPos) * (dx1 * dy2);
	const V3 value2 = env_texel(E, 0, xPos, yPos + 1) * (dx2 * dy1) + env_texel(E, 0, xPos + 1, yPos + 1) * (dx1 * dy1);
	value = (value1 + value2) * E.scale;
	pdf = (rgb_luminance(value1) * E.rowWeights[min(max(yPos, 0), H - 1)] + rgb_luminance(value2) * E.rowWeights[min(max(yPos + 1, 0), H - 1)]) * E.normalization;
	float sinPhi, cosPhi, sinTheta, cosTheta;
	sincosf(E.pixelSizeX * (px + 0.5f), &sinPhi, &cosPhi);
	sincosf(E.pixelSizeY * (py + 0.5f), &sinTheta, &cosTheta);
	d = v3(sinPhi * sinTheta, cosTheta, -cosPhi * sinTheta);
	pdf /= fmaxf(fabsf(sinTheta), MTS_EPSILON);
}
// internalPdfDirection (envmap.cpp:608-638): d in EMITTER space
__device__ inline float envmap_pdf_direction(const EnvmapDev &E, const V3 &d) {
	const float uvx = atan2f(d.x, -d.z) * (1.0f / (2.0f * ENVMAP_PI)), uvy = safe_acos(d.y) * ENVMAP_INV_PI;
	if (!isfinite(uvx) || !isfinite(uvy)) return 0.0f;
	const int W = E.w[0], H = E.h[0];
	const float u = uvx * W - 0.5f, v = uvy * H - 0.5f;
	const int xPos = (int) floorf(u), yPos = (int) floorf(v);
	const float dx1 = u - xPos, dx2 = 1.0f - dx1, dy1 = v - yPos, dy2 = 1.0f - dy1;
	const V3 value1 = env_texel(E, 0, xPos, yPos) * (dx2 * dy2) + env_texel(E, 0, xPos + 1, yPos) * (dx1 * dy2);
	const V3 value2 = env_texel(E, 0, xPos, yPos + 1) * (dx2 * dy1) + env_texel(E, 0, xPos + 1, yPos + 1) * (dx1 * dy1);
	const float sinTheta = env_safe_sqrt(1 - d.y * d.y);
	return (rgb_luminance(value1) * E.rowWeights[min(max(yPos, 0), H - 1)] + rgb_luminance(value2) * E.rowWeights[min(max(yPos + 1, 0), H - 1)])
		* E.normalization / fmaxf(fabsf(sinTheta), MTS_EPSILON);
}

} // namespace mtsb
