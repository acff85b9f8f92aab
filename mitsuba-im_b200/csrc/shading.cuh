// shading.cuh — device functions of the GPU plugin subset: BSDFs, emitters, warps, the counter-based sampler.
// Each function restates the reference plugin it stands in for (file:line given per function); all arithmetic is
// float32 on 3-channel linear RGB spectra, textures are constants.
#pragma once
#include "surface.cuh"

namespace mtsb {

#define MTS_PI        3.14159265358979323846f
#define MTS_INV_PI    0.31830988618379067154f
#define MTS_DELTA_EPS 1e-3f     /* DeltaEpsilon, include/mitsuba/core/constants.h */

// BSDF::EBSDFType bits (include/mitsuba/render/bsdf.h:224-296)
enum {
	B_NULL = 0x1, B_DIFFUSE_REFL = 0x2, B_DIFFUSE_TRANS = 0x4, B_GLOSSY_REFL = 0x8, B_GLOSSY_TRANS = 0x10,
	B_DELTA_REFL = 0x20, B_DELTA_TRANS = 0x40, B_DELTA1D_REFL = 0x80, B_DELTA1D_TRANS = 0x100,
	B_FRONT = 0x1000, B_BACK = 0x2000,
	B_SMOOTH = B_DIFFUSE_REFL | B_DIFFUSE_TRANS | B_GLOSSY_REFL | B_GLOSSY_TRANS,
	B_DELTA = B_NULL | B_DELTA_REFL | B_DELTA_TRANS,
	B_TRANSMISSION = B_DIFFUSE_TRANS | B_GLOSSY_TRANS | B_DELTA_TRANS | B_DELTA1D_TRANS | B_NULL
};
// EMeasure (include/mitsuba/render/common.h:52-66)
enum { M_SOLID_ANGLE = 1, M_LENGTH = 2, M_AREA = 3, M_DISCRETE = 4 };

// ----------------------------------------------------------------------------- sampler
// Philox4x32-10, counter {px, py, sampleIndex, dim/4}, key {seed, 0x6d747362}: the stream oracle/ctrsampler.cpp feeds
// the reference.  Replaces IndependentSampler::next1D/next2D (src/samplers/independent.cpp:51-106); the float
// conversion is Random::nextFloat's 23-bit mantissa trick (src/libcore/random.cpp:626-635).
// `sobol` != NULL selects the Sobol' sampler (src/samplers/sobol.cpp over sobolseq.h): the 1024 x 52 direction numbers; the path's
// 64-bit sequence index then travels in (py, sample) = (high, low) and `dim` is the Sobol' dimension.
struct SamplerState { uint32_t px, py, sample, dim, seed; const uint32_t *sobol; };

// sobol::sampleSingle (sobolseq.h:43-59), scramble = 0
__device__ inline float sobol_sample(const uint32_t *matrices, unsigned long long index, uint32_t dimension) {
	uint32_t result = 0;
	for (uint32_t i = min(dimension, 1023u) * 52u; index; index >>= 1, ++i)
		if (index & 1) result ^= __ldg(matrices + i);
	return fminf(result * (1.0f / 4294967296.0f), 0.999999940395355225f);
}

__host__ __device__ inline void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4]) {
	#pragma unroll
	for (int r = 0; r < 10; ++r) {
		const unsigned long long p0 = (unsigned long long) 0xD2511F53u * c0, p1 = (unsigned long long) 0xCD9E8D57u * c2;
		const uint32_t n0 = (uint32_t) (p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t) p1, n2 = (uint32_t) (p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t) p0;
		c0 = n0; c1 = n1; c2 = n2; c3 = n3;
		k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
	}
	out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// ---- `ldsampler` (src/samplers/ldsampler.cpp): per pixel and per dimension a scrambled (0,2)-sequence / van der Corput point set
// of sampleCount points in shuffled order, for the first `dimension` 1D and 2D draws of a sample; later draws are independent.
// The reference scrambles and shuffles with its sequential Random, per pixel in scheduler order (its own documentation: two runs
// differ), so the realisation cannot be shared; the point sets are the same construction (radicalInverse2Single / sample02Single,
// include/mitsuba/core/qmc.h:43-59,82-87,164-168) keyed by Philox on (pixel, dimension).  A `sobol` field holding a small
// integer instead of an address selects it: dimension | log2(sampleCount) << 8 (mtsb_render_set_ldsampler).  The 16-bit `dim`
// of the path state is split into the 1D counter (4 bits), the 2D counter (4 bits) and the count of independent draws (8 bits).
__device__ __forceinline__ uint32_t ld_tag(const uint32_t *p) { const unsigned long long v = (unsigned long long) p; return v < 0x10000ull ? (uint32_t) v : 0u; }
// Kensler's hashed permutation ("Correlated Multi-Jittered Sampling", 2013) of [0, mask] for a power-of-two length: every step is a
// bijection on the low bits, so no cycle walking is needed
__device__ __forceinline__ uint32_t ld_permute(uint32_t i, uint32_t w, uint32_t p) {
	i ^= p; i *= 0xe170893du; i ^= p >> 16; i ^= (i & w) >> 4; i ^= p >> 8; i *= 0x0929eb3fu; i ^= p >> 23; i ^= (i & w) >> 1;
	i *= 1u | p >> 27; i *= 0x6935fa69u; i ^= (i & w) >> 11; i *= 0x74dcb303u; i ^= (i & w) >> 2; i *= 0x9e501cc3u; i ^= (i & w) >> 2;
	i *= 0xc860a3dfu; i &= w; i ^= i >> 5;
	return (i + p) & w;
}
__device__ __forceinline__ float ld_radical_inverse2(uint32_t n, uint32_t scramble) {      // qmc.h:43-59
	n = (__brev(n) >> 8) ^ (scramble & 0x00ffffffu);
	return (float) n / 16777216.0f;
}
__device__ __forceinline__ float ld_sobol2(uint32_t n, uint32_t scramble) {                // qmc.h:82-87
	for (uint32_t v = 1u << 31; n != 0; n >>= 1, v ^= v >> 1)
		if (n & 1) scramble ^= v;
	return fminf((float) scramble / 4294967296.0f, 0.999999940395355225f);                  // the division rounds to 1.0f for the top 128 values
}
__device__ inline float ld_independent(SamplerState &s) {
	const uint32_t c = (s.dim >> 8) & 0xffu;
	s.dim = (s.dim & 0xffu) | (((c + 1u) & 0xffu) << 8);
	uint32_t w[4];
	philox4x32_10(s.px, s.py, s.sample, 0x80000000u | (c >> 2), s.seed, 0x6d747362u, w);
	const uint32_t word = (c & 3) == 0 ? w[0] : ((c & 3) == 1 ? w[1] : ((c & 3) == 2 ? w[2] : w[3]));
	return __uint_as_float((word >> 9) | 0x3f800000u) - 1.0f;
}
__device__ inline float ld_next1D(SamplerState &s, uint32_t tag) {                           // ldsampler.cpp:225-231
	const uint32_t d = s.dim & 0xfu;
	if (d >= (tag & 0xffu)) return ld_independent(s);
	s.dim += 1u;
	uint32_t w[4];
	philox4x32_10(s.px, s.py, d, 0x6c643144u, s.seed, 0x6d747362u, w);
	const uint32_t mask = (1u << (tag >> 8)) - 1u;
	return ld_radical_inverse2(ld_permute(s.sample & mask, mask, w[2]), w[0]);
}
__device__ inline void ld_next2D(SamplerState &s, uint32_t tag, float &a, float &b) {        // ldsampler.cpp:233-239
	const uint32_t d = (s.dim >> 4) & 0xfu;
	if (d >= (tag & 0xffu)) { a = ld_independent(s); b = ld_independent(s); return; }
	s.dim += 16u;
	uint32_t w[4];
	philox4x32_10(s.px, s.py, 16u + d, 0x6c643244u, s.seed, 0x6d747362u, w);
	const uint32_t mask = (1u << (tag >> 8)) - 1u, i = ld_permute(s.sample & mask, mask, w[2]);
	a = ld_radical_inverse2(i, w[0]); b = ld_sobol2(i, w[1]);                                // sample02Single
}

__device__ inline float next1D(SamplerState &s) {
	if (s.sobol) {
		const uint32_t tag = ld_tag(s.sobol);
		if (tag) return ld_next1D(s, tag);
		return sobol_sample(s.sobol, ((unsigned long long) s.py << 32) | s.sample, s.dim++);      // SobolSampler::next1D (sobol.cpp:218-228)
	}
	uint32_t w[4];
	philox4x32_10(s.px, s.py, s.sample, s.dim >> 2, s.seed, 0x6d747362u, w);
	const uint32_t word = (s.dim & 3) == 0 ? w[0] : ((s.dim & 3) == 1 ? w[1] : ((s.dim & 3) == 2 ? w[2] : w[3]));
	++s.dim;
	return __uint_as_float((word >> 9) | 0x3f800000u) - 1.0f;
}
__device__ inline void next2D(SamplerState &s, float &a, float &b) {
	if (s.sobol) {
		const uint32_t tag = ld_tag(s.sobol);
		if (tag) { ld_next2D(s, tag, a, b); return; }
		// SobolSampler::next2D (:230-251).  Dimensions [m_arrayStartDim, m_arrayEndDim) = [5, 5 + requests) are reserved for sample
		// arrays, and the test `m_dimension + 1 >= m_arrayStartDim && m_dimension < m_arrayEndDim` moves a pair that would start at
		// dimension 4 to 5 even when nothing is reserved.  (The pixel-relative first pair is generate_kernel's business.)
		if (s.dim == 4) s.dim = 5;
		a = next1D(s); b = next1D(s); return;
	}
	if ((s.dim & 3) < 3) {   // both words come from the same Philox block
		uint32_t w[4];
		philox4x32_10(s.px, s.py, s.sample, s.dim >> 2, s.seed, 0x6d747362u, w);
		const uint32_t k = s.dim & 3;
		const uint32_t wa = k == 0 ? w[0] : (k == 1 ? w[1] : w[2]), wb = k == 0 ? w[1] : (k == 1 ? w[2] : w[3]);
		s.dim += 2;
		a = __uint_as_float((wa >> 9) | 0x3f800000u) - 1.0f;
		b = __uint_as_float((wb >> 9) | 0x3f800000u) - 1.0f;
	} else { a = next1D(s); b = next1D(s); }
}

// ----------------------------------------------------------------------------- warps (src/libcore/warp.cpp)
__device__ inline float safe_sqrt(float x) { return sqrtf(fmaxf(x, 0.0f)); }

__device__ inline void square_to_disk_concentric(float sx, float sy, float &x, float &y) {   // warp.cpp:81-103
	const float r1 = 2.0f * sx - 1.0f, r2 = 2.0f * sy - 1.0f;
	float phi, r;
	if (r1 == 0 && r2 == 0) { r = phi = 0; }
	else if (r1 * r1 > r2 * r2) { r = r1; phi = (MTS_PI / 4.0f) * (r2 / r1); }
	else { r = r2; phi = (MTS_PI / 2.0f) - (r1 / r2) * (MTS_PI / 4.0f); }
	float s, c; sincosf(phi, &s, &c);
	x = r * c; y = r * s;
}
__device__ inline V3 square_to_cosine_hemisphere(float sx, float sy) {   // warp.cpp:43-52
	float x, y; square_to_disk_concentric(sx, sy, x, y);
	float z = safe_sqrt(1.0f - x * x - y * y);
	if (z == 0) z = 1e-10f;
	return v3(x, y, z);
}
__device__ inline V3 square_to_uniform_sphere(float sx, float sy) {      // warp.cpp:26-33
	const float z = 1.0f - 2.0f * sy, r = safe_sqrt(1.0f - z * z);
	float s, c; sincosf(2.0f * MTS_PI * sx, &s, &c);
	return v3(r * c, r * s, z);
}
__device__ inline void square_to_uniform_triangle(float sx, float sy, float &a, float &b) {  // warp.cpp:75-78
	const float t = safe_sqrt(1.0f - sx);
	a = 1 - t; b = t * sy;
}

// ----------------------------------------------------------------------------- Fresnel (src/libcore/util.cpp)
__device__ inline float fresnel_dielectric_ext(float cosThetaI_, float &cosThetaT_, float eta) {   // util.cpp:653-683
	if (eta == 1) { cosThetaT_ = -cosThetaI_; return 0.0f; }
	const float scale = (cosThetaI_ > 0) ? 1 / eta : eta;
	const float cosThetaTSqr = 1 - (1 - cosThetaI_ * cosThetaI_) * (scale * scale);
	if (cosThetaTSqr <= 0.0f) { cosThetaT_ = 0.0f; return 1.0f; }
	const float cosThetaI = fabsf(cosThetaI_), cosThetaT = sqrtf(cosThetaTSqr);
	const float Rs = (cosThetaI - eta * cosThetaT) / (cosThetaI + eta * cosThetaT);
	const float Rp = (eta * cosThetaI - cosThetaT) / (eta * cosThetaI + cosThetaT);
	cosThetaT_ = (cosThetaI_ > 0) ? -cosThetaT : cosThetaT;
	return 0.5f * (Rs * Rs + Rp * Rp);
}
__device__ inline float fresnel_dielectric_ext(float cosThetaI, float eta) { float t; return fresnel_dielectric_ext(cosThetaI, t, eta); }

__device__ inline V3 vsafe_sqrt(V3 a) { return v3(safe_sqrt(a.x), safe_sqrt(a.y), safe_sqrt(a.z)); }
__device__ inline V3 fresnel_conductor_exact(float cosThetaI, V3 eta, V3 k) {   // util.cpp:741-763
	const float cosThetaI2 = cosThetaI * cosThetaI, sinThetaI2 = 1 - cosThetaI2, sinThetaI4 = sinThetaI2 * sinThetaI2;
	const V3 temp1 = eta * eta - k * k - v3(sinThetaI2);
	const V3 a2pb2 = vsafe_sqrt(temp1 * temp1 + k * k * eta * eta * 4.0f);
	const V3 a = vsafe_sqrt((a2pb2 + temp1) * 0.5f);
	const V3 term1 = a2pb2 + v3(cosThetaI2), term2 = a * (2 * cosThetaI);
	const V3 Rs2 = (term1 - term2) / (term1 + term2);
	const V3 term3 = a2pb2 * cosThetaI2 + v3(sinThetaI4), term4 = term2 * sinThetaI2;
	const V3 Rp2 = Rs2 * (term3 - term4) / (term3 + term4);
	return 0.5f * (Rp2 + Rs2);
}

} // namespace mtsb
#include "microfacet.cuh"
#include "bsdf.cuh"
#include "emitter.cuh"
#include "texture.cuh"
