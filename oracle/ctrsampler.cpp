/*
 * oracle/ctrsampler.cpp — TEST INFRASTRUCTURE.
 *
 * A Mitsuba *sampler plugin* ("ctr") written for this repository and built against the unmodified
 * reference by oracle/Makefile.ref.  It implements the reference's Sampler interface
 * (include/mitsuba/render/sampler.h:77-173; the stock implementation it stands in for is
 * src/samplers/independent.cpp:51-106) with the SAME counter-based stream the CUDA path uses:
 *
 *     word(x, y, sampleIndex, dim) = Philox4x32-10(ctr = {x, y, sampleIndex, dim/4},
 *                                                   key = {seed, 0x6d747362})[dim % 4]
 *     float = bitcast((word >> 9) | 0x3f800000) - 1      (same 23-bit trick as Random::nextFloat,
 *                                                        src/libcore/random.cpp:626-635)
 *
 * Sample ARRAYS (Sampler::request2DArray / next2DArray, src/librender/sampler.cpp:78-93 — MIDirectIntegrator asks for them when
 * emitterSamples or bsdfSamples > 1, direct.cpp:140-145) are counter-based too: entry j of the a-th requested 2D array
 * for sample index s of pixel (x, y) is the first two words of Philox4x32-10(ctr = {x, y, s, j}, key = {seed,
 * 0x32446100 + a}); 1D arrays use key 0x31446100 + a and the first word.
 *
 * `dim` counts next1D()/next2D() draws since generate()/advance().  With it the reference's own `path`
 * plugin consumes exactly the random numbers of the CUDA wavefront, so per-sample radiance can be
 * compared directly (classic block renderer: SamplingIntegrator::renderBlock,
 * src/librender/integrator.cpp:141-189 — generate(offset) per pixel, advance() per sample).
 *
 * ldDimension > 0 turns it into the counter-based twin of the `ldsampler` (src/samplers/ldsampler.cpp), the construction the
 * CUDA path uses (csrc/shading.cuh ld_next1D / ld_next2D): the k-th 1D / 2D draw of sample s, k < ldDimension, is entry
 * permute(s) of a van der Corput / (0,2) point set — the reference's own radicalInverse2Single / sobol2Single (qmc.h) — whose
 * scramble words and permutation key are Philox4x32-10(ctr = {x, y, k or 16 + k, 'ld1D' or 'ld2D'}, key = {seed, 0x6d747362});
 * later draws are words of the block stream {x, y, s, 0x80000000 | c / 4}.  The stock ldsampler draws those from a sequential
 * Random in scheduler order, which no other implementation can share; with this twin the reference's `path` consumes the CUDA
 * path's low-discrepancy numbers draw for draw.
 */
#include <mitsuba/render/sampler.h>
#include <mitsuba/core/qmc.h>

MTS_NAMESPACE_BEGIN

static inline void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
		uint32_t k0, uint32_t k1, uint32_t out[4]) {
	for (int r = 0; r < 10; ++r) {
		uint64_t p0 = (uint64_t) 0xD2511F53u * c0, p1 = (uint64_t) 0xCD9E8D57u * c2;
		uint32_t n0 = (uint32_t) (p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t) p1,
		         n2 = (uint32_t) (p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t) p0;
		c0 = n0; c1 = n1; c2 = n2; c3 = n3;
		k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
	}
	out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

class CounterSampler : public Sampler {
public:
	CounterSampler() : Sampler(Properties()) { }
	CounterSampler(const Properties &props) : Sampler(props) {
		m_sampleCount = props.getSize("sampleCount", 4);
		m_seed = (uint32_t) props.getInteger("seed", 0);
		m_ldDim = (uint32_t) props.getInteger("ldDimension", 0);
		m_px = m_py = 0; m_dim = 0; m_block = 0xFFFFFFFFu; m_ld1 = m_ld2 = m_ldc = 0;
	}
	CounterSampler(Stream *stream, InstanceManager *manager) : Sampler(stream, manager) {
		m_seed = stream->readUInt();
		m_ldDim = stream->readUInt();
		m_px = m_py = 0; m_dim = 0; m_block = 0xFFFFFFFFu; m_ld1 = m_ld2 = m_ldc = 0;
	}
	void serialize(Stream *stream, InstanceManager *manager) const {
		Sampler::serialize(stream, manager);
		stream->writeUInt(m_seed);
		stream->writeUInt(m_ldDim);
	}
	ref<Sampler> clone() {
		ref<CounterSampler> s = new CounterSampler();
		s->m_sampleCount = m_sampleCount; s->m_seed = m_seed; s->m_ldDim = m_ldDim;
		s->m_px = m_px; s->m_py = m_py; s->m_dim = 0; s->m_block = 0xFFFFFFFFu; s->m_ld1 = s->m_ld2 = s->m_ldc = 0;
		for (size_t i=0; i<m_req1D.size(); ++i) s->request1DArray(m_req1D[i]);
		for (size_t i=0; i<m_req2D.size(); ++i) s->request2DArray(m_req2D[i]);
		return s.get();
	}
	void generate(const Point2i &offset, size_t nextSampleIdx) {
		m_px = (uint32_t) offset.x; m_py = (uint32_t) offset.y;
		if (nextSampleIdx != (size_t) ~0) m_sampleIndex = nextSampleIdx;
		m_dim = 0; m_block = 0xFFFFFFFFu; m_ld1 = m_ld2 = m_ldc = 0;
		m_dimension1DArray = m_dimension2DArray = 0;
		// the stock samplers fill their arrays for the whole pixel here as well (independent.cpp:75-90)
		uint32_t w[4];
		for (size_t a = 0; a < m_req2D.size(); ++a)
			for (size_t smp = 0; smp < m_sampleCount; ++smp)
				for (size_t j = 0; j < m_req2D[a]; ++j) {
					philox4x32_10(m_px, m_py, (uint32_t) smp, (uint32_t) j, m_seed, 0x32446100u + (uint32_t) a, w);
					m_sampleArrays2D[a][smp * m_req2D[a] + j] = Point2(toFloat(w[0]), toFloat(w[1]));
				}
		for (size_t a = 0; a < m_req1D.size(); ++a)
			for (size_t smp = 0; smp < m_sampleCount; ++smp)
				for (size_t j = 0; j < m_req1D[a]; ++j) {
					philox4x32_10(m_px, m_py, (uint32_t) smp, (uint32_t) j, m_seed, 0x31446100u + (uint32_t) a, w);
					m_sampleArrays1D[a][smp * m_req1D[a] + j] = toFloat(w[0]);
				}
	}
	static inline Float toFloat(uint32_t w) {
		union { uint32_t u; float f; } x;
		x.u = (w >> 9) | 0x3f800000u;
		return x.f - 1.0f;
	}
	void advance() { Sampler::advance(); m_dim = 0; m_block = 0xFFFFFFFFu; m_ld1 = m_ld2 = m_ldc = 0; }
	void setSampleIndex(size_t idx) { Sampler::setSampleIndex(idx); m_dim = 0; m_block = 0xFFFFFFFFu; m_ld1 = m_ld2 = m_ldc = 0; }
	// ---- the low-discrepancy twin
	uint32_t ldMask() const { uint32_t lg = 0; while (((size_t) 1 << lg) < m_sampleCount) ++lg; return (1u << lg) - 1u; }
	static uint32_t ldPermute(uint32_t i, uint32_t w, uint32_t p) {      // Kensler 2013, power-of-two length
		i ^= p; i *= 0xe170893du; i ^= p >> 16; i ^= (i & w) >> 4; i ^= p >> 8; i *= 0x0929eb3fu; i ^= p >> 23; i ^= (i & w) >> 1;
		i *= 1u | p >> 27; i *= 0x6935fa69u; i ^= (i & w) >> 11; i *= 0x74dcb303u; i ^= (i & w) >> 2; i *= 0x9e501cc3u; i ^= (i & w) >> 2;
		i *= 0xc860a3dfu; i &= w; i ^= i >> 5;
		return (i + p) & w;
	}
	Float ldIndependent() {
		const uint32_t c = m_ldc & 0xffu; m_ldc = (m_ldc + 1) & 0xffu;      // the CUDA path keeps this counter in 8 bits
		uint32_t w[4];
		philox4x32_10(m_px, m_py, (uint32_t) m_sampleIndex, 0x80000000u | (c >> 2), m_seed, 0x6d747362u, w);
		return toFloat(w[c & 3]);
	}
	Float ldNext1D() {
		if (m_ld1 >= m_ldDim) return ldIndependent();
		uint32_t w[4];
		philox4x32_10(m_px, m_py, m_ld1++, 0x6c643144u, m_seed, 0x6d747362u, w);
		const uint32_t mask = ldMask();
		return radicalInverse2Single(ldPermute((uint32_t) m_sampleIndex & mask, mask, w[2]), w[0]);
	}
	Point2 ldNext2D() {
		if (m_ld2 >= m_ldDim) { Float a = ldIndependent(); Float b = ldIndependent(); return Point2(a, b); }
		uint32_t w[4];
		philox4x32_10(m_px, m_py, 16u + m_ld2++, 0x6c643244u, m_seed, 0x6d747362u, w);
		const uint32_t mask = ldMask(), i = ldPermute((uint32_t) m_sampleIndex & mask, mask, w[2]);
		const Float b = sobol2Single(i, w[1]);
		return Point2(radicalInverse2Single(i, w[0]), b < 0.99999994f ? b : 0.99999994f);       // the CUDA path clamps the value below 1
	}
	Float next1D() {
		if (m_ldDim) return ldNext1D();
		uint32_t blk = m_dim >> 2;
		if (blk != m_block) {
			philox4x32_10(m_px, m_py, (uint32_t) m_sampleIndex, blk, m_seed, 0x6d747362u, m_words);
			m_block = blk;
		}
		uint32_t w = m_words[m_dim & 3];
		++m_dim;
		union { uint32_t u; float f; } x;
		x.u = (w >> 9) | 0x3f800000u;
		return x.f - 1.0f;
	}
	Point2 next2D() { if (m_ldDim) return ldNext2D(); Float a = next1D(); Float b = next1D(); return Point2(a, b); }
	std::string toString() const { return "CounterSampler[]"; }
	MTS_DECLARE_CLASS()
private:
	uint32_t m_seed, m_px, m_py, m_dim, m_block, m_words[4];
	uint32_t m_ldDim, m_ld1, m_ld2, m_ldc;
};

MTS_IMPLEMENT_CLASS_S(CounterSampler, false, Sampler)
MTS_EXPORT_PLUGIN(CounterSampler, "Counter-based (Philox4x32-10) sampler shared with the CUDA path");
MTS_NAMESPACE_END
