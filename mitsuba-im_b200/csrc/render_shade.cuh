// render_shade.cuh — the shade and splat kernels (included by render.cu).
#pragma once

namespace mtsb {

__device__ __forceinline__ float mi_weight(float pdfA, float pdfB) {   // path.cpp:296-300 (power heuristic)
	pdfA *= pdfA; pdfB *= pdfB;
	return pdfA / (pdfA + pdfB);
}

// What one shading step hands back to its caller: whether the path continues (its next ray is in ps.rayO/rayD[slot]) and
// the shadow-ray entry of the emitter sample, if any.  The wavefront kernel appends both to queues, the late-bounce
// kernel (render_tail.cuh) keeps them in registers.
struct ShadeOut { bool alive, shadow; float4 shO, shD, shL; };

// One shading step of one path; state lives in ps.*[slot].  Line references are to src/integrators/path/path.cpp.
template <bool LEAN = false>
__device__ __forceinline__ void shade_path(const SceneDev &S, const PathState &ps, const IntegratorDev &integ, uint32_t seed,
		uint32_t slot, ShadeOut &out) {
	// Path state touched per bounce (HBM-bound on small scenes, so every 16 bytes count): rayO rayD hit thr prev rng are read,
	// rayO rayD thr prev rng written.  Flags and depth ride in rng.w beside the sampler dimension (PATH_* below); the
	// radiance L is read and written only by the bounces that add to it (an emitter was hit); the sample position stays in
	// `misc`, written once by generate and read once by splat.
	const float4 r0 = ps.rayO[slot], r1 = ps.rayD[slot], h = ps.hit[slot];
	float4 thr4 = ps.thr[slot], prev = ps.prev[slot];
	const uint4 rg = ps.rng[slot];
	SamplerState rs; rs.px = rg.x; rs.py = rg.y; rs.sample = rg.z; rs.dim = PATH_DIM(rg.w); rs.seed = seed; rs.sobol = LEAN ? nullptr : integ.sobol;      // a render with the sobol sampler runs the full kernel
	const V3 o = v3(r0.x, r0.y, r0.z), d = v3(r1.x, r1.y, r1.z);
	V3 thr = v3(thr4.x, thr4.y, thr4.z), Li = v3(0.0f);          // Li: what THIS bounce adds
	float eta = thr4.w;
	uint32_t flags = PATH_FLAGS(rg.w), depth = PATH_DEPTH(rg.w);
	float alpha = 0.0f; bool first = false;
	const uint32_t hitSlot = __float_as_uint(h.w);
	const bool valid = hitSlot != INVALID_SLOT;
	bool alive = true;

	Surf sf;
	int emitterId = -1, bsdfId = 0;
	if (valid) {
		fill_surface(S, hitSlot, h.x, h.y, h.z, o, d, sf);
		emitterId = S.shapes[sf.shape].emitter; bsdfId = S.shapes[sf.shape].bsdf;
	}

	if (flags & PF_FROM_BSDF) {
		// ---- tail of the previous iteration (:224-286): the BSDF sample's ray has been traced
		const V3 refN = v3(prev.x, prev.y, prev.z);
		const float bsdfPdf = prev.w;
		bool hitEmitter = false; V3 value = v3(0.0f); float lumPdf = 0;
		if (valid) {
			if (emitterId >= 0) {
				value = area_emitter_eval(S, emitterId, sf.ns, -d);
				hitEmitter = true;
				if (!(flags & PF_DELTA))
					lumPdf = emitter_pdf_direct<LEAN>(S, emitterId, refN, d, h.x, sf.ns, M_SOLID_ANGLE, o) * emitter_pdf_discrete(S, emitterId);
			}
		} else if (S.envEmitter >= 0) {
			if (integ.hideEmitters && !(flags & PF_SCATTERED)) alive = false;
			else {
				const EmitterRec &env = S.emitters[S.envEmitter];
				value = env_radiance<LEAN>(S, d);
				float nearT, farT;       // fillDirectSamplingRecord (constant.cpp:238-256)
				if (!sphere_intersect(env.sphereCenter, env.sphereRadius, o, d, nearT, farT) || nearT > 0 || farT < 0) alive = false;
				else {
					hitEmitter = true;
					if (!(flags & PF_DELTA)) {
						const V3 n = normalize(env.sphereCenter - (o + d * farT));
						lumPdf = emitter_pdf_direct<LEAN>(S, S.envEmitter, refN, d, farT, n, M_SOLID_ANGLE) * emitter_pdf_discrete(S, S.envEmitter);
					}
				}
			}
		} else alive = false;
		if (alive && hitEmitter) Li += thr * value * mi_weight(bsdfPdf, lumPdf);
		if (!valid) alive = false;                      // :262 "if (!its.isValid() ...) break"
		if (alive) {
			if (depth++ >= (uint32_t) integ.rrDepth) {     // :266-286 Russian roulette
				const float q = fminf(vmax(thr) * eta * eta, 0.95f);
				if (next1D(rs) >= q) alive = false;
				else thr /= q;
			}
		}
		if (alive && !((int) depth <= integ.maxDepth || integ.maxDepth < 0)) alive = false;   // while condition (:135)
	} else {
		// ---- first iteration: the sensor ray (:129-150)
		first = true; alpha = valid ? 1.0f : 0.0f;      // RadianceQueryRecord::rayIntersect sets alpha
		if (!valid) {
			if (!integ.hideEmitters && S.envEmitter >= 0) {
				// the sensor ray carries differentials (generate_kernel left them in prev / L): a filtered lookup for an `envmap`
				V3 rx = v3(0.0f), ry = v3(0.0f);
				if (!LEAN && S.emitters[S.envEmitter].type == MTSB_EMITTER_ENVMAP) { const float4 l = ps.L[slot]; rx = v3(prev.x, prev.y, prev.z); ry = v3(l.x, l.y, l.z); }
				Li += thr * env_radiance<LEAN>(S, d, true, rx, ry);
			}
			alive = false;
		} else if (emitterId >= 0 && !integ.hideEmitters) {
			Li += thr * area_emitter_eval(S, emitterId, sf.ns, -d);
		}
	}

	bool shadow = false;
	float4 shO = make_float4(0, 0, 0, 0), shD = shO, shL = shO;
	if (alive) {
		const V3 wi = to_local(sf, -d);
		// :155-164
		if (((int) depth >= integ.maxDepth && integ.maxDepth > 0) || (integ.strictNormals && dot(d, sf.ng) * wi.z >= 0)) alive = false;
		if (alive) {
			const BsdfRec bsdf = load_bsdf<LEAN>(S, bsdfId, sf, hitSlot, o, first, prev, ps.L + slot);
			const uint32_t btype = bsdf_type(bsdf);
			const V3 refN = (btype & (B_TRANSMISSION | B_BACK)) == 0 ? sf.ns : v3(0.0f);   // DirectSamplingRecord(its), records.inl:160-164
			// ---- emitter sampling (:171-199)
			// MIDirectIntegrator draws the emitter sample before it looks at the BSDF type (direct.cpp:211-215); MIPathTracer
			// only inside the smooth branch (path.cpp:171-176) — the stream position must follow the integrator in use
			float sx = 0, sy = 0;
			const bool smooth = (btype & B_SMOOTH) && S.emitterCount > 0;
			if (smooth || integ.type == MTSB_INTEGRATOR_DIRECT) next2D(rs, sx, sy);
			if (smooth) {
				DirectSample ds;
				const V3 value = scene_sample_emitter_direct<LEAN>(S, sf.p, refN, sx, sy, ds);
				if (!is_zero(value)) {
					const V3 wo = to_local(sf, ds.d);
					const V3 bsdfVal = bsdf_eval<LEAN>(bsdf, wi, wo);
					if (!is_zero(bsdfVal) && (!integ.strictNormals || dot(sf.ng, ds.d) * wo.z > 0)) {
						// Emitter::isOnSurface(): area (area.cpp:72) and constant environment (constant.cpp:50) set EOnSurface, point does not
						const bool onSurface = S.emitters[ds.emitter].type != MTSB_EMITTER_POINT;
						const float bp = (onSurface && ds.measure == M_SOLID_ANGLE) ? bsdf_pdf<LEAN>(bsdf, wi, wo) : 0;
						const V3 c = thr * value * bsdfVal * mi_weight(ds.pdf, bp);
						shadow = true;      // visibility test of Scene::sampleEmitterDirect (scene.cpp:871-876), deferred to `connect`
						shO = make_float4(sf.p.x, sf.p.y, sf.p.z, MTS_EPSILON);
						shD = make_float4(ds.d.x, ds.d.y, ds.d.z, ds.dist * (1 - MTS_SHADOW_EPSILON));
						shL = make_float4(c.x, c.y, c.z, __uint_as_float(slot));
					}
				}
			}
			// ---- BSDF sampling (:205-222)
			next2D(rs, sx, sy);
			const float sz = (!LEAN && bsdf_uses_sampler(bsdf)) ? next1D(rs) : 0.0f;
			BsdfSample bs;
			const V3 bsdfWeight = bsdf_sample<LEAN>(bsdf, wi, sx, sy, sz, bs);
			if (is_zero(bsdfWeight)) alive = false;
			else {
				flags |= PF_SCATTERED;
				const V3 wo = to_world(sf, bs.wo);
				if (integ.strictNormals && dot(sf.ng, wo) * bs.wo.z <= 0) alive = false;
				else {
					ps.rayO[slot] = make_float4(sf.p.x, sf.p.y, sf.p.z, MTS_EPSILON);
					ps.rayD[slot] = make_float4(wo.x, wo.y, wo.z, INFINITY);
					thr *= bsdfWeight; eta *= bs.eta;           // :252-253 (applied before the trace; unused if the path ends)
					flags |= PF_FROM_BSDF;
					if (bs.type & B_DELTA) flags |= PF_DELTA; else flags &= ~PF_DELTA;
					ps.prev[slot] = make_float4(refN.x, refN.y, refN.z, bs.pdf);
				}
			}
		}
	}

	if (first) ps.L[slot] = make_float4(Li.x, Li.y, Li.z, alpha);             // the sensor ray's bounce starts the sum
	else if (!is_zero(Li)) {
		float4 L4 = ps.L[slot];
		L4.x += Li.x; L4.y += Li.y; L4.z += Li.z;
		ps.L[slot] = L4;
	}
	ps.rng[slot] = make_uint4(rg.x, rg.y, rg.z, PATH_PACK(rs.dim, flags, depth));  // always: splat reads the depth
	if (alive) ps.thr[slot] = make_float4(thr.x, thr.y, thr.z, eta);
	out.alive = alive; out.shadow = shadow; out.shO = shO; out.shD = shD; out.shL = shL;
}

// One thread per live path, grid-stride: the grid is sized to the machine (render_host.cuh), not to the chunk, so a late
// bounce with a few thousand survivors does not launch 65 K blocks that exit on the count.
#ifndef MTSB_SHADE_MIN_BLOCKS
#define MTSB_SHADE_MIN_BLOCKS 5
#endif
template <bool LEAN>
__global__ void __launch_bounds__(RENDER_BLOCK, MTSB_SHADE_MIN_BLOCKS)
shade_kernel(SceneDev S, PathState ps, IntegratorDev integ, uint32_t seed) {
	const uint32_t parity = ps.counters[CTR_PARITY] & 1u;
	const uint32_t n = ps.counters[parity];
	const uint32_t *queueIn = parity ? ps.queueB : ps.queueA;
	uint32_t *queueOut = parity ? ps.queueA : ps.queueB;
	uint32_t *countOut = ps.counters + (parity ^ 1u), *countShadow = ps.counters + CTR_SHADOW;
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
		const uint32_t slot = QUEUE_SLOT(queueIn[i]);
		// a slot whose ray the extend stage parked is not shaded yet: it travels on to the next bounce's queue, flagged
		if (__float_as_uint(ps.hit[slot].w) == SUSPENDED_SLOT) { queueOut[warp_append(countOut)] = slot | QUEUE_RESUME; continue; }
		ShadeOut so;
		shade_path<LEAN>(S, ps, integ, seed, slot, so);
		if (so.alive) queueOut[warp_append(countOut)] = slot;
		if (so.shadow) {
			const uint32_t k = warp_append(countShadow);
			ps.shO[k] = so.shO; ps.shD[k] = so.shD; ps.shL[k] = so.shL;
		}
	}
}

// ImageBlock::put (imageblock.h:161-221) with atomics (the Integrator2 path uses putAtomic, :223-318).
//
// A warp holds the 32 samples of one 8x4 pixel tile (generate_kernel's slot order), so their filter footprints overlap
// almost completely: with the Gaussian filter (radius 2) 32 x 25 pixel updates land on at most 13 x 9 pixels.  The warp
// therefore accumulates into a private shared-memory tile first (shared atomics) and flushes each touched film float
// ONCE with a global atomic: ~18 global atomics per lane instead of 125.  Warps whose footprint does not fit the tile
// (explicit sample lists, very wide filters) add straight to the film.
#define SPLAT_TILE_W 16
#define SPLAT_TILE_H 12
#define SPLAT_MAX_CH 5
#define SPLAT_STAT_STRIPES 64     // striped path-count / depth-sum counters behind ps.stats[12]

__global__ void __launch_bounds__(RENDER_BLOCK)
splat_kernel(PathState ps, FilmDev film, uint32_t count, float *samplesOut, float *posOut) {
	__shared__ float shTile[RENDER_BLOCK / 32][SPLAT_TILE_W * SPLAT_TILE_H * SPLAT_MAX_CH];
	const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
	const unsigned lane = threadIdx.x & 31u, FULL = 0xffffffffu;
	float *tile = shTile[threadIdx.x >> 5];
	unsigned depth = 0, one = 0;
	bool ok = false;
	float value[SPLAT_MAX_CH] = { 0, 0, 0, 0, 1.0f };
	float px = 0, py = 0;
	int x0 = 1 << 30, y0 = 1 << 30, x1 = -(1 << 30), y1 = -(1 << 30);
	const int sizeX = film.width + 2 * film.border, sizeY = film.height + 2 * film.border;
	if (j < count && ps.rng[j].x != INVALID_SLOT) {
		one = 1;
		const float4 L = ps.L[j], misc = ps.misc[j];
		depth = PATH_DEPTH(ps.rng[j].w);
		if (samplesOut) {
			((float4 *) samplesOut)[j] = L;
			if (posOut) { posOut[2 * j] = misc.x; posOut[2 * j + 1] = misc.y; }
		} else {
			value[0] = L.x; value[1] = L.y; value[2] = L.z; value[3] = L.w;
			ok = true;
			for (int c = 0; c < film.channels; ++c) if (!isfinite(value[c]) || value[c] < 0) ok = false;
			if (ok) {
				px = misc.x - 0.5f - (float) (film.offX - film.border); py = misc.y - 0.5f - (float) (film.offY - film.border);
				x0 = max((int) ceilf(px - film.radius), 0); y0 = max((int) ceilf(py - film.radius), 0);
				x1 = min((int) floorf(px + film.radius), sizeX - 1); y1 = min((int) floorf(py + film.radius), sizeY - 1);
				if (x1 < x0 || y1 < y0) ok = false;
			}
		}
	}
	if (!samplesOut) {
		// warp footprint
		int bx0 = ok ? x0 : (1 << 30), by0 = ok ? y0 : (1 << 30), bx1 = ok ? x1 : -(1 << 30), by1 = ok ? y1 : -(1 << 30);
		#pragma unroll
		for (int o = 16; o > 0; o >>= 1) {
			bx0 = min(bx0, __shfl_xor_sync(FULL, bx0, o)); by0 = min(by0, __shfl_xor_sync(FULL, by0, o));
			bx1 = max(bx1, __shfl_xor_sync(FULL, bx1, o)); by1 = max(by1, __shfl_xor_sync(FULL, by1, o));
		}
		const bool any = bx1 >= bx0;
		const int tw = bx1 - bx0 + 1, th = by1 - by0 + 1;
		const bool fits = any && tw <= SPLAT_TILE_W && th <= SPLAT_TILE_H;
		const int ch = film.channels;
		if (fits) {
			const int n = tw * th * ch;
			for (int i = lane; i < n; i += 32) tile[i] = 0.0f;
			__syncwarp();
		}
		if (ok) {
			for (int y = y0; y <= y1; ++y) {
				const float wy = film.table[min((int) fabsf((y - py) * film.scale), 31)];
				for (int x = x0; x <= x1; ++x) {
					const float w = film.table[min((int) fabsf((x - px) * film.scale), 31)] * wy;
					if (fits) {
						float *dest = tile + ((y - by0) * tw + (x - bx0)) * ch;
						for (int c = 0; c < ch; ++c) atomicAdd(dest + c, w * value[c]);
					} else {
						float *dest = film.data + ((size_t) y * sizeX + x) * ch;
						for (int c = 0; c < ch; ++c) atomicAdd(dest + c, w * value[c]);
					}
				}
			}
		}
		if (fits) {
			__syncwarp();
			const int rowFloats = tw * ch, n = rowFloats * th;
			for (int i = lane; i < n; i += 32) {
				const float v = tile[i];
				if (v != 0.0f) {
					const int y = i / rowFloats, r = i - y * rowFloats;
					atomicAdd(film.data + ((size_t) (by0 + y) * sizeX + bx0) * ch + r, v);
				}
			}
		}
	}
	// avgPathLength += rRec.depth (path.cpp:290-291)
	unsigned sum = depth;
	#pragma unroll
	for (int o = 16; o > 0; o >>= 1) { sum += __shfl_xor_sync(FULL, sum, o); one += __shfl_xor_sync(FULL, one, o); }
	// one pair of counters per 64 blocks' worth of stripes: half a million warps adding to ONE address cost more than the
	// film atomics (the same-address adds serialise in L2: 1 ms of the 12 ms Cornell frame); mtsb_render_stats sums the stripes
	if (lane == 0 && one) {
		unsigned long long *stripe = ps.stats + 12 + 2 * (blockIdx.x & (SPLAT_STAT_STRIPES - 1));
		atomicAdd(stripe + 1, (unsigned long long) sum); atomicAdd(stripe, (unsigned long long) one);
	}
}

// Headless stand-in for the stacking arithmetic of the im-mts preview (src/im-mts/preview_opengl2.cpp:188-270): the GUI
// uploads every worker framebuffer (un-normalised RGBA sums) as a texture and blends them into one target with
// glBlendFunc(GL_CONSTANT_COLOR, GL_ONE), blend colour (s, s, s, 0), s = 1 / max(sum of the buffers' sample counts, 1), over
// a clear colour of alpha 1 — i.e. rgb = s * sum_k film_k.rgb, a = 1.  Same arithmetic here, on the device films directly.
struct StackIn { const float *film[8]; int n, channels; float scale; };
__global__ void preview_stack_kernel(StackIn in, size_t pixels, float4 *out) {
	const size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x;
	if (i >= pixels) return;
	float r = 0, g = 0, b = 0;
	for (int k = 0; k < in.n; ++k) {
		const float *p = in.film[k] + i * in.channels;
		r += p[0]; g += p[1]; b += p[2];
	}
	out[i] = make_float4(r * in.scale, g * in.scale, b * in.scale, 1.0f);
}

// ---- pointwise plugin math for parity tests (mtsb_eval_bsdf / mtsb_eval_emitter) ----
// The sampler's draws for `n` consecutive samples of one pixel: `draws` next2D calls, then `draws` next1D calls per sample,
// written as [n][draws][3] (pointwise checks of the samplers: stratification of the ldsampler's point sets, Sobol' dimensions)
__global__ void sampler_dump_kernel(IntegratorDev integ, uint32_t seed, uint32_t px, uint32_t py, uint32_t first, uint32_t n, uint32_t draws, float *out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	SamplerState rs; rs.px = px; rs.py = py; rs.sample = first + i; rs.dim = 0; rs.seed = seed; rs.sobol = integ.sobol;
	float *o = out + (size_t) i * draws * 3;
	for (uint32_t k = 0; k < draws; ++k) next2D(rs, o[k * 3], o[k * 3 + 1]);
	for (uint32_t k = 0; k < draws; ++k) o[k * 3 + 2] = next1D(rs);
}

__global__ void eval_bsdf_kernel(SceneDev S, int bsdfId, const float *in, const float *s3, size_t n, float *out) {
	const size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x;
	if (i >= n) return;
	const BsdfRec b = S.bsdfs[bsdfId];
	const float *r = in + i * 8; float *o = out + i * 16;
	const V3 wi = v3(r[0], r[1], r[2]), wo = v3(r[3], r[4], r[5]);
	const V3 f = bsdf_eval(b, wi, wo);
	o[0] = f.x; o[1] = f.y; o[2] = f.z; o[3] = bsdf_pdf(b, wi, wo);
	BsdfSample bs;
	const V3 w = bsdf_sample(b, wi, r[6], r[7], s3 ? s3[i] : 0.5f, bs);
	o[4] = w.x; o[5] = w.y; o[6] = w.z; o[7] = bs.wo.x; o[8] = bs.wo.y; o[9] = bs.wo.z;
	o[10] = bs.pdf; o[11] = bs.eta; o[12] = (float) bs.type; o[13] = o[14] = o[15] = 0;
}

__global__ void eval_emitter_kernel(SceneDev S, const float *in, size_t n, float *out) {
	const size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x;
	if (i >= n) return;
	const float *r = in + i * 8; float *o = out + i * 12;
	const V3 p = v3(r[0], r[1], r[2]), refN = v3(r[3], r[4], r[5]);
	DirectSample ds;
	const V3 value = scene_sample_emitter_direct(S, p, refN, r[6], r[7], ds);
	o[0] = value.x; o[1] = value.y; o[2] = value.z; o[3] = ds.pdf;
	o[4] = ds.d.x; o[5] = ds.d.y; o[6] = ds.d.z; o[7] = ds.dist; o[8] = (float) ds.measure;
	o[9] = (ds.pdf > 0 && !is_zero(value)) ? emitter_pdf_direct(S, ds.emitter, refN, ds.d, ds.dist, ds.n, ds.measure, p) * emitter_pdf_discrete(S, ds.emitter) : 0.0f;
	o[10] = (float) ds.emitter; o[11] = 0;
}

} // namespace mtsb
