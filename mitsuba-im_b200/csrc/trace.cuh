// trace.cuh — device-side BVH8 traversal + primitive tests (closest hit and any hit).
//
// Replaces SAHKDTree3D::rayIntersectHavran<shadowRay> (include/mitsuba/render/sahkdtree3.h:178-308) and
// TriAccel::rayIntersect (include/mitsuba/render/triaccel.h:96-158).  The acceleration structure is a
// GPU-built 8-wide BVH with quantised child boxes (80-byte nodes read as five 16-byte vector loads);
// the primitive test is the reference's own Wald projection test on 48-byte records so that accepted
// hits, t and barycentrics agree with the CPU path to rounding.
//
// Node layout (5 x uint4):
//   n0 = { p.x, p.y, p.z (float bits), ex | ey<<8 | ez<<16 | imask<<24 }     quantisation frame + internal mask
//   n1 = { childBase, triBase, meta[0..3], meta[4..7] }
//   n2 = { qlo.x[0..3], qlo.x[4..7], qlo.y[0..3], qlo.y[4..7] }
//   n3 = { qlo.z[0..3], qlo.z[4..7], qhi.x[0..3], qhi.x[4..7] }
//   n4 = { qhi.y[0..3], qhi.y[4..7], qhi.z[0..3], qhi.z[4..7] }
// meta[i]: 0 = empty; internal child = (1<<5) | (24 + slot); leaf = (unary triangle count << 5) | triangle
// offset (0..23).  Children sit in octant-ordered slots (slot bit set = child lies towards +axis), so the
// traversal order for a ray is "slot XOR (7 - octant)", highest first — no per-node distance sort.
#pragma once
#include "common.cuh"

namespace mtsb {

#define MTSB_STACK_SHARED 8      // stack entries per thread kept in shared memory
#define MTSB_STACK_LOCAL  56     // overflow entries in local memory
#define MTSB_TRACE_BLOCK  128    // threads per block of every traversal kernel
#define MTSB_SUSP_STACK   32     // deepest stack a suspended ray can carry (deeper ones keep running)
#define MTSB_SUSP_WORDS   (5 + MTSB_SUSP_STACK)   // uint2 words of a suspension record (Trav::save)
#ifndef MTSB_TRACE_BLOCKS_PER_SM
#define MTSB_TRACE_BLOCKS_PER_SM 8   // resident CTAs per SM of the persistent traversal grids (bounds the register budget)
#endif

struct HitRec { float t, u, v; uint32_t slot; };

__device__ __forceinline__ float byte_to_float(uint32_t word, int j) {
	// PRMT picks byte j into the low byte of 0x4B000000 -> 2^23 + byte, exact subtraction afterwards
	return __uint_as_float(__byte_perm(word, 0x4B000000u, 0x7440u | (uint32_t) j)) - 8388608.0f;
}

// Reference Wald test, evaluated without FMA contraction in source order (triaccel.h:96-158).
__device__ __forceinline__ bool tri_test(const float4 a, const float4 b, const float4 c,
		const V3 &o, const V3 &d, float mint, float maxt, float &t, float &u, float &v) {
	const uint32_t k = __float_as_uint(a.x);
	if (k > 2u) return false;                       // degenerate triangle (k = 3, triaccel.h:77-80)
	// axis permutation as selects, not as three divergent branches (each ran with ~3 lanes, profiles/r01b)
	const bool k0 = k == 0, k1 = k == 1;
	const float o_u = k0 ? o.y : (k1 ? o.z : o.x), o_v = k0 ? o.z : (k1 ? o.x : o.y), o_k = k0 ? o.x : (k1 ? o.y : o.z);
	const float d_u = k0 ? d.y : (k1 ? d.z : d.x), d_v = k0 ? d.z : (k1 ? d.x : d.y), d_k = k0 ? d.x : (k1 ? d.y : d.z);
	const float n_u = a.y, n_v = a.z, n_d = a.w;
	float num = __fsub_rn(__fsub_rn(__fsub_rn(n_d, __fmul_rn(o_u, n_u)), __fmul_rn(o_v, n_v)), o_k);
	float den = __fadd_rn(__fadd_rn(__fmul_rn(d_u, n_u), __fmul_rn(d_v, n_v)), d_k);
	t = __fdiv_rn(num, den);
	if (t < mint || t > maxt) return false;   // NaN passes, exactly as on the CPU; u/v test below rejects it
	const float hu = __fsub_rn(__fadd_rn(o_u, __fmul_rn(t, d_u)), b.x);
	const float hv = __fsub_rn(__fadd_rn(o_v, __fmul_rn(t, d_v)), b.y);
	u = __fadd_rn(__fmul_rn(hv, b.z), __fmul_rn(hu, b.w));
	v = __fadd_rn(__fmul_rn(hu, c.x), __fmul_rn(hv, c.y));
	return u >= 0 && v >= 0 && __fadd_rn(u, v) <= 1.0f;
}

// Rectangle::rayIntersect (src/shapes/rectangle.cpp:125-147): object-space plane z=0, |x|,|y| <= 1.
__device__ __forceinline__ bool rect_test(const RectRec &r, const V3 &o, const V3 &d,
		float mint, float maxt, float &t, float &u, float &v) {
	const float *m = r.toObject;
	float ox = m[0]*o.x + m[1]*o.y + m[2]*o.z + m[3];
	float oy = m[4]*o.x + m[5]*o.y + m[6]*o.z + m[7];
	float oz = m[8]*o.x + m[9]*o.y + m[10]*o.z + m[11];
	float dx = m[0]*d.x + m[1]*d.y + m[2]*d.z;
	float dy = m[4]*d.x + m[5]*d.y + m[6]*d.z;
	float dz = m[8]*d.x + m[9]*d.y + m[10]*d.z;
	float hit = -oz / dz;
	if (!(hit >= mint && hit <= maxt)) return false;
	float lx = ox + dx * hit, ly = oy + dy * hit;
	if (fabsf(lx) <= 1 && fabsf(ly) <= 1) { t = hit; u = lx; v = ly; return true; }
	return false;
}

// Disk::rayIntersect (src/shapes/disk.cpp:139-165): object-space plane z = 0, x^2 + y^2 <= 1; u, v = the local x, y.
__device__ __forceinline__ bool disk_test(const RectRec &r, const V3 &o, const V3 &d,
		float mint, float maxt, float &t, float &u, float &v) {
	const float *m = r.toObject;
	float ox = m[0]*o.x + m[1]*o.y + m[2]*o.z + m[3];
	float oy = m[4]*o.x + m[5]*o.y + m[6]*o.z + m[7];
	float oz = m[8]*o.x + m[9]*o.y + m[10]*o.z + m[11];
	float dx = m[0]*d.x + m[1]*d.y + m[2]*d.z;
	float dy = m[4]*d.x + m[5]*d.y + m[6]*d.z;
	float dz = m[8]*d.x + m[9]*d.y + m[10]*d.z;
	float hit = -oz / dz;
	if (!(hit >= mint && hit <= maxt)) return false;
	float lx = ox + dx * hit, ly = oy + dy * hit;
	if (lx * lx + ly * ly <= 1) { t = hit; u = lx; v = ly; return true; }
	return false;
}

// Sphere::rayIntersect (src/shapes/sphere.cpp:158-183) with solveQuadraticDouble (src/libcore/util.cpp:489-527): the
// reference solves the quadratic in double precision, so does this (B200 has full-rate-enough FP64 for one leaf test).
__device__ __forceinline__ bool sphere_test(const RectRec &r, const V3 &o, const V3 &d,
		float mint, float maxt, float &t, float &u, float &v) {
	const double ox = (double) o.x - (double) r.dpdu.x, oy = (double) o.y - (double) r.dpdu.y, oz = (double) o.z - (double) r.dpdu.z;
	const double dx = d.x, dy = d.y, dz = d.z, rad = r.dpdv.x;
	const double A = dx*dx + dy*dy + dz*dz, B = 2 * (ox*dx + oy*dy + oz*dz), C = ox*ox + oy*oy + oz*oz - rad * rad;
	double nearT, farT;
	if (A == 0) {
		if (B == 0) return false;
		nearT = farT = -C / B;
	} else {
		const double discrim = B*B - 4.0 * A * C;
		if (discrim < 0) return false;
		const double sq = sqrt(discrim);
		const double temp = B < 0 ? -0.5 * (B - sq) : -0.5 * (B + sq);
		nearT = temp / A; farT = C / temp;
		if (nearT > farT) { const double tmp = nearT; nearT = farT; farT = tmp; }
	}
	if (!(nearT <= (double) maxt && farT >= (double) mint)) return false;
	if (nearT < (double) mint) {
		if (farT > (double) maxt) return false;
		t = (float) farT;
	} else t = (float) nearT;
	u = 0; v = 0;
	return true;
}

// disk and sphere behind a real call: their registers (FP64 quadratic) must not weigh on the traversal loop
static __device__ __noinline__ bool disk_or_sphere_test(const RectRec *r, uint32_t type, float ox, float oy, float oz, float dx, float dy, float dz,
		float mint, float maxt, float *tuv) {
	const V3 o = v3(ox, oy, oz), d = v3(dx, dy, dz);
	float t, u, v;
	const bool h = type == SHAPE_DISK ? disk_test(*r, o, d, mint, maxt, t, u, v) : sphere_test(*r, o, d, mint, maxt, t, u, v);
	if (h) { tuv[0] = t; tuv[1] = u; tuv[2] = v; }
	return h;
}

// leaf test of a non-triangle slot: the shape's own rayIntersect
__device__ __forceinline__ bool analytic_test(const SceneDev &S, uint32_t shapeIdx, const V3 &o, const V3 &d,
		float mint, float maxt, float &t, float &u, float &v) {
	const ShapeRec &sh = S.shapes[shapeIdx];
	const RectRec &r = S.rects[sh.rect];
	if (sh.type == SHAPE_RECT) return rect_test(r, o, d, mint, maxt, t, u, v);
	float tuv[3];
	const bool h = disk_or_sphere_test(&r, sh.type, o.x, o.y, o.z, d.x, d.y, d.z, mint, maxt, tuv);
	if (h) { t = tuv[0]; u = tuv[1]; v = tuv[2]; }
	return h;
}

// Byte j of `word` placed in bits 8..15 of the mantissa of 1.0f: 1 + q * 2^-15, exact, ONE PRMT and no arithmetic.
// The slab distance q * step * idir + (p - o) * idir is then a single FMA with the 2^15 folded into the scale and
// the "1 +" folded into the offset (node_test).  `one` = 0x3F800000 comes from the kernel parameter bank: PRMT has a single immediate slot, and with both the
// selector and the pattern constant ptxas re-materialises one of them before every PRMT (48 extra moves per node).
__device__ __forceinline__ float q_to_unit(uint32_t word, uint32_t one, int j) {
	return __uint_as_float(__byte_perm(word, one, 0x7604u | ((uint32_t) j << 4)));
}

// Tests the 8 quantised child boxes of a node; returns the hit mask (bits 24..31 internal children in
// traversal priority order, bits 0..23 triangles).  Conservative by construction: every near distance is lowered and
// every far distance raised by a bound on the rounding error of the offset (a few ulps of |scale| + |offset|).
__device__ __forceinline__ uint32_t node_test(const uint4 n0, const uint4 n1, const uint4 n2, const uint4 n3,
		const uint4 n4, const V3 &o, const V3 &idir, uint32_t octinv4, float tmin, float tmax, uint32_t one) {
	const uint32_t e = n0.w;
	// 2^(e + 15) * idir: the scale per unit of q_to_unit's mantissa step
	const float sx = __uint_as_float(((e & 0xffu) << 23) + (15u << 23)) * idir.x;
	const float sy = __uint_as_float((((e >> 8) & 0xffu) << 23) + (15u << 23)) * idir.y;
	const float sz = __uint_as_float((((e >> 16) & 0xffu) << 23) + (15u << 23)) * idir.z;
	const float px = (__uint_as_float(n0.x) - o.x) * idir.x, py = (__uint_as_float(n0.y) - o.y) * idir.y,
	            pz = (__uint_as_float(n0.z) - o.z) * idir.z;
	const float mx = (fabsf(sx) + fabsf(px)) * 4.8e-7f, my = (fabsf(sy) + fabsf(py)) * 4.8e-7f, mz = (fabsf(sz) + fabsf(pz)) * 4.8e-7f;
	const float bnx = (px - sx) - mx, bny = (py - sy) - my, bnz = (pz - sz) - mz;
	const float bfx = (px - sx) + mx, bfy = (py - sy) + my, bfz = (pz - sz) + mz;
	uint32_t hitmask = 0;
	#pragma unroll
	for (int g = 0; g < 2; ++g) {
		const uint32_t meta4 = g ? n1.w : n1.z;
		const uint32_t isInner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
		const uint32_t innerMask4 = (isInner4 >> 4) * 0xffu;
		// per child byte: [unary triangle count : 3 | bit index : 5]; SHF.L.W uses the low 5 bits of the shift amount
		const uint32_t m4 = meta4 ^ (octinv4 & innerMask4 & 0x07070707u);
		const uint32_t qlox = g ? n2.y : n2.x, qloy = g ? n2.w : n2.z, qloz = g ? n3.y : n3.x;
		const uint32_t qhix = g ? n3.w : n3.z, qhiy = g ? n4.y : n4.x, qhiz = g ? n4.w : n4.z;
		const uint32_t xn = idir.x < 0 ? qhix : qlox, xf = idir.x < 0 ? qlox : qhix;
		const uint32_t yn = idir.y < 0 ? qhiy : qloy, yf = idir.y < 0 ? qloy : qhiy;
		const uint32_t zn = idir.z < 0 ? qhiz : qloz, zf = idir.z < 0 ? qloz : qhiz;
		#pragma unroll
		for (int j = 0; j < 4; ++j) {
			const float tnx = fmaf(q_to_unit(xn, one, j), sx, bnx), tfx = fmaf(q_to_unit(xf, one, j), sx, bfx);
			const float tny = fmaf(q_to_unit(yn, one, j), sy, bny), tfy = fmaf(q_to_unit(yf, one, j), sy, bfy);
			const float tnz = fmaf(q_to_unit(zn, one, j), sz, bnz), tfz = fmaf(q_to_unit(zf, one, j), sz, bfz);
			const float cmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, tmin));
			const float cmax = fminf(fminf(tfx, tfy), fminf(tfz, tmax));
			if (cmin <= cmax) {
				const uint32_t m = (m4 >> (8 * j)) & 0xffu;
				hitmask |= (m >> 5) << (m & 31u);
			}
		}
	}
	return hitmask;
}

// Per-thread traversal stack: first MTSB_STACK_SHARED entries in shared memory (interleaved by thread),
// the rest in local memory.
struct TravStack {
	uint2 *sh;            // &shared[threadIdx.x], stride blockDim.x
	uint2 *loc;           // MTSB_STACK_LOCAL entries of local memory, declared by the kernel (kept out of this struct so
	                      // that the scalar members stay in registers)
	int n;
	__device__ __forceinline__ void push(uint2 v) {
		if (n < MTSB_STACK_SHARED) sh[n * MTSB_TRACE_BLOCK] = v;
		else if (n < MTSB_STACK_SHARED + MTSB_STACK_LOCAL) loc[n - MTSB_STACK_SHARED] = v;
		else return;  // cannot happen for trees built by bvh.cu (depth bounded well below)
		++n;
	}
	__device__ __forceinline__ uint2 pop() {
		--n;
		return n < MTSB_STACK_SHARED ? sh[n * MTSB_TRACE_BLOCK] : loc[n - MTSB_STACK_SHARED];
	}
};

// The scene-AABB clip and adaptive epsilon that precede traversal in the three ShapeKDTree::rayIntersect
// overloads (src/librender/skdtree.cpp:112-142, 207-226; AABB::rayIntersect include/mitsuba/core/aabb.h:308-338).
// floorEps: the closest-hit overload floors the epsilon scale at Epsilon, the other two do not.
__device__ __forceinline__ bool clip_ray(const SceneDev &S, const V3 &o, const V3 &d, float rayMint, float rayMaxt,
		bool floorEps, float &mint, float &maxt, V3 &rcpOut) {
	float nearT = -INFINITY, farT = INFINITY;
	float rcps[3] = { 0.0f, 0.0f, 0.0f };
	#pragma unroll
	for (int i = 0; i < 3; ++i) {
		const float origin = comp(o, i), dir = comp(d, i);
		const float minVal = comp(S.aabbMin, i), maxVal = comp(S.aabbMax, i);
		if (dir == 0) {
			if (origin < minVal || origin > maxVal) return false;
		} else {
			const float rcp = __frcp_rn(dir);          // correctly rounded, bit-identical to 1.0f / dir without the division's slow path
			rcps[i] = rcp;
			float t1 = __fmul_rn(__fsub_rn(minVal, origin), rcp);
			float t2 = __fmul_rn(__fsub_rn(maxVal, origin), rcp);
			if (t1 > t2) { float tmp = t1; t1 = t2; t2 = tmp; }
			nearT = fmaxf(t1, nearT);
			farT = fminf(t2, farT);
			if (!(nearT <= farT)) return false;
		}
	}
	float rayMinT = rayMint;
	if (rayMinT == MTS_EPSILON) {
		float m = fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z));
		if (floorEps) m = fmaxf(m, MTS_EPSILON);
		rayMinT = __fmul_rn(rayMinT, m);
	}
	rcpOut = v3(rcps[0], rcps[1], rcps[2]);
	mint = nearT; maxt = farT;
	if (rayMinT > mint) mint = rayMinT;
	if (rayMaxt < maxt) maxt = rayMaxt;
	return maxt > mint;
}

// Per-lane traversal state + one warp-convergent traversal iteration, shared by the persistent batch kernels
// (trace_persistent) and by the late-bounce kernel of the path tracer (render_tail.cuh: trace_warp).
struct Trav {
	V3 o, d, idir;
	float tmin;
	HitRec hit;
	uint2 ng, tg;                 // pending node group {child base, hit mask | imask} / triangle group {slot base, mask}
	uint32_t octinv, octinv4;
	bool active, found;
	TravStack st;

	__device__ __forceinline__ void init(uint2 *shStack, uint2 *locStack) {
		active = false; found = false;
		o = v3(0.0f); d = v3(0.0f); idir = v3(0.0f); tmin = 0;
		hit.t = 0; hit.u = hit.v = 0; hit.slot = 0xffffffffu;
		ng = make_uint2(0u, 0u); tg = make_uint2(0u, 0u); octinv = 0; octinv4 = 0;
		st.sh = shStack; st.loc = locStack; st.n = 0;
	}
	// o, d, tmin already set; rcp = the exact reciprocals clip_ray computed; axis-parallel components are clamped
	__device__ __forceinline__ void begin(float tmax, const V3 &rcp) {
		idir = rcp;
		const float tiny = 1e-18f, big = 1e18f;
		if (!(fabsf(d.x) > tiny)) idir.x = copysignf(big, d.x);
		if (!(fabsf(d.y) > tiny)) idir.y = copysignf(big, d.y);
		if (!(fabsf(d.z) > tiny)) idir.z = copysignf(big, d.z);
		octinv = ((idir.x < 0 ? 1u : 0u) | (idir.y < 0 ? 2u : 0u) | (idir.z < 0 ? 4u : 0u)) ^ 7u;
		octinv4 = octinv * 0x01010101u;
		ng = make_uint2(0u, 0x80000000u); tg = make_uint2(0u, 0u);
		hit.t = tmax; hit.u = hit.v = 0; hit.slot = 0xffffffffu;
		found = false; st.n = 0; active = true;
	}

	// Suspension (extend stage of the wavefront, render.cu): a ray that is still traversing when its launch has run out of
	// other work parks its whole state — pending groups, best hit, stack — in MTSB_SUSP_WORDS 8-byte words and is picked up
	// again by the next launch, so that a launch lasts as long as its typical rays, not as its longest one.
	__device__ __forceinline__ void save(uint2 *rec) const {
		rec[0] = ng; rec[1] = tg;
		rec[2] = make_uint2(__float_as_uint(hit.t), __float_as_uint(hit.u));
		rec[3] = make_uint2(__float_as_uint(hit.v), hit.slot);
		rec[4] = make_uint2((uint32_t) st.n | (found ? 0x80000000u : 0u), __float_as_uint(tmin));
		for (int k = 0; k < st.n; ++k) rec[5 + k] = k < MTSB_STACK_SHARED ? st.sh[k * MTSB_TRACE_BLOCK] : st.loc[k - MTSB_STACK_SHARED];
	}
	// o and d already set (from the path's ray); everything else comes back from the record
	__device__ __forceinline__ void restore(const uint2 *rec) {
		const V3 rcp = v3(d.x != 0 ? __frcp_rn(d.x) : 0.0f, d.y != 0 ? __frcp_rn(d.y) : 0.0f, d.z != 0 ? __frcp_rn(d.z) : 0.0f);   // as clip_ray
		begin(0.0f, rcp);
		ng = rec[0]; tg = rec[1];
		const uint2 h0 = rec[2], h1 = rec[3], m = rec[4];
		hit.t = __uint_as_float(h0.x); hit.u = __uint_as_float(h0.y); hit.v = __uint_as_float(h1.x); hit.slot = h1.y;
		found = (m.x & 0x80000000u) != 0; tmin = __uint_as_float(m.y);
		const int n = (int) (m.x & 0xffffu);
		st.n = 0;
		for (int k = 0; k < n; ++k) st.push(rec[5 + k]);
	}

	// One iteration: node phase, triangle phase, pop-or-finish.  Must be called by all 32 lanes of the warp (it ballots).
	// Returns true for the lanes whose ray finished in this iteration (`active` is already false for them).
	// CURVED: the scene holds analytic disks / spheres.  Their leaf test sits behind a real call whose register traffic cost
	// the loop 3 % (config 2: 2946 -> 2863 Mrays/s) even when it never ran, so scenes without them get kernels without it.
	template <bool ANY, bool COUNT, bool CURVED>
	__device__ __forceinline__ bool step(const SceneDev &S, uint32_t one, unsigned &nNodes, unsigned &nTris) {
		const unsigned FULL = 0xffffffffu;
		// ---- node phase (lanes still holding untested triangles sit this one out: they drain one per iteration)
		if (active && tg.y == 0) {
			if (ng.y > 0x00ffffffu) {
				const uint32_t bit = 31u - __clz(ng.y);
				const uint32_t imask = ng.y & 0xffu;
				ng.y &= ~(1u << bit);
				if (ng.y > 0x00ffffffu) st.push(ng);
				const uint32_t slot = (bit - 24u) ^ octinv;
				const uint32_t rel = __popc(imask & ~(0xffffffffu << slot));
				const uint4 *np = S.nodes + (size_t) (ng.x + rel) * 5;
				const uint4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
				if (COUNT) ++nNodes;
				const uint32_t hm = node_test(n0, n1, n2, n3, n4, o, idir, octinv4, tmin, hit.t, one);
				ng = make_uint2(n1.x, (hm & 0xff000000u) | (n0.w >> 24));
				tg = make_uint2(n1.y, hm & 0x00ffffffu);
			} else {
				tg = ng; ng = make_uint2(0u, 0u);
			}
		}

		// ---- triangle phase: ONE triangle per lane per round.  The first round runs whenever any lane has a triangle;
		// further rounds only while at least triMinLanes lanes still have one — draining a long leaf with two or three
		// lanes active (the old while-loop: 2.5 of 32 threads, profiles/r01) costs more issue slots than letting those
		// lanes skip a node step.  Round count is warp-uniform, so the loop stays convergent.
		for (bool first = true; ; first = false) {
			const bool want = active && tg.y;
			const unsigned wm = __ballot_sync(FULL, want);
			if (wm == 0 || (!first && __popc(wm) < S.triMinLanes)) break;
			if (want) {
				const uint32_t bit = __ffs(tg.y) - 1;
				tg.y &= tg.y - 1;
				const uint32_t slot = tg.x + bit;
				const float4 *tp = S.tris + (size_t) slot * 3;
				const float4 a = __ldg(tp), b = __ldg(tp + 1), c = __ldg(tp + 2);
				if (COUNT) ++nTris;
				float t, u, v; bool h;
				if (__float_as_uint(a.x) != K_NO_TRIANGLE)
					h = tri_test(a, b, c, o, d, tmin, hit.t, t, u, v);
				else
					h = CURVED ? analytic_test(S, __float_as_uint(c.z), o, d, tmin, hit.t, t, u, v)
					           : rect_test(S.rects[S.shapes[__float_as_uint(c.z)].rect], o, d, tmin, hit.t, t, u, v);
				if (h) {
					hit.t = t; hit.u = u; hit.v = v; hit.slot = slot; found = true;
					if (ANY) { tg.y = 0; ng = make_uint2(0u, 0u); st.n = 0; }
				}
			}
		}

		// ---- pop or finish
		if (active && tg.y == 0 && ng.y <= 0x00ffffffu) {
			if (st.n == 0) { active = false; return true; }
			ng = st.pop();
		}
		return false;
	}
};

// Persistent, warp-synchronous traversal of a whole ray batch.
//
// Every warp keeps 32 rays in flight.  Each loop iteration is warp-convergent (it starts with a full-mask
// ballot): (1) lanes whose ray has finished have written their record; once fetchMinIdle lanes are idle the
// warp pulls that many new rays from the global work counter with ONE atomic (lane 0) broadcast by shuffle, each
// idle lane taking the slot given by its rank in the ballot; (2) every lane with a pending node group pops its
// nearest child and tests the 8 quantised boxes; (3) lanes with leaf triangles test them; (4) lanes whose group is
// exhausted pop the stack or finish.  ANY=false: closest hit (shrinks hit.t); ANY=true: first accepted hit ends the ray.
//
// IO supplies the rays and consumes the records:
//   bool IO::load(i, o, d, tmin, tmax, rcp) -> rcp = 1/d from clip_ray; false if the ray needs no traversal (IO has already written a miss)
//   void IO::store(i, found, hit, o, d)
template <bool ANY, bool COUNT, bool CURVED, class IO>
__device__ __forceinline__ void trace_persistent(const SceneDev &S, IO &io, unsigned long long n,
		unsigned long long *__restrict__ workCounter, uint2 *shStack, unsigned suspendAfter = 0) {
	const unsigned lane = threadIdx.x & 31u, ltMask = (1u << lane) - 1u;
	const unsigned FULL = 0xffffffffu;
	bool exhausted = false;
	uint2 locStack[MTSB_STACK_LOCAL];
	Trav tv; tv.init(shStack, locStack);
	unsigned long long idx = 0;
	unsigned nNodes = 0, nTris = 0, sinceExhausted = 0;
	const uint32_t one = S.oneBits;                 // 1.0f bit pattern from the parameter bank (see q_to_unit)

	for (;;) {
		// ---- (1) refill idle lanes
		const unsigned idleMask = __ballot_sync(FULL, !tv.active);
		const unsigned nIdle = __popc(idleMask);
		if (!exhausted && nIdle >= (unsigned) S.fetchMinIdle) {
			unsigned long long base = 0;
			if (lane == 0) base = atomicAdd(workCounter, (unsigned long long) nIdle);
			base = __shfl_sync(FULL, base, 0);
			if (!tv.active) {
				idx = base + __popc(idleMask & ltMask);
				if (idx < n) {
					float tmax; V3 rcp;
					const int how = io.load(idx, tv.o, tv.d, tv.tmin, tmax, rcp);     // 0: no traversal, 1: fresh ray, 2: suspended ray
					if (how == 1) tv.begin(tmax, rcp);
					else if constexpr (IO::CAN_SUSPEND) { if (how == 2) io.resume(idx, tv); }
				}
			}
			if (base + nIdle >= n) exhausted = true;
		}
		if (!__any_sync(FULL, tv.active)) { if (exhausted) break; else continue; }
		// ---- suspension: the queue is empty and this warp has spent its allowance on what is left -> park it
		if constexpr (IO::CAN_SUSPEND) {
			if (suspendAfter && exhausted && ++sinceExhausted > suspendAfter) {
				if (tv.active && tv.st.n <= MTSB_SUSP_STACK) { io.suspend(idx, tv); tv.active = false; }
				// (lanes with a deeper stack run on until it fits, or they finish)
			}
		}
		// ---- (2)-(4)
		if (tv.template step<ANY, COUNT, CURVED>(S, one, nNodes, nTris)) io.store(idx, tv.found, tv.hit, tv.o, tv.d);
	}
	if (COUNT && S.counters) {
		atomicAdd(S.counters, (unsigned long long) nNodes);
		atomicAdd(S.counters + 1, (unsigned long long) nTris);
	}
}

// Traces the ONE ray each lane holds in `tv` (tv.begin called by the lanes that have one) until every lane of the warp
// is done; no refill.  Used where a warp owns 32 paths and alternates traversal with shading (render_tail.cuh).
template <bool ANY, bool CURVED>
__device__ __forceinline__ void trace_warp(const SceneDev &S, Trav &tv) {
	const uint32_t one = S.oneBits;
	unsigned nNodes = 0, nTris = 0;
	while (__any_sync(0xffffffffu, tv.active)) tv.template step<ANY, false, CURVED>(S, one, nNodes, nTris);
}

} // namespace mtsb
