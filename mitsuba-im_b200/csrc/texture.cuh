// texture.cuh — `bitmap` textures on BSDF parameters (src/textures/bitmap.cpp; the MIP lookups live in envmap.cuh).
//   Intersection::getBSDF(ray) -> computePartials          include/mitsuba/render/records.inl:69-75, src/librender/intersection.cpp:5-76
//   dpdu / dpdv of the hit                                   TriMesh::computeUVTangents (src/librender/trimesh.cpp:683-733),
//                                                            skdtree.h:376-383 (no texture coordinates: the triangle's edges), rectangle.cpp:156-157
// The only textured slot of the subset is the reflectance of `diffuse`: the looked-up colour replaces p[0..2] of the thread's
// copy of the BSDF record before eval / pdf / sample run, which is where SmoothDiffuse::eval / sample call m_reflectance->eval(its).
#pragma once

namespace mtsb {

// the diffuse reflectance of `diffuse` (p[0..2]) or of `plastic` (p[1..3]) comes from a texture: p[12] = 1, p[13..14] = its record
__device__ __forceinline__ bool bsdf_textured(const BsdfRec &b) { const int t = b.type & 0xff; return (t == MTSB_BSDF_DIFFUSE || t == MTSB_BSDF_PLASTIC) && b.p[12] != 0.0f; }

// its.dpdu / its.dpdv for a hit on a triangle mesh or a rectangle
__device__ inline void surface_tangents(const SceneDev &S, uint32_t slot, const ShapeRec &sh, uint32_t prim, V3 &dpdu, V3 &dpdv) {
	if (sh.type == SHAPE_RECT) { const RectRec &r = S.rects[sh.rect]; dpdu = r.dpdu; dpdv = r.dpdv; return; }
	const float4 *vr = S.verts + (size_t) slot * 5;
	const float4 q0 = __ldg(vr), q1 = __ldg(vr + 1), q2 = __ldg(vr + 2);
	const V3 p0 = v3(q0.x, q0.y, q0.z), p1 = v3(q0.w, q1.x, q1.y), p2 = v3(q1.z, q1.w, q2.x);
	const V3 dP1 = p1 - p0, dP2 = p2 - p0;
	dpdu = dP1; dpdv = dP2;
	if (!(sh.flags & SHAPE_HAS_UV)) return;
	const uint32_t *ix = S.indices + (size_t) (sh.triBase + prim) * 3;
	const uint32_t i0 = sh.vtxBase + ix[0], i1 = sh.vtxBase + ix[1], i2 = sh.vtxBase + ix[2];
	const float *T = S.texcoords;
	const float du1 = T[i1*2] - T[i0*2], dv1 = T[i1*2+1] - T[i0*2+1], du2 = T[i2*2] - T[i0*2], dv2 = T[i2*2+1] - T[i0*2+1];
	const V3 n = cross(dP1, dP2);
	const float len = length(n);
	if (len == 0) { dpdu = v3(0.0f); dpdv = v3(0.0f); return; }          // the memset of computeUVTangents
	const float det = du1 * dv2 - dv1 * du2;
	if (det == 0) coordinate_system(n / len, dpdu, dpdv);
	else {
		const float invDet = 1.0f / det;
		dpdu = (dv2 * dP1 - dv1 * dP2) * invDet;
		dpdv = (-du2 * dP1 + du1 * dP2) * invDet;
	}
}

__device__ inline bool solve_linear_2x2(const float a[2][2], const float b[2], float x[2]) {     // src/libcore/util.cpp: solveLinearSystem2x2
	const float det = a[0][0] * a[1][1] - a[0][1] * a[1][0];
	if (fabsf(det) <= 2.93873587705571876e-39f) return false;    // RCPOVERFLOW
	const float inverse = 1.0f / det;
	x[0] = (a[1][1] * b[0] - a[0][1] * b[1]) * inverse;
	x[1] = (a[0][0] * b[1] - a[1][0] * b[0]) * inverse;
	return true;
}

// Intersection::computePartials for a ray whose differentials share its origin (perspective / thinlens sensor rays)
__device__ inline void compute_partials(const V3 &n, const V3 &p, const V3 &o, const V3 &rxD, const V3 &ryD, const V3 &dpdu, const V3 &dpdv,
		float &dudx, float &dvdx, float &dudy, float &dvdy) {
	dudx = dvdx = dudy = dvdy = 0.0f;
	if (is_zero(dpdu) && is_zero(dpdv)) return;
	const float pp = dot(n, p), pox = dot(n, o), prx = dot(n, rxD), pry = dot(n, ryD);
	if (prx == 0 || pry == 0) return;
	const float tx = (pp - pox) / prx, ty = (pp - pox) / pry;
	const float absX = fabsf(n.x), absY = fabsf(n.y), absZ = fabsf(n.z);
	int a0, a1;
	if (absX > absY && absX > absZ) { a0 = 1; a1 = 2; } else if (absY > absZ) { a0 = 0; a1 = 2; } else { a0 = 0; a1 = 1; }
	float A[2][2] = { { comp(dpdu, a0), comp(dpdv, a0) }, { comp(dpdu, a1), comp(dpdv, a1) } };
	const V3 px = o + rxD * tx, py = o + ryD * ty;
	const float Bx[2] = { comp(px, a0) - comp(p, a0), comp(px, a1) - comp(p, a1) }, By[2] = { comp(py, a0) - comp(p, a0), comp(py, a1) - comp(p, a1) };
	float x[2];
	if (solve_linear_2x2(A, Bx, x)) { dudx = x[0]; dvdx = x[1]; } else { dudx = 1; dvdx = 0; }
	if (solve_linear_2x2(A, By, x)) { dudy = x[0]; dvdy = x[1]; } else { dudy = 1; dvdy = 0; }      // (":74 dudy = 0; dudy = 1;" as written)
}

// Replaces the diffuse reflectance of a textured `diffuse` / `plastic` record by the texture's value at this hit.  hasDifferentials: the ray is a
// sensor ray (rx, ry = its differential directions); kept out of line, like the environment lookup.
static __device__ __noinline__ void resolve_bsdf_texture(const SceneDev *S, BsdfRec *b, const Surf *sf, uint32_t slot, bool hasDifferentials,
		float ox, float oy, float oz, float rxx, float rxy, float rxz, float ryx, float ryy, float ryz) {
	const unsigned long long addr = (unsigned long long) __float_as_uint(b->p[13]) | ((unsigned long long) __float_as_uint(b->p[14]) << 32);
	const EnvmapDev &T = *(const EnvmapDev *) addr;
	float dudx = 0, dvdx = 0, dudy = 0, dvdy = 0;
	if (hasDifferentials) {
		V3 dpdu, dpdv;
		surface_tangents(*S, slot, S->shapes[sf->shape], sf->prim, dpdu, dpdv);
		compute_partials(sf->ng, sf->p, v3(ox, oy, oz), v3(rxx, rxy, rxz), v3(ryx, ryy, ryz), dpdu, dpdv, dudx, dvdx, dudy, dvdy);
	}
	const V3 c = texture_eval(T, sf->u, sf->v, hasDifferentials, dudx, dvdx, dudy, dvdy);
	const int off = (b->type & 0xff) == MTSB_BSDF_PLASTIC ? 1 : 0;
	b->p[off] = c.x; b->p[off + 1] = c.y; b->p[off + 2] = c.z;
}

// the BSDF record of a hit, textures resolved; `first`: the path's sensor ray, whose differentials generate_kernel left in prev / L
template <bool LEAN = false>
__device__ __forceinline__ BsdfRec load_bsdf(const SceneDev &S, int bsdfId, const Surf &sf, uint32_t hitSlot, const V3 &o, bool first,
		const float4 &prev, const float4 *L) {
	BsdfRec b = S.bsdfs[bsdfId];
	if (!LEAN && bsdf_textured(b)) {
		V3 rx = v3(0.0f), ry = v3(0.0f);
		if (first) { const float4 l = *L; rx = v3(prev.x, prev.y, prev.z); ry = v3(l.x, l.y, l.z); }
		resolve_bsdf_texture(&S, &b, &sf, hitSlot, first, o.x, o.y, o.z, rx.x, rx.y, rx.z, ry.x, ry.y, ry.z);
	}
	return b;
}

} // namespace mtsb
