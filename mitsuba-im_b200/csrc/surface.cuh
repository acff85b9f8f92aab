// surface.cuh — hit record -> shading geometry.
// Follows ShapeKDTree::fillIntersectionRecord<true> (include/mitsuba/render/skdtree.h:343-428),
// Rectangle::fillIntersectionRecord (src/shapes/rectangle.cpp:154-168), TriMesh::computeUVTangents
// (src/librender/trimesh.cpp:683-733), computeShadingFrame / coordinateSystem (src/libcore/util.cpp:594-610).
#pragma once
#include "trace.cuh"

namespace mtsb {

struct Surf {
	V3 p, ng, ns, s, t;
	float u, v;          // its.uv
	uint32_t shape, prim;
};

__host__ __device__ inline void coordinate_system(const V3 &a, V3 &b, V3 &c) {
	if (fabsf(a.x) > fabsf(a.y)) {
		float invLen = 1.0f / sqrtf(a.x * a.x + a.z * a.z);
		c = v3(a.z * invLen, 0.0f, -a.x * invLen);
	} else {
		float invLen = 1.0f / sqrtf(a.y * a.y + a.z * a.z);
		c = v3(0.0f, a.z * invLen, -a.y * invLen);
	}
	b = cross(c, a);
}

__host__ __device__ inline V3 to_local(const Surf &s, const V3 &v) { return v3(dot(v, s.s), dot(v, s.t), dot(v, s.ns)); }
__host__ __device__ inline V3 to_world(const Surf &s, const V3 &v) { return s.s * v.x + s.t * v.y + s.ns * v.z; }

// `slot` indexes the TriAccel array; o,d are the ray (needed for the rectangle's p = ray(t)).
__device__ inline void fill_surface(const SceneDev &S, uint32_t slot, float t, float bu, float bv,
		const V3 &o, const V3 &d, Surf &out) {
	// one 80-byte record per slot (SceneDev::verts) instead of TriAccel tail -> shape -> indices -> 6 vertex gathers
	const float4 *vr = S.verts + (size_t) slot * 5;
	const float4 q0 = __ldg(vr), q1 = __ldg(vr + 1), q2 = __ldg(vr + 2), q3 = __ldg(vr + 3), q4 = __ldg(vr + 4);
	const uint32_t shapeIdx = __float_as_uint(q4.z), prim = __float_as_uint(q4.w);
	const ShapeRec sh = S.shapes[shapeIdx];
	out.shape = shapeIdx; out.prim = prim;
	V3 dpdu;
	if (sh.type == SHAPE_MESH) {
		const V3 p0 = v3(q0.x, q0.y, q0.z), p1 = v3(q0.w, q1.x, q1.y), p2 = v3(q1.z, q1.w, q2.x);
		const float b0 = 1 - bu - bv, b1 = bu, b2 = bv;
		out.p = p0 * b0 + p1 * b1 + p2 * b2;
		const V3 side1 = p1 - p0, side2 = p2 - p0;
		V3 faceNormal = cross(side1, side2);
		const float len = length(faceNormal);
		if (!is_zero(faceNormal)) faceNormal /= len;
		dpdu = side1;
		if (sh.flags & SHAPE_HAS_UV) {
			const uint32_t *ix = S.indices + (size_t) (sh.triBase + prim) * 3;
			const uint32_t i0 = sh.vtxBase + ix[0], i1 = sh.vtxBase + ix[1], i2 = sh.vtxBase + ix[2];
			const float *T = S.texcoords;
			const float u0 = T[i0*2], v0 = T[i0*2+1], u1 = T[i1*2], v1 = T[i1*2+1], u2 = T[i2*2], v2 = T[i2*2+1];
			out.u = u0 * b0 + u1 * b1 + u2 * b2; out.v = v0 * b0 + v1 * b1 + v2 * b2;
			// per-triangle UV tangent (TangentSpace::dpdu)
			const float du1 = u1 - u0, dv1 = v1 - v0, du2 = u2 - u0, dv2 = v2 - v0;
			const float det = du1 * dv2 - dv1 * du2;
			if (len == 0) dpdu = v3(0.0f);
			else if (det == 0) { V3 dpdv; coordinate_system(faceNormal, dpdu, dpdv); }
			else { const float invDet = 1.0f / det; dpdu = (dv2 * side1 - dv1 * side2) * invDet; }
		} else {
			out.u = b1; out.v = b2;
		}
		if (sh.flags & SHAPE_HAS_NORMALS) {
			const V3 n0 = v3(q2.y, q2.z, q2.w), n1 = v3(q3.x, q3.y, q3.z), n2 = v3(q3.w, q4.x, q4.y);
			out.ns = normalize(n0 * b0 + n1 * b1 + n2 * b2);
			if (dot(faceNormal, out.ns) < 0) faceNormal = -faceNormal;
		} else {
			out.ns = faceNormal;
		}
		out.ng = faceNormal;
	} else if (sh.type == SHAPE_SPHERE) {                   // Sphere::fillIntersectionRecord, src/shapes/sphere.cpp:205-252
		const RectRec &r = S.rects[sh.rect];
		const V3 center = r.dpdu; const float radius = r.dpdv.x;
		out.p = center + normalize((o + d * t) - center) * radius;         // re-projection (SINGLE_PRECISION branch)
		const V3 rel = out.p - center;
		const float *m = r.toObject;
		const V3 local = v3(m[0]*rel.x + m[1]*rel.y + m[2]*rel.z, m[4]*rel.x + m[5]*rel.y + m[6]*rel.z, m[8]*rel.x + m[9]*rel.y + m[10]*rel.z);
		const float theta = acosf(fminf(1.0f, fmaxf(-1.0f, local.z / radius)));   // math::safe_acos
		float phi = atan2f(local.y, local.x);
		if (phi < 0) phi += 2 * 3.14159265358979323846f;
		out.u = phi * (0.5f * 0.31830988618379067154f); out.v = theta * 0.31830988618379067154f;
		const float *w = r.toWorld;
		const V3 lu = v3(-local.y, local.x, 0.0f) * (2 * 3.14159265358979323846f);
		dpdu = v3(w[0]*lu.x + w[1]*lu.y + w[2]*lu.z, w[4]*lu.x + w[5]*lu.y + w[6]*lu.z, w[8]*lu.x + w[9]*lu.y + w[10]*lu.z);
		out.ng = normalize(out.p - center);
		if (r.pad[0]) out.ng = -out.ng;                                     // m_flipNormals
		out.ns = out.ng;
	} else if (sh.type == SHAPE_DISK) {                     // Disk::fillIntersectionRecord, src/shapes/disk.cpp:168-197
		const RectRec &r = S.rects[sh.rect];
		const float rr = sqrtf(bu * bu + bv * bv), invR = rr == 0 ? 0.0f : 1.0f / rr;
		float phi = atan2f(bv, bu);
		if (phi < 0) phi += 2 * 3.14159265358979323846f;
		const float *w = r.toWorld;
		const V3 lu = rr != 0 ? v3(bu * invR, bv * invR, 0.0f) : v3(1.0f, 0.0f, 0.0f);
		dpdu = v3(w[0]*lu.x + w[1]*lu.y, w[4]*lu.x + w[5]*lu.y, w[8]*lu.x + w[9]*lu.y);
		out.ns = r.n; out.ng = r.n;           // the reference leaves geoFrame unset here; the shading normal is the sensible stand-in
		out.u = rr; out.v = phi * (0.5f * 0.31830988618379067154f);
		out.p = o + d * t;
	} else {
		const RectRec &r = S.rects[sh.rect];
		out.ng = r.n; out.ns = r.n;
		dpdu = r.dpdu;
		out.u = 0.5f * (bu + 1); out.v = 0.5f * (bv + 1);
		out.p = o + d * t;
	}
	// computeShadingFrame(its.shFrame.n, its.dpdu, its.shFrame)
	out.s = normalize(dpdu - out.ns * dot(out.ns, dpdu));
	out.t = cross(out.ns, out.s);
}

} // namespace mtsb
