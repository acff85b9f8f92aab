// emitter.cuh — direct-illumination sampling for the emitter subset.
//   Scene::sampleEmitterDirect / pdfEmitterDirect   src/librender/scene.cpp:860-884, 981-984 (without the shadow ray)
//   DiscreteDistribution::sampleReuse                 include/mitsuba/core/pmf.h:124-188
//   AreaLight::{eval,sampleDirect,pdfDirect}          src/emitters/area.cpp:106-111,160-184
//   Shape::sampleDirect / pdfDirect                   src/librender/shape.cpp:102-126
//   TriMesh::samplePosition, Triangle::sample         src/librender/trimesh.cpp:413-425, src/libcore/triangle.cpp:24-60
//   Rectangle::samplePosition                         src/shapes/rectangle.cpp:210-216
//   Disk::samplePosition                              src/shapes/disk.cpp:247-255
//   Sphere::{sampleDirect,pdfDirect}                  src/shapes/sphere.cpp:286-385 (cone sampling from outside)
//   PointEmitter::{sampleDirect,pdfDirect}            src/emitters/point.cpp:133-153
//   ConstantBackgroundEmitter::{sampleDirect,pdfDirect,fillDirectSamplingRecord}  src/emitters/constant.cpp:175-256
#pragma once
#include "envmap.cuh"

namespace mtsb {

struct DirectSample {
	V3 value;        // radiance / pdf (already divided by the emitter selection probability)
	V3 d, n;         // direction ref -> emitter, emitter normal
	V3 p;            // dRec.p, the sampled position (the far end of Scene::evalTransmittance; render_vol.cuh)
	float dist, pdf;
	int measure, emitter;
};

// DiscreteDistribution::sample on a normalised cdf with count+1 entries (pmf.h:124-136)
__device__ inline uint32_t cdf_sample(const float *cdf, uint32_t count, float value) {
	uint32_t lo = 0, hi = count + 1;            // std::lower_bound over [0, count+1)
	while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (cdf[mid] < value) lo = mid + 1; else hi = mid; }
	long idx = (long) lo - 1;
	if (idx < 0) idx = 0;
	if (idx > (long) count - 1) idx = (long) count - 1;
	uint32_t index = (uint32_t) idx;
	while (cdf[index + 1] - cdf[index] == 0 && index < count) ++index;
	return index;
}
__device__ inline uint32_t cdf_sample_reuse(const float *cdf, uint32_t count, float &value, float &pdf) {
	const uint32_t index = cdf_sample(cdf, count, value);
	pdf = cdf[index + 1] - cdf[index];
	value = (value - cdf[index]) / (cdf[index + 1] - cdf[index]);
	return index;
}

__device__ inline bool sphere_intersect(const V3 &center, float radius, const V3 &ro, const V3 &rd, float &nearT, float &farT) {
	const V3 o = ro - center;                     // BSphere::rayIntersect + solveQuadratic (util.cpp:449-487)
	const float A = dot(rd, rd), Bq = 2 * dot(o, rd), C = dot(o, o) - radius * radius;
	if (A == 0) { if (Bq != 0) { nearT = farT = -C / Bq; return true; } return false; }
	const float discrim = Bq * Bq - 4.0f * A * C;
	if (discrim < 0) return false;
	const float sq = sqrtf(discrim);
	const float temp = Bq < 0 ? -0.5f * (Bq - sq) : -0.5f * (Bq + sq);
	float x0 = temp / A, x1 = C / temp;
	if (x0 > x1) { const float t = x0; x0 = x1; x1 = t; }
	nearT = x0; farT = x1;
	return true;
}

// Shape::samplePosition for the shape an area light is attached to.
// LEAN (here and below): the scene has no environment map and no area light on an analytic sphere / disk (scene.cu), so the
// kernels of the benchmark configurations are compiled without those branches — as for the BSDFs (bsdf.cuh)
template <bool LEAN = false>
__device__ inline void shape_sample_position(const SceneDev &S, uint32_t shapeIdx, float sx, float sy, V3 &p, V3 &n, float &pdf) {
	const ShapeRec sh = S.shapes[shapeIdx];
	if (sh.type == SHAPE_RECT) {
		const RectRec &r = S.rects[sh.rect];
		const float x = sx * 2 - 1, y = sy * 2 - 1;
		const float *m = r.toWorld;
		p = v3(m[0] * x + m[1] * y + m[3], m[4] * x + m[5] * y + m[7], m[8] * x + m[9] * y + m[11]);
		n = r.n; pdf = r.invSurfaceArea;
		return;
	}
	if (!LEAN && sh.type == SHAPE_DISK) {
		const RectRec &r = S.rects[sh.rect];
		float x, y; square_to_disk_concentric(sx, sy, x, y);
		const float *m = r.toWorld;
		p = v3(m[0] * x + m[1] * y + m[3], m[4] * x + m[5] * y + m[7], m[8] * x + m[9] * y + m[11]);
		n = r.n; pdf = r.invSurfaceArea;
		return;
	}
	const float *cdf = S.triAreaCdf + S.triAreaCdfOffset[shapeIdx];
	float dummy;
	const uint32_t tri = cdf_sample_reuse(cdf, sh.primCount, sy, dummy);
	const uint32_t *ix = S.indices + (size_t) (sh.triBase + tri) * 3;
	const uint32_t i0 = sh.vtxBase + ix[0], i1 = sh.vtxBase + ix[1], i2 = sh.vtxBase + ix[2];
	const float *P = S.positions;
	const V3 p0 = v3(P[i0*3], P[i0*3+1], P[i0*3+2]), p1 = v3(P[i1*3], P[i1*3+1], P[i1*3+2]), p2 = v3(P[i2*3], P[i2*3+1], P[i2*3+2]);
	float bx, by; square_to_uniform_triangle(sx, sy, bx, by);
	const V3 sideA = p1 - p0, sideB = p2 - p0;
	p = p0 + sideA * bx + sideB * by;
	if (sh.flags & SHAPE_HAS_NORMALS) {
		const float *N = S.normals;
		const V3 n0 = v3(N[i0*3], N[i0*3+1], N[i0*3+2]), n1 = v3(N[i1*3], N[i1*3+1], N[i1*3+2]), n2 = v3(N[i2*3], N[i2*3+1], N[i2*3+2]);
		n = normalize(n0 * (1.0f - bx - by) + n1 * bx + n2 * by);
	} else {
		n = normalize(cross(sideA, sideB));
	}
	pdf = sh.invArea;
}

// solveQuadratic (src/libcore/util.cpp:449-487)
__device__ inline bool solve_quadratic(float a, float b, float c, float &x0, float &x1) {
	if (a == 0) { if (b != 0) { x0 = x1 = -c / b; return true; } return false; }
	const float discrim = b * b - 4.0f * a * c;
	if (discrim < 0) return false;
	const float sq = sqrtf(discrim);
	const float temp = b < 0 ? -0.5f * (b - sq) : -0.5f * (b + sq);
	x0 = temp / a; x1 = c / temp;
	if (x0 > x1) { const float t = x0; x0 = x1; x1 = t; }
	return true;
}

// Sphere::sampleDirect (sphere.cpp:286-355): the cone that contains the sphere when the reference point is outside,
// uniform over the sphere otherwise.  center = r.dpdu, radius = r.dpdv.x, flipNormals = r.pad[0] (scene.cu)
__device__ inline void sphere_sample_direct(const RectRec &r, const V3 &ref, float sx, float sy, DirectSample &out) {
	const V3 center = r.dpdu; const float radius = r.dpdv.x;
	const V3 refToCenter = center - ref;
	const float refDist2 = dot(refToCenter, refToCenter);
	const float invRefDist = 1.0f / sqrtf(refDist2);
	const float sinAlpha = radius * invRefDist;
	if (sinAlpha < 1 - MTS_EPSILON) {
		const float cosAlpha = safe_sqrt(1.0f - sinAlpha * sinAlpha);
		const float cosTheta = (1 - sx) + sx * cosAlpha;                      // warp::squareToUniformCone (warp.cpp:54-63)
		const float sinTheta = safe_sqrt(1.0f - cosTheta * cosTheta);
		float sinPhi, cosPhi; sincosf(2.0f * MTS_PI * sy, &sinPhi, &cosPhi);
		const V3 axis = refToCenter * invRefDist;
		V3 s, t; coordinate_system(axis, s, t);
		out.d = s * (cosPhi * sinTheta) + t * (sinPhi * sinTheta) + axis * cosTheta;
		out.pdf = (1.0f / (2.0f * MTS_PI)) / (1 - cosAlpha);
		const float projDist = dot(refToCenter, out.d);
		const float baseT = refDist2 / projDist;
		const V3 query = ref + out.d * baseT;
		const V3 queryToCenter = center - query;
		const float queryDist2 = dot(queryToCenter, queryToCenter), queryProjDist = dot(queryToCenter, out.d);
		float nearT, farT;
		if (!solve_quadratic(1.0f, -2 * queryProjDist, queryDist2 - radius * radius, nearT, farT)) nearT = queryProjDist;
		out.dist = baseT + nearT;
		out.n = normalize(out.d * nearT - queryToCenter);
		out.p = center + out.n * radius;
	} else {
		const V3 d = square_to_uniform_sphere(sx, sy);
		out.p = center + d * radius;
		out.n = d;
		V3 dd = out.p - ref;
		const float dist2 = dot(dd, dd);
		out.dist = sqrtf(dist2);
		out.d = dd / out.dist;
		out.pdf = r.invSurfaceArea * dist2 / fabsf(dot(out.d, out.n));
	}
	if (r.pad[0]) out.n = v3(-out.n.x, -out.n.y, -out.n.z);
}
// Sphere::pdfDirect (sphere.cpp:357-385), solid-angle measure
__device__ inline float sphere_pdf_direct(const RectRec &r, const V3 &ref, const V3 &d, float dist, const V3 &n) {
	const V3 refToCenter = r.dpdu - ref;
	const float invRefDist = 1.0f / length(refToCenter);
	const float sinAlpha = r.dpdv.x * invRefDist;
	if (sinAlpha < 1 - MTS_EPSILON) {
		const float cosAlpha = safe_sqrt(1 - sinAlpha * sinAlpha);
		return (1.0f / (2.0f * MTS_PI)) / (1 - cosAlpha);
	}
	return r.invSurfaceArea * dist * dist / fabsf(dot(d, n));
}

// Emitter::sampleDirect for emitter `e`
template <bool LEAN = false>
__device__ inline V3 emitter_sample_direct(const SceneDev &S, uint32_t e, const V3 &ref, const V3 &refN, float sx, float sy, DirectSample &out) {
	const EmitterRec &em = S.emitters[e];
	out.emitter = (int) e; out.pdf = 0; out.n = v3(0.0f); out.d = v3(0.0f); out.dist = 0; out.measure = M_SOLID_ANGLE; out.p = ref;
	if (em.type == MTSB_EMITTER_AREA) {
		V3 d; float pdf;
		const ShapeRec &esh = S.shapes[em.shape];
		if (!LEAN && esh.type == SHAPE_SPHERE) {              // the sphere overrides Shape::sampleDirect
			sphere_sample_direct(S.rects[esh.rect], ref, sx, sy, out);
			d = out.d; pdf = out.pdf;
		} else {
			V3 p;
			shape_sample_position<LEAN>(S, em.shape, sx, sy, p, out.n, pdf);
			out.p = p;
			d = p - ref;
			const float distSquared = dot(d, d);
			out.dist = sqrtf(distSquared);
			d /= out.dist;
			const float dp = fabsf(dot(d, out.n));
			pdf *= dp != 0 ? (distSquared / dp) : 0.0f;
		}
		out.d = d; out.pdf = pdf;
		if (dot(d, refN) >= 0 && dot(d, out.n) < 0 && pdf != 0) return em.radiance / pdf;
		out.pdf = 0.0f;
		return v3(0.0f);
	} else if (em.type == MTSB_EMITTER_POINT) {
		V3 d = em.pos - ref;
		out.dist = length(d);
		const float invDist = 1.0f / out.dist;
		out.d = d * invDist; out.pdf = 1; out.measure = M_DISCRETE; out.p = em.pos;
		return em.radiance * (invDist * invDist);
	} else if (!LEAN && em.type == MTSB_EMITTER_ENVMAP) {        // EnvironmentMap::sampleDirect (envmap.cpp:525-553)
		const EnvmapDev &E = *envmap_of(em);
		V3 value, dl; float pdf;
		envmap_sample_direction(E, sx, sy, dl, value, pdf);
		const V3 d = mat3_mul(E.toWorld, dl);
		float nearT, farT;
		if (is_zero(value) || pdf == 0 || !sphere_intersect(em.sphereCenter, em.sphereRadius, ref, d, nearT, farT) || nearT >= 0 || farT <= 0) return v3(0.0f);
		const V3 p = ref + d * farT;
		out.n = normalize(em.sphereCenter - p);
		out.d = d; out.dist = farT; out.pdf = pdf; out.p = p;
		return value / pdf;
	} else {
		V3 d; float pdf;
		if (!is_zero(refN)) {
			const V3 l = square_to_cosine_hemisphere(sx, sy);
			pdf = MTS_INV_PI * l.z;
			V3 s, t; coordinate_system(refN, s, t);
			d = s * l.x + t * l.y + refN * l.z;
		} else {
			d = square_to_uniform_sphere(sx, sy);
			pdf = 1.0f / (4.0f * MTS_PI);
		}
		float nearT, farT;
		if (!sphere_intersect(em.sphereCenter, em.sphereRadius, ref, d, nearT, farT)) return v3(0.0f);
		if (!(nearT < 0 && farT > 0)) return v3(0.0f);
		const V3 p = ref + d * farT;
		out.n = normalize(em.sphereCenter - p);
		out.d = d; out.dist = farT; out.pdf = pdf; out.p = p;
		if (!is_zero(refN) && dot(d, refN) <= 0) return v3(0.0f);
		return em.radiance / pdf;
	}
}

// Emitter::pdfDirect for a direction that hit emitter `e` (d, dist, n of the hit; measure = solid angle).  `ref` (dRec.ref) is
// read by sphere emitters only.
template <bool LEAN = false>
__device__ inline float emitter_pdf_direct(const SceneDev &S, uint32_t e, const V3 &refN, const V3 &d, float dist, const V3 &n, int measure,
		const V3 &ref = V3()) {
	const EmitterRec &em = S.emitters[e];
	if (em.type == MTSB_EMITTER_AREA) {
		if (dot(d, refN) >= 0 && dot(d, n) < 0) {
			const ShapeRec &esh = S.shapes[em.shape];
			if (!LEAN && esh.type == SHAPE_SPHERE) {
				const float pdfSA = sphere_pdf_direct(S.rects[esh.rect], ref, d, dist, n);
				if (measure == M_SOLID_ANGLE) return pdfSA;
				if (measure == M_AREA) {
					const RectRec &r = S.rects[esh.rect];
					const float sinAlpha = r.dpdv.x / length(r.dpdu - ref);
					return sinAlpha < 1 - MTS_EPSILON ? pdfSA * fabsf(dot(d, n)) / (dist * dist) : r.invSurfaceArea;
				}
				return 0.0f;
			}
			if (measure == M_SOLID_ANGLE) return em.invSurfaceArea * (dist * dist) / fabsf(dot(d, n));
			if (measure == M_AREA) return em.invSurfaceArea;
		}
		return 0.0f;
	} else if (em.type == MTSB_EMITTER_POINT) {
		return measure == M_DISCRETE ? 1.0f : 0.0f;
	} else if (!LEAN && em.type == MTSB_EMITTER_ENVMAP) {        // EnvironmentMap::pdfDirect (envmap.cpp:555-565)
		const EnvmapDev &E = *envmap_of(em);
		const float pdfSA = envmap_pdf_direction(E, mat3_mul(E.toLocal, d));
		if (measure == M_SOLID_ANGLE) return pdfSA;
		if (measure == M_AREA) return pdfSA * fabsf(dot(d, n)) / (dist * dist);
		return 0.0f;
	} else {
		const float pdfSA = !is_zero(refN) ? MTS_INV_PI * fmaxf(0.0f, dot(d, refN)) : 1.0f / (4.0f * MTS_PI);
		if (measure == M_SOLID_ANGLE) return pdfSA;
		if (measure == M_AREA) return pdfSA * fabsf(dot(d, n)) / (dist * dist);
		return 0.0f;
	}
}

// Scene::pdfEmitterDiscrete (include/mitsuba/render/scene.h): samplingWeight * m_emitterPDF.getNormalization()
__device__ inline float emitter_pdf_discrete(const SceneDev &S, uint32_t e) { return S.emitters[e].samplingWeight * S.emitterNorm; }

// Scene::sampleEmitterDirect(dRec, sample, testVisibility = false)
template <bool LEAN = false>
__device__ inline V3 scene_sample_emitter_direct(const SceneDev &S, const V3 &ref, const V3 &refN, float sx, float sy, DirectSample &out) {
	float emPdf;
	const uint32_t index = cdf_sample_reuse(S.emitterCdf, S.emitterCount, sx, emPdf);
	V3 value = emitter_sample_direct<LEAN>(S, index, ref, refN, sx, sy, out);
	if (out.pdf != 0) { out.pdf *= emPdf; return value / emPdf; }
	return v3(0.0f);
}

// Scene::evalEnvironment(ray) for the scene's environment emitter (S.envEmitter >= 0): constant.cpp:211-213 / envmap.cpp:383-414.
// rx, ry: the differential directions of a sensor ray (hasDifferentials); kept out of line — the EWA lookup is large and rare.
static __device__ __noinline__ V3 envmap_radiance(const EmitterRec *env, float dx, float dy, float dz, bool hasDifferentials,
		float rxx, float rxy, float rxz, float ryx, float ryy, float ryz) {
	return envmap_eval(*envmap_of(*env), v3(dx, dy, dz), hasDifferentials, v3(rxx, rxy, rxz), v3(ryx, ryy, ryz));
}
template <bool LEAN = false>
__device__ __forceinline__ V3 env_radiance(const SceneDev &S, const V3 &d, bool hasDifferentials = false, const V3 &rx = V3(), const V3 &ry = V3()) {
	const EmitterRec &env = S.emitters[S.envEmitter];
	if (!LEAN && env.type == MTSB_EMITTER_ENVMAP) return envmap_radiance(&env, d.x, d.y, d.z, hasDifferentials, rx.x, rx.y, rx.z, ry.x, ry.y, ry.z);
	return env.radiance;
}

// Emitter::eval(its, d) for an area light hit from direction -d (area.cpp:106-111)
__device__ inline V3 area_emitter_eval(const SceneDev &S, uint32_t e, const V3 &ns, const V3 &d) {
	return dot(ns, d) <= 0 ? v3(0.0f) : S.emitters[e].radiance;
}

} // namespace mtsb
