/*
 * oracle/restate.c — TEST INFRASTRUCTURE, not product code.
 *
 * Plain-C, single-threaded CPU restatement of the reference's intersection path, written from the reference's
 * algorithm (each function cites the file:line it follows).  It exists so that tests can check the CUDA path on
 * arbitrary seeded inputs; it is itself pinned against the UNMODIFIED reference (oracle/_ref, built from
 * /root/reference by oracle/Makefile.ref) through the golden vectors in tests/golden/ (tests/test_oracle.py).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load it.
 *
 * The acceleration structure is deliberately NOT restated: the kd-tree only prunes; the accepted hit is defined by
 * the primitive tests, the [mint,maxt] rules and the "later test wins ties" shrink rule, which a brute-force loop
 * over primitives in index order reproduces (sahkdtree3.h:275-291) — O(N) per ray, so small scenes only.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_API __attribute__((visibility("default")))
#define EPSILON 1e-4f            /* include/mitsuba/core/constants.h:26 */
#define KD_AABB_EPSILON 1e-3f    /* include/mitsuba/render/gkdtree.h:34 */

typedef struct { uint32_t k; float n_u, n_v, n_d, a_u, a_v, b_nu, b_nv, c_nu, c_nv; } TriAccel;

/* TriAccel::load — include/mitsuba/render/triaccel.h:61-94 */
static int triaccel_load(TriAccel *ta, const float *A, const float *B, const float *C) {
	static const int waldModulo[4] = { 1, 2, 0, 1 };
	float b[3], c[3], N[3];
	for (int i = 0; i < 3; ++i) { b[i] = C[i] - A[i]; c[i] = B[i] - A[i]; }
	N[0] = c[1]*b[2] - c[2]*b[1]; N[1] = c[2]*b[0] - c[0]*b[2]; N[2] = c[0]*b[1] - c[1]*b[0];
	int k = 0;
	for (int j = 0; j < 3; ++j) if (fabsf(N[j]) > fabsf(N[k])) k = j;
	int u = waldModulo[k], v = waldModulo[k+1];
	float n_k = N[k], denom = b[u]*c[v] - b[v]*c[u];
	ta->k = (uint32_t) k;
	if (denom == 0) { ta->k = 3; return 1; }
	ta->n_u = N[u] / n_k; ta->n_v = N[v] / n_k;
	ta->n_d = (A[0]*N[0] + A[1]*N[1] + A[2]*N[2]) / n_k;
	ta->b_nu = b[u] / denom; ta->b_nv = -b[v] / denom;
	ta->a_u = A[u]; ta->a_v = A[v];
	ta->c_nu = c[v] / denom; ta->c_nv = -c[u] / denom;
	return 0;
}

/* TriAccel::rayIntersect — include/mitsuba/render/triaccel.h:96-158 */
static int triaccel_intersect(const TriAccel *ta, const float *o, const float *d, float mint, float maxt,
		float *u, float *v, float *t) {
	float o_u, o_v, o_k, d_u, d_v, d_k;
	switch (ta->k) {
		case 0: o_u = o[1]; o_v = o[2]; o_k = o[0]; d_u = d[1]; d_v = d[2]; d_k = d[0]; break;
		case 1: o_u = o[2]; o_v = o[0]; o_k = o[1]; d_u = d[2]; d_v = d[0]; d_k = d[1]; break;
		case 2: o_u = o[0]; o_v = o[1]; o_k = o[2]; d_u = d[0]; d_v = d[1]; d_k = d[2]; break;
		default: return 0;
	}
	*t = (ta->n_d - o_u*ta->n_u - o_v*ta->n_v - o_k) / (d_u * ta->n_u + d_v * ta->n_v + d_k);
	if (*t < mint || *t > maxt) return 0;
	const float hu = o_u + *t * d_u - ta->a_u;
	const float hv = o_v + *t * d_v - ta->a_v;
	*u = hv * ta->b_nu + hu * ta->b_nv;
	*v = hu * ta->c_nu + hv * ta->c_nv;
	return *u >= 0 && *v >= 0 && *u + *v <= 1.0f;
}

/* Rectangle::rayIntersect — src/shapes/rectangle.cpp:125-147 (toObject: 3x4 row-major affine) */
static int rect_intersect(const float *m, const float *o, const float *d, float mint, float maxt,
		float *u, float *v, float *t) {
	float lo[3], ld[3];
	for (int r = 0; r < 3; ++r) {
		lo[r] = m[r*4]*o[0] + m[r*4+1]*o[1] + m[r*4+2]*o[2] + m[r*4+3];
		ld[r] = m[r*4]*d[0] + m[r*4+1]*d[1] + m[r*4+2]*d[2];
	}
	float hit = -lo[2] / ld[2];
	if (!(hit >= mint && hit <= maxt)) return 0;
	float lx = lo[0] + ld[0]*hit, ly = lo[1] + ld[1]*hit;
	if (fabsf(lx) <= 1 && fabsf(ly) <= 1) { *t = hit; *u = lx; *v = ly; return 1; }
	return 0;
}

/* AABB::rayIntersect — include/mitsuba/core/aabb.h:308-338 */
static int aabb_intersect(const float *mn, const float *mx, const float *o, const float *d, float *nearT, float *farT) {
	*nearT = -INFINITY; *farT = INFINITY;
	for (int i = 0; i < 3; ++i) {
		if (d[i] == 0) {
			if (o[i] < mn[i] || o[i] > mx[i]) return 0;
		} else {
			float rcp = 1.0f / d[i];
			float t1 = (mn[i] - o[i]) * rcp, t2 = (mx[i] - o[i]) * rcp;
			if (t1 > t2) { float tmp = t1; t1 = t2; t2 = tmp; }
			if (t1 > *nearT) *nearT = t1;
			if (t2 < *farT) *farT = t2;
			if (!(*nearT <= *farT)) return 0;
		}
	}
	return 1;
}

typedef struct {
	uint32_t nprims;
	TriAccel *accel;       /* per global primitive; k = 0xFFFFFFFF marks a rectangle */
	uint32_t *rectIndex;   /* per primitive: index into rects (rectangles only) */
	float *rects;          /* 12 floats (world->object) per rectangle */
	float aabbMin[3], aabbMax[3];
} OracleScene;

/* Builds the primitive table in ShapeKDTree::build order (src/librender/skdtree.cpp:68-110).
 * kinds[s] = 0 mesh (positions/triangles concatenated, vtxBase/triBase/triCount per shape) or 1 rectangle
 * (rectToWorld 16 floats row-major each, consumed in order).  Scene AABB as gkdtree.h:1213-1220. */
ORACLE_API OracleScene *oracle_scene_create(uint32_t nshapes, const int32_t *kinds, const float *positions,
		const uint32_t *triangles, const uint32_t *vtxBase, const uint32_t *triBase, const uint32_t *triCount,
		const float *rectToWorld, const float *rectToObject) {
	OracleScene *s = (OracleScene *) calloc(1, sizeof(OracleScene));
	uint32_t n = 0, nrect = 0;
	for (uint32_t i = 0; i < nshapes; ++i) { if (kinds[i] == 0) n += triCount[i]; else { n += 1; ++nrect; } }
	s->nprims = n;
	s->accel = (TriAccel *) calloc(n, sizeof(TriAccel));
	s->rectIndex = (uint32_t *) calloc(n, sizeof(uint32_t));
	s->rects = (float *) calloc(nrect ? nrect * 12 : 1, sizeof(float));
	float mn[3] = { INFINITY, INFINITY, INFINITY }, mx[3] = { -INFINITY, -INFINITY, -INFINITY };
	uint32_t idx = 0, r = 0;
	for (uint32_t i = 0; i < nshapes; ++i) {
		if (kinds[i] == 0) {
			for (uint32_t j = 0; j < triCount[i]; ++j) {
				const uint32_t *tri = triangles + (size_t) (triBase[i] + j) * 3;
				const float *A = positions + (size_t) (vtxBase[i] + tri[0]) * 3;
				const float *B = positions + (size_t) (vtxBase[i] + tri[1]) * 3;
				const float *C = positions + (size_t) (vtxBase[i] + tri[2]) * 3;
				triaccel_load(&s->accel[idx], A, B, C);
				for (int a = 0; a < 3; ++a) {
					mn[a] = fminf(mn[a], fminf(A[a], fminf(B[a], C[a])));
					mx[a] = fmaxf(mx[a], fmaxf(A[a], fmaxf(B[a], C[a])));
				}
				++idx;
			}
		} else {
			const float *m = rectToWorld + (size_t) r * 16;
			memcpy(s->rects + (size_t) r * 12, rectToObject + (size_t) r * 12, 12 * sizeof(float));
			s->accel[idx].k = 0xFFFFFFFFu; s->rectIndex[idx] = r;
			static const float corners[4][2] = { {-1,-1}, {1,-1}, {1,1}, {-1,1} };   /* Rectangle::getAABB, rectangle.cpp:111-118 */
			for (int c = 0; c < 4; ++c)
				for (int a = 0; a < 3; ++a) {
					float p = m[a*4]*corners[c][0] + m[a*4+1]*corners[c][1] + m[a*4+3];
					mn[a] = fminf(mn[a], p); mx[a] = fmaxf(mx[a], p);
				}
			++idx; ++r;
		}
	}
	for (int a = 0; a < 3; ++a) {
		float lo = mn[a], hi = mx[a];
		lo -= (hi - lo) * KD_AABB_EPSILON + KD_AABB_EPSILON;
		hi += (hi - lo) * KD_AABB_EPSILON + KD_AABB_EPSILON;
		s->aabbMin[a] = lo; s->aabbMax[a] = hi;
	}
	return s;
}

ORACLE_API void oracle_scene_destroy(OracleScene *s) {
	if (!s) return;
	free(s->accel); free(s->rectIndex); free(s->rects); free(s);
}

ORACLE_API void oracle_scene_aabb(const OracleScene *s, float *mn, float *mx) {
	memcpy(mn, s->aabbMin, 12); memcpy(mx, s->aabbMax, 12);
}

/* The prologue shared by the ShapeKDTree::rayIntersect overloads (src/librender/skdtree.cpp:112-142, 207-226). */
static int prologue(const OracleScene *s, const float *ray, int floorEps, float *mint, float *maxt) {
	const float *o = ray, *d = ray + 4;
	float nearT, farT;
	if (!aabb_intersect(s->aabbMin, s->aabbMax, o, d, &nearT, &farT)) return 0;
	float rayMinT = ray[3];
	if (rayMinT == EPSILON) {
		float m = fmaxf(fmaxf(fabsf(o[0]), fabsf(o[1])), fabsf(o[2]));
		if (floorEps) m = fmaxf(m, EPSILON);
		rayMinT *= m;
	}
	*mint = nearT; *maxt = farT;
	if (rayMinT > *mint) *mint = rayMinT;
	if (ray[7] < *maxt) *maxt = ray[7];
	return *maxt > *mint;
}

/* rays: n x 8 {o, mint, d, maxt}; hits: n x 4 {t, u, v, prim(bits)} — closest hit (skdtree.cpp:112-142 with the
 * leaf loop of sahkdtree3.h:268-292 run over all primitives). */
ORACLE_API void oracle_closest(const OracleScene *s, const float *rays, size_t n, float *hits) {
	for (size_t i = 0; i < n; ++i) {
		const float *ray = rays + i * 8;
		float mint, maxt, bestU = 0, bestV = 0; uint32_t best = 0xFFFFFFFFu; int found = 0;
		if (prologue(s, ray, 1, &mint, &maxt)) {
			for (uint32_t p = 0; p < s->nprims; ++p) {
				float t, u, v; int h;
				if (s->accel[p].k != 0xFFFFFFFFu) h = triaccel_intersect(&s->accel[p], ray, ray + 4, mint, maxt, &u, &v, &t);
				else h = rect_intersect(s->rects + (size_t) s->rectIndex[p] * 12, ray, ray + 4, mint, maxt, &u, &v, &t);
				if (h) { maxt = t; bestU = u; bestV = v; best = p; found = 1; }
			}
		}
		float *o = hits + i * 4;
		o[0] = found ? maxt : INFINITY; o[1] = bestU; o[2] = bestV; memcpy(o + 3, &best, 4);
	}
}

/* any hit (skdtree.cpp:207-226): epsilon scale without the Epsilon floor */
ORACLE_API void oracle_any(const OracleScene *s, const float *rays, size_t n, uint8_t *out) {
	for (size_t i = 0; i < n; ++i) {
		const float *ray = rays + i * 8;
		float mint, maxt; int found = 0;
		if (prologue(s, ray, 0, &mint, &maxt)) {
			for (uint32_t p = 0; p < s->nprims && !found; ++p) {
				float t, u, v;
				if (s->accel[p].k != 0xFFFFFFFFu) found = triaccel_intersect(&s->accel[p], ray, ray + 4, mint, maxt, &u, &v, &t);
				else found = rect_intersect(s->rects + (size_t) s->rectIndex[p] * 12, ray, ray + 4, mint, maxt, &u, &v, &t);
			}
		}
		out[i] = (uint8_t) found;
	}
}
