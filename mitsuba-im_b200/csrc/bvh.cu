// bvh.cu — GPU construction of the 8-wide compressed BVH that replaces the reference's SAH kd-tree
// (GenericKDTree::buildInternal, include/mitsuba/render/gkdtree.h:958-1263) and of the TriAccel records
// (ShapeKDTree::build, src/librender/skdtree.cpp:68-110; TriAccel::load, include/mitsuba/render/triaccel.h:61-94).
//
// Pipeline (all on the context stream; the PLOC rounds and the collapse levels read their counts from the device, the host
// looks at them once per batch of rounds / levels):
//   prim_bounds   : AABB + centroid of every primitive (triangle or analytic rectangle), scene AABB reduction
//   morton + sort : 63-bit Morton codes of centroids, cub::DeviceRadixSort
//   binary tree   : LBVH (Karras 2012) or PLOC (Meister & Bittner 2018) over the sorted primitives
//   refit         : bottom-up boxes + primitive counts
//   collapse      : top-down, level by level, binary -> 8-wide; children are octant-ordered and their boxes
//                   quantised to 8 bits relative to the node frame (Ylitie et al. 2017 layout)
//   triaccel      : Wald records in leaf order
// The tree does not have to match the kd-tree — only the hits do.
#include "bvh.cuh"
#include <cub/cub.cuh>
#include <stdlib.h>
#include <algorithm>
#include <string.h>

namespace mtsb {

// ---------------------------------------------------------------------------------------------------
// helpers
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t float_flip(float f) {   // order-preserving float -> uint
	uint32_t u = __float_as_uint(f);
	return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float float_unflip(uint32_t u) {
	return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

__device__ __forceinline__ V3 xform_point(const float *m, float x, float y, float z) {
	return v3(m[0]*x + m[1]*y + m[2]*z + m[3], m[4]*x + m[5]*y + m[6]*z + m[7], m[8]*x + m[9]*y + m[10]*z + m[11]);
}

__device__ __forceinline__ uint32_t find_shape(const ShapeRec *shapes, uint32_t shapeCount, uint32_t prim) {
	uint32_t lo = 0, hi = shapeCount;       // last shape with primOffset <= prim
	while (hi - lo > 1) {
		uint32_t mid = (lo + hi) >> 1;
		if (shapes[mid].primOffset <= prim) lo = mid; else hi = mid;
	}
	return lo;
}

__device__ __forceinline__ void load_tri(const SceneDev &S, const ShapeRec &sh, uint32_t local, V3 &A, V3 &B, V3 &C) {
	const uint32_t *ix = S.indices + (size_t) (sh.triBase + local) * 3;
	const float *p0 = S.positions + (size_t) (sh.vtxBase + ix[0]) * 3;
	const float *p1 = S.positions + (size_t) (sh.vtxBase + ix[1]) * 3;
	const float *p2 = S.positions + (size_t) (sh.vtxBase + ix[2]) * 3;
	A = v3(p0[0], p0[1], p0[2]); B = v3(p1[0], p1[1], p1[2]); C = v3(p2[0], p2[1], p2[2]);
}

// ---------------------------------------------------------------------------------------------------
// 1. primitive bounds
// ---------------------------------------------------------------------------------------------------
__global__ void prim_bounds_kernel(SceneDev S, uint32_t primCount, float4 *lo, float4 *hi, uint32_t *sceneBox) {
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	V3 mn = v3(INFINITY), mx = v3(-INFINITY);
	if (i < primCount) {
		uint32_t s = find_shape(S.shapes, S.shapeCount, i);
		const ShapeRec sh = S.shapes[s];
		if (sh.type == SHAPE_MESH) {
			V3 A, B, C; load_tri(S, sh, i - sh.primOffset, A, B, C);
			mn = vmin3(A, vmin3(B, C)); mx = vmax3(A, vmax3(B, C));
		} else if (sh.type == SHAPE_SPHERE) {                // Sphere::getAABB, src/shapes/sphere.cpp:147-152
			const RectRec &r = S.rects[sh.rect];
			mn = r.dpdu - v3(r.dpdv.x); mx = r.dpdu + v3(r.dpdv.x);
		} else if (sh.type == SHAPE_DISK) {                  // Disk::getAABB, src/shapes/disk.cpp:117-130
			const float *m = S.rects[sh.rect].toWorld;
			V3 c0 = xform_point(m, 1, 0, 0), c1 = xform_point(m, -1, 0, 0), c2 = xform_point(m, 0, 1, 0), c3 = xform_point(m, 0, -1, 0);
			mn = vmin3(vmin3(c0, c1), vmin3(c2, c3)); mx = vmax3(vmax3(c0, c1), vmax3(c2, c3));
		} else {
			const float *m = S.rects[sh.rect].toWorld;   // Rectangle::getAABB, src/shapes/rectangle.cpp:111-118
			V3 c0 = xform_point(m, -1, -1, 0), c1 = xform_point(m, 1, -1, 0), c2 = xform_point(m, 1, 1, 0), c3 = xform_point(m, -1, 1, 0);
			mn = vmin3(vmin3(c0, c1), vmin3(c2, c3)); mx = vmax3(vmax3(c0, c1), vmax3(c2, c3));
		}
		lo[i] = make_float4(mn.x, mn.y, mn.z, 0); hi[i] = make_float4(mx.x, mx.y, mx.z, 0);
	}
	// warp reduce then one atomic per warp
	#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		mn.x = fminf(mn.x, __shfl_xor_sync(0xffffffffu, mn.x, o)); mn.y = fminf(mn.y, __shfl_xor_sync(0xffffffffu, mn.y, o));
		mn.z = fminf(mn.z, __shfl_xor_sync(0xffffffffu, mn.z, o)); mx.x = fmaxf(mx.x, __shfl_xor_sync(0xffffffffu, mx.x, o));
		mx.y = fmaxf(mx.y, __shfl_xor_sync(0xffffffffu, mx.y, o)); mx.z = fmaxf(mx.z, __shfl_xor_sync(0xffffffffu, mx.z, o));
	}
	if ((threadIdx.x & 31) == 0 && mn.x <= mx.x) {
		atomicMin(sceneBox + 0, float_flip(mn.x)); atomicMin(sceneBox + 1, float_flip(mn.y)); atomicMin(sceneBox + 2, float_flip(mn.z));
		atomicMax(sceneBox + 3, float_flip(mx.x)); atomicMax(sceneBox + 4, float_flip(mx.y)); atomicMax(sceneBox + 5, float_flip(mx.z));
	}
}

__global__ void init_scene_box_kernel(uint32_t *sceneBox) {
	if (threadIdx.x < 3) sceneBox[threadIdx.x] = 0xffffffffu; else if (threadIdx.x < 6) sceneBox[threadIdx.x] = 0u;
}

// ---------------------------------------------------------------------------------------------------
// 2. Morton codes
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t expand21(uint32_t v) {   // spread 21 bits to every third bit
	uint64_t x = v & 0x1fffffu;
	x = (x | x << 32) & 0x1f00000000ffffull;
	x = (x | x << 16) & 0x1f0000ff0000ffull;
	x = (x | x << 8)  & 0x100f00f00f00f00full;
	x = (x | x << 4)  & 0x10c30c30c30c30c3ull;
	x = (x | x << 2)  & 0x1249249249249249ull;
	return x;
}

__global__ void morton_kernel(uint32_t primCount, const float4 *lo, const float4 *hi, const uint32_t *sceneBox,
		uint64_t *keys, uint32_t *vals) {
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= primCount) return;
	V3 smin = v3(float_unflip(sceneBox[0]), float_unflip(sceneBox[1]), float_unflip(sceneBox[2]));
	V3 smax = v3(float_unflip(sceneBox[3]), float_unflip(sceneBox[4]), float_unflip(sceneBox[5]));
	V3 ext = smax - smin;
	float4 l = lo[i], h = hi[i];
	float cx = ext.x > 0 ? (0.5f * (l.x + h.x) - smin.x) / ext.x : 0.5f;
	float cy = ext.y > 0 ? (0.5f * (l.y + h.y) - smin.y) / ext.y : 0.5f;
	float cz = ext.z > 0 ? (0.5f * (l.z + h.z) - smin.z) / ext.z : 0.5f;
	const float scale = 2097152.0f;   // 2^21
	uint32_t ix = (uint32_t) fminf(fmaxf(cx * scale, 0.0f), scale - 1.0f);
	uint32_t iy = (uint32_t) fminf(fmaxf(cy * scale, 0.0f), scale - 1.0f);
	uint32_t iz = (uint32_t) fminf(fmaxf(cz * scale, 0.0f), scale - 1.0f);
	keys[i] = (expand21(ix) << 2) | (expand21(iy) << 1) | expand21(iz);
	vals[i] = i;
}

// ---------------------------------------------------------------------------------------------------
// 3a. LBVH (Karras 2012).  Binary node ids: [0, N-1) internal, [N-1, 2N-1) leaves in sorted order.
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ int delta(const uint64_t *keys, int n, int i, int j) {
	if (j < 0 || j >= n) return -1;
	uint64_t a = keys[i], b = keys[j];
	if (a == b) return 64 + __clz((uint32_t) i ^ (uint32_t) j);
	return __clzll((long long) (a ^ b));
}

__global__ void lbvh_kernel(int n, const uint64_t *keys, int2 *children, int *parent) {
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n - 1) return;
	int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
	int dmin = delta(keys, n, i, i - d);
	int lmax = 2;
	while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
	int l = 0;
	for (int t = lmax >> 1; t >= 1; t >>= 1)
		if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
	int j = i + l * d;
	int dnode = delta(keys, n, i, j);
	int s = 0;
	for (int t = (l + 1) >> 1; ; t = (t + 1) >> 1) {
		if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
		if (t == 1) break;
	}
	int gamma = i + s * d + min(d, 0);
	int left = (min(i, j) == gamma) ? (n - 1 + gamma) : gamma;
	int right = (max(i, j) == gamma + 1) ? (n - 1 + gamma + 1) : gamma + 1;
	children[i] = make_int2(left, right);
	parent[left] = i; parent[right] = i;
	if (i == 0) parent[0] = -1;
}

// bottom-up refit: boxes and primitive counts (count in lo.w as bits)
__global__ void refit_kernel(int n, const uint32_t *sortedPrim, const float4 *primLo, const float4 *primHi,
		const int2 *children, const int *parent, float4 *blo, float4 *bhi, uint32_t *flags) {
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	uint32_t prim = sortedPrim[i];
	float4 l = primLo[prim], h = primHi[prim];
	l.w = __uint_as_float(1u);
	int node = n - 1 + i;
	blo[node] = l; bhi[node] = h;
	if (n == 1) return;
	int p = parent[node];
	while (p >= 0) {
		__threadfence();
		if (atomicAdd(flags + p, 1u) == 0) return;     // first arrival leaves, second continues
		int2 c = children[p];
		float4 l0 = __ldcg(blo + c.x), h0 = __ldcg(bhi + c.x), l1 = __ldcg(blo + c.y), h1 = __ldcg(bhi + c.y);
		float4 lo4 = make_float4(fminf(l0.x, l1.x), fminf(l0.y, l1.y), fminf(l0.z, l1.z),
			__uint_as_float(__float_as_uint(l0.w) + __float_as_uint(l1.w)));
		float4 hi4 = make_float4(fmaxf(h0.x, h1.x), fmaxf(h0.y, h1.y), fmaxf(h0.z, h1.z), 0);
		blo[p] = lo4; bhi[p] = hi4;
		p = parent[p];
	}
}

// ---------------------------------------------------------------------------------------------------
// 3b. PLOC (Meister & Bittner 2018): iterative merging of mutually-nearest clusters within a window of
// the Morton-sorted order.  Produces the same binary representation as LBVH (ids 0..N-2 internal).
// ---------------------------------------------------------------------------------------------------
#ifndef PLOC_RADIUS
#define PLOC_RADIUS 16
#endif
__device__ __forceinline__ float box_area(float4 l, float4 h) {
	float dx = h.x - l.x, dy = h.y - l.y, dz = h.z - l.z;
	return 2.0f * (dx * dy + dy * dz + dz * dx);
}

// The merge rounds are driven from the device: every kernel reads the live cluster count m from mPtr and is launched over
// an upper bound `mBound` the host knew a few rounds ago, so a round needs no host round trip (the host reads m back once
// per PLOC_BATCH rounds to shrink the grids); below PLOC_TAIL clusters one CTA finishes the tree by itself (ploc_tail_kernel).
#define PLOC_BATCH 4
#define PLOC_TAIL 1024

__device__ __forceinline__ int ploc_nearest(int i, int m, const int *cluster, const float4 *blo, const float4 *bhi) {
	float4 l = blo[cluster[i]], h = bhi[cluster[i]];
	float best = INFINITY; int bestJ = -1;
	int j0 = max(0, i - PLOC_RADIUS), j1 = min(m - 1, i + PLOC_RADIUS);
	for (int j = j0; j <= j1; ++j) {
		if (j == i) continue;
		float4 l2 = blo[cluster[j]], h2 = bhi[cluster[j]];
		float4 ml = make_float4(fminf(l.x, l2.x), fminf(l.y, l2.y), fminf(l.z, l2.z), 0);
		float4 mh = make_float4(fmaxf(h.x, h2.x), fmaxf(h.y, h2.y), fmaxf(h.z, h2.z), 0);
		float a = box_area(ml, mh);
		if (a < best || (a == best && j < bestJ)) { best = a; bestJ = j; }
	}
	return bestJ;
}

// merge mutual pairs (lower index of the pair creates the node); returns the cluster id that stays at position i or -1
__device__ __forceinline__ int ploc_merge(int i, const int *cluster, const int *nearest, float4 *blo, float4 *bhi,
		int2 *children, int *parent, int *nodeCounter) {
	int j = nearest[i];
	int res = cluster[i];
	if (j >= 0 && nearest[j] == i) {
		if (i < j) {
			int a = cluster[i], b = cluster[j];
			int id = atomicSub(nodeCounter, 1) - 1;    // internal ids handed out from N-2 downwards; root ends at 0
			float4 l0 = blo[a], h0 = bhi[a], l1 = blo[b], h1 = bhi[b];
			blo[id] = make_float4(fminf(l0.x, l1.x), fminf(l0.y, l1.y), fminf(l0.z, l1.z),
				__uint_as_float(__float_as_uint(l0.w) + __float_as_uint(l1.w)));
			bhi[id] = make_float4(fmaxf(h0.x, h1.x), fmaxf(h0.y, h1.y), fmaxf(h0.z, h1.z), 0);
			children[id] = make_int2(a, b);
			parent[a] = id; parent[b] = id;
			res = id;
		} else {
			res = -1;
		}
	}
	return res;
}

__global__ void ploc_nearest_kernel(const int *mPtr, const int *cluster, const float4 *blo, const float4 *bhi, int *nearest) {
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	const int m = *mPtr;
	if (i >= m) return;
	nearest[i] = ploc_nearest(i, m, cluster, blo, bhi);
}

// positions m .. mBound-1 are dead: their flag is cleared so the scan over mBound items counts the live ones only
__global__ void ploc_merge_kernel(const int *mPtr, int mBound, const int *cluster, const int *nearest, float4 *blo, float4 *bhi,
		int2 *children, int *parent, int *nodeCounter, int *out, int *validFlag) {
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= mBound) return;
	if (i >= *mPtr) { validFlag[i] = 0; return; }
	int res = ploc_merge(i, cluster, nearest, blo, bhi, children, parent, nodeCounter);
	out[i] = res;
	validFlag[i] = res >= 0 ? 1 : 0;
}

// the last position also publishes the next round's cluster count (mNext is the other slot of the ping-pong pair, no
// kernel of this round reads it)
__global__ void ploc_compact_kernel(const int *mPtr, int mBound, const int *in, const int *validFlag, const int *scan, int *outCluster, int *mNext) {
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= mBound) return;
	if (i == mBound - 1) *mNext = scan[i] + validFlag[i];
	if (i < *mPtr && validFlag[i]) outCluster[scan[i]] = in[i];
}

// one CTA runs every remaining round over <= PLOC_TAIL clusters: same nearest / merge rules, the compaction is a block scan
__global__ void __launch_bounds__(PLOC_TAIL) ploc_tail_kernel(const int *mPtr, const int *cluster, float4 *blo, float4 *bhi,
		int2 *children, int *parent, int *nodeCounter) {
	__shared__ int cl[2][PLOC_TAIL];
	__shared__ int nearest[PLOC_TAIL];
	__shared__ int warpOffset[PLOC_TAIL / 32];
	__shared__ int total;
	const int i = threadIdx.x, lane = i & 31, warp = i >> 5;
	int m = min(*mPtr, PLOC_TAIL);
	if (i < m) cl[0][i] = cluster[i];
	int cur = 0;
	__syncthreads();
	while (m > 1) {
		if (i < m) nearest[i] = ploc_nearest(i, m, cl[cur], blo, bhi);
		__syncthreads();
		const int res = i < m ? ploc_merge(i, cl[cur], nearest, blo, bhi, children, parent, nodeCounter) : -1;
		const unsigned keep = __ballot_sync(0xffffffffu, res >= 0);
		if (lane == 0) warpOffset[warp] = __popc(keep);
		__syncthreads();
		if (warp == 0) {
			const int c = warpOffset[lane];
			int incl = c;
			for (int d = 1; d < 32; d <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += t; }
			warpOffset[lane] = incl - c;
			if (lane == 31) total = incl;
		}
		__syncthreads();
		if (res >= 0) cl[cur ^ 1][warpOffset[warp] + __popc(keep & ((1u << lane) - 1u))] = res;
		m = total; cur ^= 1;
		__syncthreads();      // also orders this round's blo / bhi / cl writes before the next round's reads
	}
}

__global__ void ploc_init_kernel(int n, const uint32_t *sortedPrim, const float4 *primLo, const float4 *primHi,
		float4 *blo, float4 *bhi, int *cluster) {
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	uint32_t prim = sortedPrim[i];
	float4 l = primLo[prim], h = primHi[prim];
	l.w = __uint_as_float(1u);
	blo[n - 1 + i] = l; bhi[n - 1 + i] = h;
	cluster[i] = n - 1 + i;
}

// ---------------------------------------------------------------------------------------------------
// 4. collapse binary -> BVH8, one level per launch
// ---------------------------------------------------------------------------------------------------
#define LEAF_MAX 3   // triangles per leaf child (unary count in 3 bits)

// Cost-driven collapse (after Ylitie, Karras, Laine 2017, sec. 4.2): bottom-up dynamic programme over the binary tree.
//   C(n,1) = min( leaf: A(n) P(n) c_prim  [P(n) <= LEAF_MAX],  wide node: A(n) c_node + D(n,8) )
//   D(n,j) = min over 0<k<j of C(left,k) + C(right,j-k)            ("n's subtree shown as j roots")
//   C(n,i) = min( D(n,i), C(n,i-1) )                               i = 2..7
// cost[n*7 + i-1] = C(n,i); split[n*8 + j-1] = the k of D(n,j) (0 = "use j-1 roots" for j < 8; split[..][0]: 0 leaf, 1 node).
// Same second-arrival climb as refit_kernel.
__global__ void collapse_dp_kernel(int n, const int2 *children, const int *parent, const float4 *blo, const float4 *bhi,
		uint32_t *flags, float *cost, uint8_t *split, float cNode, float cPrim) {
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	int node = n - 1 + i;
	{
		const float a = box_area(blo[node], bhi[node]) * cPrim;
		for (int j = 0; j < 7; ++j) cost[(size_t) node * 7 + j] = a;
		for (int j = 0; j < 8; ++j) split[(size_t) node * 8 + j] = 0;
	}
	if (n == 1) return;
	int p = parent[node];
	while (p >= 0) {
		__threadfence();
		if (atomicAdd(flags + p, 1u) == 0) return;
		const int2 c = children[p];
		float cl[8], cr[8];
		for (int j = 1; j <= 7; ++j) { cl[j] = __ldcg(cost + (size_t) c.x * 7 + j - 1); cr[j] = __ldcg(cost + (size_t) c.y * 7 + j - 1); }
		const float4 l = blo[p], h = bhi[p];
		const float A = box_area(l, h);
		const uint32_t P = __float_as_uint(l.w);
		float D[9]; uint8_t K[9];
		for (int j = 2; j <= 8; ++j) {
			float best = INFINITY; int bk = 1;
			for (int k = 1; k < j; ++k) {
				if (k > 7 || j - k > 7) continue;
				const float v = cl[k] + cr[j - k];
				if (v < best) { best = v; bk = k; }
			}
			D[j] = best; K[j] = (uint8_t) bk;
		}
		float C[8];
		uint8_t *sp = split + (size_t) p * 8;
		if (P <= LEAF_MAX) { C[1] = A * (float) P * cPrim; sp[0] = 0; }
		else { C[1] = A * cNode + D[8]; sp[0] = 1; }
		for (int j = 2; j <= 7; ++j) {
			if (D[j] < C[j - 1]) { C[j] = D[j]; sp[j - 1] = K[j]; } else { C[j] = C[j - 1]; sp[j - 1] = 0; }
		}
		sp[7] = K[8];
		for (int j = 1; j <= 7; ++j) cost[(size_t) p * 7 + j - 1] = C[j];
		p = parent[p];
	}
}


struct CollapseItem { int bnode; uint32_t wnode; };

__device__ __forceinline__ int enumerate_leaf(int n, int node, const int2 *children, int *outLeaves) {
	// collects the (<= LEAF_MAX) binary leaves below `node`
	int stack[8]; int sp = 0, cnt = 0;
	stack[sp++] = node;
	while (sp) {
		int x = stack[--sp];
		if (x >= n - 1) { if (cnt < LEAF_MAX) outLeaves[cnt] = x; ++cnt; }
		else { int2 c = children[x]; stack[sp++] = c.y; stack[sp++] = c.x; }
	}
	return cnt;
}

// The queue length of a level is read from the device (inCountPtr); the host launches a level over the bound 8^level
// (a wide node has at most 8 children) and looks at the counts once per batch of levels only.
__global__ void collapse_kernel(int n, int rootIsLeaf, const CollapseItem *inQ, const uint32_t *inCountPtr, CollapseItem *outQ, uint32_t *outCount,
		const int2 *children, const float4 *blo, const float4 *bhi, const uint32_t *sortedPrim,
		uint4 *nodes, uint32_t *nodeCounter, uint32_t *slotCounter, uint32_t *slotPrim, float *sahAccum, uint32_t *leafCounter,
		const uint8_t *dpSplit, uint32_t *depthCounter, uint32_t level) {
	uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t inCount = *inCountPtr;
	if (w >= inCount) return;
	if (w == 0) *depthCounter = level + 1;        // levels run in stream order: the last non-empty one leaves the depth
	const CollapseItem item = inQ[w];

	int kids[8]; int nk = 0;
	if (rootIsLeaf || item.bnode >= n - 1) {
		kids[nk++] = item.bnode;        // whole tree is a single small leaf
	} else if (dpSplit) {
		// follow the dynamic programme: this node is a wide node, its subtree is shown as (up to) 8 roots
		int stN[16]; int stI[16]; int sp = 0;
		{
			const int2 c = children[item.bnode];
			const int k = dpSplit[(size_t) item.bnode * 8 + 7];
			stN[sp] = c.y; stI[sp++] = 8 - k; stN[sp] = c.x; stI[sp++] = k;
		}
		while (sp) {
			const int c = stN[--sp]; int j = stI[sp];
			if (c >= n - 1) { kids[nk++] = c; continue; }
			const uint8_t *ds = dpSplit + (size_t) c * 8;
			while (j > 1 && ds[j - 1] == 0) --j;
			if (j == 1 || nk + sp + 2 > 8) { kids[nk++] = c; continue; }     // one root: leaf (P <= LEAF_MAX) or wide node
			const int2 cc = children[c];
			const int k = ds[j - 1];
			stN[sp] = cc.y; stI[sp++] = j - k; stN[sp] = cc.x; stI[sp++] = k;
		}
	} else {
		int2 c = children[item.bnode];
		kids[nk++] = c.x; kids[nk++] = c.y;
		while (nk < 8) {
			float best = -1; int bi = -1;
			for (int i = 0; i < nk; ++i) {
				int k = kids[i];
				if (k >= n - 1) continue;
				if (__float_as_uint(blo[k].w) <= LEAF_MAX) continue;
				float a = box_area(blo[k], bhi[k]);
				if (a > best) { best = a; bi = i; }
			}
			if (bi < 0) break;
			int2 cc = children[kids[bi]];
			kids[bi] = cc.x; kids[nk++] = cc.y;
		}
#ifndef MTSB_NO_LEAF_SPLIT
		// free slots left: open multi-triangle leaves (largest box first) — tighter leaf boxes at no extra node
		while (nk < 8) {
			float best = -1; int bi = -1;
			for (int i = 0; i < nk; ++i) {
				int k = kids[i];
				if (k >= n - 1) continue;
				float a = box_area(blo[k], bhi[k]);
				if (a > best) { best = a; bi = i; }
			}
			if (bi < 0) break;
			int2 cc = children[kids[bi]];
			kids[bi] = cc.x; kids[nk++] = cc.y;
		}
#endif
	}

	// node frame
	float4 nl = blo[item.bnode], nh = bhi[item.bnode];
	V3 lo = v3(nl.x, nl.y, nl.z), hi = v3(nh.x, nh.y, nh.z);
	uint32_t ebits[3]; float escale[3];
	#pragma unroll
	for (int a = 0; a < 3; ++a) {
		float ext = comp(hi, a) - comp(lo, a);
		// smallest power of two s with ext / s <= 255, kept conservative against rounding
		int e = 0;
		float m = frexpf(fmaxf(ext, 1e-30f) / 255.0f, &e);   // value = m * 2^e, m in [0.5,1)
		(void) m;
		int be = e + 127;
		be = max(1, min(254, be));
		float s = __uint_as_float((uint32_t) be << 23);
		while (ext / s > 255.0f && be < 254) { ++be; s = __uint_as_float((uint32_t) be << 23); }
		ebits[a] = (uint32_t) be; escale[a] = s;
	}

	// octant slot assignment (greedy on cost = dot(child centre - node centre, slot direction))
	V3 nc = 0.5f * (lo + hi);
	float cost[8][8];
	for (int i = 0; i < nk; ++i) {
		float4 l = blo[kids[i]], h = bhi[kids[i]];
		V3 cc = v3(0.5f * (l.x + h.x), 0.5f * (l.y + h.y), 0.5f * (l.z + h.z)) - nc;
		for (int s = 0; s < 8; ++s) {
			V3 ds = v3((s & 1) ? -1.0f : 1.0f, (s & 2) ? -1.0f : 1.0f, (s & 4) ? -1.0f : 1.0f);
			cost[i][s] = dot(cc, ds);
		}
	}
	int slotOf[8]; int kidAt[8];
	for (int i = 0; i < 8; ++i) { slotOf[i] = -1; kidAt[i] = -1; }
	for (int r = 0; r < nk; ++r) {
		float best = INFINITY; int bi = -1, bs = -1;
		for (int i = 0; i < nk; ++i) {
			if (slotOf[i] >= 0) continue;
			for (int s = 0; s < 8; ++s) {
				if (kidAt[s] >= 0) continue;
				if (cost[i][s] < best) { best = cost[i][s]; bi = i; bs = s; }
			}
		}
		slotOf[bi] = bs; kidAt[bs] = bi;
	}

	// classify, allocate
	uint32_t imask = 0, nInner = 0, nTri = 0;
	uint32_t triCount[8];
	for (int s = 0; s < 8; ++s) {
		triCount[s] = 0;
		if (kidAt[s] < 0) continue;
		int k = kids[kidAt[s]];
		uint32_t cnt = __float_as_uint(blo[k].w);
		if (k < n - 1 && cnt > LEAF_MAX) { imask |= 1u << s; ++nInner; }
		else { triCount[s] = cnt; nTri += cnt; }
	}
	uint32_t childBase = nInner ? atomicAdd(nodeCounter, nInner) : 0;
	uint32_t triBase = nTri ? atomicAdd(slotCounter, nTri) : 0;
	uint32_t qBase = nInner ? atomicAdd(outCount, nInner) : 0;

	uint32_t meta[8], q[6][8];
	uint32_t triOff = 0, innerIdx = 0;
	float nodeArea = box_area(nl, nh);
	float sah = 0;
	for (int s = 0; s < 8; ++s) {
		meta[s] = 0;
		for (int a = 0; a < 6; ++a) q[a][s] = 0;
		if (kidAt[s] < 0) continue;
		int k = kids[kidAt[s]];
		float4 l = blo[k], h = bhi[k];
		const float cl[3] = { l.x, l.y, l.z }, ch[3] = { h.x, h.y, h.z };
		for (int a = 0; a < 3; ++a) {
			float base = comp(lo, a), sc = escale[a];
			int ql = (int) floorf((cl[a] - base) / sc), qh = (int) ceilf((ch[a] - base) / sc);
			ql = max(0, min(255, ql)); qh = max(0, min(255, qh));
			while (ql > 0 && base + ql * sc > cl[a]) --ql;       // keep the dequantised box conservative
			while (qh < 255 && base + qh * sc < ch[a]) ++qh;
			q[a][s] = (uint32_t) ql; q[3 + a][s] = (uint32_t) qh;
		}
		if (imask & (1u << s)) {
			meta[s] = (1u << 5) | (24u + s);
			outQ[qBase + innerIdx].bnode = k; outQ[qBase + innerIdx].wnode = childBase + innerIdx;
			++innerIdx;
			sah += box_area(l, h);
		} else {
			uint32_t cnt = triCount[s];
			meta[s] = (((1u << cnt) - 1u) << 5) | triOff;
			int leaves[LEAF_MAX];
			enumerate_leaf(n, k, children, leaves);
			for (uint32_t t = 0; t < cnt; ++t)
				slotPrim[triBase + triOff + t] = sortedPrim[leaves[t] - (n - 1)];
			triOff += cnt;
			sah += box_area(l, h) * cnt;
		}
	}
	if (sahAccum) atomicAdd(sahAccum, sah);   // children's area-weighted cost (normalised on the host)
	(void) nodeArea;
	if (leafCounter) atomicAdd(leafCounter, __popc(~imask & 0xffu) - (8 - nk));

	uint4 *np = nodes + (size_t) item.wnode * 5;
	np[0] = make_uint4(__float_as_uint(lo.x), __float_as_uint(lo.y), __float_as_uint(lo.z),
		ebits[0] | (ebits[1] << 8) | (ebits[2] << 16) | (imask << 24));
	#define PACK4(arr, o) ((arr)[o] | ((arr)[o+1] << 8) | ((arr)[o+2] << 16) | ((arr)[o+3] << 24))
	np[1] = make_uint4(childBase, triBase, PACK4(meta, 0), PACK4(meta, 4));
	np[2] = make_uint4(PACK4(q[0], 0), PACK4(q[0], 4), PACK4(q[1], 0), PACK4(q[1], 4));
	np[3] = make_uint4(PACK4(q[2], 0), PACK4(q[2], 4), PACK4(q[3], 0), PACK4(q[3], 4));
	np[4] = make_uint4(PACK4(q[4], 0), PACK4(q[4], 4), PACK4(q[5], 0), PACK4(q[5], 4));
	#undef PACK4
}

// ---------------------------------------------------------------------------------------------------
// 5. TriAccel records in slot order (TriAccel::load, triaccel.h:61-94, evaluated without FMA contraction)
// ---------------------------------------------------------------------------------------------------
__global__ void triaccel_kernel(SceneDev S, uint32_t slotCount, const uint32_t *slotPrim, float4 *tris) {
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= slotCount) return;
	uint32_t prim = slotPrim[i];
	uint32_t s = find_shape(S.shapes, S.shapeCount, prim);
	const ShapeRec sh = S.shapes[s];
	float4 r0 = make_float4(0, 0, 0, 0), r1 = r0, r2 = r0;
	if (sh.type != SHAPE_MESH) {
		r0.x = __uint_as_float(K_NO_TRIANGLE);
		r2.z = __uint_as_float(s); r2.w = __uint_as_float(0u);
	} else {
		uint32_t local = prim - sh.primOffset;
		V3 A, B, C; load_tri(S, sh, local, A, B, C);
		V3 b = v3(__fsub_rn(C.x, A.x), __fsub_rn(C.y, A.y), __fsub_rn(C.z, A.z));
		V3 c = v3(__fsub_rn(B.x, A.x), __fsub_rn(B.y, A.y), __fsub_rn(B.z, A.z));
		V3 N = v3(__fsub_rn(__fmul_rn(c.y, b.z), __fmul_rn(c.z, b.y)),
		          __fsub_rn(__fmul_rn(c.z, b.x), __fmul_rn(c.x, b.z)),
		          __fsub_rn(__fmul_rn(c.x, b.y), __fmul_rn(c.y, b.x)));
		int k = 0;
		for (int j = 0; j < 3; ++j) if (fabsf(comp(N, j)) > fabsf(comp(N, k))) k = j;
		const int u = (k + 1) % 3, v = (k + 2) % 3;
		const float n_k = comp(N, k);
		const float denom = __fsub_rn(__fmul_rn(comp(b, u), comp(c, v)), __fmul_rn(comp(b, v), comp(c, u)));
		if (denom == 0) {
			k = 3;
		} else {
			r0.y = __fdiv_rn(comp(N, u), n_k);
			r0.z = __fdiv_rn(comp(N, v), n_k);
			float dotAN = __fadd_rn(__fadd_rn(__fmul_rn(A.x, N.x), __fmul_rn(A.y, N.y)), __fmul_rn(A.z, N.z));
			r0.w = __fdiv_rn(dotAN, n_k);
			r1.x = comp(A, u); r1.y = comp(A, v);
			r1.z = __fdiv_rn(comp(b, u), denom);          // b_nu
			r1.w = __fdiv_rn(-comp(b, v), denom);         // b_nv
			r2.x = __fdiv_rn(comp(c, v), denom);          // c_nu
			r2.y = __fdiv_rn(-comp(c, u), denom);         // c_nv
		}
		r0.x = __uint_as_float((uint32_t) k);
		r2.z = __uint_as_float(s); r2.w = __uint_as_float(local);
	}
	tris[(size_t) i * 3] = r0; tris[(size_t) i * 3 + 1] = r1; tris[(size_t) i * 3 + 2] = r2;
}

// ---------------------------------------------------------------------------------------------------
// 6. de-indexed shading records in slot order (layout in common.cuh: SceneDev::verts)
// ---------------------------------------------------------------------------------------------------
__global__ void shading_record_kernel(SceneDev S, uint32_t slotCount, const uint32_t *slotPrim, float4 *verts) {
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= slotCount) return;
	const uint32_t prim = slotPrim[i];
	const uint32_t s = find_shape(S.shapes, S.shapeCount, prim);
	const ShapeRec sh = S.shapes[s];
	float4 r0 = make_float4(0, 0, 0, 0), r1 = r0, r2 = r0, r3 = r0, r4 = r0;
	uint32_t local = 0;
	if (sh.type == SHAPE_MESH) {
		local = prim - sh.primOffset;
		const uint32_t *ix = S.indices + (size_t) (sh.triBase + local) * 3;
		const uint32_t i0 = sh.vtxBase + ix[0], i1 = sh.vtxBase + ix[1], i2 = sh.vtxBase + ix[2];
		const float *P = S.positions, *N = S.normals;
		r0 = make_float4(P[i0*3], P[i0*3+1], P[i0*3+2], P[i1*3]);
		r1 = make_float4(P[i1*3+1], P[i1*3+2], P[i2*3], P[i2*3+1]);
		r2.x = P[i2*3+2];
		if (sh.flags & SHAPE_HAS_NORMALS) {
			r2.y = N[i0*3]; r2.z = N[i0*3+1]; r2.w = N[i0*3+2];
			r3 = make_float4(N[i1*3], N[i1*3+1], N[i1*3+2], N[i2*3]);
			r4.x = N[i2*3+1]; r4.y = N[i2*3+2];
		}
	}
	r4.z = __uint_as_float(s); r4.w = __uint_as_float(local);
	float4 *o = verts + (size_t) i * 5;
	o[0] = r0; o[1] = r1; o[2] = r2; o[3] = r3; o[4] = r4;
}

// ---------------------------------------------------------------------------------------------------
// host driver
// ---------------------------------------------------------------------------------------------------
static inline unsigned div_up(size_t a, unsigned b) { return (unsigned) ((a + b - 1) / b); }

int build_bvh(mtsb_ctx *ctx, SceneDev &S, uint32_t primCount, int builder, BvhResult &out) {
	cudaStream_t st = ctx->stream;
	const int n = (int) primCount;
	if (n <= 0) MTSB_FAIL(MTSB_EINVAL, "build_bvh: scene has no primitives");
	cudaEvent_t ev0, ev1;
	MTSB_CUDA(cudaEventCreate(&ev0)); MTSB_CUDA(cudaEventCreate(&ev1));
	MTSB_CUDA(cudaEventRecord(ev0, st));

	DevBuf<float4> primLo, primHi, blo, bhi;
	DevBuf<uint32_t> sceneBox, vals, sortedPrim, flags;
	DevBuf<uint64_t> keys, sortedKeys;
	DevBuf<int2> children; DevBuf<int> parent;
	MTSB_CHECK(primLo.alloc(n)); MTSB_CHECK(primHi.alloc(n));
	MTSB_CHECK(sceneBox.alloc(8));
	MTSB_CHECK(keys.alloc(n)); MTSB_CHECK(sortedKeys.alloc(n)); MTSB_CHECK(vals.alloc(n)); MTSB_CHECK(sortedPrim.alloc(n));
	MTSB_CHECK(blo.alloc(2 * (size_t) n)); MTSB_CHECK(bhi.alloc(2 * (size_t) n));
	MTSB_CHECK(children.alloc(n)); MTSB_CHECK(parent.alloc(2 * (size_t) n)); MTSB_CHECK(flags.alloc(n));

	init_scene_box_kernel<<<1, 32, 0, st>>>(sceneBox.ptr);
	prim_bounds_kernel<<<div_up(n, 256), 256, 0, st>>>(S, primCount, primLo.ptr, primHi.ptr, sceneBox.ptr);
	morton_kernel<<<div_up(n, 256), 256, 0, st>>>(primCount, primLo.ptr, primHi.ptr, sceneBox.ptr, keys.ptr, vals.ptr);
	{
		size_t tmpBytes = 0;
		MTSB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmpBytes, keys.ptr, sortedKeys.ptr, vals.ptr, sortedPrim.ptr, n, 0, 63, st));
		DevBuf<uint8_t> tmp; MTSB_CHECK(tmp.alloc(tmpBytes));
		MTSB_CUDA(cub::DeviceRadixSort::SortPairs(tmp.ptr, tmpBytes, keys.ptr, sortedKeys.ptr, vals.ptr, sortedPrim.ptr, n, 0, 63, st));
		MTSB_CUDA(cudaStreamSynchronize(st));
	}

	int root = 0;
	if (n == 1) {
		MTSB_CUDA(cudaMemsetAsync(flags.ptr, 0, sizeof(uint32_t), st));
		refit_kernel<<<1, 32, 0, st>>>(n, sortedPrim.ptr, primLo.ptr, primHi.ptr, children.ptr, parent.ptr, blo.ptr, bhi.ptr, flags.ptr);
		root = 0;   // node id n-1+0 == 0
	} else if (builder == MTSB_BUILDER_PLOC) {
		DevBuf<int> clA, clB, nearest, merged, valid, scan, counter;
		MTSB_CHECK(clA.alloc(n)); MTSB_CHECK(clB.alloc(n)); MTSB_CHECK(nearest.alloc(n)); MTSB_CHECK(merged.alloc(n));
		MTSB_CHECK(valid.alloc(n)); MTSB_CHECK(scan.alloc(n)); MTSB_CHECK(counter.alloc(1));
		size_t tmpBytes = 0;
		MTSB_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmpBytes, valid.ptr, scan.ptr, n, st));
		DevBuf<uint8_t> tmp; MTSB_CHECK(tmp.alloc(tmpBytes));
		ploc_init_kernel<<<div_up(n, 256), 256, 0, st>>>(n, sortedPrim.ptr, primLo.ptr, primHi.ptr, blo.ptr, bhi.ptr, clA.ptr);
		int nodeCounter = n - 1;
		MTSB_CUDA(cudaMemcpyAsync(counter.ptr, &nodeCounter, sizeof(int), cudaMemcpyHostToDevice, st));
		DevBuf<int> mDev; MTSB_CHECK(mDev.alloc(2));          // ping-pong pair: round r reads slot r & 1, publishes into the other
		int m = n, round = 0; int *cur = clA.ptr, *nxt = clB.ptr;
		MTSB_CUDA(cudaMemcpyAsync(mDev.ptr, &m, sizeof(int), cudaMemcpyHostToDevice, st));
		while (m > PLOC_TAIL) {
			for (int b = 0; b < PLOC_BATCH; ++b, ++round) {
				const int *mIn = mDev.ptr + (round & 1); int *mOut = mDev.ptr + ((round + 1) & 1);
				ploc_nearest_kernel<<<div_up(m, 128), 128, 0, st>>>(mIn, cur, blo.ptr, bhi.ptr, nearest.ptr);
				ploc_merge_kernel<<<div_up(m, 256), 256, 0, st>>>(mIn, m, cur, nearest.ptr, blo.ptr, bhi.ptr, children.ptr, parent.ptr, counter.ptr, merged.ptr, valid.ptr);
				MTSB_CUDA(cub::DeviceScan::ExclusiveSum(tmp.ptr, tmpBytes, valid.ptr, scan.ptr, m, st));
				ploc_compact_kernel<<<div_up(m, 256), 256, 0, st>>>(mIn, m, merged.ptr, valid.ptr, scan.ptr, nxt, mOut);
				std::swap(cur, nxt);
			}
			MTSB_CUDA(cudaMemcpyAsync(&m, mDev.ptr + (round & 1), sizeof(int), cudaMemcpyDeviceToHost, st));
			MTSB_CUDA(cudaStreamSynchronize(st));
		}
		if (m > 1) ploc_tail_kernel<<<1, PLOC_TAIL, 0, st>>>(mDev.ptr + (round & 1), cur, blo.ptr, bhi.ptr, children.ptr, parent.ptr, counter.ptr);
		root = 0;    // last id handed out is 0
	} else {
		lbvh_kernel<<<div_up(n - 1, 256), 256, 0, st>>>(n, sortedKeys.ptr, children.ptr, parent.ptr);
		MTSB_CUDA(cudaMemsetAsync(flags.ptr, 0, (size_t) n * sizeof(uint32_t), st));
		refit_kernel<<<div_up(n, 256), 256, 0, st>>>(n, sortedPrim.ptr, primLo.ptr, primHi.ptr, children.ptr, parent.ptr, blo.ptr, bhi.ptr, flags.ptr);
		root = 0;
	}
	MTSB_CUDA(cudaGetLastError());

	// collapse
	const size_t maxNodes = (size_t) std::max(1, n);   // a wide node has >= 2 children unless the tree is a single leaf
	DevBuf<uint4> nodesTmp; MTSB_CHECK(nodesTmp.alloc(maxNodes * 5));
	DevBuf<uint32_t> slotPrim; MTSB_CHECK(slotPrim.alloc(n));
	DevBuf<CollapseItem> qA, qB; MTSB_CHECK(qA.alloc(maxNodes)); MTSB_CHECK(qB.alloc(maxNodes));
	DevBuf<uint32_t> counters; MTSB_CHECK(counters.alloc(8));   // 0 nodeCounter, 1 slotCounter, 2 | 5 queue lengths (ping-pong), 3 sah(float), 4 leaves, 6 depth
	uint32_t initCounters[8] = { 1, 0, 1, 0, 0, 0, 0, 0 };
	MTSB_CUDA(cudaMemcpyAsync(counters.ptr, initCounters, sizeof(initCounters), cudaMemcpyHostToDevice, st));
	// cost-driven collapse (default; MTSB_COLLAPSE_DP=0 falls back to the greedy largest-area collapse).  c_prim / c_node
	// = 0.7 measured best on configs 1-3 (0.4 ... 1.5 within 1 %): a triangle round costs fewer instructions than a node
	// step but runs at a quarter of its SIMD width
	DevBuf<float> dpCost; DevBuf<uint8_t> dpSplit; DevBuf<uint32_t> dpFlags;
	const uint8_t *dpSplitPtr = nullptr;
	{
		if (ctx->collapseDp && n > LEAF_MAX) {
			const float cNode = ctx->dpCNode, cPrim = ctx->dpCPrim;
			MTSB_CHECK(dpCost.alloc((size_t) (2 * n) * 7)); MTSB_CHECK(dpSplit.alloc((size_t) (2 * n) * 8)); MTSB_CHECK(dpFlags.alloc(n));
			MTSB_CUDA(cudaMemsetAsync(dpFlags.ptr, 0, (size_t) n * sizeof(uint32_t), st));
			const int minusOne = -1;     // PLOC never writes the root's parent
			MTSB_CUDA(cudaMemcpyAsync(parent.ptr + root, &minusOne, sizeof(int), cudaMemcpyHostToDevice, st));
			collapse_dp_kernel<<<div_up(n, 256), 256, 0, st>>>(n, children.ptr, parent.ptr, blo.ptr, bhi.ptr, dpFlags.ptr, dpCost.ptr, dpSplit.ptr, cNode, cPrim);
			dpSplitPtr = dpSplit.ptr;
		}
	}
	CollapseItem rootItem = { root, 0 };
	MTSB_CUDA(cudaMemcpyAsync(qA.ptr, &rootItem, sizeof(rootItem), cudaMemcpyHostToDevice, st));
	int rootIsLeaf = n <= LEAF_MAX ? 1 : 0;
	uint32_t live = 1; int level = 0; size_t bound = 1;
	int firstBatch = 4; while (firstBatch < 12 && ((size_t) 1 << (3 * (firstBatch - 3))) < (size_t) n) ++firstBatch;   // a balanced tree's depth + 3
	CollapseItem *qi = qA.ptr, *qo = qB.ptr;
	const int slot[2] = { 2, 5 };
	while (live) {
		for (int b = level == 0 ? firstBatch : 8; b > 0; --b, ++level) {
			uint32_t *inCount = counters.ptr + slot[level & 1], *outCount = counters.ptr + slot[(level + 1) & 1];
			MTSB_CUDA(cudaMemsetAsync(outCount, 0, sizeof(uint32_t), st));
			collapse_kernel<<<div_up(bound, 64), 64, 0, st>>>(n, rootIsLeaf, qi, inCount, qo, outCount, children.ptr, blo.ptr, bhi.ptr,
				sortedPrim.ptr, nodesTmp.ptr, counters.ptr + 0, counters.ptr + 1, slotPrim.ptr, (float *) (counters.ptr + 3), counters.ptr + 4, dpSplitPtr,
				counters.ptr + 6, (uint32_t) level);
			std::swap(qi, qo);
			bound = std::min(bound * 8, maxNodes);
		}
		MTSB_CUDA(cudaMemcpyAsync(&live, counters.ptr + slot[level & 1], sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
		MTSB_CUDA(cudaStreamSynchronize(st));
		if (level > 256) MTSB_FAIL(MTSB_EINVAL, "build_bvh: wide tree is more than 256 levels deep, the traversal stack holds 64 (degenerate geometry: many coincident primitives? the LBVH builder copes with them)");
	}
	uint32_t fin[8];
	MTSB_CUDA(cudaMemcpy(fin, counters.ptr, sizeof(fin), cudaMemcpyDeviceToHost));
	const uint32_t nodeCount = fin[0], slotCount = fin[1];
	const int depth = (int) fin[6];
	if (slotCount != primCount) MTSB_FAIL(MTSB_EINVAL, "build_bvh: %u primitives but %u leaf slots", primCount, slotCount);

	// the traversal stack holds one entry per level (MTSB_STACK_SHARED + MTSB_STACK_LOCAL); a deeper tree (thousands of
	// coincident centroids under duplicate Morton keys) would silently drop entries, so it is refused here instead
	if (depth > 64) MTSB_FAIL(MTSB_EINVAL, "build_bvh: wide tree is %d levels deep, the traversal stack holds 64 (degenerate geometry: many coincident primitives?)", depth);
	MTSB_CHECK(out.accel.alloc((size_t) nodeCount * 5 + (size_t) slotCount * 3));
	out.nodes = out.accel.ptr; out.tris = (float4 *) (out.accel.ptr + (size_t) nodeCount * 5);
	out.nodeBytes = (size_t) nodeCount * 5 * sizeof(uint4); out.triBytes = (size_t) slotCount * 3 * sizeof(float4);
	MTSB_CUDA(cudaMemcpyAsync(out.nodes, nodesTmp.ptr, out.nodeBytes, cudaMemcpyDeviceToDevice, st));
	triaccel_kernel<<<div_up(slotCount, 256), 256, 0, st>>>(S, slotCount, slotPrim.ptr, out.tris);
	MTSB_CHECK(out.verts.alloc((size_t) slotCount * 5));
	shading_record_kernel<<<div_up(slotCount, 256), 256, 0, st>>>(S, slotCount, slotPrim.ptr, out.verts.ptr);
	MTSB_CUDA(cudaGetLastError());

	uint32_t box[6];
	MTSB_CUDA(cudaMemcpyAsync(box, sceneBox.ptr, sizeof(box), cudaMemcpyDeviceToHost, st));
	MTSB_CUDA(cudaEventRecord(ev1, st));
	MTSB_CUDA(cudaStreamSynchronize(st));
	float ms = 0; cudaEventElapsedTime(&ms, ev0, ev1);
	cudaEventDestroy(ev0); cudaEventDestroy(ev1);

	auto unflip = [](uint32_t u) { uint32_t v = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u; float f; memcpy(&f, &v, 4); return f; };
	for (int a = 0; a < 3; ++a) { out.tightMin[a] = unflip(box[a]); out.tightMax[a] = unflip(box[3 + a]); }
	out.nodeCount = nodeCount; out.slotCount = slotCount; out.buildMs = ms; out.depth = depth;
	float sahSum; memcpy(&sahSum, &fin[3], 4);
	float ex = out.tightMax[0] - out.tightMin[0], ey = out.tightMax[1] - out.tightMin[1], ez = out.tightMax[2] - out.tightMin[2];
	float rootArea = 2.0f * (ex * ey + ey * ez + ez * ex);
	out.sahCost = rootArea > 0 ? sahSum / rootArea : 0;
	out.leafCount = fin[4];
	return 0;
}

} // namespace mtsb
