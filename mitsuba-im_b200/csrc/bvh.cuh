// bvh.cuh — host interface of the GPU BVH builder (bvh.cu).
#pragma once
#include "common.cuh"

namespace mtsb {

struct BvhResult {
	// nodes and TriAccel records share ONE allocation (nodes first) so that a single cudaAccessPolicyWindow can keep the
	// hot part of the acceleration structure resident in L2 while rays and hit records stream past it
	DevBuf<uint4>  accel;
	uint4  *nodes = nullptr;   // 5 x uint4 per wide node
	float4 *tris = nullptr;    // 3 x float4 per leaf slot (TriAccel record)
	size_t nodeBytes = 0, triBytes = 0;
	DevBuf<float4> verts;   // 5 x float4 per leaf slot: the triangle's de-indexed vertices + normals + (shape, prim), see surface.cuh
	uint32_t nodeCount = 0, slotCount = 0, leafCount = 0;
	int depth = 0;
	float buildMs = 0, sahCost = 0;
	float tightMin[3] = {0,0,0}, tightMax[3] = {0,0,0};   // GenericKDTree::m_tightAABB
};

// Builds the BVH over the primitives described by S (shapes/rects/positions/indices must be valid device
// pointers; S.nodes/S.tris are filled in by the caller from `out`).
int build_bvh(mtsb_ctx *ctx, SceneDev &S, uint32_t primCount, int builder, BvhResult &out);

} // namespace mtsb
