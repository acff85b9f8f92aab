// scene.cuh — host-side scene object behind the mtsb_scene handle.
#pragma once
#include "common.cuh"
#include "bvh.cuh"
#include <mutex>
#include <algorithm>

namespace mtsb { struct TracePipe; }

struct mtsb_scene {
	mtsb_ctx *ctx = nullptr;
	bool committed = false;

	// host staging (filled by mtsb_scene_add_*)
	std::vector<mtsb::ShapeRec> shapes;
	std::vector<mtsb::RectRec> rects;
	std::vector<mtsb::BsdfRec> bsdfs;
	std::vector<mtsb::EmitterRec> emitters;
	std::vector<float> emitterWeights;
	std::vector<float> positions, normals, texcoords;
	std::vector<uint32_t> indices;
	std::vector<uint32_t> shapeOffsets;   // n+1 prefix sums of primitive counts
	uint32_t primCount = 0;

	// device copies
	mtsb::DevBuf<mtsb::ShapeRec> dShapes;
	mtsb::DevBuf<mtsb::RectRec> dRects;
	mtsb::DevBuf<mtsb::BsdfRec> dBsdfs;
	mtsb::DevBuf<mtsb::EmitterRec> dEmitters;
	mtsb::DevBuf<float> dPositions, dNormals, dTexcoords, dEmitterCdf, dTriAreaCdf;
	mtsb::DevBuf<uint32_t> dIndices, dTriAreaCdfOffset;
	mtsb::DevBuf<unsigned long long> dCounters;
	mtsb::BvhResult bvh;

	mtsb::SceneDev dev;                   // kernel parameter block
	float aabbMin[3], aabbMax[3];         // enlarged (ShapeKDTree::getAABB)

	// staging ring + streams of the host-pointer ray queries (trace.cu), created on first use
	std::mutex pipeMutex;
	mtsb::TracePipe *pipe = nullptr;
};

namespace mtsb {

// Launch attribute that keeps the acceleration structure resident in L2 while a traversal kernel runs: a window over
// the head of the `accel` allocation (nodes first, then TriAccel records) whose hits are marked persisting — as large a
// share of it as the device's persisting set-aside holds — while everything outside the window (rays, hit records, path
// state) takes the normal, evictable path.  Returns 0 attributes when the set-aside is unavailable or switched off.
inline int l2_window(const mtsb_scene *scene, cudaLaunchAttribute *attr) {
	const mtsb_ctx *ctx = scene->ctx;
	const size_t bytes = std::min(scene->bvh.accel.bytes(), ctx->l2WindowMax);
	if (ctx->l2Persist == 0 || bytes == 0) return 0;
	attr->id = cudaLaunchAttributeAccessPolicyWindow;
	cudaAccessPolicyWindow &w = attr->val.accessPolicyWindow;
	w.base_ptr = (void *) scene->bvh.accel.ptr;
	w.num_bytes = bytes;
	w.hitRatio = (float) std::min(1.0, (double) ctx->l2Persist / (double) bytes);
	w.hitProp = cudaAccessPropertyPersisting;
	w.missProp = cudaAccessPropertyStreaming;
	return 1;
}

} // namespace mtsb
