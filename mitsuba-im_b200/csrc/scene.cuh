// scene.cuh — host-side scene object behind the mtsb_scene handle.
#pragma once
#include "common.cuh"
#include "bvh.cuh"

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
};
