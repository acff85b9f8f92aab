// trace_all.cu — Scene::rayIntersectAll for batches (src/librender/scene.cpp:716-730, 769-792): the kd-tree query of trace.cu
// followed by a linear scan over the scene's special shapes.
#include "scene.cuh"
#include "trace.cuh"
#include <string.h>
#include <vector>

using namespace mtsb;

enum { MODE_CLOSEST_ = 0, MODE_ANY_ = 1 };

template <bool ANY>
__global__ void special_kernel(const RectRec *recs, const uint32_t *types, uint32_t count, const float4 *__restrict__ rays, size_t n, void *out) {
	const size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x;
	if (i >= n) return;
	const float4 r0 = rays[2 * i], r1 = rays[2 * i + 1];
	const V3 o = v3(r0.x, r0.y, r0.z), d = v3(r1.x, r1.y, r1.z);
	float mint = r0.w;
	if (mint == MTS_EPSILON) mint = __fmul_rn(mint, fmaxf(fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z)), MTS_EPSILON));
	if (ANY) {
		uint8_t *a = (uint8_t *) out;
		if (a[i]) return;                                        // "if (rayIntersect(ray)) return true"
		for (uint32_t k = 0; k < count; ++k) {
			float t, u, v;
			const bool h = types[k] == SHAPE_DISK ? disk_test(recs[k], o, d, mint, r1.w, t, u, v) : sphere_test(recs[k], o, d, mint, r1.w, t, u, v);
			if (h) { a[i] = 1; return; }
		}
	} else {
		mtsb_hit *hits = (mtsb_hit *) out;
		mtsb_hit hit = hits[i];
		float maxt = hit.prim != MTSB_NO_HIT ? hit.t : r1.w;     // :776
		bool changed = false;
		for (uint32_t k = 0; k < count; ++k) {
			float t, u, v;
			const bool h = types[k] == SHAPE_DISK ? disk_test(recs[k], o, d, mint, maxt, t, u, v) : sphere_test(recs[k], o, d, mint, maxt, t, u, v);
			if (h) { hit.t = t; hit.u = u; hit.v = v; hit.prim = MTSB_SPECIAL_PRIM + k; maxt = t; changed = true; }
		}
		if (changed) hits[i] = hit;
	}
}

// caller-added aperture disks, then the bounding sphere of every constant environment emitter (the order of
// Scene::initializeBidirectional: sensor shape, then the emitters' shapes)
static int upload_specials(mtsb_scene *scene, cudaStream_t st, uint32_t *countOut) {
	const size_t nUser = scene->specialRecs.size();
	std::vector<RectRec> recs(scene->specialRecs); std::vector<uint32_t> types(scene->specialTypes);
	for (const EmitterRec &e : scene->emitters) {
		if (e.type != MTSB_EMITTER_CONSTANT && e.type != MTSB_EMITTER_ENVMAP) continue;
		RectRec rr; memset(&rr, 0, sizeof(rr));
		rr.dpdu = e.sphereCenter; rr.dpdv = v3(e.sphereRadius, 0.0f, 0.0f); rr.pad[0] = 1u;      // constant.cpp:82-91: a flipped sphere
		recs.push_back(rr); types.push_back(SHAPE_SPHERE);
	}
	*countOut = (uint32_t) recs.size();
	if (scene->specialUploaded == nUser || recs.empty()) { scene->specialUploaded = nUser; return 0; }
	MTSB_CHECK(scene->dSpecialRecs.upload(recs, st));
	MTSB_CHECK(scene->dSpecialTypes.upload(types, st));
	MTSB_CUDA(cudaStreamSynchronize(st));
	scene->specialUploaded = nUser;
	return 0;
}

static int trace_all(mtsb_scene *scene, int mode, const mtsb_ray *rays, size_t n, void *out, int onDevice, const char *who) {
	if (!scene || !scene->committed) MTSB_FAIL(MTSB_EINVAL, "%s: scene not committed", who);
	if (n == 0) return 0;
	if (!rays || !out) MTSB_FAIL(MTSB_EINVAL, "%s: NULL buffer", who);
	mtsb_ctx *ctx = scene->ctx;
	MTSB_CUDA(cudaSetDevice(ctx->device));
	cudaStream_t st = ctx->stream;
	uint32_t count = 0;
	MTSB_CHECK(upload_specials(scene, st, &count));
	const size_t ob = mode == MODE_ANY_ ? 1 : sizeof(mtsb_hit);
	// the special-shape pass reads the rays and rewrites the records on the device: host buffers are staged whole
	DevBuf<mtsb_ray> dRays; DevBuf<uint8_t> dOut;
	const mtsb_ray *src = rays; void *dst = out;
	if (!onDevice) {
		MTSB_CHECK(dRays.upload(rays, n, st)); MTSB_CHECK(dOut.alloc(n * ob));
		src = dRays.ptr; dst = dOut.ptr;
	}
	// the kd-tree half: the ordinary batched query on device pointers (asynchronous on the context stream)
	if (mode == MODE_ANY_) MTSB_CHECK(mtsb_trace_any(scene, src, n, (uint8_t *) dst, 1));
	else MTSB_CHECK(mtsb_trace_closest(scene, src, n, (mtsb_hit *) dst, 1));
	if (count) {
		const unsigned grid = (unsigned) ((n + 255) / 256);
		if (mode == MODE_ANY_) special_kernel<true><<<grid, 256, 0, st>>>(scene->dSpecialRecs.ptr, scene->dSpecialTypes.ptr, count, (const float4 *) src, n, dst);
		else special_kernel<false><<<grid, 256, 0, st>>>(scene->dSpecialRecs.ptr, scene->dSpecialTypes.ptr, count, (const float4 *) src, n, dst);
		MTSB_CUDA(cudaGetLastError());
	}
	if (!onDevice) {
		MTSB_CUDA(cudaMemcpyAsync(out, dOut.ptr, n * ob, cudaMemcpyDeviceToHost, st));
		MTSB_CUDA(cudaStreamSynchronize(st));
	}
	return 0;
}

extern "C" {

int mtsb_trace_closest_all(mtsb_scene *scene, const mtsb_ray *rays, size_t n, mtsb_hit *hits, int on_device) {
	return trace_all(scene, MODE_CLOSEST_, rays, n, hits, on_device, "mtsb_trace_closest_all");
}
int mtsb_trace_any_all(mtsb_scene *scene, const mtsb_ray *rays, size_t n, uint8_t *out, int on_device) {
	return trace_all(scene, MODE_ANY_, rays, n, out, on_device, "mtsb_trace_any_all");
}

} // extern "C"
