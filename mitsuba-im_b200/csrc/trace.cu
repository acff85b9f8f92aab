// trace.cu — batched ray queries of the C ABI: mtsb_trace_closest / mtsb_trace_any / mtsb_trace_surface.
//
// Batched equivalents of the three ShapeKDTree::rayIntersect overloads (src/librender/skdtree.cpp:112-226).
// Kernels are persistent: one grid sized to the SM count, each warp pulls batches of 32 rays from a global
// counter (one atomic per warp, broadcast by shuffle), rays and hits move as 16-byte vectors.
// Host-pointer calls are pipelined in 2M-ray chunks (smaller towards the end) over four streams — H2D, two alternating
// traversal streams, D2H — and a four-buffer ring so that PCIe traffic in both directions overlaps traversal.
#include "scene.cuh"
#include "trace.cuh"
#include "surface.cuh"
#include <algorithm>
#include <mutex>
#include <stdlib.h>
#include <string.h>

namespace mtsb {

enum { MODE_CLOSEST = 0, MODE_ANY = 1, MODE_SURFACE = 2, MODE_LIGHT = 3 };

// Ray source / record sink of the batched C-ABI queries: AoS mtsb_ray in, mtsb_hit / byte / mtsb_surface out.
template <int MODE>
struct BatchIO {
	static constexpr bool CAN_SUSPEND = false;
	SceneDev S;
	const float4 *__restrict__ rays;
	void *__restrict__ out;

	__device__ __forceinline__ void write_miss(unsigned long long i) const {
		if (MODE == MODE_ANY) ((uint8_t *) out)[i] = 0;
		else if (MODE == MODE_CLOSEST) __stcs((float4 *) out + i, make_float4(INFINITY, 0.0f, 0.0f, __uint_as_float(MTSB_NO_HIT)));
		else if (MODE == MODE_LIGHT) {
			((float4 *) out)[2 * i] = make_float4(INFINITY, __uint_as_float(MTSB_NO_HIT), 0.0f, 0.0f);
			((float4 *) out)[2 * i + 1] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
		} else {
			mtsb_surface rec; memset(&rec, 0, sizeof(rec));
			rec.t = INFINITY; rec.prim = MTSB_NO_HIT; rec.shape = MTSB_NO_HIT;
			((mtsb_surface *) out)[i] = rec;
		}
	}
	__device__ __forceinline__ bool load(unsigned long long i, V3 &o, V3 &d, float &tmin, float &tmax, V3 &rcp) const {
		// rays are read once and records written once: evict-first (ld/st.global.cs) so that the stream does not push the
		// acceleration structure out of L2
		const float4 r0 = __ldcs(rays + 2 * i), r1 = __ldcs(rays + 2 * i + 1);
		o = v3(r0.x, r0.y, r0.z); d = v3(r1.x, r1.y, r1.z);
		// the closest-hit overload floors the epsilon scale at Epsilon, the shadow-ray and light-record overloads do not
		if (clip_ray(S, o, d, r0.w, r1.w, MODE == MODE_CLOSEST || MODE == MODE_SURFACE, tmin, tmax, rcp)) return true;
		write_miss(i);
		return false;
	}
	__device__ __forceinline__ void store(unsigned long long i, bool found, const HitRec &hit, const V3 &o, const V3 &d) const {
		if (!found) { write_miss(i); return; }
		if (MODE == MODE_ANY) {
			((uint8_t *) out)[i] = 1;
		} else if (MODE == MODE_CLOSEST) {
			const float4 c = __ldg(S.tris + (size_t) hit.slot * 3 + 2);
			const uint32_t prim = S.shapes[__float_as_uint(c.z)].primOffset + __float_as_uint(c.w);
			__stcs((float4 *) out + i, make_float4(hit.t, hit.u, hit.v, __uint_as_float(prim)));
		} else if (MODE == MODE_LIGHT) {
			Surf sf; fill_surface(S, hit.slot, hit.t, hit.u, hit.v, o, d, sf);
			// a mesh without texture coordinates reports uv = 0 here (skdtree.cpp:182-184), not the barycentrics of its.uv
			const ShapeRec &sr = S.shapes[sf.shape];
			const bool noUv = sr.type == SHAPE_MESH && !(sr.flags & SHAPE_HAS_UV);
			((float4 *) out)[2 * i] = make_float4(hit.t, __uint_as_float(sf.shape), sf.ng.x, sf.ng.y);
			((float4 *) out)[2 * i + 1] = make_float4(sf.ng.z, noUv ? 0.0f : sf.u, noUv ? 0.0f : sf.v, 0.0f);
		} else {
			mtsb_surface rec; memset(&rec, 0, sizeof(rec));
			Surf sf; fill_surface(S, hit.slot, hit.t, hit.u, hit.v, o, d, sf);
			rec.t = hit.t; rec.u = hit.u; rec.v = hit.v;
			rec.shape = sf.shape; rec.prim_local = sf.prim; rec.prim = S.shapes[sf.shape].primOffset + sf.prim;
			rec.p[0] = sf.p.x; rec.p[1] = sf.p.y; rec.p[2] = sf.p.z;
			rec.ng[0] = sf.ng.x; rec.ng[1] = sf.ng.y; rec.ng[2] = sf.ng.z;
			rec.ns[0] = sf.ns.x; rec.ns[1] = sf.ns.y; rec.ns[2] = sf.ns.z;
			rec.s[0] = sf.s.x; rec.s[1] = sf.s.y; rec.s[2] = sf.s.z;
			rec.tt[0] = sf.t.x; rec.tt[1] = sf.t.y; rec.tt[2] = sf.t.z;
			rec.uv[0] = sf.u; rec.uv[1] = sf.v;
			((mtsb_surface *) out)[i] = rec;
		}
	}
};

template <int MODE, bool COUNT, bool CURVED>
__global__ void __launch_bounds__(MTSB_TRACE_BLOCK, MTSB_TRACE_BLOCKS_PER_SM)
trace_kernel(SceneDev S, const float4 *__restrict__ rays, unsigned long long n, void *__restrict__ out,
		unsigned long long *__restrict__ workCounter) {
	__shared__ uint2 shStack[MTSB_STACK_SHARED * MTSB_TRACE_BLOCK];
	BatchIO<MODE> io; io.S = S; io.rays = rays; io.out = out;
	trace_persistent<MODE == MODE_ANY, COUNT, CURVED>(S, io, n, workCounter, shStack + threadIdx.x);
	if (COUNT && threadIdx.x == 0 && blockIdx.x == 0 && S.counters) atomicAdd(S.counters + 2, n);
}

static size_t out_stride(int mode) {
	return mode == MODE_ANY ? 1 : (mode == MODE_CLOSEST ? sizeof(mtsb_hit) : (mode == MODE_LIGHT ? sizeof(mtsb_light_hit) : sizeof(mtsb_surface)));
}

// launches one persistent grid over n rays that already live on the device
int launch_trace(mtsb_scene *scene, int mode, const mtsb_ray *dRays, size_t n, void *dOut, unsigned long long *dWork,
		cudaStream_t st, bool count) {
	if (n == 0) return 0;
	mtsb_ctx *ctx = scene->ctx;
	l2_phase(scene, true);
	MTSB_CUDA(cudaMemsetAsync(dWork, 0, sizeof(unsigned long long), st));
	const int blocksPerSm = MTSB_TRACE_BLOCKS_PER_SM;
	unsigned grid = (unsigned) std::min<size_t>((size_t) ctx->smCount * blocksPerSm, (n + MTSB_TRACE_BLOCK - 1) / MTSB_TRACE_BLOCK);
	SceneDev S = scene->dev;
	S.counters = count ? scene->dCounters.ptr : nullptr;
	const float4 *r = (const float4 *) dRays;
	const unsigned long long nn = n;
	cudaLaunchAttribute attr[1];
	cudaLaunchConfig_t cfg; memset(&cfg, 0, sizeof(cfg));
	cfg.gridDim = dim3(grid); cfg.blockDim = dim3(MTSB_TRACE_BLOCK); cfg.stream = st;
	cfg.attrs = attr; cfg.numAttrs = l2_window(scene, attr);
	const bool curved = scene->curved;
	#define LAUNCH(M) do { if (count) MTSB_CUDA(cudaLaunchKernelEx(&cfg, trace_kernel<M, true, true>, S, r, nn, dOut, dWork)); \
		else if (curved) MTSB_CUDA(cudaLaunchKernelEx(&cfg, trace_kernel<M, false, true>, S, r, nn, dOut, dWork)); \
		else MTSB_CUDA(cudaLaunchKernelEx(&cfg, trace_kernel<M, false, false>, S, r, nn, dOut, dWork)); } while (0)
	if (mode == MODE_CLOSEST) LAUNCH(MODE_CLOSEST);
	else if (mode == MODE_ANY) LAUNCH(MODE_ANY);
	else if (mode == MODE_LIGHT) LAUNCH(MODE_LIGHT);
	else LAUNCH(MODE_SURFACE);
	#undef LAUNCH
	return 0;
}

#define MTSB_PIPE_BUFS 4
struct TracePipe {
	// device staging ring + streams for host-pointer calls: with four buffers the H2D of chunk c+3 never waits for the
	// D2H of chunk c, so the steady-state period is max(H2D, kernel, D2H) of one chunk, not their sum over two
	cudaStream_t sIn = nullptr, sOut = nullptr;
	cudaStream_t sK2 = nullptr;   // odd chunks traverse on their own stream (see trace_common)
	cudaEvent_t inDone[MTSB_PIPE_BUFS] = {}, kDone[MTSB_PIPE_BUFS] = {}, outDone[MTSB_PIPE_BUFS] = {};
	DevBuf<mtsb_ray> rays[MTSB_PIPE_BUFS];
	DevBuf<uint8_t> out[MTSB_PIPE_BUFS];
	DevBuf<unsigned long long> work;
	size_t chunk = 0, outBytes = 0;
	bool ready = false;
};

} // namespace mtsb

using namespace mtsb;

// one pipe per scene, created on first use and owned by the scene
static int get_pipe(mtsb_scene *scene, size_t chunk, size_t outBytesPerRay, TracePipe **outPipe) {
	std::lock_guard<std::mutex> lock(scene->pipeMutex);
	TracePipe *&p = scene->pipe;
	if (!p) p = new TracePipe();
	if (!p->ready) {
		MTSB_CUDA(cudaStreamCreateWithFlags(&p->sIn, cudaStreamNonBlocking));
		MTSB_CUDA(cudaStreamCreateWithFlags(&p->sOut, cudaStreamNonBlocking));
		MTSB_CUDA(cudaStreamCreateWithFlags(&p->sK2, cudaStreamNonBlocking));
		for (int i = 0; i < MTSB_PIPE_BUFS; ++i) {
			MTSB_CUDA(cudaEventCreateWithFlags(&p->inDone[i], cudaEventDisableTiming));
			MTSB_CUDA(cudaEventCreateWithFlags(&p->kDone[i], cudaEventDisableTiming));
			MTSB_CUDA(cudaEventCreateWithFlags(&p->outDone[i], cudaEventDisableTiming));
		}
		MTSB_CHECK(p->work.alloc(MTSB_PIPE_BUFS + 1));
		p->ready = true;
	}
	if (p->chunk < chunk) {
		for (int i = 0; i < MTSB_PIPE_BUFS; ++i) MTSB_CHECK(p->rays[i].alloc(chunk));
		p->chunk = chunk;
	}
	if (p->outBytes < chunk * outBytesPerRay) {
		for (int i = 0; i < MTSB_PIPE_BUFS; ++i) MTSB_CHECK(p->out[i].alloc(chunk * outBytesPerRay));
		p->outBytes = chunk * outBytesPerRay;
	}
	*outPipe = p;
	return 0;
}

void mtsb_release_pipe(mtsb_scene *scene) {
	std::lock_guard<std::mutex> lock(scene->pipeMutex);
	TracePipe *p = scene->pipe;
	if (!p) return;
	if (p->ready) {
		cudaStreamDestroy(p->sIn); cudaStreamDestroy(p->sOut); cudaStreamDestroy(p->sK2);
		for (int i = 0; i < MTSB_PIPE_BUFS; ++i) { cudaEventDestroy(p->inDone[i]); cudaEventDestroy(p->kDone[i]); cudaEventDestroy(p->outDone[i]); }
	}
	delete p;
	scene->pipe = nullptr;
}

static int trace_common(mtsb_scene *scene, int mode, const mtsb_ray *rays, size_t n, void *out, int onDevice, const char *who) {
	if (!scene || !scene->committed) MTSB_FAIL(MTSB_EINVAL, "%s: scene not committed", who);
	if (n == 0) return 0;
	if (!rays || !out) MTSB_FAIL(MTSB_EINVAL, "%s: NULL buffer", who);
	mtsb_ctx *ctx = scene->ctx;
	MTSB_CUDA(cudaSetDevice(ctx->device));
	const size_t ob = out_stride(mode);
	TracePipe *p = nullptr;
	if (onDevice) {
		MTSB_CHECK(get_pipe(scene, 1, ob, &p));
		return launch_trace(scene, mode, rays, n, out, p->work.ptr, ctx->stream, false);
	}
	// host path: chunks of 2M rays (64 MB in, 32 MB out; uniform 1M and smaller measured slower), H2D on sIn, kernel on ctx->stream, D2H on sOut
	const size_t hostChunk = ctx->hostChunk;
	const size_t chunk = std::min<size_t>(n, hostChunk);
	MTSB_CHECK(get_pipe(scene, chunk, ob, &p));
	// The H2D engine is the bottleneck and is busy from the first byte on; what the pipeline adds to the bare copy time is
	// the tail — kernel + D2H of the LAST chunk.  So the last two chunks' worth of rays goes in smaller pieces (1M, 1M,
	// 1M, 512K, 512K; measured +12 % end to end, pieces below 256K lose it again to launch overheads).
	const size_t minChunk = ctx->hostTaper;
	size_t off = 0;
	// Even and odd chunks traverse on two streams: a persistent grid ends with a tail in which the last warps finish
	// their longest rays on a mostly idle GPU (≈ 0.17 ms per launch, measured: 2M rays 1.0 ms, 512K rays 0.36 ms); with
	// the next chunk's kernel already resident on the other stream its CTAs fill the SMs as the tail frees them.
	// (The H2D copies stay on ONE stream: on two they would share the link and every chunk would arrive later.)
	MTSB_CUDA(cudaEventRecord(p->kDone[0], ctx->stream));        // work queued earlier on the context stream comes first
	MTSB_CUDA(cudaStreamWaitEvent(p->sK2, p->kDone[0], 0));
	for (size_t c = 0; off < n; ++c) {
		const int b = (int) (c % MTSB_PIPE_BUFS);
		cudaStream_t sIn = p->sIn, sK = (c & 1) ? p->sK2 : ctx->stream;
		const size_t remaining = n - off;
		size_t cnt = std::min(chunk, remaining);
		if (remaining <= chunk && remaining > minChunk) cnt = std::max(remaining / 2, minChunk);
		else if (remaining <= 2 * chunk && remaining > chunk && minChunk < chunk) cnt = std::max(chunk / 2, minChunk);
		// buffer b is free once the D2H of chunk c - MTSB_PIPE_BUFS is done
		if (c >= MTSB_PIPE_BUFS) MTSB_CUDA(cudaStreamWaitEvent(sIn, p->outDone[b], 0));
		MTSB_CUDA(cudaMemcpyAsync(p->rays[b].ptr, rays + off, cnt * sizeof(mtsb_ray), cudaMemcpyHostToDevice, sIn));
		MTSB_CUDA(cudaEventRecord(p->inDone[b], sIn));
		MTSB_CUDA(cudaStreamWaitEvent(sK, p->inDone[b], 0));
		if (c >= MTSB_PIPE_BUFS) MTSB_CUDA(cudaStreamWaitEvent(sK, p->outDone[b], 0));
		MTSB_CHECK(launch_trace(scene, mode, p->rays[b].ptr, cnt, p->out[b].ptr, p->work.ptr + b, sK, false));
		MTSB_CUDA(cudaEventRecord(p->kDone[b], sK));
		MTSB_CUDA(cudaStreamWaitEvent(p->sOut, p->kDone[b], 0));
		MTSB_CUDA(cudaMemcpyAsync((uint8_t *) out + off * ob, p->out[b].ptr, cnt * ob, cudaMemcpyDeviceToHost, p->sOut));
		MTSB_CUDA(cudaEventRecord(p->outDone[b], p->sOut));
		off += cnt;
	}
	MTSB_CUDA(cudaStreamSynchronize(p->sOut));
	MTSB_CUDA(cudaStreamSynchronize(p->sK2));
	MTSB_CUDA(cudaStreamSynchronize(ctx->stream));
	return 0;
}

extern "C" {

int mtsb_trace_closest(mtsb_scene *scene, const mtsb_ray *rays, size_t n, mtsb_hit *hits, int on_device) {
	return trace_common(scene, MODE_CLOSEST, rays, n, hits, on_device, "mtsb_trace_closest");
}
int mtsb_trace_any(mtsb_scene *scene, const mtsb_ray *rays, size_t n, uint8_t *out, int on_device) {
	return trace_common(scene, MODE_ANY, rays, n, out, on_device, "mtsb_trace_any");
}
int mtsb_trace_surface(mtsb_scene *scene, const mtsb_ray *rays, size_t n, mtsb_surface *out, int on_device) {
	return trace_common(scene, MODE_SURFACE, rays, n, out, on_device, "mtsb_trace_surface");
}
int mtsb_trace_light(mtsb_scene *scene, const mtsb_ray *rays, size_t n, mtsb_light_hit *out, int on_device) {
	return trace_common(scene, MODE_LIGHT, rays, n, out, on_device, "mtsb_trace_light");
}

// ---- allocation-free closest-hit query (no counterpart in the reference) ----
//
// Everything the ordinary entry point takes from cudaMalloc'ed memory — the persistent grid's work counter and, for host
// pointers, the device staging of the rays and of the hit records (TracePipe above) — lives in ONE file-scope device array
// of fixed size instead:
//
//     [0, 256)                                the work counter (8 bytes used)
//     [256, 256 + 32 * MAX)                   rays staged from a host pointer
//     [256 + 32 * MAX, 256 + 48 * MAX)        hit records on their way to a host pointer
//
// MAX = MTSB_NOALLOC_MAX_RAYS = 65536 rays a call (3 MiB + 256 bytes per device).  The array belongs to the module, so
// every call on a device — whatever the scene or the context — shares it, and it keeps the last call's contents: the
// counter is cleared in stream order before each launch, the staging areas are overwritten by the copy before they are
// read.  Two calls that overlap on the GPU (two contexts = two streams on one device, or two host threads) race on it;
// calls on one context are ordered by its stream.
#define MTSB_NOALLOC_BYTES (256 + (size_t) MTSB_NOALLOC_MAX_RAYS * (sizeof(mtsb_ray) + sizeof(mtsb_hit)))
static __device__ __align__(256) unsigned char g_noallocScratch[MTSB_NOALLOC_BYTES];

// device address of the array, looked up once per device with that device current
static std::mutex g_noallocMutex;
static unsigned char *g_noallocPtr[64];

static int noalloc_scratch(int device, unsigned char **out) {
	if (device < 0 || device >= 64) MTSB_FAIL(MTSB_EINVAL, "mtsb_trace_closest_noalloc: device %d outside the table (64 devices)", device);
	std::lock_guard<std::mutex> lock(g_noallocMutex);
	if (!g_noallocPtr[device]) {
		MTSB_CUDA(cudaSetDevice(device));
		unsigned char *ptr = nullptr;
		MTSB_CUDA(cudaGetSymbolAddress((void **) &ptr, g_noallocScratch));
		if (!ptr) MTSB_FAIL(MTSB_ECUDA, "mtsb_trace_closest_noalloc: cudaGetSymbolAddress returned a null address on device %d", device);
		g_noallocPtr[device] = ptr;
	}
	*out = g_noallocPtr[device];
	return 0;
}

int mtsb_trace_closest_noalloc(mtsb_scene *scene, const mtsb_ray *rays, size_t n, mtsb_hit *hits, int on_device) {
	if (!scene || !scene->committed) MTSB_FAIL(MTSB_EINVAL, "mtsb_trace_closest_noalloc: scene not committed");
	if (n > MTSB_NOALLOC_MAX_RAYS)
		MTSB_FAIL(MTSB_EINVAL, "mtsb_trace_closest_noalloc: %zu rays exceed the fixed scratch of %u rays a call — split the batch or use mtsb_trace_closest",
			n, (unsigned) MTSB_NOALLOC_MAX_RAYS);
	if (n == 0) return 0;
	if (!rays || !hits) MTSB_FAIL(MTSB_EINVAL, "mtsb_trace_closest_noalloc: NULL buffer");
	mtsb_ctx *ctx = scene->ctx;
	MTSB_CUDA(cudaSetDevice(ctx->device));
	unsigned char *scratch = nullptr;
	MTSB_CHECK(noalloc_scratch(ctx->device, &scratch));
	cudaStream_t st = ctx->stream;
	unsigned long long *dWork = (unsigned long long *) scratch;
	mtsb_ray *dRays = (mtsb_ray *) (scratch + 256);
	mtsb_hit *dHits = (mtsb_hit *) (scratch + 256 + (size_t) MTSB_NOALLOC_MAX_RAYS * sizeof(mtsb_ray));
	const mtsb_ray *src = rays;
	mtsb_hit *dst = hits;
	if (!on_device) {
		MTSB_CUDA(cudaMemcpyAsync(dRays, rays, n * sizeof(mtsb_ray), cudaMemcpyHostToDevice, st));
		src = dRays; dst = dHits;
	}
	MTSB_CUDA(cudaMemsetAsync(dWork, 0, sizeof(unsigned long long), st));      // the counter must start at zero
	const unsigned grid = (unsigned) std::min<size_t>((size_t) ctx->smCount * MTSB_TRACE_BLOCKS_PER_SM, (n + MTSB_TRACE_BLOCK - 1) / MTSB_TRACE_BLOCK);
	SceneDev S = scene->dev;
	S.counters = nullptr;
	const float4 *r = (const float4 *) src;
	const unsigned long long nn = n;
	void *o = dst;
	if (scene->curved) trace_kernel<MODE_CLOSEST, false, true><<<grid, MTSB_TRACE_BLOCK, 0, st>>>(S, r, nn, o, dWork);
	else trace_kernel<MODE_CLOSEST, false, false><<<grid, MTSB_TRACE_BLOCK, 0, st>>>(S, r, nn, o, dWork);
	MTSB_CUDA(cudaGetLastError());
	if (!on_device) {
		MTSB_CUDA(cudaMemcpyAsync(hits, dHits, n * sizeof(mtsb_hit), cudaMemcpyDeviceToHost, st));
		MTSB_CUDA(cudaStreamSynchronize(st));
	}
	return 0;
}

// counting variant: re-traces a device or host batch with node / primitive-test counters enabled
int mtsb_trace_count(mtsb_scene *scene, const mtsb_ray *rays, size_t n, int any_hit, int on_device, uint64_t out[4]) {
	if (!scene || !scene->committed) MTSB_FAIL(MTSB_EINVAL, "mtsb_trace_count: scene not committed");
	mtsb_ctx *ctx = scene->ctx;
	MTSB_CUDA(cudaSetDevice(ctx->device));
	cudaStream_t st = ctx->stream;
	DevBuf<mtsb_ray> dRays; DevBuf<uint8_t> dOut; DevBuf<unsigned long long> work;
	const int mode = any_hit ? MODE_ANY : MODE_CLOSEST;
	MTSB_CHECK(dOut.alloc(n * out_stride(mode))); MTSB_CHECK(work.alloc(1));
	const mtsb_ray *src = rays;
	if (!on_device) { MTSB_CHECK(dRays.upload(rays, n, st)); src = dRays.ptr; }
	MTSB_CUDA(cudaMemsetAsync(scene->dCounters.ptr, 0, 4 * sizeof(unsigned long long), st));
	MTSB_CHECK(launch_trace(scene, mode, src, n, dOut.ptr, work.ptr, st, true));
	unsigned long long h[4];
	MTSB_CUDA(cudaMemcpyAsync(h, scene->dCounters.ptr, sizeof(h), cudaMemcpyDeviceToHost, st));
	MTSB_CUDA(cudaStreamSynchronize(st));
	for (int i = 0; i < 4; ++i) out[i] = h[i];
	return 0;
}

int mtsb_trace_counters(mtsb_scene *scene, uint64_t out[4]) {
	if (!scene || !scene->committed) MTSB_FAIL(MTSB_EINVAL, "mtsb_trace_counters: scene not committed");
	unsigned long long h[4];
	MTSB_CUDA(cudaSetDevice(scene->ctx->device));
	MTSB_CUDA(cudaMemcpy(h, scene->dCounters.ptr, sizeof(h), cudaMemcpyDeviceToHost));
	for (int i = 0; i < 4; ++i) out[i] = h[i];
	return 0;
}

// pinned host memory for callers that want the pipelined copies to run at PCIe speed
void *mtsb_host_alloc(size_t bytes) {
	void *p = nullptr;
	if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); mtsb::set_error("cudaHostAlloc(%zu) failed", bytes); return nullptr; }
	return p;
}
void mtsb_host_free(void *p) { if (p) cudaFreeHost(p); }

} // extern "C"
