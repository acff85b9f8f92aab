// render_host.cuh — host orchestration of the wavefront passes + the rendering half of the C ABI (included by render.cu).
#pragma once

using namespace mtsb;

static const size_t MAX_PATHS_IN_FLIGHT = (size_t) 1 << 23;

static int ensure_capacity(mtsb_render *r, size_t paths) {
	if (paths <= r->capacity) return 0;
	size_t cap = std::max<size_t>(paths, 1024);
	MTSB_CHECK(r->bufF4.alloc(cap * 10));
	MTSB_CHECK(r->bufRng.alloc(cap));
	MTSB_CHECK(r->bufQueue.alloc(cap * 2));
	PathState &ps = r->ps;
	float4 *f = r->bufF4.ptr;
	ps.rayO = f; ps.rayD = f + cap; ps.hit = f + 2 * cap; ps.thr = f + 3 * cap; ps.L = f + 4 * cap; ps.prev = f + 5 * cap;
	ps.misc = f + 6 * cap; ps.shO = f + 7 * cap; ps.shD = f + 8 * cap; ps.shL = f + 9 * cap;
	ps.rng = r->bufRng.ptr;
	ps.queueA = r->bufQueue.ptr; ps.queueB = r->bufQueue.ptr + cap;
	r->capacity = cap;
	return 0;
}

static inline unsigned blocks_for(size_t n, unsigned b) { return (unsigned) std::max<size_t>(1, (n + b - 1) / b); }

// cudaLaunchKernelEx with the scene's L2 access-policy window attached (traversal kernels only)
template <typename... KArgs, typename... Args>
static int launch_trav(mtsb_scene *scene, void (*kernel)(KArgs...), unsigned grid, cudaStream_t st, Args... args) {
	cudaLaunchAttribute attr[1];
	cudaLaunchConfig_t cfg; memset(&cfg, 0, sizeof(cfg));
	cfg.gridDim = dim3(grid); cfg.blockDim = dim3(MTSB_TRACE_BLOCK); cfg.stream = st;
	cfg.attrs = attr; cfg.numAttrs = l2_window(scene, attr);
	MTSB_CUDA(cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...));
	return 0;
}

// Runs generate -> bounces -> splat for one chunk of `count` path slots.  Nothing here waits for the device: every
// kernel reads its counts from device memory, the wavefront stages run for a number of bounces chosen from the
// previous chunks' feedback word, and the tail kernel (render_tail.cuh) finishes whatever is alive after them — so the
// call returns as soon as the launches are queued (mtsb_render_pass is asynchronous, as include/mtsb200.h says).
static int run_chunk(mtsb_render *r, const GenParams &g, float *samplesOut, float *posOut = nullptr) {
	mtsb_scene *scene = r->scene;
	mtsb_ctx *ctx = scene->ctx;
	cudaStream_t st = ctx->stream;
	PathState &ps = r->ps;
	uint32_t *cnt = ps.counters;       // [0] queue A count, [1] queue B count, [2] shadow count, [4..7] tail hand-over
	const unsigned traceGrid = (unsigned) ctx->smCount * MTSB_TRACE_BLOCKS_PER_SM;
	const unsigned shadeGrid = std::min(blocks_for(g.count, RENDER_BLOCK), (unsigned) ctx->smCount * MTSB_SHADE_MIN_BLOCKS * 2u);
	SceneDev S = scene->dev; S.counters = nullptr;

	MTSB_CUDA(cudaMemsetAsync(cnt, 0, 8 * sizeof(uint32_t), st));
	generate_kernel<<<blocks_for(g.count, RENDER_BLOCK), RENDER_BLOCK, 0, st>>>(g, r->cam, ps, ps.queueA);
	r->launches += 1;
	uint32_t *qIn = ps.queueA, *qOut = ps.queueB, *cIn = cnt + 0, *cOut = cnt + 1;
	const int maxDepth = r->integ.maxDepth;
	const int bounded = maxDepth > 0 ? maxDepth : 1 << 30;
	const uint32_t tailBelow = (uint32_t) std::max(0, r->tailPaths);
	// wavefront bounces to enqueue: all of them without the tail kernel, else what the feedback of earlier chunks suggests
	if (tailBelow && r->hostFeedback) {
		const uint32_t fb = r->hostFeedback[0], live = r->hostFeedback[1];
		if (fb != 0xffffffffu) {
			const bool forced = (fb & 0x80000000u) != 0;
			const int at = (int) (fb & 0x7fffffffu);
			if (!forced) r->waveBounces = at;
			else if (live > 8u * tailBelow) r->waveBounces = at + 2;
			else if (live > 2u * tailBelow) r->waveBounces = at + 1;
			r->hostFeedback[0] = 0xffffffffu;
		}
	}
	const int waves = tailBelow ? std::min(bounded, std::max(0, r->waveBounces)) : bounded;
	uint32_t parity = 0;
	int bounce = 0;
	for (; bounce < waves; ++bounce) {
		prep_kernel<<<1, 1, 0, st>>>(ps, cIn, cOut, cnt + 2, parity, (uint32_t) bounce, tailBelow, 0u);
		MTSB_CHECK(launch_trav(scene, extend_kernel, traceGrid, st, S, ps, (const uint32_t *) qIn, (const uint32_t *) cIn));
		shade_kernel<<<shadeGrid, RENDER_BLOCK, 0, st>>>(S, ps, r->integ, r->seed, qIn, qOut, cIn, cOut, cnt + 2);
		MTSB_CHECK(launch_trav(scene, connect_kernel, traceGrid, st, S, ps, (const uint32_t *) (cnt + 2)));
		r->launches += 4;
		std::swap(qIn, qOut); std::swap(cIn, cOut); parity ^= 1u;
		// debugging configuration only (tail kernel switched off AND unbounded depth): the host has to look at the count
		if (!tailBelow && maxDepth < 0 && (bounce & 7) == 7) {
			uint32_t live = 0;
			MTSB_CUDA(cudaMemcpyAsync(&live, cIn, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
			MTSB_CUDA(cudaStreamSynchronize(st));
			if (live == 0) break;
			if (bounce > 100000) MTSB_FAIL(MTSB_EINVAL, "render: path depth exceeded 100000 bounces");
		}
	}
	if (tailBelow && bounce < bounded) {
		// hand whatever is alive to the tail kernel (a no-op if the hand-over already happened or nothing is left)
		prep_kernel<<<1, 1, 0, st>>>(ps, cIn, nullptr, cnt + 2, parity, (uint32_t) bounce, tailBelow, 1u);
		MTSB_CHECK(launch_trav(scene, tail_kernel, (unsigned) ctx->smCount * MTSB_TAIL_MIN_BLOCKS, st, S, ps, r->integ, r->seed));
		r->launches += 2;
	}
	splat_kernel<<<blocks_for(g.count, RENDER_BLOCK), RENDER_BLOCK, 0, st>>>(ps, r->film, g.count, samplesOut, posOut);
	r->launches += 1;
	MTSB_CUDA(cudaGetLastError());
	return 0;
}

static void fill_camera(CameraDev &c, const mtsb_camera *cam) {
	memcpy(c.s2c, cam->sample_to_camera, sizeof(c.s2c));
	memcpy(c.c2w, cam->camera_to_world, sizeof(c.c2w));
	c.nearClip = cam->near_clip; c.farClip = cam->far_clip;
	c.invResX = cam->inv_resolution[0]; c.invResY = cam->inv_resolution[1];
	c.apertureRadius = cam->aperture_radius; c.focusDistance = cam->focus_distance;
}

extern "C" {

int mtsb_render_create(mtsb_scene *scene, const mtsb_camera *camera, const mtsb_film_desc *film,
		const mtsb_integrator_desc *integrator, uint64_t seed, mtsb_render **out) {
	if (!scene || !scene->committed || !camera || !film || !integrator || !out)
		MTSB_FAIL(MTSB_EINVAL, "mtsb_render_create: NULL argument or scene not committed");
	if (film->width <= 0 || film->height <= 0 || (film->channels != 4 && film->channels != 5) || film->border < 0 || film->filter_radius <= 0)
		MTSB_FAIL(MTSB_EINVAL, "mtsb_render_create: bad film description (%dx%d, %d channels)", film->width, film->height, film->channels);
	if (integrator->type != MTSB_INTEGRATOR_PATH && integrator->type != MTSB_INTEGRATOR_DIRECT)
		MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_render_create: integrator type %d is outside the GPU subset", integrator->type);
	if (integrator->type == MTSB_INTEGRATOR_DIRECT && (integrator->emitter_samples != 1 || integrator->bsdf_samples != 1))
		MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_render_create: direct integrator supported with emitterSamples = bsdfSamples = 1 only");
	MTSB_CUDA(cudaSetDevice(scene->ctx->device));
	mtsb_render *r = new mtsb_render();
	r->scene = scene;
	fill_camera(r->cam, camera);
	FilmDev &f = r->film;
	f.width = film->width; f.height = film->height; f.offX = film->offset[0]; f.offY = film->offset[1];
	f.border = film->border; f.channels = film->channels; f.radius = film->filter_radius;
	f.scale = 31.0f / film->filter_radius;           // m_scaleFactor = MTS_FILTER_RESOLUTION / m_radius (rfilter.cpp:48)
	memcpy(f.table, film->filter_table, sizeof(f.table));
	f.cropW = film->crop_size[0] > 0 ? film->crop_size[0] : film->width;
	f.cropH = film->crop_size[1] > 0 ? film->crop_size[1] : film->height;
	IntegratorDev &ig = r->integ;
	ig.type = integrator->type; ig.maxDepth = integrator->max_depth; ig.rrDepth = integrator->rr_depth;
	ig.strictNormals = integrator->strict_normals; ig.hideEmitters = integrator->hide_emitters;
	ig.emitterSamples = integrator->emitter_samples; ig.bsdfSamples = integrator->bsdf_samples;
	if (ig.type == MTSB_INTEGRATOR_DIRECT) { ig.maxDepth = 2; ig.rrDepth = 1 << 30; }   // direct(1,1) == one-bounce path
	if (ig.maxDepth == 0) { delete r; MTSB_FAIL(MTSB_EINVAL, "mtsb_render_create: maxDepth must be -1 or > 0"); }
	r->seed = (uint32_t) seed;
	const size_t filmFloats = (size_t) (f.width + 2 * f.border) * (f.height + 2 * f.border) * f.channels;
	int rc = r->dFilm.alloc(filmFloats);
	if (rc == 0) rc = r->bufCounters.alloc(16);
	if (rc == 0) rc = r->bufWork.alloc(16 + 2 * SPLAT_STAT_STRIPES);
	if (rc < 0) { delete r; return rc; }
	f.data = r->dFilm.ptr;
	r->ps.counters = r->bufCounters.ptr;
	r->ps.work = r->bufWork.ptr; r->ps.stats = r->bufWork.ptr + 4;
	r->tailPaths = scene->ctx->tailPaths; r->waveBounces = 3;
	r->ps.feedback = nullptr;
	if (cudaHostAlloc((void **) &r->hostFeedback, 2 * sizeof(uint32_t), cudaHostAllocMapped) == cudaSuccess) {
		r->hostFeedback[0] = 0xffffffffu; r->hostFeedback[1] = 0;
		uint32_t *dp = nullptr;
		if (cudaHostGetDevicePointer((void **) &dp, r->hostFeedback, 0) == cudaSuccess) r->ps.feedback = dp;
	} else { cudaGetLastError(); r->hostFeedback = nullptr; }     // stats[12 ..): striped path / depth counters of splat_kernel
	cudaStream_t st = scene->ctx->stream;
	MTSB_CUDA(cudaMemsetAsync(r->dFilm.ptr, 0, filmFloats * sizeof(float), st));
	MTSB_CUDA(cudaMemsetAsync(r->bufWork.ptr, 0, (16 + 2 * SPLAT_STAT_STRIPES) * sizeof(unsigned long long), st));
	MTSB_CUDA(cudaStreamSynchronize(st));
	*out = r;
	return 0;
}

void mtsb_render_destroy(mtsb_render *r) {
	if (!r) return;
	cudaSetDevice(r->scene->ctx->device);
	cudaStreamSynchronize(r->scene->ctx->stream);     // passes are asynchronous: nothing of this render may still be queued
	if (r->hostStage) cudaFreeHost(r->hostStage);
	if (r->hostFeedback) cudaFreeHost(r->hostFeedback);
	delete r;
}

int mtsb_render_set_camera(mtsb_render *r, const mtsb_camera *camera) {
	if (!r || !camera) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_set_camera: NULL argument");
	fill_camera(r->cam, camera);       // kernel parameter only: no device traffic, no rebuild
	return 0;
}

int mtsb_render_set_tail(mtsb_render *r, int paths) {
	if (!r) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_set_tail: NULL argument");
	r->tailPaths = paths < 0 ? r->scene->ctx->tailPaths : paths;
	r->waveBounces = 3;
	if (r->hostFeedback) r->hostFeedback[0] = 0xffffffffu;
	return 0;
}

int mtsb_render_clear(mtsb_render *r) {
	if (!r) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_clear: NULL argument");
	MTSB_CUDA(cudaSetDevice(r->scene->ctx->device));
	cudaStream_t st = r->scene->ctx->stream;
	MTSB_CUDA(cudaMemsetAsync(r->dFilm.ptr, 0, r->dFilm.bytes(), st));
	MTSB_CUDA(cudaMemsetAsync(r->ps.stats, 0, (12 + 2 * SPLAT_STAT_STRIPES) * sizeof(unsigned long long), st));
	r->launches = 0;
	return 0;
}

int mtsb_render_pass(mtsb_render *r, uint32_t spp_begin, uint32_t spp_count, uint32_t spp_total, const int32_t rect[4],
		uint32_t sample_stride, uint32_t sample_phase) {
	(void) spp_total;   // only scales ray differentials in the reference (integrator.cpp:145-146); constant textures ignore them
	if (!r) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_pass: NULL argument");
	if (sample_stride == 0 || sample_phase >= sample_stride) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_pass: bad sample stride/phase");
	if (spp_count == 0) return 0;
	MTSB_CUDA(cudaSetDevice(r->scene->ctx->device));
	int x0 = 0, y0 = 0, x1 = r->film.cropW, y1 = r->film.cropH;
	if (rect) { x0 = rect[0]; y0 = rect[1]; x1 = rect[2]; y1 = rect[3]; }
	if (x0 < 0 || y0 < 0 || x1 > r->film.cropW || y1 > r->film.cropH || x1 <= x0 || y1 <= y0)
		MTSB_FAIL(MTSB_EINVAL, "mtsb_render_pass: rectangle [%d,%d)x[%d,%d) outside the %dx%d crop window", x0, x1, y0, y1, r->film.cropW, r->film.cropH);
	GenParams g; memset(&g, 0, sizeof(g));
	g.x0 = x0; g.y0 = y0; g.w = x1 - x0; g.h = y1 - y0;
	g.tilesX = (g.w + 7) / 8;
	g.pixPadded = g.tilesX * ((g.h + 3) / 4) * 32;
	g.sppBegin = spp_begin; g.sppStride = sample_stride; g.sppPhase = sample_phase; g.seed = r->seed;
	g.samples = nullptr;
	const unsigned long long total = (unsigned long long) g.pixPadded * spp_count;
	// chunks hold whole sample planes when possible so that a warp covers an 8x4 pixel tile of one sample index
	size_t chunk = std::min<unsigned long long>(total, MAX_PATHS_IN_FLIGHT);
	if (chunk > g.pixPadded) chunk = (chunk / g.pixPadded) * g.pixPadded;
	MTSB_CHECK(ensure_capacity(r, chunk));
	for (unsigned long long first = 0; first < total; first += chunk) {
		g.first = first; g.count = (uint32_t) std::min<unsigned long long>(chunk, total - first);
		MTSB_CHECK(run_chunk(r, g, nullptr));
	}
	return 0;
}

int mtsb_render_sync(mtsb_render *r) {
	if (!r) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_sync: NULL argument");
	MTSB_CUDA(cudaSetDevice(r->scene->ctx->device));
	MTSB_CUDA(cudaStreamSynchronize(r->scene->ctx->stream));
	return 0;
}

int mtsb_film_device_ptr(mtsb_render *r, void **ptr_out, size_t *float_count_out) {
	if (!r || !ptr_out || !float_count_out) MTSB_FAIL(MTSB_EINVAL, "mtsb_film_device_ptr: NULL argument");
	*ptr_out = r->dFilm.ptr; *float_count_out = r->dFilm.count;
	return 0;
}

int mtsb_film_to_host(mtsb_render *r, float *dst, size_t float_count, int accumulate) {
	if (!r || !dst) MTSB_FAIL(MTSB_EINVAL, "mtsb_film_to_host: NULL argument");
	if (float_count != r->dFilm.count) MTSB_FAIL(MTSB_EINVAL, "mtsb_film_to_host: %zu floats given, film has %zu", float_count, r->dFilm.count);
	MTSB_CUDA(cudaSetDevice(r->scene->ctx->device));
	cudaStream_t st = r->scene->ctx->stream;
	if (!accumulate) {
		MTSB_CUDA(cudaMemcpyAsync(dst, r->dFilm.ptr, float_count * sizeof(float), cudaMemcpyDeviceToHost, st));
		MTSB_CUDA(cudaStreamSynchronize(st));
		return 0;
	}
	// accumulate (ResponsiveIntegrator contract: the caller's target holds un-normalised sums of earlier passes): the film
	// lands in a page-locked staging buffer kept with the render — a fresh pageable vector per frame costs more in page
	// faults than the copy itself
	if (!r->hostStage) MTSB_CUDA(cudaMallocHost((void **) &r->hostStage, float_count * sizeof(float)));
	MTSB_CUDA(cudaMemcpyAsync(r->hostStage, r->dFilm.ptr, float_count * sizeof(float), cudaMemcpyDeviceToHost, st));
	MTSB_CUDA(cudaStreamSynchronize(st));
	const float *src = r->hostStage;
	for (size_t i = 0; i < float_count; ++i) dst[i] += src[i];
	return 0;
}

int mtsb_film_stage(mtsb_render *r, const float **stage_out, size_t *float_count_out) {
	if (!r || !stage_out) MTSB_FAIL(MTSB_EINVAL, "mtsb_film_stage: NULL argument");
	MTSB_CUDA(cudaSetDevice(r->scene->ctx->device));
	cudaStream_t st = r->scene->ctx->stream;
	const size_t n = r->dFilm.count;
	if (!r->hostStage) MTSB_CUDA(cudaMallocHost((void **) &r->hostStage, n * sizeof(float)));
	MTSB_CUDA(cudaMemcpyAsync(r->hostStage, r->dFilm.ptr, n * sizeof(float), cudaMemcpyDeviceToHost, st));
	MTSB_CUDA(cudaStreamSynchronize(st));
	*stage_out = r->hostStage;
	if (float_count_out) *float_count_out = n;
	return 0;
}

int mtsb_render_stats(mtsb_render *r, uint64_t out[8]) {
	if (!r || !out) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_stats: NULL argument");
	MTSB_CUDA(cudaSetDevice(r->scene->ctx->device));
	unsigned long long h[12 + 2 * SPLAT_STAT_STRIPES];
	MTSB_CUDA(cudaMemcpyAsync(h, r->ps.stats, sizeof(h), cudaMemcpyDeviceToHost, r->scene->ctx->stream));
	MTSB_CUDA(cudaStreamSynchronize(r->scene->ctx->stream));
	for (int i = 0; i < SPLAT_STAT_STRIPES; ++i) { h[0] += h[12 + 2 * i]; h[3] += h[12 + 2 * i + 1]; }
	for (int i = 0; i < 8; ++i) out[i] = 0;
	out[0] = h[0]; out[1] = h[1]; out[2] = h[2]; out[3] = h[3]; out[4] = r->launches;
	return 0;
}

int mtsb_render_samples(mtsb_render *r, const uint32_t *samples, size_t n, uint32_t spp_total, float *out) {
	return mtsb_render_samples_ex(r, samples, n, spp_total, out, nullptr);
}

int mtsb_render_samples_ex(mtsb_render *r, const uint32_t *samples, size_t n, uint32_t spp_total, float *out, float *pos_out) {
	(void) spp_total;
	if (!r || !samples || !out) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_samples: NULL argument");
	if (n == 0) return 0;
	if (n > MAX_PATHS_IN_FLIGHT) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_samples: at most %zu samples per call", MAX_PATHS_IN_FLIGHT);
	MTSB_CUDA(cudaSetDevice(r->scene->ctx->device));
	cudaStream_t st = r->scene->ctx->stream;
	MTSB_CHECK(ensure_capacity(r, n));
	DevBuf<uint32_t> dS; DevBuf<float> dO, dP;
	MTSB_CHECK(dS.upload(samples, n * 4, st)); MTSB_CHECK(dO.alloc(n * 4));
	if (pos_out) MTSB_CHECK(dP.alloc(n * 2));
	GenParams g; memset(&g, 0, sizeof(g));
	g.count = (uint32_t) n; g.samples = dS.ptr; g.seed = r->seed; g.pixPadded = 1; g.tilesX = 1; g.sppStride = 1;
	MTSB_CHECK(run_chunk(r, g, dO.ptr, pos_out ? dP.ptr : nullptr));
	MTSB_CUDA(cudaMemcpyAsync(out, dO.ptr, n * 4 * sizeof(float), cudaMemcpyDeviceToHost, st));
	if (pos_out) MTSB_CUDA(cudaMemcpyAsync(pos_out, dP.ptr, n * 2 * sizeof(float), cudaMemcpyDeviceToHost, st));
	MTSB_CUDA(cudaStreamSynchronize(st));
	return 0;
}

int mtsb_eval_bsdf_s3(mtsb_scene *scene, int bsdf, const float *in, const float *s3, size_t n, float *out) {
	if (!scene || !scene->committed || !in || !out) MTSB_FAIL(MTSB_EINVAL, "mtsb_eval_bsdf: NULL argument or scene not committed");
	if (bsdf < 0 || bsdf >= (int) scene->bsdfs.size()) MTSB_FAIL(MTSB_EINVAL, "mtsb_eval_bsdf: bsdf id %d unknown", bsdf);
	if (n == 0) return 0;
	MTSB_CUDA(cudaSetDevice(scene->ctx->device));
	cudaStream_t st = scene->ctx->stream;
	DevBuf<float> dI, dO, dS;
	MTSB_CHECK(dI.upload(in, n * 8, st)); MTSB_CHECK(dO.alloc(n * 16));
	if (s3) MTSB_CHECK(dS.upload(s3, n, st));
	eval_bsdf_kernel<<<blocks_for(n, 128), 128, 0, st>>>(scene->dev, bsdf, dI.ptr, s3 ? dS.ptr : nullptr, n, dO.ptr);
	MTSB_CUDA(cudaMemcpyAsync(out, dO.ptr, n * 16 * sizeof(float), cudaMemcpyDeviceToHost, st));
	MTSB_CUDA(cudaStreamSynchronize(st));
	return 0;
}

int mtsb_eval_bsdf(mtsb_scene *scene, int bsdf, const float *in, size_t n, float *out) {
	return mtsb_eval_bsdf_s3(scene, bsdf, in, nullptr, n, out);
}

int mtsb_eval_emitter(mtsb_scene *scene, const float *in, size_t n, float *out) {
	if (!scene || !scene->committed || !in || !out) MTSB_FAIL(MTSB_EINVAL, "mtsb_eval_emitter: NULL argument or scene not committed");
	if (scene->emitters.empty()) MTSB_FAIL(MTSB_EINVAL, "mtsb_eval_emitter: scene has no emitters");
	if (n == 0) return 0;
	MTSB_CUDA(cudaSetDevice(scene->ctx->device));
	cudaStream_t st = scene->ctx->stream;
	DevBuf<float> dI, dO;
	MTSB_CHECK(dI.upload(in, n * 8, st)); MTSB_CHECK(dO.alloc(n * 12));
	eval_emitter_kernel<<<blocks_for(n, 128), 128, 0, st>>>(scene->dev, dI.ptr, n, dO.ptr);
	MTSB_CUDA(cudaMemcpyAsync(out, dO.ptr, n * 12 * sizeof(float), cudaMemcpyDeviceToHost, st));
	MTSB_CUDA(cudaStreamSynchronize(st));
	return 0;
}

} // extern "C"
