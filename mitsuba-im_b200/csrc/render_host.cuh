// render_host.cuh — host orchestration of the wavefront passes + the rendering half of the C ABI (included by render.cu).
#pragma once

using namespace mtsb;

static const size_t MAX_PATHS_IN_FLIGHT = (size_t) 1 << 23;

// `direct` with emitterSamples / bsdfSamples other than 1/1 fans one camera path out into N shadow + M BSDF rays
// (render_direct.cuh); everything else runs one shadow ray and one continuation ray per path slot.
static inline bool fan_mode(const mtsb_render *r) {
	return r->integ.type == MTSB_INTEGRATOR_DIRECT && (r->integ.emitterSamples != 1 || r->integ.bsdfSamples != 1);
}

// SimpleVolumetricPathTracer (render_vol.cuh): its own shade / connect kernels, no tail kernel, no ray suspension
static inline bool vol_mode(const mtsb_render *r) { return r->integ.type == MTSB_INTEGRATOR_VOLPATH_SIMPLE || r->integ.type == MTSB_INTEGRATOR_VOLPATH; }
static inline bool volmis_mode(const mtsb_render *r) { return r->integ.type == MTSB_INTEGRATOR_VOLPATH; }

// camera paths per chunk: the fan-out of `direct` shares the same budget of rays in flight
static inline size_t max_paths(const mtsb_render *r) {
	if (!fan_mode(r)) return MAX_PATHS_IN_FLIGHT;
	return MAX_PATHS_IN_FLIGHT / (size_t) (1 + std::max(r->integ.emitterSamples, r->integ.bsdfSamples));
}

// warp iterations after which the extend stage parks the rays it is still working on (0: never).  Only the graph loop
// uses it: its WHILE node runs as many extra bounces as the parked rays need; the fixed launch sequences do not.
static inline uint32_t suspend_after(const mtsb_render *r) {
	const mtsb_ctx *ctx = r->scene->ctx;
	return (fan_mode(r) || vol_mode(r) || !ctx->useGraph || r->graphFailed) ? 0u : (uint32_t) ctx->suspendAfter;
}

static int ensure_capacity(mtsb_render *r, size_t paths) {
	if (paths <= r->capacity) return 0;
	const size_t cap = std::max<size_t>(paths, 1024);
	// connection-queue entries per camera slot: N emitter samples in fan mode; 2 for volpath_simple (the opacity query of
	// the sensor ray travels through the same queue as its emitter sample, render_vol.cuh)
	const size_t M = fan_mode(r) ? (size_t) r->integ.bsdfSamples : 0, N = fan_mode(r) ? (size_t) std::max(1, r->integ.emitterSamples) : (vol_mode(r) ? 2 : 1);
	const size_t wide = cap * (1 + M);            // camera slots + secondary slots (one per BSDF sample)
	MTSB_CHECK(r->bufF4.alloc(wide * 6 + cap + cap * N * (volmis_mode(r) ? 4 : 3)));
	MTSB_CHECK(r->bufRng.alloc(cap));
	MTSB_CHECK(r->bufQueue.alloc(cap + std::max(cap, cap * M)));
	if (suspend_after(r)) MTSB_CHECK(r->bufSusp.alloc(cap * MTSB_SUSP_WORDS));
	r->ps.susp = r->bufSusp.ptr;
	PathState &ps = r->ps;
	float4 *f = r->bufF4.ptr;
	ps.rayO = f; ps.rayD = f + wide; ps.hit = f + 2 * wide; ps.thr = f + 3 * wide; ps.prev = f + 4 * wide; ps.misc = f + 5 * wide;
	f += 6 * wide;
	ps.L = f; f += cap;
	ps.shO = f; ps.shD = f + cap * N; ps.shL = f + 2 * cap * N;
	ps.shX = volmis_mode(r) ? f + 3 * cap * N : nullptr;
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

struct LoopShape { unsigned traceGrid, shadeGrid, tailGrid; };

// which instantiation of the volumetric shade kernel a render runs (render_vol.cuh): MIS = `volpath`, PLAIN / LEAN by the scene
typedef void (*VolShadeKernel)(SceneDev, PathState, IntegratorDev, uint32_t);
static VolShadeKernel volshade_kernel_for(const mtsb_render *r, bool mis) {
	const mtsb_scene *scene = r->scene;
	const bool lean = scene->volLean && !r->integ.sobol;
	if (mis) return lean ? volshade_kernel<1, true, true> : (scene->plain ? volshade_kernel<1, true, false> : volshade_kernel<1, false, false>);
	return lean ? volshade_kernel<0, true, true> : (scene->plain ? volshade_kernel<0, true, false> : volshade_kernel<0, false, false>);
}

static void destroy_bounce_graph(mtsb_render *r) {
	if (r->bounceExec) cudaGraphExecDestroy(r->bounceExec);
	if (r->bounceGraph) cudaGraphDestroy(r->bounceGraph);
	r->bounceExec = nullptr; r->bounceGraph = nullptr; r->graphKey = nullptr; r->graphTail = -1;
}

// Adds one kernel node (optionally with the L2 access-policy window of the traversal kernels) behind `dep`.
static cudaError_t add_kernel(cudaGraph_t g, cudaGraphNode_t *node, cudaGraphNode_t dep, void *func, unsigned grid, unsigned block,
		void **args, mtsb_scene *windowOf) {
	cudaKernelNodeParams kp; memset(&kp, 0, sizeof(kp));
	kp.func = func; kp.gridDim = dim3(grid); kp.blockDim = dim3(block); kp.sharedMemBytes = 0; kp.kernelParams = args; kp.extra = nullptr;
	cudaError_t e = cudaGraphAddKernelNode(node, g, dep ? &dep : nullptr, dep ? 1 : 0, &kp);
	if (e != cudaSuccess || !windowOf) return e;
	cudaLaunchAttribute attr[1];
	if (l2_window(windowOf, attr) == 1) e = cudaGraphKernelNodeSetAttribute(*node, cudaLaunchAttributeAccessPolicyWindow, &attr[0].val);
	return e;
}

// The bounce loop of one chunk as ONE graph:   begin -> WHILE { extend -> shade -> connect -> advance } -> tail
// The WHILE condition is written by begin_kernel / advance_kernel on the device (render_tail.cuh).  All kernel arguments
// are constants of the render (path-state pointers, integrator, seed): per-chunk quantities travel through ps.counters.
static int build_bounce_graph(mtsb_render *r, const LoopShape &sh, uint32_t tailBelow, uint32_t maxBounces) {
	destroy_bounce_graph(r);
	mtsb_scene *scene = r->scene;
	SceneDev S = scene->dev; S.counters = nullptr;
	PathState ps = r->ps;
	IntegratorDev integ = r->integ;
	uint32_t seed = r->seed;
	cudaGraph_t g = nullptr;
	MTSB_CUDA(cudaGraphCreate(&g, 0));
	r->bounceGraph = g;
	LoopCtl ctl; ctl.tailBelow = tailBelow; ctl.maxBounces = maxBounces; ctl.inGraph = 1;
	MTSB_CUDA(cudaGraphConditionalHandleCreate(&ctl.cond, g, 0, cudaGraphCondAssignDefault));
	cudaGraphNode_t nBegin, nWhile, nTail, n1, n2, n3, n4;
	void *aCtl[] = { &ps, &ctl };
	MTSB_CUDA(add_kernel(g, &nBegin, nullptr, (void *) begin_kernel, 1, 1, aCtl, nullptr));
	cudaGraphNodeParams wp = { cudaGraphNodeTypeConditional };
	wp.conditional.handle = ctl.cond; wp.conditional.type = cudaGraphCondTypeWhile; wp.conditional.size = 1;
	MTSB_CUDA(cudaGraphAddNode(&nWhile, g, &nBegin, 1, &wp));
	cudaGraph_t body = wp.conditional.phGraph_out[0];
	uint32_t suspendAfter = suspend_after(r);
	void *aTrace[] = { &S, &ps };
	void *aExtend[] = { &S, &ps, &suspendAfter };
	void *aShade[] = { &S, &ps, &integ, &seed };
	MTSB_CUDA(add_kernel(body, &n1, nullptr, scene->curved ? (void *) extend_kernel<true> : (void *) extend_kernel<false>, sh.traceGrid, MTSB_TRACE_BLOCK, aExtend, scene));
	if (vol_mode(r)) {
		MTSB_CUDA(add_kernel(body, &n2, n1, (void *) volshade_kernel_for(r, volmis_mode(r)), sh.shadeGrid, RENDER_BLOCK, aShade, nullptr));
		MTSB_CUDA(add_kernel(body, &n3, n2, scene->curved ? (void *) connect_vol_kernel<true> : (void *) connect_vol_kernel<false>, sh.tailGrid, MTSB_TRACE_BLOCK, aTrace, scene));
	} else {
		MTSB_CUDA(add_kernel(body, &n2, n1, (scene->lean && !r->integ.sobol) ? (void *) shade_kernel<true> : (void *) shade_kernel<false>, sh.shadeGrid, RENDER_BLOCK, aShade, nullptr));
		MTSB_CUDA(add_kernel(body, &n3, n2, scene->curved ? (void *) connect_kernel<true> : (void *) connect_kernel<false>, sh.traceGrid, MTSB_TRACE_BLOCK, aTrace, scene));
	}
	MTSB_CUDA(add_kernel(body, &n4, n3, (void *) advance_kernel, 1, 1, aCtl, nullptr));
	if (tailBelow) MTSB_CUDA(add_kernel(g, &nTail, nWhile, scene->curved ? (void *) tail_kernel<true> : (void *) tail_kernel<false>, sh.tailGrid, MTSB_TRACE_BLOCK, aShade, scene));
	MTSB_CUDA(cudaGraphInstantiate(&r->bounceExec, g, 0));
	r->graphKey = r->bufF4.ptr; r->graphTail = (int) tailBelow;
	return 0;
}

// Runs generate -> bounces -> splat for one chunk of `count` path slots.  Nothing here waits for the device: the bounce
// loop is a graph whose WHILE node is steered from the device, and the tail kernel (render_tail.cuh) finishes whatever is
// alive when the wavefront stops paying — so the call returns as soon as four launches are queued (mtsb_render_pass is
// asynchronous, as include/mtsb200.h says).
static int run_chunk(mtsb_render *r, const GenParams &g, float *samplesOut, float *posOut = nullptr) {
	mtsb_scene *scene = r->scene;
	mtsb_ctx *ctx = scene->ctx;
	cudaStream_t st = r->stream;
	PathState &ps = r->ps;
	uint32_t *cnt = ps.counters;
	LoopShape sh;
	sh.traceGrid = (unsigned) ctx->smCount * MTSB_TRACE_BLOCKS_PER_SM;
	// sized to the machine and to the chunk the render was made for (r->capacity), not to this chunk: graph nodes are fixed
	sh.shadeGrid = blocks_for(r->capacity, RENDER_BLOCK);
	if (ctx->shadeWaves > 0) sh.shadeGrid = std::min(sh.shadeGrid, (unsigned) ctx->smCount * MTSB_SHADE_MIN_BLOCKS * (unsigned) ctx->shadeWaves);
	sh.tailGrid = (unsigned) ctx->smCount * MTSB_TAIL_MIN_BLOCKS;
	SceneDev S = scene->dev; S.counters = nullptr;
	l2_phase(scene, false);

	MTSB_CUDA(cudaMemsetAsync(cnt, 0, CTR_WORDS * sizeof(uint32_t), st));
	generate_kernel<<<blocks_for(g.count, RENDER_BLOCK), RENDER_BLOCK, 0, st>>>(g, r->cam, ps, ps.queueA);
	r->launches += 1;
	const int maxDepth = r->integ.maxDepth;
	const uint32_t tailBelow = vol_mode(r) ? 0u : (uint32_t) std::max(0, r->tailPaths);
	const uint32_t bounded = maxDepth > 0 ? (uint32_t) maxDepth : 0x7fffffffu;

	if (fan_mode(r)) {
		// MIDirectIntegrator with N emitter / M BSDF samples: a fixed sequence, no loop (render_direct.cuh)
		const int N = r->integ.emitterSamples, M = r->integ.bsdfSamples;
		DirectFan fan; fan.capacity = (uint32_t) r->capacity;
		fan.weightBSDF = 1.0f / (float) M; fan.weightLum = 1.0f / (float) N;                 // direct.cpp:128-137 (inf for a count of 0, never used)
		fan.fracBSDF = (float) M / (float) (N + M); fan.fracLum = (float) N / (float) (N + M);
		LoopCtl ctl; ctl.tailBelow = 0; ctl.maxBounces = 2; ctl.inGraph = 0; ctl.cond = 0;
		const unsigned fanGrid = std::min(blocks_for(g.count, RENDER_BLOCK), (unsigned) ctx->smCount * 16u);
		begin_kernel<<<1, 1, 0, st>>>(ps, ctl);
		MTSB_CHECK(launch_trav(scene, scene->curved ? extend_kernel<true> : extend_kernel<false>, sh.traceGrid, st, S, ps, 0u));
		direct_head_kernel<<<fanGrid, RENDER_BLOCK, 0, st>>>(S, ps, r->integ, r->seed, fan);
		MTSB_CHECK(launch_trav(scene, scene->curved ? connect_fan_kernel<true> : connect_fan_kernel<false>, sh.traceGrid, st, S, ps));
		if (M > 0) {
			fan_flip_kernel<<<1, 1, 0, st>>>(ps);
			MTSB_CHECK(launch_trav(scene, scene->curved ? extend_kernel<true> : extend_kernel<false>, sh.traceGrid, st, S, ps, 0u));
			direct_tail_kernel<<<fanGrid, RENDER_BLOCK, 0, st>>>(S, ps, r->integ, fan);
		}
		r->launches += M > 0 ? 7 : 4;
	} else if (ctx->useGraph && !r->graphFailed) {
		if (!r->bounceExec || r->graphKey != (const void *) r->bufF4.ptr || r->graphTail != (int) tailBelow) {
			// parked rays cost their paths extra trips round the loop; it ends when nothing is alive, whatever the bound
			const uint32_t loopBound = suspend_after(r) && bounded < 0x10000000u ? bounded * 4u + 16u : bounded;
			if (build_bounce_graph(r, sh, tailBelow, loopBound) < 0) {
				// a driver without conditional graph nodes: the same kernels as plain launches (still no host round trip)
				cudaGetLastError();
				destroy_bounce_graph(r);
				r->graphFailed = true;
			}
		}
	}
	if (fan_mode(r)) {
		// queued above
	} else if (r->bounceExec) {
		MTSB_CUDA(cudaGraphLaunch(r->bounceExec, st));
		r->launches += tailBelow ? 2 : 1;               // begin + tail; the kernels inside the WHILE node are counted on the device
	} else {
		// plain launches: every bounded bounce is enqueued (an unbounded integrator gets 16, the tail kernel takes the rest)
		const bool poll = !tailBelow && maxDepth < 0;   // debugging configuration only: no tail kernel AND unbounded depth
		LoopCtl ctl; ctl.tailBelow = tailBelow; ctl.inGraph = 0; ctl.cond = 0;
		ctl.maxBounces = (maxDepth > 0 || poll) ? bounded : 16u;
		begin_kernel<<<1, 1, 0, st>>>(ps, ctl);
		for (uint32_t bounce = 0; bounce < ctl.maxBounces; ++bounce) {
			MTSB_CHECK(launch_trav(scene, scene->curved ? extend_kernel<true> : extend_kernel<false>, sh.traceGrid, st, S, ps, 0u));
			if (vol_mode(r)) {
				volshade_kernel_for(r, volmis_mode(r))<<<sh.shadeGrid, RENDER_BLOCK, 0, st>>>(S, ps, r->integ, r->seed);
				MTSB_CHECK(launch_trav(scene, scene->curved ? connect_vol_kernel<true> : connect_vol_kernel<false>, sh.tailGrid, st, S, ps));
			} else {
				if (scene->lean && !r->integ.sobol) shade_kernel<true><<<sh.shadeGrid, RENDER_BLOCK, 0, st>>>(S, ps, r->integ, r->seed);
				else shade_kernel<false><<<sh.shadeGrid, RENDER_BLOCK, 0, st>>>(S, ps, r->integ, r->seed);
				MTSB_CHECK(launch_trav(scene, scene->curved ? connect_kernel<true> : connect_kernel<false>, sh.traceGrid, st, S, ps));
			}
			advance_kernel<<<1, 1, 0, st>>>(ps, ctl);
			if (poll && (bounce & 7) == 7) {
				uint32_t live[2] = { 0, 0 };
				MTSB_CUDA(cudaMemcpyAsync(live, cnt, sizeof(live), cudaMemcpyDeviceToHost, st));
				MTSB_CUDA(cudaStreamSynchronize(st));
				if (live[0] + live[1] == 0) break;
				if (bounce > 100000) MTSB_FAIL(MTSB_EINVAL, "render: path depth exceeded 100000 bounces");
			}
		}
		r->launches += 1;
		if (tailBelow) {
			MTSB_CHECK(launch_trav(scene, scene->curved ? tail_kernel<true> : tail_kernel<false>, sh.tailGrid, st, S, ps, r->integ, r->seed));
			r->launches += 1;
		}
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
	if (integrator->type != MTSB_INTEGRATOR_PATH && integrator->type != MTSB_INTEGRATOR_DIRECT && integrator->type != MTSB_INTEGRATOR_VOLPATH_SIMPLE &&
			integrator->type != MTSB_INTEGRATOR_VOLPATH)
		MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_render_create: integrator type %d is outside the GPU subset", integrator->type);
	if (integrator->type == MTSB_INTEGRATOR_DIRECT && (integrator->emitter_samples < 0 || integrator->bsdf_samples < 0 ||
			integrator->emitter_samples + integrator->bsdf_samples <= 0 || integrator->emitter_samples > 4096 || integrator->bsdf_samples > 4096))
		MTSB_FAIL(MTSB_EINVAL, "mtsb_render_create: direct integrator needs emitterSamples + bsdfSamples > 0 (direct.cpp:109)");
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
	ig.cameraMedium = -1; ig.sobol = nullptr;
	if (ig.type == MTSB_INTEGRATOR_DIRECT) { ig.maxDepth = 2; ig.rrDepth = 1 << 30; }   // direct(1,1) == one-bounce path
	if (ig.maxDepth == 0) { delete r; MTSB_FAIL(MTSB_EINVAL, "mtsb_render_create: maxDepth must be -1 or > 0"); }
	r->seed = (uint32_t) seed;
	const size_t filmFloats = (size_t) (f.width + 2 * f.border) * (f.height + 2 * f.border) * f.channels;
	int rc = r->dFilm.alloc(filmFloats);
	if (rc == 0) rc = r->bufCounters.alloc(CTR_WORDS);
	if (rc == 0) rc = r->bufWork.alloc(16 + 2 * SPLAT_STAT_STRIPES);
	if (rc < 0) { delete r; return rc; }
	f.data = r->dFilm.ptr;
	r->ps.counters = r->bufCounters.ptr;
	r->ps.work = r->bufWork.ptr; r->ps.stats = r->bufWork.ptr + 4;
	r->tailPaths = scene->ctx->tailPaths;
	r->stream = scene->ctx->stream;
	cudaStream_t st = r->stream;
	MTSB_CUDA(cudaMemsetAsync(r->dFilm.ptr, 0, filmFloats * sizeof(float), st));
	MTSB_CUDA(cudaMemsetAsync(r->bufWork.ptr, 0, (16 + 2 * SPLAT_STAT_STRIPES) * sizeof(unsigned long long), st));
	MTSB_CUDA(cudaStreamSynchronize(st));
	*out = r;
	return 0;
}

void mtsb_render_destroy(mtsb_render *r) {
	if (!r) return;
	cudaSetDevice(r->scene->ctx->device);
	cudaStreamSynchronize(r->stream);                 // passes are asynchronous: nothing of this render may still be queued
	if (r->ownStream) cudaStreamDestroy(r->stream);
	if (r->hostStage) cudaFreeHost(r->hostStage);
	if (r->flushDone) cudaEventDestroy(r->flushDone);
	if (r->previewDev) cudaFree(r->previewDev);
	if (r->previewHost) cudaFreeHost(r->previewHost);
	destroy_bounce_graph(r);
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
	return 0;
}

int mtsb_render_private_stream(mtsb_render *r) {
	if (!r) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_private_stream: NULL argument");
	if (r->ownStream) return 0;
	MTSB_CUDA(cudaSetDevice(r->scene->ctx->device));
	MTSB_CUDA(cudaStreamSynchronize(r->stream));
	cudaStream_t s = nullptr;
	MTSB_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
	r->stream = s; r->ownStream = true;
	return 0;
}

int mtsb_render_set_sobol(mtsb_render *r, const uint32_t *tables, size_t words) {
	if (!r || !tables) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_set_sobol: NULL argument");
	if (words != MTSB_SOBOL_TABLE_WORDS) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_set_sobol: %zu table words (%d expected)", words, (int) MTSB_SOBOL_TABLE_WORDS);
	if (fan_mode(r)) MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_render_set_sobol: sample arrays (direct with emitterSamples / bsdfSamples != 1) are not available with the sobol sampler");
	MTSB_CUDA(cudaSetDevice(r->scene->ctx->device));
	MTSB_CHECK(r->dSobol.upload(tables, words, r->stream));
	MTSB_CUDA(cudaStreamSynchronize(r->stream));
	// setFilmResolution(film->getCropSize(), bucketed) (sobol.cpp:147-158): math::roundToPowerOfTwo + log2i
	uint32_t res = 1, lg = 0;
	while (res < (uint32_t) std::max(r->film.cropW, r->film.cropH)) { res <<= 1; ++lg; }
	if (lg > 16) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_set_sobol: film larger than 65536 pixels a side");
	r->sobolLogRes = (int) lg;
	r->integ.sobol = r->dSobol.ptr;
	destroy_bounce_graph(r);                     // the integrator block is a kernel argument of the graph's nodes
	return 0;
}

static inline uint32_t ld_tag_host(const uint32_t *p) { const uintptr_t v = (uintptr_t) p; return v < 0x10000u ? (uint32_t) v : 0u; }

int mtsb_render_set_ldsampler(mtsb_render *r, int dimension, size_t sample_count) {
	if (!r) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_set_ldsampler: NULL argument");
	if (dimension < 1 || dimension > 15) MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_render_set_ldsampler: dimension %d (1..15: the counters ride in the path state)", dimension);
	if (sample_count == 0 || sample_count > ((size_t) 1 << 24)) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_set_ldsampler: sample count %zu", sample_count);
	if (fan_mode(r)) MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_render_set_ldsampler: sample arrays (direct with emitterSamples / bsdfSamples != 1) are not available with the ldsampler");
	uint32_t lg = 0;
	while (((size_t) 1 << lg) < sample_count) ++lg;          // math::roundToPowerOfTwo (ldsampler.cpp:83-87)
	r->dSobol.release();
	r->integ.sobol = (const uint32_t *) (uintptr_t) ((uint32_t) dimension | (lg << 8));      // shading.cuh ld_tag
	destroy_bounce_graph(r);
	return 0;
}

int mtsb_render_set_camera_medium(mtsb_render *r, int medium) {
	if (!r) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_set_camera_medium: NULL argument");
	if (medium < -1 || medium >= (int) r->scene->media.size()) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_set_camera_medium: medium %d unknown", medium);
	r->cameraMedium = medium;
	if (r->integ.cameraMedium != medium) { r->integ.cameraMedium = medium; destroy_bounce_graph(r); }     // a kernel argument of the graph's nodes
	return 0;
}

int mtsb_render_clear(mtsb_render *r) {
	if (!r) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_clear: NULL argument");
	MTSB_CUDA(cudaSetDevice(r->scene->ctx->device));
	cudaStream_t st = r->stream;
	MTSB_CUDA(cudaMemsetAsync(r->dFilm.ptr, 0, r->dFilm.bytes(), st));
	MTSB_CUDA(cudaMemsetAsync(r->ps.stats, 0, (12 + 2 * SPLAT_STAT_STRIPES) * sizeof(unsigned long long), st));
	r->launches = 0;
	return 0;
}

int mtsb_render_pass(mtsb_render *r, uint32_t spp_begin, uint32_t spp_count, uint32_t spp_total, const int32_t rect[4],
		uint32_t sample_stride, uint32_t sample_phase) {
	// spp_total only scales ray differentials in the reference (integrator.cpp:145-146); an `envmap` seen by sensor rays reads them
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
	g.sobol = r->integ.sobol; g.sobolLogRes = (uint32_t) r->sobolLogRes;
	g.diffScale = (r->scene->envmapEmitter >= 0 || !r->scene->texRecords.empty()) ? 1.0f / sqrtf((float) std::max(1u, spp_total)) : 0.0f;
	const unsigned long long total = (unsigned long long) g.pixPadded * spp_count;
	// chunks hold whole sample planes when possible so that a warp covers an 8x4 pixel tile of one sample index
	size_t chunk = std::min<unsigned long long>(total, max_paths(r));
	if (chunk > g.pixPadded) chunk = (chunk / g.pixPadded) * g.pixPadded;
	if (chunk > r->capacity) MTSB_CHECK(ensure_capacity(r, chunk));
	for (unsigned long long first = 0; first < total; first += chunk) {
		g.first = first; g.count = (uint32_t) std::min<unsigned long long>(chunk, total - first);
		MTSB_CHECK(run_chunk(r, g, nullptr));
	}
	return 0;
}

int mtsb_render_reserve(mtsb_render *r, uint32_t spp_per_pass) {
	if (!r) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_reserve: NULL argument");
	MTSB_CUDA(cudaSetDevice(r->scene->ctx->device));
	const unsigned long long pixPadded = (unsigned long long) ((r->film.cropW + 7) / 8) * ((r->film.cropH + 3) / 4) * 32;
	const unsigned long long total = pixPadded * std::max(1u, spp_per_pass);
	size_t chunk = (size_t) std::min<unsigned long long>(total, max_paths(r));
	if (chunk > pixPadded) chunk = (chunk / pixPadded) * pixPadded;
	MTSB_CHECK(ensure_capacity(r, chunk));
	// one 8x4-pixel pass: loads every kernel of the pipeline, instantiates the bounce graph for this capacity
	const int32_t rect[4] = { 0, 0, std::min(8, r->film.cropW), std::min(4, r->film.cropH) };
	MTSB_CHECK(mtsb_render_pass(r, 0, 1, 1, rect, 1, 0));
	MTSB_CHECK(mtsb_render_clear(r));
	MTSB_CUDA(cudaStreamSynchronize(r->stream));
	return 0;
}

int mtsb_render_sync(mtsb_render *r) {
	if (!r) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_sync: NULL argument");
	MTSB_CUDA(cudaSetDevice(r->scene->ctx->device));
	MTSB_CUDA(cudaStreamSynchronize(r->stream));
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
	cudaStream_t st = r->stream;
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
	cudaStream_t st = r->stream;
	const size_t n = r->dFilm.count;
	if (!r->hostStage) MTSB_CUDA(cudaMallocHost((void **) &r->hostStage, n * sizeof(float)));
	MTSB_CUDA(cudaMemcpyAsync(r->hostStage, r->dFilm.ptr, n * sizeof(float), cudaMemcpyDeviceToHost, st));
	MTSB_CUDA(cudaStreamSynchronize(st));
	*stage_out = r->hostStage;
	if (float_count_out) *float_count_out = n;
	return 0;
}

int mtsb_film_flush_begin(mtsb_render *r) {
	if (!r) MTSB_FAIL(MTSB_EINVAL, "mtsb_film_flush_begin: NULL argument");
	MTSB_CUDA(cudaSetDevice(r->scene->ctx->device));
	cudaStream_t st = r->stream;
	const size_t n = r->dFilm.count;
	if (!r->hostStage) MTSB_CUDA(cudaMallocHost((void **) &r->hostStage, n * sizeof(float)));
	if (!r->flushDone) MTSB_CUDA(cudaEventCreateWithFlags(&r->flushDone, cudaEventDisableTiming));
	// in stream order: [passes queued so far] -> film to the staging buffer -> film cleared -> [passes queued from now on]
	MTSB_CUDA(cudaMemcpyAsync(r->hostStage, r->dFilm.ptr, n * sizeof(float), cudaMemcpyDeviceToHost, st));
	MTSB_CUDA(cudaEventRecord(r->flushDone, st));
	MTSB_CUDA(cudaMemsetAsync(r->dFilm.ptr, 0, r->dFilm.bytes(), st));
	return 0;
}

int mtsb_film_flush_wait(mtsb_render *r, const float **stage_out, size_t *float_count_out) {
	if (!r || !stage_out || !r->flushDone) MTSB_FAIL(MTSB_EINVAL, "mtsb_film_flush_wait: NULL argument or no flush begun");
	MTSB_CUDA(cudaSetDevice(r->scene->ctx->device));
	MTSB_CUDA(cudaEventSynchronize(r->flushDone));       // the copy only — passes queued after flush_begin keep running
	*stage_out = r->hostStage;
	if (float_count_out) *float_count_out = r->dFilm.count;
	return 0;
}

int mtsb_preview_stack(mtsb_render *const *renders, const double *spp, int n, float *rgba_out, size_t pixel_count) {
	if (!renders || !spp || !rgba_out || n < 1 || n > 8) MTSB_FAIL(MTSB_EINVAL, "mtsb_preview_stack: bad argument (1..8 renders)");
	mtsb_render *r0 = renders[0];
	if (!r0) MTSB_FAIL(MTSB_EINVAL, "mtsb_preview_stack: NULL render");
	const size_t pixels = r0->dFilm.count / (size_t) r0->film.channels;
	if (pixel_count != pixels) MTSB_FAIL(MTSB_EINVAL, "mtsb_preview_stack: %zu pixels given, the film (with its border) has %zu", pixel_count, pixels);
	StackIn in; memset(&in, 0, sizeof(in));
	double total = 0;
	for (int k = 0; k < n; ++k) {
		if (!renders[k] || renders[k]->scene->ctx != r0->scene->ctx || renders[k]->dFilm.count != r0->dFilm.count)
			MTSB_FAIL(MTSB_EINVAL, "mtsb_preview_stack: renders must share one context and one film size");
		in.film[k] = renders[k]->dFilm.ptr; total += spp[k];
	}
	in.n = n; in.channels = r0->film.channels; in.scale = 1.0f / std::max((float) total, 1.0f);
	MTSB_CUDA(cudaSetDevice(r0->scene->ctx->device));
	cudaStream_t st = r0->stream;
	// every film must be complete before it is read: the renders may run on streams of their own
	for (int k = 1; k < n; ++k) if (renders[k]->stream != st) MTSB_CUDA(cudaStreamSynchronize(renders[k]->stream));
	if (r0->previewCount < pixels) {
		if (r0->previewDev) cudaFree(r0->previewDev);
		if (r0->previewHost) cudaFreeHost(r0->previewHost);
		MTSB_CUDA(cudaMalloc((void **) &r0->previewDev, pixels * sizeof(float4)));
		MTSB_CUDA(cudaMallocHost((void **) &r0->previewHost, pixels * sizeof(float4)));
		r0->previewCount = pixels;
	}
	preview_stack_kernel<<<blocks_for(pixels, 256), 256, 0, st>>>(in, pixels, r0->previewDev);
	MTSB_CUDA(cudaMemcpyAsync(r0->previewHost, r0->previewDev, pixels * sizeof(float4), cudaMemcpyDeviceToHost, st));
	MTSB_CUDA(cudaStreamSynchronize(st));
	memcpy(rgba_out, r0->previewHost, pixels * sizeof(float4));
	return 0;
}

int mtsb_render_stats(mtsb_render *r, uint64_t out[8]) {
	if (!r || !out) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_stats: NULL argument");
	MTSB_CUDA(cudaSetDevice(r->scene->ctx->device));
	unsigned long long h[12 + 2 * SPLAT_STAT_STRIPES];
	MTSB_CUDA(cudaMemcpyAsync(h, r->ps.stats, sizeof(h), cudaMemcpyDeviceToHost, r->stream));
	MTSB_CUDA(cudaStreamSynchronize(r->stream));
	for (int i = 0; i < SPLAT_STAT_STRIPES; ++i) { h[0] += h[12 + 2 * i]; h[3] += h[12 + 2 * i + 1]; }
	for (int i = 0; i < 8; ++i) out[i] = 0;
	out[0] = h[0]; out[1] = h[1]; out[2] = h[2]; out[3] = h[3]; out[4] = r->launches + h[5];
	return 0;
}

int mtsb_render_samples(mtsb_render *r, const uint32_t *samples, size_t n, uint32_t spp_total, float *out) {
	return mtsb_render_samples_ex(r, samples, n, spp_total, out, nullptr);
}

int mtsb_render_samples_ex(mtsb_render *r, const uint32_t *samples, size_t n, uint32_t spp_total, float *out, float *pos_out) {
	if (!r || !samples || !out) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_samples: NULL argument");
	if (n == 0) return 0;
	if (n > MAX_PATHS_IN_FLIGHT) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_samples: at most %zu samples per call", MAX_PATHS_IN_FLIGHT);
	MTSB_CUDA(cudaSetDevice(r->scene->ctx->device));
	cudaStream_t st = r->stream;
	MTSB_CHECK(ensure_capacity(r, n));
	DevBuf<uint32_t> dS; DevBuf<float> dO, dP;
	MTSB_CHECK(dS.upload(samples, n * 4, st)); MTSB_CHECK(dO.alloc(n * 4));
	if (pos_out) MTSB_CHECK(dP.alloc(n * 2));
	GenParams g; memset(&g, 0, sizeof(g));
	g.count = (uint32_t) n; g.samples = dS.ptr; g.seed = r->seed; g.pixPadded = 1; g.tilesX = 1; g.sppStride = 1;
	g.sobol = r->integ.sobol; g.sobolLogRes = (uint32_t) r->sobolLogRes;
	g.diffScale = (r->scene->envmapEmitter >= 0 || !r->scene->texRecords.empty()) ? 1.0f / sqrtf((float) std::max(1u, spp_total)) : 0.0f;
	MTSB_CHECK(run_chunk(r, g, dO.ptr, pos_out ? dP.ptr : nullptr));
	MTSB_CUDA(cudaMemcpyAsync(out, dO.ptr, n * 4 * sizeof(float), cudaMemcpyDeviceToHost, st));
	if (pos_out) MTSB_CUDA(cudaMemcpyAsync(pos_out, dP.ptr, n * 2 * sizeof(float), cudaMemcpyDeviceToHost, st));
	MTSB_CUDA(cudaStreamSynchronize(st));
	return 0;
}

// ---- Scene::evalTransmittance for a batch of segments: the connection walk of the volumetric integrators on its own
__global__ void transmittance_setup_kernel(const mtsb_segment *segs, uint32_t n, int maxInteractions, PathState ps) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i == 0) { ps.counters[CTR_SHADOW] = n; ps.work[1] = 0ull; }
	if (i >= n) return;
	const mtsb_segment s = segs[i];
	const uint32_t maxi = maxInteractions < 0 ? 0xffffu : (uint32_t) min(maxInteractions, 0xfffe);
	const uint32_t bits = VQ_PACK(s.flags & 3u, s.medium, maxi);
	ps.shO[i] = make_float4(s.p1[0], s.p1[1], s.p1[2], __uint_as_float(bits));
	ps.shD[i] = make_float4(s.p2[0], s.p2[1], s.p2[2], 0.0f);
	ps.shL[i] = make_float4(1.0f, 1.0f, 1.0f, __uint_as_float(i));
	ps.L[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
}

int mtsb_eval_transmittance(mtsb_scene *scene, const mtsb_segment *segments, size_t n, int max_interactions, float *out_rgb) {
	if (!scene || !scene->committed || !segments || !out_rgb) MTSB_FAIL(MTSB_EINVAL, "mtsb_eval_transmittance: NULL argument or scene not committed");
	if (n == 0) return 0;
	if (n > 0x7fffffffull) MTSB_FAIL(MTSB_EINVAL, "mtsb_eval_transmittance: at most 2^31 segments a call");
	for (size_t i = 0; i < n; ++i)
		if (segments[i].medium < -1 || segments[i].medium >= (int) scene->media.size()) MTSB_FAIL(MTSB_EINVAL, "mtsb_eval_transmittance: segment %zu names medium %d", i, segments[i].medium);
	mtsb_ctx *ctx = scene->ctx;
	MTSB_CUDA(cudaSetDevice(ctx->device));
	cudaStream_t st = ctx->stream;
	DevBuf<mtsb_segment> dS; DevBuf<float4> q; DevBuf<uint32_t> counters; DevBuf<unsigned long long> work;
	MTSB_CHECK(dS.upload(segments, n, st)); MTSB_CHECK(q.alloc(4 * n)); MTSB_CHECK(counters.alloc(CTR_WORDS)); MTSB_CHECK(work.alloc(8));
	MTSB_CUDA(cudaMemsetAsync(counters.ptr, 0, CTR_WORDS * sizeof(uint32_t), st));
	MTSB_CUDA(cudaMemsetAsync(work.ptr, 0, 8 * sizeof(unsigned long long), st));
	PathState ps; memset(&ps, 0, sizeof(ps));
	ps.shO = q.ptr; ps.shD = q.ptr + n; ps.shL = q.ptr + 2 * n; ps.L = q.ptr + 3 * n;
	ps.counters = counters.ptr; ps.work = work.ptr; ps.stats = work.ptr + 4;      // stats[2] = shadow rays lands in work[6]
	transmittance_setup_kernel<<<blocks_for(n, 256), 256, 0, st>>>(dS.ptr, (uint32_t) n, max_interactions, ps);
	l2_phase(scene, false);
	const unsigned grid = (unsigned) ctx->smCount * MTSB_TAIL_MIN_BLOCKS;
	MTSB_CHECK(launch_trav(scene, scene->curved ? connect_vol_kernel<true> : connect_vol_kernel<false>, grid, st, scene->dev, ps));
	MTSB_CUDA(cudaGetLastError());
	std::vector<float4> host;
	try { host.resize(n); } catch (const std::exception &) { MTSB_FAIL(MTSB_ENOMEM, "mtsb_eval_transmittance: %zu bytes of host memory", n * sizeof(float4)); }
	MTSB_CUDA(cudaMemcpyAsync(host.data(), ps.L, n * sizeof(float4), cudaMemcpyDeviceToHost, st));
	MTSB_CUDA(cudaStreamSynchronize(st));
	for (size_t i = 0; i < n; ++i) { out_rgb[3 * i] = host[i].x; out_rgb[3 * i + 1] = host[i].y; out_rgb[3 * i + 2] = host[i].z; }
	return 0;
}

int mtsb_render_sampler_dump(mtsb_render *r, uint32_t px, uint32_t py, uint32_t first, uint32_t n, uint32_t draws, float *out) {
	if (!r || !out) MTSB_FAIL(MTSB_EINVAL, "mtsb_render_sampler_dump: NULL argument");
	if (n == 0 || draws == 0) return 0;
	if (ld_tag_host(r->integ.sobol) == 0 && r->integ.sobol) MTSB_FAIL(MTSB_EUNSUPPORTED, "mtsb_render_sampler_dump: the sobol sampler's sequence index is per path (use mtsb_render_samples_ex)");
	MTSB_CUDA(cudaSetDevice(r->scene->ctx->device));
	cudaStream_t st = r->stream;
	DevBuf<float> dO; MTSB_CHECK(dO.alloc((size_t) n * draws * 3));
	sampler_dump_kernel<<<blocks_for(n, 128), 128, 0, st>>>(r->integ, r->seed, px, py, first, n, draws, dO.ptr);
	MTSB_CUDA(cudaGetLastError());
	MTSB_CUDA(cudaMemcpyAsync(out, dO.ptr, (size_t) n * draws * 3 * sizeof(float), cudaMemcpyDeviceToHost, st));
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
