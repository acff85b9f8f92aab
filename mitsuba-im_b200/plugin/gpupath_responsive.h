// gpupath_responsive.h — ResponsiveIntegrator (Integrator2) adapter + plugin exports (included by gpupath.cpp).
//
// Contract served (include/mitsuba/render/integrator2.h:49-100; callers src/mitsuba/im_render.cpp:103-266 and
// src/im-mts/scene.cpp): render() is entered concurrently by `threadCount` mitsuba::Threads that may share their
// ImageBlock; each accumulates un-normalised sums into it, reports its own sample count through progress() — the caller
// divides the summed targets by the summed counts — and polls abort / continu.
//
// How it maps to the GPU: thread t owns the sample indices s = t (mod threadCount) of every pixel and renders them on
// GPU t % gpus, one asynchronous mtsb_render_pass per pass into ITS OWN device film, which keeps the running sum of the
// passes not yet handed to the host.  A flush (D2H into page-locked staging + add into the caller's target under that
// target's lock + device clear) happens after the first pass — an interactive viewer sees the first frame at once — and
// then whenever the unflushed passes reach half of what has already landed (or after 250 ms), so an N-pass render costs
// O(log N) film read-backs instead of N.  progress() is only ever told about samples that HAVE landed in the target.
// Locks: one per GPU, held only while a pass is being queued (microseconds: nothing waits for the device there), and
// one per distinct target, held for the host-side add.  Threads on different GPUs never wait for one another.
#pragma once

MTS_NAMESPACE_BEGIN

GPUResponsiveIntegrator::GPUResponsiveIntegrator(GPUPathTracer *parent) : ResponsiveIntegrator(parent->getProperties()), m_parent(parent) { }

GPUResponsiveIntegrator::~GPUResponsiveIntegrator() {
	for (size_t i = 0; i < m_renders.size(); ++i) if (m_renders[i]) mtsb_render_destroy(m_renders[i]);
}

/// Called before EVERY frame / restart (src/im-mts/shell.cpp:240-242, src/mitsuba/im_render.cpp:272): the device scene
/// is cached on scene identity, so this is a no-op after the first call — a camera move never rebuilds the BVH.
bool GPUResponsiveIntegrator::preprocess(const Scene *scene, const Sensor *, const Sampler *) {
	for (int i = 0; i < m_parent->m_gpus; ++i) m_parent->device(i, scene);
	return true;
}

bool GPUResponsiveIntegrator::allocate(const Scene &, Sampler *const *, ImageBlock *const *targets, int threadCount) {
	for (size_t i = 0; i < m_renders.size(); ++i) if (m_renders[i]) mtsb_render_destroy(m_renders[i]);
	m_renders.assign(threadCount, (mtsb_render *) NULL);
	// one lock per distinct target (the CLI hands one ImageBlock to every 8 threads, the GUI one to all)
	m_targetLocks.clear();
	for (int i = 0; i < threadCount; ++i)
		if (targets && targets[i] && !m_targetLocks.count(targets[i])) m_targetLocks[targets[i]] = new Mutex();
	return true;
}

Mutex *GPUResponsiveIntegrator::targetLock(const ImageBlock *target) {
	LockGuard lock(m_parent->m_mutex);
	ref<Mutex> &m = m_targetLocks[target];
	if (!m) m = new Mutex();
	return m.get();
}

int GPUResponsiveIntegrator::render(const Scene &scene, const Sensor &sensor, Sampler &sampler, ImageBlock &target,
		Controls controls, int threadIdx, int threadCount) {
	GPUPathTracer *p = m_parent.get();
	const int gpu = threadIdx % p->m_gpus;
	GPUDevice &dev = p->device(gpu, &scene);
	const uint32_t sampleCount = (uint32_t) sampler.getSampleCount();
	mtsb_camera cam; describeCamera(&sensor, cam);
	{
		LockGuard lock(p->m_mutex);
		if ((int) m_renders.size() <= threadIdx) m_renders.resize(threadIdx + 1, (mtsb_render *) NULL);
		if (!m_renders[threadIdx]) {
			mtsb_film_desc fd; describeFilm(target, scene.getFilm()->getReconstructionFilter(), scene.getFilm(), fd);
			// Integrator2 renders over its target: sample positions run over the target's own pixel grid and go through the
			// sensor as they are (ImageOrderIntegrator::render, integrator.cpp:338-390), whatever the film's crop window says
			fd.crop_size[0] = target.getSize().x; fd.crop_size[1] = target.getSize().y;
			mtsb_integrator_desc id; p->describeIntegrator(id);
			MTSB_CALL(mtsb_render_create(dev.scene, &cam, &fd, &id, p->seedFor(&sampler), &m_renders[threadIdx]));
		}
	}
	mtsb_render *r = m_renders[threadIdx];
	MTSB_CALL(mtsb_render_set_camera(r, &cam));     // camera change = new kernel parameters, nothing else
	float *dst = target.getBitmap()->getFloatData();
	const size_t n = (size_t) target.getBitmap()->getSize().x * target.getBitmap()->getSize().y * target.getChannelCount();
	Mutex *tlock = targetLock(&target);

	// this thread's share: sample indices threadIdx, threadIdx + threadCount, ...
	uint32_t mine = 0;
	for (uint32_t s = (uint32_t) threadIdx; s < sampleCount; s += (uint32_t) threadCount) ++mine;
	const uint32_t pass = p->m_sppPerPass > 0 ? (uint32_t) p->m_sppPerPass : 1u;   // 1 spp per pass keeps abort latency at one wave
	uint32_t queued = 0, landed = 0;          // samples per pixel handed to the GPU / present in `target`
	ref<Timer> sinceFlush = new Timer();
	MTSB_CALL(mtsb_render_clear(r));          // the device film holds only what has not landed yet
	while (true) {
		if (controls.abort && *controls.abort) return -1;
		if (controls.continu && !*controls.continu) return -2;
		// flush policy: first pass at once, then geometrically (or every 250 ms), and always at the end of the share
		const uint32_t unflushed = queued - landed;
		if (unflushed && (queued >= mine || unflushed >= std::max(1u, landed / 2) || sinceFlush->getMilliseconds() >= 250)) {
			const float *src = NULL; size_t cnt = 0;
			MTSB_CALL(mtsb_film_stage(r, &src, &cnt));     // waits for this thread's passes only up to here
			if (cnt != n) Log(EError, "gpupath: target holds %zu floats, the device film %zu", n, cnt);
			if (threadIdx == 0) {
				uint64_t st[8]; mtsb_render_stats(r, st);
				char buf[160];
				snprintf(buf, sizeof(buf), "GPU wavefront: %.2f rays/sample, avg path length %.2f", st[0] ? (double) (st[1] + st[2]) / st[0] : 0.0,
					st[0] ? (double) st[3] / st[0] : 0.0);
				m_stats = buf;
			}
			MTSB_CALL(mtsb_render_clear(r));
			{
				LockGuard lock(tlock);
				for (size_t i = 0; i < n; ++i) dst[i] += src[i];
			}
			landed = queued;
			sinceFlush->reset();
		}
		if (controls.interrupt) {
			const int rc = controls.interrupt->progress(this, scene, sensor, sampler, target, (double) landed, controls, threadIdx, threadCount);
			if (rc != 0) return rc;
		}
		if (queued >= mine) return 0;                    // "sampler exhausted"
		const uint32_t cnt = std::min(pass, mine - queued);
		{
			LockGuard lock(dev.submit);                  // queueing only: mtsb_render_pass does not wait for the device
			MTSB_CALL(mtsb_render_pass(r, queued * threadCount, cnt, sampleCount, NULL, (uint32_t) threadCount, (uint32_t) threadIdx));
		}
		queued += cnt;
		// keep at most a few passes in flight per thread so that abort / camera changes take effect within one wave
		if (queued - landed >= 4 && (queued - landed) % 4 == 0) MTSB_CALL(mtsb_render_sync(r));
	}
}

MTS_IMPLEMENT_CLASS(GPUResponsiveIntegrator, false, ResponsiveIntegrator)
MTS_IMPLEMENT_CLASS_S(GPUPathTracer, false, MonteCarloIntegrator)
MTS_EXPORT_PLUGIN(GPUPathTracer, "B200 wavefront path tracer (libmtsb200)");
MTS_NAMESPACE_END
