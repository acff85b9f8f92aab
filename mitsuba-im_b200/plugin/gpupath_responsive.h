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
// Every thread's render owns a CUDA stream (mtsb_render_private_stream), so the passes of the threads that share a GPU
// overlap on it.  Only lock: one per distinct target, held for the host-side add.
#pragma once

MTS_NAMESPACE_BEGIN

GPUResponsiveIntegrator::GPUResponsiveIntegrator(GPUPathTracer *parent) : ResponsiveIntegrator(parent->getProperties()), m_parent(parent),
	m_active(0), m_roundSamples(0) { m_roundTimer = new Timer(); }

GPUResponsiveIntegrator::~GPUResponsiveIntegrator() {
	if (m_parent->m_responsive == this) m_parent->m_responsive = NULL;
	for (size_t i = 0; i < m_renders.size(); ++i) if (m_renders[i]) mtsb_render_destroy(m_renders[i]);
}

/// Called before EVERY frame / restart (src/im-mts/shell.cpp:240-242, src/mitsuba/im_render.cpp:272): the device scene
/// is cached on scene identity, so this is a no-op after the first call — a camera move never rebuilds the BVH.
bool GPUResponsiveIntegrator::preprocess(const Scene *scene, const Sensor *, const Sampler *) {
	for (int i = 0; i < m_parent->m_gpus; ++i) m_parent->device(i, scene);
	return true;
}

/// The render of thread `threadIdx` (created on first use): own film on GPU threadIdx % gpus, own stream, path state
/// reserved and kernels loaded — all of it before the first timed pass.
mtsb_render *GPUResponsiveIntegrator::renderFor(int threadIdx, const Scene &scene, const Sensor &sensor, const Sampler &sampler, const ImageBlock &target) {
	GPUPathTracer *p = m_parent.get();
	GPUDevice &dev = p->device(threadIdx % p->m_gpus, &scene);
	{
		LockGuard lock(p->m_mutex);
		if ((int) m_renders.size() <= threadIdx) m_renders.resize(threadIdx + 1, (mtsb_render *) NULL);
		if (m_renders[threadIdx]) return m_renders[threadIdx];
	}
	mtsb_camera cam; describeCamera(&sensor, cam);
	mtsb_film_desc fd; describeFilm(target, scene.getFilm()->getReconstructionFilter(), scene.getFilm(), fd);
	// Integrator2 renders over its target: sample positions run over the target's own pixel grid and go through the
	// sensor as they are (ImageOrderIntegrator::render, integrator.cpp:338-390), whatever the film's crop window says
	fd.crop_size[0] = target.getSize().x; fd.crop_size[1] = target.getSize().y;
	mtsb_integrator_desc id; p->describeIntegrator(id);
	mtsb_render *r = NULL;
	MTSB_CALL(mtsb_render_create(dev.scene, &cam, &fd, &id, p->seedFor(&sampler), &r));
	MTSB_CALL(mtsb_render_private_stream(r));       // the threads' passes overlap on the device
	MTSB_CALL(mtsb_render_reserve(r, passSize(target)));
	LockGuard lock(p->m_mutex);
	m_renders[threadIdx] = r;
	return r;
}

/// Samples per pixel of one pass: about a million paths (1 spp from 1024^2 up) — abort / camera changes take effect
/// within one wave of the wavefront, while small frames do not pay the per-bounce launch floor once per sample.
uint32_t GPUResponsiveIntegrator::passSize(const ImageBlock &target) const {
	if (m_parent->m_sppPerPass > 0) return (uint32_t) m_parent->m_sppPerPass;
	const size_t pixelCount = (size_t) target.getSize().x * target.getSize().y;
	return (uint32_t) std::max<size_t>(1, ((size_t) 1 << 20) / std::max<size_t>(pixelCount, 1));
}

bool GPUResponsiveIntegrator::allocate(const Scene &scene, Sampler *const *samplers, ImageBlock *const *targets, int threadCount) {
	for (size_t i = 0; i < m_renders.size(); ++i) if (m_renders[i]) mtsb_render_destroy(m_renders[i]);
	m_renders.assign(threadCount, (mtsb_render *) NULL);
	// one lock per distinct target (the CLI hands one ImageBlock to every 8 threads, the GUI one to all)
	m_targetLocks.clear();
	for (int i = 0; i < threadCount; ++i)
		if (targets && targets[i] && !m_targetLocks.count(targets[i])) m_targetLocks[targets[i]] = new Mutex();
	// remembered for prepare(): the scene is not initialised yet when the CLI calls this (im_render.cpp:52-57 runs in the
	// process constructor, Scene::initialize() in Scene::preprocess), so nothing touches the device here
	m_allocTargets.assign(threadCount, (ImageBlock *) NULL); m_allocSamplers.assign(threadCount, (Sampler *) NULL);
	for (int i = 0; i < threadCount; ++i) {
		if (targets) m_allocTargets[i] = targets[i];
		if (samplers) m_allocSamplers[i] = samplers[i];
	}
	return true;
}

/// Called from GPUPathTracer::preprocess (i.e. from Scene::preprocess, after Scene::initialize and before the threads
/// enter render()): device-side allocation of every thread's render happens here, not inside the first timed round.
void GPUResponsiveIntegrator::prepare(const Scene &scene) {
	if (!scene.getSensor()) return;
	for (size_t i = 0; i < m_allocTargets.size(); ++i)
		if (m_allocTargets[i] && m_allocSamplers[i]) renderFor((int) i, scene, *scene.getSensor(), *m_allocSamplers[i], *m_allocTargets[i]);
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
	const uint32_t sampleCount = (uint32_t) sampler.getSampleCount();
	mtsb_camera cam; describeCamera(&sensor, cam);
	mtsb_render *r = renderFor(threadIdx, scene, sensor, sampler, target);
	MTSB_CALL(mtsb_render_set_camera(r, &cam));     // camera change = new kernel parameters, nothing else
	// wall time of one render() round over all threads, logged by the last thread out (the responsive CLI path prints no
	// "Render time" of its own: src/mitsuba/mitsuba.cpp:387-400)
	struct Round {
		GPUResponsiveIntegrator *self; const uint32_t *landed; size_t pixels;
		Round(GPUResponsiveIntegrator *s, const uint32_t *l, size_t px) : self(s), landed(l), pixels(px) {
			LockGuard lock(self->m_parent->m_mutex);
			if (self->m_active++ == 0) { self->m_roundTimer->reset(); self->m_roundSamples = 0; }
		}
		~Round() {
			LockGuard lock(self->m_parent->m_mutex);
			self->m_roundSamples += (double) *landed * (double) pixels;
			if (--self->m_active == 0) {
				const double sec = self->m_roundTimer->getMicroseconds() * 1e-6;
				SLog(EInfo, "gpupath: responsive round: %.0f samples in %.4f s, %.1f M samples/s", self->m_roundSamples, sec,
					self->m_roundSamples / std::max(sec, 1e-9) / 1e6);
			}
		}
	};
	float *dst = target.getBitmap()->getFloatData();
	const size_t n = (size_t) target.getBitmap()->getSize().x * target.getBitmap()->getSize().y * target.getChannelCount();
	Mutex *tlock = targetLock(&target);

	// this thread's share: sample indices threadIdx, threadIdx + threadCount, ...
	uint32_t mine = 0;
	for (uint32_t s = (uint32_t) threadIdx; s < sampleCount; s += (uint32_t) threadCount) ++mine;
	const uint32_t pass = passSize(target);
	uint32_t queued = 0, landed = 0;          // samples per pixel handed to the GPU / present in `target`
	int sinceSync = 0;
	Round round(this, &landed, (size_t) target.getSize().x * target.getSize().y);
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
		// queueing only: mtsb_render_pass does not wait for the device, and touches nothing but this thread's render
		MTSB_CALL(mtsb_render_pass(r, queued * threadCount, cnt, sampleCount, NULL, (uint32_t) threadCount, (uint32_t) threadIdx));
		queued += cnt;
		// keep at most a few passes in flight per thread so that abort / camera changes take effect within a few waves
		if (++sinceSync >= 4) { MTSB_CALL(mtsb_render_sync(r)); sinceSync = 0; }
	}
}

MTS_IMPLEMENT_CLASS(GPUResponsiveIntegrator, false, ResponsiveIntegrator)
MTS_IMPLEMENT_CLASS_S(GPUPathTracer, false, MonteCarloIntegrator)
MTS_EXPORT_PLUGIN(GPUPathTracer, "B200 wavefront path tracer (libmtsb200)");
MTS_NAMESPACE_END
