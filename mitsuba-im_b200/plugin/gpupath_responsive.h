// gpupath_responsive.h — ResponsiveIntegrator (Integrator2) adapter + plugin exports (included by gpupath.cpp).
//
// Contract served (include/mitsuba/render/integrator2.h:49-100; callers src/mitsuba/im_render.cpp:103-266 and
// src/im-mts/scene.cpp): render() is entered concurrently by `threadCount` mitsuba::Threads that may share their
// ImageBlock; each accumulates un-normalised sums into its target, reports a sample count through progress() — the
// caller divides the SUM of the targets by the SUM of the counts (InteractiveSceneProcess::develop) — and polls
// abort / continu.
//
// How it maps to the GPU.  The host threads exist to keep CPU cores busy; a GPU wants ONE submitter and large passes.
// The threads are therefore grouped by device (thread t -> GPU t % G, G = min(gpus, threadCount)); the lowest thread of a
// group is its LEADER and renders the group's whole share — sample indices s = g (mod G) of every pixel — into one device
// film, in asynchronous passes that grow from one sample per pixel (an interactive viewer sees a first frame at once) to
// a full chunk of paths in flight.  The film keeps the running sum of the passes not yet handed to the host; a flush (D2H
// into page-locked staging + add into the leader's target + device clear) happens after the first pass and then whenever
// the unflushed passes reach half of what has already landed (or after 250 ms): O(log N) read-backs for N passes.
// The other threads of the group are FOLLOWERS: they render nothing, watch the group's landed count and report their
// equal share of it (landed / group size — progress() takes a double) so that the caller's per-thread accounting
// (sum of counts; `spp * threadCount >= sampleCount` as the stop test) holds exactly as if every thread had rendered.
// Measured: 16 threads each driving 1/16 of a 512x512x64 frame through a stream of their own took 44 ms for 11 ms of
// device work (every small pass pays the per-bounce latency floor of the wavefront); one leader runs at the classic
// path's speed.
#pragma once
#include <unistd.h>

MTS_NAMESPACE_BEGIN

static inline double self_now() {      // milliseconds on a process-wide clock (timing log only)
	struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts);
	return (double) (ts.tv_sec % 1000) * 1e3 + ts.tv_nsec * 1e-6;
}

GPUResponsiveIntegrator::GPUResponsiveIntegrator(GPUPathTracer *parent) : ResponsiveIntegrator(parent->getProperties()), m_parent(parent),
	m_active(0), m_roundSamples(0) { m_roundTimer = new Timer(); m_addLock = new Mutex(); m_groups.resize(std::max(1, parent->m_gpus)); }

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

/// The render of device group `gpu` (created on first use): own film, own stream, path state reserved and kernels
/// loaded — all of it before the first timed pass.
mtsb_render *GPUResponsiveIntegrator::renderFor(int gpu, const Scene &scene, const Sensor &sensor, const Sampler &sampler, const ImageBlock &target) {
	GPUPathTracer *p = m_parent.get();
	GPUDevice &dev = p->device(gpu, &scene);
	{
		LockGuard lock(p->m_mutex);
		if ((int) m_renders.size() <= gpu) m_renders.resize(gpu + 1, (mtsb_render *) NULL);
		if (m_renders[gpu]) return m_renders[gpu];
	}
	mtsb_camera cam; describeCamera(&sensor, cam);
	mtsb_film_desc fd; describeFilm(target, scene.getFilm()->getReconstructionFilter(), scene.getFilm(), fd);
	// Integrator2 renders over its target: sample positions run over the target's own pixel grid and go through the
	// sensor as they are (ImageOrderIntegrator::render, integrator.cpp:338-390), whatever the film's crop window says
	fd.crop_size[0] = target.getSize().x; fd.crop_size[1] = target.getSize().y;
	mtsb_integrator_desc id; p->describeIntegrator(id);
	mtsb_render *r = NULL;
	MTSB_CALL(mtsb_render_create(dev.scene, &cam, &fd, &id, p->seedFor(&sampler), &r));
	MTSB_CALL(mtsb_render_set_camera_medium(r, mediumIndex(collectMedia(&scene), sensor.getMedium())));
	configureDeviceSampler(r, &sampler);
	MTSB_CALL(mtsb_render_private_stream(r));
	MTSB_CALL(mtsb_render_reserve(r, maxPass(target)));
	LockGuard lock(p->m_mutex);
	m_renders[gpu] = r;
	return r;
}

/// Largest pass, in samples per pixel: one full chunk of paths in flight (8 M) — or the `sppPerPass` property.
uint32_t GPUResponsiveIntegrator::maxPass(const ImageBlock &target) const {
	if (m_parent->m_sppPerPass > 0) return (uint32_t) m_parent->m_sppPerPass;
	const size_t pixelCount = (size_t) target.getSize().x * target.getSize().y;
	return (uint32_t) std::max<size_t>(1, ((size_t) 1 << 23) / std::max<size_t>(pixelCount, 1));
}

bool GPUResponsiveIntegrator::allocate(const Scene &, Sampler *const *samplers, ImageBlock *const *targets, int threadCount) {
	for (size_t i = 0; i < m_renders.size(); ++i) if (m_renders[i]) mtsb_render_destroy(m_renders[i]);
	m_renders.assign(m_parent->m_gpus, (mtsb_render *) NULL);
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
/// enter render()): device-side allocation of every group's render happens here, not inside the first timed round.
void GPUResponsiveIntegrator::prepare(const Scene &scene) {
	if (!scene.getSensor()) return;
	const int G = std::min<int>(m_parent->m_gpus, (int) m_allocTargets.size());
	for (int g = 0; g < G; ++g)        // the leader of group g is thread g
		if (m_allocTargets[g] && m_allocSamplers[g]) renderFor(g, scene, *scene.getSensor(), *m_allocSamplers[g], *m_allocTargets[g]);
}

int GPUResponsiveIntegrator::render(const Scene &scene, const Sensor &sensor, Sampler &sampler, ImageBlock &target,
		Controls controls, int threadIdx, int threadCount) {
	GPUPathTracer *p = m_parent.get();
	const int G = std::min(p->m_gpus, threadCount), gpu = threadIdx % G;
	const bool leader = threadIdx == gpu;
	int groupSize = 0;
	for (int t = gpu; t < threadCount; t += G) ++groupSize;
	Group &grp = m_groups[gpu];
	const uint32_t sampleCount = (uint32_t) sampler.getSampleCount();
	const size_t pixels = (size_t) target.getSize().x * target.getSize().y;
	mtsb_render *r = NULL;
	if (leader) {
		mtsb_camera cam; describeCamera(&sensor, cam);
		r = renderFor(gpu, scene, sensor, sampler, target);
		MTSB_CALL(mtsb_render_set_camera(r, &cam));     // camera change = new kernel parameters, nothing else
	}

	// One render() round over all threads: group state is reset by whoever enters first, the wall time of the round is
	// logged by the last thread out (the responsive CLI path prints no "Render time": src/mitsuba/mitsuba.cpp:387-400).
	struct Round {
		GPUResponsiveIntegrator *self; Group *grp; const uint32_t *landed; size_t pixels;
		Round(GPUResponsiveIntegrator *s, Group *g, int groupSize, const uint32_t *l, size_t px) : self(s), grp(g), landed(l), pixels(px) {
			LockGuard lock(self->m_parent->m_mutex);
			if (self->m_active++ == 0) { self->m_roundTimer->reset(); self->m_roundSamples = 0; }
			// every thread of the group enters every round exactly once (the caller joins them all before the next one), so
			// the first entrant is recognised by count — not by "nobody is inside", which a fast leader would defeat
			if (grp->entered == grp->roundSize) grp->entered = 0;
			if (grp->entered == 0) { grp->landed = 0; grp->done = false; grp->rc = 0; grp->roundSize = groupSize; }
			++grp->entered;
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
	uint32_t queued = 0, landed = 0;          // samples per pixel handed to the GPU / present in `target` (leader only)
	Round round(this, &grp, groupSize, &landed, pixels);

	if (!leader) {
		// ---- follower: report an equal share of what the group's leader has landed
		double seen = -1;
		while (true) {
			if (controls.abort && *controls.abort) return -1;
			if (controls.continu && !*controls.continu) return -2;
			double now; bool done; int rc;
			{ LockGuard lock(p->m_mutex); now = grp.landed; done = grp.done; rc = grp.rc; }
			if (now != seen && controls.interrupt) {
				seen = now;
				const int prc = controls.interrupt->progress(this, scene, sensor, sampler, target, now / groupSize, controls, threadIdx, threadCount);
				if (prc != 0) return prc;
			}
			if (done) return rc;
			usleep(200);
		}
	}

	// ---- leader
	struct Finish {      // whatever way the leader leaves, its followers are released with its return code
		GPUPathTracer *p; Group *grp; int rc;
		~Finish() { LockGuard lock(p->m_mutex); grp->done = true; grp->rc = rc; }
	} fin = { p, &grp, 0 };
	float *dst = target.getBitmap()->getFloatData();
	const size_t n = (size_t) target.getBitmap()->getSize().x * target.getBitmap()->getSize().y * target.getChannelCount();
	// the group's share: sample indices gpu, gpu + G, ...
	uint32_t mine = 0;
	for (uint32_t s = (uint32_t) gpu; s < sampleCount; s += (uint32_t) G) ++mine;
	const uint32_t passCap = maxPass(target);
	uint32_t pass = p->m_sppPerPass > 0 ? passCap : 1u;     // grows 1, 4, 16, ... up to a full chunk of paths
	int sinceSync = 0;
	ref<Timer> sinceFlush = new Timer();
	static const bool timing = getenv("MTSB_PLUGIN_TIMING") != NULL;      // breakdown of a round in the log
	ref<Timer> tt = new Timer();
	double tSubmit = 0, tStage = 0, tAdd = 0, tProgress = 0; const double tEnter = self_now();
	MTSB_CALL(mtsb_render_clear(r));          // the device film holds only what has not landed yet
	while (true) {
		if (controls.abort && *controls.abort) return fin.rc = -1;
		if (controls.continu && !*controls.continu) return fin.rc = -2;
		// flush policy: first pass at once, then geometrically (or every 250 ms), and always at the end of the share
		const uint32_t unflushed = queued - landed;
		const bool flush = unflushed && (queued >= mine || unflushed >= std::max(1u, landed / 2) || sinceFlush->getMilliseconds() >= 250);
		if (flush) {
			// in stream order: film -> staging buffer, film cleared; whatever is queued next accumulates into the cleared film
			MTSB_CALL(mtsb_film_flush_begin(r));
		}
		const uint32_t flushedUpTo = queued;
		// queue the next pass BEFORE the host's share of the flush, so that the device renders while the host adds
		if (queued < mine) {
			const uint32_t cnt = std::min(pass, mine - queued);
			tt->reset();
			MTSB_CALL(mtsb_render_pass(r, queued * G, cnt, sampleCount, NULL, (uint32_t) G, (uint32_t) gpu));
			tSubmit += tt->getMicroseconds() * 1e-3;
			queued += cnt;
			pass = std::min(passCap, pass * 4);
			++sinceSync;
		}
		if (flush) {
			const float *src = NULL; size_t cnt = 0;
			tt->reset();
			MTSB_CALL(mtsb_film_flush_wait(r, &src, &cnt));    // waits for the copy, not for the pass queued above
			tStage += tt->getMicroseconds() * 1e-3;
			if (cnt != n) Log(EError, "gpupath: target holds %zu floats, the device film %zu", n, cnt);
			tt->reset();
			{
				// only leaders write targets; they can meet in one ImageBlock when the caller shares it between GPUs' threads
				LockGuard lock(m_addLock);
				for (size_t i = 0; i < n; ++i) dst[i] += src[i];
			}
			tAdd += tt->getMicroseconds() * 1e-3;
			landed = flushedUpTo;
			{ LockGuard lock(p->m_mutex); grp.landed = (double) landed; }
			sinceFlush->reset();
			sinceSync = 0;
		} else if (sinceSync >= 4) {
			// keep at most a few passes in flight so that abort / camera changes take effect within a few waves
			MTSB_CALL(mtsb_render_sync(r)); sinceSync = 0;
		}
		if (controls.interrupt) {
			tt->reset();
			const int rc = controls.interrupt->progress(this, scene, sensor, sampler, target, (double) landed / groupSize, controls, threadIdx, threadCount);
			tProgress += tt->getMicroseconds() * 1e-3;
			if (rc != 0) { mtsb_render_sync(r); return fin.rc = rc; }
		}
		if (landed >= mine) {                            // "sampler exhausted" — and everything is in the target
			if (threadIdx == 0) {
				uint64_t st[8]; mtsb_render_stats(r, st);
				char buf[160];
				snprintf(buf, sizeof(buf), "GPU wavefront: %.2f rays/sample, avg path length %.2f", st[0] ? (double) (st[1] + st[2]) / st[0] : 0.0,
					st[0] ? (double) st[3] / st[0] : 0.0);
				m_stats = buf;
			}
			if (timing) SLog(EInfo, "gpupath: leader %d (group of %d): %.2f ms in render(); submit %.2f, flush wait %.2f, add %.2f, progress %.2f ms",
				threadIdx, groupSize, self_now() - tEnter, tSubmit, tStage, tAdd, tProgress);
			return fin.rc = 0;
		}
	}
}

MTS_IMPLEMENT_CLASS(GPUResponsiveIntegrator, false, ResponsiveIntegrator)
MTS_IMPLEMENT_CLASS_S(GPUPathTracer, false, MonteCarloIntegrator)
MTS_EXPORT_PLUGIN(GPUPathTracer, "B200 wavefront path tracer (libmtsb200)");
MTS_NAMESPACE_END
