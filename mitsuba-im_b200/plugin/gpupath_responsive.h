// gpupath_responsive.h — ResponsiveIntegrator (Integrator2) adapter + plugin exports (included by gpupath.cpp).
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

bool GPUResponsiveIntegrator::allocate(const Scene &scene, Sampler *const *, ImageBlock *const *targets, int threadCount) {
	for (size_t i = 0; i < m_renders.size(); ++i) if (m_renders[i]) mtsb_render_destroy(m_renders[i]);
	m_renders.assign(threadCount, (mtsb_render *) NULL);
	const Film *film = scene.getFilm();
	if (film->getCropSize() != film->getSize() || targets[0]->getSize() != film->getSize())
		Log(EError, "gpupath: the responsive path expects full-frame targets (no crop window)");
	return true;
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
			mtsb_integrator_desc id; p->describeIntegrator(id);
			MTSB_CALL(mtsb_render_create(dev.scene, &cam, &fd, &id, p->seedFor(&sampler), &m_renders[threadIdx]));
		}
	}
	mtsb_render *r = m_renders[threadIdx];
	MTSB_CALL(mtsb_render_set_camera(r, &cam));     // camera change = new kernel parameters, nothing else
	float *dst = target.getBitmap()->getFloatData();
	const size_t n = (size_t) target.getBitmap()->getSize().x * target.getBitmap()->getSize().y * target.getChannelCount();

	// this thread's share: sample indices threadIdx, threadIdx + threadCount, ...
	uint32_t mine = 0;
	for (uint32_t s = (uint32_t) threadIdx; s < sampleCount; s += (uint32_t) threadCount) ++mine;
	const uint32_t pass = p->m_sppPerPass > 0 ? (uint32_t) p->m_sppPerPass : 1u;   // 1 spp per pass keeps abort latency at one wave
	uint32_t done = 0;
	int rc = 0;
	while (true) {
		if (controls.abort && *controls.abort) return -1;
		if (controls.continu && !*controls.continu) return -2;
		if (controls.interrupt) {
			rc = controls.interrupt->progress(this, scene, sensor, sampler, target, (double) done, controls, threadIdx, threadCount);
			if (rc != 0) return rc;
		}
		if (done >= mine) return 0;                      // "sampler exhausted"
		const uint32_t cnt = std::min(pass, mine - done);
		{
			// one submission at a time per GPU; the film of this pass is added into the (possibly shared) ImageBlock
			LockGuard lock(p->m_mutex);
			MTSB_CALL(mtsb_render_clear(r));
			MTSB_CALL(mtsb_render_pass(r, done * threadCount, cnt, sampleCount, NULL, (uint32_t) threadCount, (uint32_t) threadIdx));
			MTSB_CALL(mtsb_film_to_host(r, dst, n, 1));
			if (threadIdx == 0) {
				uint64_t st[8]; mtsb_render_stats(r, st);
				char buf[160];
				snprintf(buf, sizeof(buf), "GPU wavefront: %.2f rays/sample, avg path length %.2f", st[0] ? (double) (st[1] + st[2]) / st[0] : 0.0,
					st[0] ? (double) st[3] / st[0] : 0.0);
				m_stats = buf;
			}
		}
		done += cnt;
	}
}

MTS_IMPLEMENT_CLASS(GPUResponsiveIntegrator, false, ResponsiveIntegrator)
MTS_IMPLEMENT_CLASS_S(GPUPathTracer, false, MonteCarloIntegrator)
MTS_EXPORT_PLUGIN(GPUPathTracer, "B200 wavefront path tracer (libmtsb200)");
MTS_NAMESPACE_END
