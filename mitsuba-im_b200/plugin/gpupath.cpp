// gpupath.cpp — Mitsuba integrator plugin that routes the `path` / `direct` hot path through libmtsb200 (B200 CUDA).
//
// Drop-in at the reference's plugin boundary: exports CreateInstance / GetDescription
// (include/mitsuba/core/cobject.h:102-110) and the `mitsuba_integrator_plugin` marker
// (src/integrators/mark_integrator.cpp), derives from MonteCarloIntegrator (src/librender/integrator.cpp:191-226, same
// XML properties: maxDepth, rrDepth, strictNormals, hideEmitters) and serves both schedulers:
//   classic     Integrator::render   (include/mitsuba/render/integrator.h:61-120; caller RenderJob::run, renderjob.cpp:86-117)
//   responsive  ResponsiveIntegrator (include/mitsuba/render/integrator2.h:49-100; callers src/mitsuba/im_render.cpp, src/im-mts/scene.cpp)
// Built as gpupath.so; installed under the name path.so / direct.so in an overlay plugin directory it replaces the stock
// integrators for unmodified scene files.  Extra properties: gpus, builder (lbvh|ploc), sppPerPass, seed.
#include <mitsuba/render/integrator.h>
#include <mitsuba/render/integrator2.h>
#include <mitsuba/render/renderjob.h>
#include <mitsuba/render/imageblock.h>
#include <mitsuba/core/statistics.h>
#include <mitsuba/core/lock.h>
#include "extract.h"
#include <map>

MTS_NAMESPACE_BEGIN

/// Per-GPU state shared by both scheduler paths; cached on scene identity so that Integrator2's per-frame preprocess
/// (src/im-mts/shell.cpp:240-242) does not rebuild the BVH.
struct GPUDevice {
	mtsb_ctx *ctx; mtsb_scene *scene; const Scene *owner;
	ref<Mutex> submit;     // held while one pass is being QUEUED on this GPU (the responsive path, several host threads per GPU)
	GPUDevice() : ctx(NULL), scene(NULL), owner(NULL) { submit = new Mutex(); }
	void ensure(int device, const Scene *s, int builder) {
		if (!ctx) MTSB_CALL(mtsb_ctx_create(device, &ctx));
		if (scene && owner == s) return;
		if (scene) mtsb_scene_destroy(scene);
		scene = uploadScene(ctx, s, builder);
		owner = s;
		double st[8]; mtsb_scene_build_stats(scene, st);
		SLog(EInfo, "gpupath: device %d: BVH8 over %u primitives built in %.1f ms (%d nodes, SAH cost %.1f)", device,
			(unsigned) mtsb_scene_primitive_count(scene), st[0], (int) st[1], st[4]);
	}
	~GPUDevice() { if (scene) mtsb_scene_destroy(scene); if (ctx) mtsb_ctx_destroy(ctx); }
};

class GPUPathTracer;

/// Responsive (Integrator2) adapter: thread t renders the sample indices s = t (mod threadCount) of every pixel on GPU
/// t % gpus, accumulates un-normalised sums into the caller's ImageBlock and reports its own spp share, which is what
/// InteractiveSceneProcess::develop sums up (src/mitsuba/im_render.cpp:225-247).
class GPUResponsiveIntegrator : public ResponsiveIntegrator {
public:
	GPUResponsiveIntegrator(GPUPathTracer *parent);
	bool preprocess(const Scene *scene, const Sensor *sensor, const Sampler *sampler);
	bool allocate(const Scene &scene, Sampler *const *samplers, ImageBlock *const *targets, int threadCount);
	int render(const Scene &scene, const Sensor &sensor, Sampler &sampler, ImageBlock &target, Controls controls, int threadIdx, int threadCount);
	const char *getRealtimeStatistics() { return m_stats.c_str(); }
	MTS_DECLARE_CLASS()
protected:
	virtual ~GPUResponsiveIntegrator();
	mtsb_render *renderFor(int gpu, const Scene &scene, const Sensor &sensor, const Sampler &sampler, const ImageBlock &target);
	uint32_t maxPass(const ImageBlock &target) const;
	/// state of one device group during a render() round (guarded by the parent's mutex)
	struct Group { int entered, roundSize; double landed; bool done; int rc; Group() : entered(0), roundSize(0), landed(0), done(false), rc(0) { } };
public:
	void prepare(const Scene &scene);
protected:
	ref<GPUPathTracer> m_parent;
	std::vector<mtsb_render *> m_renders;      // one per device group
	std::vector<Group> m_groups;
	ref<Mutex> m_addLock;                      // host-side add into a target (leaders only)
	std::string m_stats;
	std::vector<ImageBlock *> m_allocTargets; std::vector<Sampler *> m_allocSamplers;   // from allocate(), for prepare()
	int m_active; double m_roundSamples; ref<Timer> m_roundTimer;   // one render() round over all threads (guarded by the parent's mutex)
};

class GPUPathTracer : public MonteCarloIntegrator {
public:
	GPUPathTracer(const Properties &props) : MonteCarloIntegrator(props), m_cancel(false), m_responsive(NULL) {
		m_gpus = props.getInteger("gpus", 1);
		m_sppPerPass = props.getInteger("sppPerPass", 0);
		m_seed = props.getInteger("seed", 0);
		const std::string b = to_lower_copy(props.getString("builder", "ploc"));
		if (b != "ploc" && b != "lbvh") Log(EError, "gpupath: builder must be \"lbvh\" or \"ploc\"");
		m_builder = b == "ploc" ? MTSB_BUILDER_PLOC : MTSB_BUILDER_LBVH;
		// `direct` semantics when installed as direct.so or asked for explicitly
		m_direct = props.getPluginName() == "direct" || props.getBoolean("direct", false);
		// SimpleVolumetricPathTracer semantics (participating media) when installed as volpath_simple.so
		m_volpath = props.getPluginName() == "volpath_simple" || props.getBoolean("volpath", false);
		// VolumetricPathTracer (the variant with multiple importance sampling) when installed as volpath.so
		m_volmis = props.getPluginName() == "volpath" || props.getBoolean("volpathMIS", false);
		m_emitterSamples = props.getInteger("emitterSamples", props.getInteger("shadingSamples", 1));
		m_bsdfSamples = props.getInteger("bsdfSamples", props.getInteger("shadingSamples", 1));
		const int n = mtsb_device_count();
		if (n == 0) Log(EError, "gpupath: no CUDA device — this integrator has no CPU fallback");
		if (m_gpus < 1 || m_gpus > n) Log(EError, "gpupath: gpus=%d but %d device(s) are visible", m_gpus, n);
		m_mutex = new Mutex();
		m_devices.resize(m_gpus);
	}
	GPUPathTracer(Stream *stream, InstanceManager *manager) : MonteCarloIntegrator(stream, manager) {
		Log(EError, "gpupath: network rendering is not supported (the GPU path replaces the use of remote workers)");
	}

	/// Per-ray entry point other integrators (adaptive, irrcache) call; the batch pipeline has no per-ray mode.
	Spectrum Li(const RayDifferential &, RadianceQueryRecord &) const {
		Log(EError, "gpupath: Li() per ray is not available — use the stock `path` plugin for integrators that call it");
		return Spectrum(0.0f);
	}

	void describeIntegrator(mtsb_integrator_desc &d) const {
		memset(&d, 0, sizeof(d));
		d.type = m_direct ? MTSB_INTEGRATOR_DIRECT : (m_volmis ? MTSB_INTEGRATOR_VOLPATH : (m_volpath ? MTSB_INTEGRATOR_VOLPATH_SIMPLE : MTSB_INTEGRATOR_PATH));
		d.max_depth = m_maxDepth; d.rr_depth = m_rrDepth;
		d.strict_normals = m_strictNormals ? 1 : 0; d.hide_emitters = m_hideEmitters ? 1 : 0;
		d.emitter_samples = m_emitterSamples; d.bsdf_samples = m_bsdfSamples;
	}

	uint64_t seedFor(const Sampler *sampler) const {
		// share the stream of oracle/ctrsampler.cpp when the scene uses it, so renders are comparable sample by sample
		if (sampler && sampler->getProperties().hasProperty("seed")) return (uint64_t) sampler->getProperties().getInteger("seed");
		return (uint64_t) m_seed;
	}

	GPUDevice &device(int i, const Scene *scene) {
		LockGuard lock(m_mutex);
		m_devices[i].ensure(i, scene, m_builder);
		return m_devices[i];
	}

	bool preprocess(const Scene *scene, RenderQueue *, const RenderJob *, int, int, int) {
		for (int i = 0; i < m_gpus; ++i) device(i, scene);
		if (m_responsive) m_responsive->prepare(*scene);
		return true;
	}

	/// Classic scheduler path: the whole image in this thread, sample passes split over the GPUs.
	bool render(Scene *scene, RenderQueue *queue, const RenderJob *job, int, int, int) {
		ref<Sensor> sensor = scene->getSensor();
		ref<Film> film = sensor->getFilm();
		const Sampler *sampler = scene->getSampler();
		const uint32_t sampleCount = (uint32_t) sampler->getSampleCount();
		m_cancel = false;
		Log(EInfo, "Starting render job (%ix%i, %u %s, %d GPU%s) ..", film->getCropSize().x, film->getCropSize().y,
			sampleCount, sampleCount == 1 ? "sample" : "samples", m_gpus, m_gpus == 1 ? "" : "s");
		ref<ImageBlock> block = new ImageBlock(Bitmap::ESpectrumAlphaWeight, film->getCropSize(), film->getReconstructionFilter());
		block->setOffset(Point2i(0, 0));
		block->clear();
		mtsb_camera cam; describeCamera(sensor, cam);
		mtsb_film_desc fd; describeFilm(*block, film->getReconstructionFilter(), film, fd);
		mtsb_integrator_desc id; describeIntegrator(id);
		std::vector<mtsb_render *> renders(m_gpus, (mtsb_render *) NULL);
		for (int g = 0; g < m_gpus; ++g) {
			MTSB_CALL(mtsb_render_create(device(g, scene).scene, &cam, &fd, &id, seedFor(sampler), &renders[g]));
			MTSB_CALL(mtsb_render_set_camera_medium(renders[g], mediumIndex(collectMedia(scene), sensor->getMedium())));
			configureDeviceSampler(renders[g], sampler);
		}
		const size_t pixels = (size_t) film->getCropSize().x * film->getCropSize().y;
		const uint32_t pass = m_sppPerPass > 0 ? (uint32_t) m_sppPerPass : (uint32_t) std::max<size_t>(1, ((size_t) 1 << 23) / std::max<size_t>(pixels, 1));
		// path state, kernels and the bounce graph are set up before the clock starts (the reference's kd-tree build and
		// its preprocess() are outside its render timer too)
		for (int g = 0; g < m_gpus; ++g) MTSB_CALL(mtsb_render_reserve(renders[g], pass));
		// so is the NCCL communicator of the film reduce (about a second of bootstrap, once per integrator)
		if (m_gpus > 1 && m_comms.empty()) {
			std::vector<mtsb_ctx *> ctxs(m_gpus);
			for (int g = 0; g < m_gpus; ++g) ctxs[g] = m_devices[g].ctx;
			m_comms.assign(m_gpus, (mtsb_comm *) NULL);
			MTSB_CALL(mtsb_comm_create_all(ctxs.data(), m_gpus, m_comms.data()));
			// first collective on a fresh communicator sets up its channels: run it on the still empty films
			MTSB_CALL(mtsb_film_reduce_all(renders.data(), m_comms.data(), m_gpus, 0));
			for (int g = 0; g < m_gpus; ++g) MTSB_CALL(mtsb_render_sync(renders[g]));
		}
		ref<Timer> timer = new Timer();
		// GPU g renders sample indices s = g (mod gpus).  mtsb_render_pass only QUEUES work (no host wait inside), so this one
		// thread keeps all devices busy: pass k goes to every GPU before anything is waited for.  A pass covers about one
		// chunk of paths in flight (8 M) so that cancel() is looked at every few tens of milliseconds; the host waits for
		// the devices only every fourth pass.
		const uint32_t perGpu = (sampleCount + m_gpus - 1) / m_gpus;
		uint32_t passIdx = 0;
		for (uint32_t done = 0; done < perGpu && !m_cancel; done += pass, ++passIdx) {
			for (int g = 0; g < m_gpus; ++g) {
				const uint32_t first = done, last = std::min(perGpu, done + pass);
				// the k-th local sample of GPU g has index g + k * gpus; drop indices beyond sampleCount
				uint32_t cnt = 0;
				for (uint32_t k = first; k < last; ++k) if ((uint32_t) g + k * m_gpus < sampleCount) ++cnt;
				if (cnt) MTSB_CALL(mtsb_render_pass(renders[g], first * m_gpus, cnt, sampleCount, NULL, m_gpus, g));
			}
			if ((passIdx & 3) == 3) for (int g = 0; g < m_gpus; ++g) MTSB_CALL(mtsb_render_sync(renders[g]));
		}
		for (int g = 0; g < m_gpus; ++g) MTSB_CALL(mtsb_render_sync(renders[g]));
		float *dst = block->getBitmap()->getFloatData();
		const size_t n = (size_t) block->getBitmap()->getSize().x * block->getBitmap()->getSize().y * block->getChannelCount();
		uint64_t st[8] = { 0 }, tot[8] = { 0 };
		// the per-GPU films meet on the device: one NCCL reduce into GPU 0's film, then a single D2H into the ImageBlock
		if (m_gpus > 1) MTSB_CALL(mtsb_film_reduce_all(renders.data(), m_comms.data(), m_gpus, 0));
		MTSB_CALL(mtsb_film_to_host(renders[0], dst, n, 0));
		// the clock stops with the film in host memory; what follows is the reference's own CPU code (Film::put, develop)
		const double s = timer->getMicroseconds() * 1e-6;
		for (int g = 0; g < m_gpus; ++g) {
			mtsb_render_stats(renders[g], st);
			for (int i = 0; i < 8; ++i) tot[i] += st[i];
			mtsb_render_destroy(renders[g]);
		}
		film->put(block);
		if (queue && job) queue->signalRefresh(job);
		Log(EInfo, "gpupath: %llu paths, %llu closest-hit + %llu shadow rays, average path length %.2f, %.4f s to the film in host memory, %.1f M samples/s",
			(unsigned long long) tot[0], (unsigned long long) tot[1], (unsigned long long) tot[2],
			tot[0] ? (double) tot[3] / tot[0] : 0.0, s, tot[0] / std::max(s, 1e-9) / 1e6);
		return !m_cancel;
	}

	void cancel() { m_cancel = true; }
	ref<ResponsiveIntegrator> makeResponsiveIntegrator() { m_responsive = new GPUResponsiveIntegrator(this); return m_responsive; }
	void serialize(Stream *stream, InstanceManager *manager) const { MonteCarloIntegrator::serialize(stream, manager); }
	std::string toString() const {
		std::ostringstream oss;
		oss << "GPUPathTracer[maxDepth = " << m_maxDepth << ", rrDepth = " << m_rrDepth << ", strictNormals = " << m_strictNormals
			<< ", gpus = " << m_gpus << ", direct = " << m_direct << "]";
		return oss.str();
	}
	MTS_DECLARE_CLASS()

	int m_gpus, m_sppPerPass, m_seed, m_builder, m_emitterSamples, m_bsdfSamples;
	bool m_direct, m_volpath, m_volmis;
	volatile bool m_cancel;
	GPUResponsiveIntegrator *m_responsive;      // the adapter made by makeResponsiveIntegrator (not owned; clears itself)
	ref<Mutex> m_mutex;
	std::vector<GPUDevice> m_devices;
	std::vector<mtsb_comm *> m_comms;
	virtual ~GPUPathTracer() { for (size_t i = 0; i < m_comms.size(); ++i) mtsb_comm_destroy(m_comms[i]); }
};

MTS_NAMESPACE_END
#include "gpupath_responsive.h"
