// skdtree_overlay.cpp — per-ray drop-in for the three ShapeKDTree::rayIntersect overloads.
//
//     LD_PRELOAD=libmtsb_overlay.so mitsuba scene.xml
//
// Scene::rayIntersect* are inline forwarders (include/mitsuba/render/scene.h:187-243) to three out-of-line members of
// ShapeKDTree exported by libmitsuba-render.so (src/librender/skdtree.cpp:112-142 closest hit + record, :144-204 light
// record, :207-226 shadow ray).  This library defines the same three symbols; loaded first, the dynamic linker binds
// every caller to them — the stock `path`, `direct`, `bdpt`, photon mappers, Scene::sampleEmitterDirect's visibility
// test: all of them, unmodified — while the rest of ShapeKDTree (addShape, build, the kd-tree itself) stays the
// reference's.  The overloads here do not traverse anything on the CPU:
//
//   caller thread                              service thread (one per process)
//   ------------------------------------      ---------------------------------------------------------------
//   take a ticket, write the ray into its      gather the contiguous run of READY slots of a queue, hand the batch to
//   slot of the request ring, mark it READY,   mtsb_trace_closest / _light / _any (host pointers: H2D, BVH8 traversal
//   sleep on the slot (futex)                  on the GPU, D2H), scatter the records, mark DONE, wake the sleepers
//   wake up, turn (t, u, v, primitive) into
//   an Intersection with the reference's OWN fillIntersectionRecord<true> (skdtree.h:343-428)
//
// The fourth interposed member is ShapeKDTree::build(): the CPU SAH build is skipped (a 10 M-triangle scene
// loads without minutes of kd-tree construction), see skdtree_overlay_impl.h.
// There is no flush timer: while one batch is on the device the next one accumulates, so the batch size adapts to the
// number of threads that are blocked.  A per-ray synchronous interface cannot feed a GPU from 16 threads — run the
// renderer with hundreds of threads (`mitsuba -p 512`: blocked threads cost nothing) and the batches grow with them.
// The device scene is built on the first query of each kd-tree from ShapeKDTree::getShapes() (triangle meshes and
// `rectangle`; any other shape is refused loudly — no CPU fallback).
#include <mitsuba/render/skdtree.h>
#include <mitsuba/render/trimesh.h>
#include <mitsuba/core/statistics.h>
#include <mitsuba/core/properties.h>
#include <mitsuba/core/timer.h>
#include <dlfcn.h>
#include <stdlib.h>
#include "../../include/mtsb200.h"
#include <atomic>
#include <algorithm>
#include <map>
#include <mutex>
#include <thread>
#include <vector>
#include <string.h>
#include <unistd.h>
#include <sys/syscall.h>
#include <linux/futex.h>

MTS_NAMESPACE_BEGIN

namespace {

static StatsCounter ovRays("GPU overlay", "Closest-hit rays (ShapeKDTree::rayIntersect(ray, its))");
static StatsCounter ovShadow("GPU overlay", "Shadow rays (ShapeKDTree::rayIntersect(ray))");
static StatsCounter ovLight("GPU overlay", "Light-record rays (ShapeKDTree::rayIntersect(ray, t, shape, n, uv))");

inline void futexWait(std::atomic<int> *a, int expected) { syscall(SYS_futex, (int *) a, FUTEX_WAIT_PRIVATE, expected, NULL, NULL, 0); }
inline void futexWake(std::atomic<int> *a, int n) { syscall(SYS_futex, (int *) a, FUTEX_WAKE_PRIVATE, n, NULL, NULL, 0); }

enum { Q_CLOSEST = 0, Q_LIGHT = 1, Q_ANY = 2, Q_COUNT = 3 };
enum { S_FREE = 0, S_READY = 1, S_DONE = 2 };
static const unsigned RING = 1u << 14;          // slots per queue (>= threads that can be blocked at once)

struct Slot {
	mtsb_ray ray;
	union { mtsb_hit hit; mtsb_light_hit light; uint8_t any; } out;
	std::atomic<int> state;
};

struct Queue {
	std::vector<Slot> ring;
	std::atomic<unsigned long long> head;       // tickets issued
	unsigned long long tail;                    // next ticket the service thread looks at (service thread only)
	void *rays, *outs;                          // pinned staging (mtsb_host_alloc)
	unsigned long long batches, requests;
	Queue() : ring(RING), head(0), tail(0), rays(NULL), outs(NULL), batches(0), requests(0) {
		for (unsigned i = 0; i < RING; ++i) ring[i].state.store(S_FREE);
	}
};

struct Overlay {
	mtsb_scene *scene;
	Queue q[Q_COUNT];
	std::atomic<int> pending;                   // requests not yet picked up, all queues (the service thread sleeps on it)
	std::thread service;
	std::atomic<bool> stop;
	Overlay() : scene(NULL), pending(0), stop(false) { }
};

std::mutex g_mutex;
std::map<const ShapeKDTree *, Overlay *> g_overlays;
mtsb_ctx *g_ctx = NULL;

void serviceLoop(Overlay *ov) {
	static const size_t outBytes[Q_COUNT] = { sizeof(mtsb_hit), sizeof(mtsb_light_hit), 1 };
	for (int k = 0; k < Q_COUNT; ++k) {
		ov->q[k].rays = mtsb_host_alloc((size_t) RING * sizeof(mtsb_ray));
		ov->q[k].outs = mtsb_host_alloc((size_t) RING * sizeof(mtsb_light_hit));
	}
	while (!ov->stop.load(std::memory_order_relaxed)) {
		bool any = false;
		for (int k = 0; k < Q_COUNT; ++k) {
			Queue &q = ov->q[k];
			// the contiguous run of READY slots behind the tail (a ticket taken but not yet written ends the run)
			const unsigned long long head = q.head.load(std::memory_order_acquire);
			unsigned n = 0;
			while (q.tail + n < head && n < RING && q.ring[(q.tail + n) % RING].state.load(std::memory_order_acquire) == S_READY) ++n;
			if (n == 0) continue;
			any = true;
			mtsb_ray *rays = (mtsb_ray *) q.rays;
			for (unsigned i = 0; i < n; ++i) rays[i] = q.ring[(q.tail + i) % RING].ray;
			ov->pending.fetch_sub((int) n, std::memory_order_relaxed);
			// Small batches: the staging buffers are page-locked (mtsb_host_alloc) and therefore addressable by the device
			// under unified addressing, so the kernel reads the rays and writes the records straight through PCIe
			// ("on_device" pointers): one launch + one sync, 15-20 us a batch, instead of the chunked copy pipeline of the
			// host-pointer path (~70 us), which only pays from a few thousand rays on.
			const int zeroCopy = n <= 4096 ? 1 : 0;
			int rc;
			if (k == Q_CLOSEST) rc = mtsb_trace_closest(ov->scene, rays, n, (mtsb_hit *) q.outs, zeroCopy);
			else if (k == Q_LIGHT) rc = mtsb_trace_light(ov->scene, rays, n, (mtsb_light_hit *) q.outs, zeroCopy);
			else rc = mtsb_trace_any(ov->scene, rays, n, (uint8_t *) q.outs, zeroCopy);
			if (rc == 0 && zeroCopy) rc = mtsb_ctx_sync(g_ctx);
			if (rc < 0) { fprintf(stderr, "mtsb overlay: %s\n", mtsb_last_error()); abort(); }     // no CPU fallback
			for (unsigned i = 0; i < n; ++i) {
				Slot &s = q.ring[(q.tail + i) % RING];
				memcpy(&s.out, (const uint8_t *) q.outs + (size_t) i * outBytes[k], outBytes[k]);
				s.state.store(S_DONE, std::memory_order_release);
				futexWake(&s.state, 1);
			}
			q.tail += n; q.batches += 1; q.requests += n;
		}
		if (!any) {
			// nothing ready: sleep until a caller publishes a request (a short spin first — the next request is
			// usually microseconds away while a render is running)
			for (int spin = 0; spin < 200 && ov->pending.load(std::memory_order_acquire) == 0; ++spin) __builtin_ia32_pause();
			if (ov->pending.load(std::memory_order_acquire) == 0) futexWait(&ov->pending, 0);
		}
	}
}

/// One request through queue `k`: blocks until the service thread has filled the slot.
inline Slot &submit(Overlay *ov, int k, const Ray &ray) {
	Queue &q = ov->q[k];
	const unsigned long long ticket = q.head.fetch_add(1, std::memory_order_acq_rel);
	Slot &s = q.ring[ticket % RING];
	while (s.state.load(std::memory_order_acquire) != S_FREE) sched_yield();      // ring wrapped onto a slot still being read
	s.ray.o[0] = ray.o.x; s.ray.o[1] = ray.o.y; s.ray.o[2] = ray.o.z; s.ray.mint = ray.mint;
	s.ray.d[0] = ray.d.x; s.ray.d[1] = ray.d.y; s.ray.d[2] = ray.d.z; s.ray.maxt = ray.maxt;
	s.state.store(S_READY, std::memory_order_release);
	if (ov->pending.fetch_add(1, std::memory_order_acq_rel) == 0) futexWake(&ov->pending, 1);
	for (int spin = 0; spin < 100 && s.state.load(std::memory_order_acquire) != S_DONE; ++spin) __builtin_ia32_pause();
	while (s.state.load(std::memory_order_acquire) != S_DONE) futexWait(&s.state, S_READY);
	return s;
}

inline void release(Slot &s) { s.state.store(S_FREE, std::memory_order_release); }

} // namespace

MTS_NAMESPACE_END
#include "skdtree_overlay_impl.h"
