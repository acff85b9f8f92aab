// comm.cu — multi-GPU film reduction of the C ABI (mtsb_comm_*, mtsb_film_reduce*).
//
// The path shards by sample index: every GPU holds the whole scene + BVH and renders the sample indices
// s = rank (mod nranks) of every pixel into its own full-size device film (mtsb_render_pass stride/phase).  Nothing
// is exchanged while tracing; the ONE exchange step is the sum of the per-GPU films into the root's film, which is
// the device-side analogue of what the reference does on the host when its workers hand their ImageBlocks to the
// film: Film::put(block) -> ImageBlock::put(const ImageBlock*) -> Bitmap::accumulate
// (src/films/hdrfilm.cpp:391-393, include/mitsuba/render/imageblock.h:140-143) and, for the responsive scheduler,
// the caller summing per-thread targets (src/mitsuba/im_render.cpp:225-247).
//
// NCCL is bound at run time (dlopen of libnccl.so.2) so that libmtsb200.so carries no link-time dependency on it:
// inside a PyTorch process the already-loaded bundled NCCL is picked up by SONAME, inside the reference's own
// `mitsuba` executable the system one is.  Without NCCL the calls fail with MTSB_EUNSUPPORTED (no host-side sum).
#include "render.cuh"
#include <dlfcn.h>
#include <mutex>
#include <nccl.h>
#include <string.h>

namespace mtsb {

struct NcclApi {
	void *handle = nullptr;
	ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
	ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
	ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
	ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
	ncclResult_t (*Reduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
	ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
	ncclResult_t (*GroupStart)() = nullptr;
	ncclResult_t (*GroupEnd)() = nullptr;
	const char *(*GetErrorString)(ncclResult_t) = nullptr;
	ncclResult_t (*GetVersion)(int *) = nullptr;
};

static NcclApi g_nccl;
static std::once_flag g_ncclOnce;
static bool g_ncclOk = false;

static void load_nccl() {
	const char *names[] = { "libnccl.so.2", "libnccl.so" };
	for (const char *n : names) {
		g_nccl.handle = dlopen(n, RTLD_NOW | RTLD_LOCAL);
		if (g_nccl.handle) break;
	}
	if (!g_nccl.handle) return;
	#define SYM(field, name) do { *(void **) (&g_nccl.field) = dlsym(g_nccl.handle, name); if (!g_nccl.field) return; } while (0)
	SYM(GetUniqueId, "ncclGetUniqueId"); SYM(CommInitRank, "ncclCommInitRank"); SYM(CommInitAll, "ncclCommInitAll");
	SYM(CommDestroy, "ncclCommDestroy"); SYM(Reduce, "ncclReduce"); SYM(AllReduce, "ncclAllReduce");
	SYM(GroupStart, "ncclGroupStart"); SYM(GroupEnd, "ncclGroupEnd"); SYM(GetErrorString, "ncclGetErrorString");
	SYM(GetVersion, "ncclGetVersion");
	#undef SYM
	g_ncclOk = true;
}

static int need_nccl(const char *who) {
	std::call_once(g_ncclOnce, load_nccl);
	if (!g_ncclOk) MTSB_FAIL(MTSB_EUNSUPPORTED, "%s: NCCL (libnccl.so.2) could not be bound at run time", who);
	return 0;
}

#define MTSB_NCCL(call) do { ncclResult_t r__ = (call); if (r__ != ncclSuccess) { \
	mtsb::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString(r__)); return MTSB_ECUDA; } } while (0)

} // namespace mtsb

struct mtsb_comm {
	mtsb_ctx *ctx = nullptr;
	ncclComm_t comm = nullptr;
	int nranks = 0, rank = 0;
};

using namespace mtsb;

static_assert(sizeof(ncclUniqueId) == MTSB_COMM_ID_BYTES, "ncclUniqueId size changed");

extern "C" {

int mtsb_comm_version(void) {
	if (need_nccl("mtsb_comm_version") < 0) return MTSB_EUNSUPPORTED;
	int v = 0;
	MTSB_NCCL(g_nccl.GetVersion(&v));
	return v;
}

int mtsb_comm_unique_id(uint8_t id[MTSB_COMM_ID_BYTES]) {
	if (!id) MTSB_FAIL(MTSB_EINVAL, "mtsb_comm_unique_id: NULL argument");
	MTSB_CHECK(need_nccl("mtsb_comm_unique_id"));
	ncclUniqueId u;
	MTSB_NCCL(g_nccl.GetUniqueId(&u));
	memcpy(id, &u, sizeof(u));
	return 0;
}

int mtsb_comm_create(mtsb_ctx *ctx, int nranks, int rank, const uint8_t id[MTSB_COMM_ID_BYTES], mtsb_comm **out) {
	if (!ctx || !id || !out || nranks < 1 || rank < 0 || rank >= nranks) MTSB_FAIL(MTSB_EINVAL, "mtsb_comm_create: bad argument (rank %d of %d)", rank, nranks);
	MTSB_CHECK(need_nccl("mtsb_comm_create"));
	MTSB_CUDA(cudaSetDevice(ctx->device));
	ncclUniqueId u; memcpy(&u, id, sizeof(u));
	mtsb_comm *c = new mtsb_comm();
	c->ctx = ctx; c->nranks = nranks; c->rank = rank;
	ncclResult_t r = g_nccl.CommInitRank(&c->comm, nranks, u, rank);
	if (r != ncclSuccess) { delete c; MTSB_FAIL(MTSB_ECUDA, "mtsb_comm_create: ncclCommInitRank -> %s", g_nccl.GetErrorString(r)); }
	*out = c;
	return 0;
}

int mtsb_comm_create_all(mtsb_ctx *const *ctxs, int n, mtsb_comm **out) {
	if (!ctxs || !out || n < 1 || n > 64) MTSB_FAIL(MTSB_EINVAL, "mtsb_comm_create_all: bad argument");
	MTSB_CHECK(need_nccl("mtsb_comm_create_all"));
	int devs[64]; ncclComm_t comms[64];
	for (int i = 0; i < n; ++i) { if (!ctxs[i]) MTSB_FAIL(MTSB_EINVAL, "mtsb_comm_create_all: NULL context %d", i); devs[i] = ctxs[i]->device; }
	MTSB_NCCL(g_nccl.CommInitAll(comms, n, devs));
	for (int i = 0; i < n; ++i) {
		out[i] = new mtsb_comm();
		out[i]->ctx = ctxs[i]; out[i]->comm = comms[i]; out[i]->nranks = n; out[i]->rank = i;
	}
	return 0;
}

void mtsb_comm_destroy(mtsb_comm *c) {
	if (!c) return;
	if (c->comm && g_ncclOk) { cudaSetDevice(c->ctx->device); g_nccl.CommDestroy(c->comm); }
	delete c;
}

static int film_collective(mtsb_render *r, mtsb_comm *c, int root, const char *who) {
	if (!r || !c) MTSB_FAIL(MTSB_EINVAL, "%s: NULL argument", who);
	if (r->scene->ctx != c->ctx) MTSB_FAIL(MTSB_EINVAL, "%s: render and communicator live on different contexts", who);
	if (root >= c->nranks) MTSB_FAIL(MTSB_EINVAL, "%s: root %d of %d ranks", who, root, c->nranks);
	MTSB_CUDA(cudaSetDevice(c->ctx->device));
	float *film = r->dFilm.ptr;
	if (root < 0) MTSB_NCCL(g_nccl.AllReduce(film, film, r->dFilm.count, ncclFloat, ncclSum, c->comm, r->stream));
	else MTSB_NCCL(g_nccl.Reduce(film, film, r->dFilm.count, ncclFloat, ncclSum, root, c->comm, r->stream));
	return 0;
}

int mtsb_film_reduce(mtsb_render *r, mtsb_comm *c, int root) {
	MTSB_CHECK(need_nccl("mtsb_film_reduce"));
	return film_collective(r, c, root, "mtsb_film_reduce");
}

int mtsb_film_reduce_all(mtsb_render *const *renders, mtsb_comm *const *comms, int n, int root) {
	if (!renders || !comms || n < 1) MTSB_FAIL(MTSB_EINVAL, "mtsb_film_reduce_all: bad argument");
	MTSB_CHECK(need_nccl("mtsb_film_reduce_all"));
	for (int i = 1; i < n; ++i)
		if (!renders[i] || !renders[0] || renders[i]->dFilm.count != renders[0]->dFilm.count)
			MTSB_FAIL(MTSB_EINVAL, "mtsb_film_reduce_all: films differ in size");
	MTSB_NCCL(g_nccl.GroupStart());
	int rc = 0;
	for (int i = 0; i < n && rc == 0; ++i) rc = film_collective(renders[i], comms[i], root, "mtsb_film_reduce_all");
	ncclResult_t e = g_nccl.GroupEnd();
	if (rc < 0) return rc;
	if (e != ncclSuccess) MTSB_FAIL(MTSB_ECUDA, "mtsb_film_reduce_all: ncclGroupEnd -> %s", g_nccl.GetErrorString(e));
	return 0;
}

} // extern "C"
