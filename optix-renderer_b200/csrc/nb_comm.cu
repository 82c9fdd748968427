/* nb_comm.cu — the path's only collective, inside the C ABI: one NCCL sum-reduce of the film accumulators to rank 0.
 *
 * The film is linear (ImageBlock stores sum w*rgb and sum w, src/utils/block.cpp:120-122; the division happens once at
 * the end, block.cpp:80, color.h:113-118), so G GPUs that render disjoint sample slices of the same image produce the
 * one-GPU film of all samples by summing their accumulators. Two ways to set the communicator up:
 *   - one process per GPU (torchrun, MPI): rank 0 calls nb_comm_unique_id, the launcher's own channel carries the 128
 *     bytes to the other ranks, every rank calls nb_comm_init_rank on its context;
 *   - one process, n GPUs (the reference's caller is ONE render thread, src/utils/render.cpp:462-574):
 *     nb_create_multi makes a context per device and one communicator clique (ncclCommInitAll); because
 *     nb_render_subframe is asynchronous, that one host thread keeps all devices busy.
 * NCCL is looked up at run time (dlopen of libnccl.so.2 — in a process that already loaded PyTorch that is the copy
 * PyTorch loaded) so that the library itself has no link-time dependency on it; every other entry works without it.
 * The reduce runs on the context's own stream, in order after the subframes queued there and before any later
 * nb_film_download, with no host synchronisation. */
#include "nb_internal.h"
#include <dlfcn.h>
#include <cstdlib>
#include <cstring>
#include <mutex>

typedef struct { char internal[128]; } NbNcclUniqueId;        /* ncclUniqueId, NCCL_UNIQUE_ID_BYTES == 128 */
static_assert(sizeof(NbNcclUniqueId) == NB_COMM_ID_BYTES, "nccl id size");
typedef void *NbNcclComm;
enum { NB_NCCL_SUCCESS = 0, NB_NCCL_FLOAT = 7, NB_NCCL_SUM = 0 };

struct NbNccl {
    void *lib = nullptr;
    int (*GetUniqueId)(NbNcclUniqueId *) = nullptr;
    int (*CommInitRank)(NbNcclComm *, int, NbNcclUniqueId, int) = nullptr;
    int (*CommInitAll)(NbNcclComm *, int, const int *) = nullptr;
    int (*CommDestroy)(NbNcclComm) = nullptr;
    int (*Reduce)(const void *, void *, size_t, int, int, int, NbNcclComm, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int *) = nullptr;
    std::string why;
};

struct NbGroup { std::vector<nb_ctx *> members; };

static std::mutex g_nccl_mutex;
static NbNccl g_nccl;

static NbNccl *nb_nccl(nb_ctx *ctx) {
    std::lock_guard<std::mutex> lock(g_nccl_mutex);
    NbNccl &n = g_nccl;
    if (n.lib) return &n;
    const char *names[] = { getenv("NORIB200_NCCL"), "libnccl.so.2", "libnccl.so" };
    for (const char *nm : names) {
        if (!nm || !*nm) continue;
        n.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (n.lib) break;
    }
    if (!n.lib) {
        nb_fail(ctx, NB_E_UNSUPPORTED, "NCCL is not available (%s): multi-GPU film reduce needs libnccl.so.2", dlerror());
        return nullptr;
    }
    bool ok = true;
    auto sym = [&](const char *name) { void *p = dlsym(n.lib, name); if (!p) { ok = false; n.why = name; } return p; };
    n.GetUniqueId = (decltype(n.GetUniqueId)) sym("ncclGetUniqueId");
    n.CommInitRank = (decltype(n.CommInitRank)) sym("ncclCommInitRank");
    n.CommInitAll = (decltype(n.CommInitAll)) sym("ncclCommInitAll");
    n.CommDestroy = (decltype(n.CommDestroy)) sym("ncclCommDestroy");
    n.Reduce = (decltype(n.Reduce)) sym("ncclReduce");
    n.GroupStart = (decltype(n.GroupStart)) sym("ncclGroupStart");
    n.GroupEnd = (decltype(n.GroupEnd)) sym("ncclGroupEnd");
    n.GetErrorString = (decltype(n.GetErrorString)) sym("ncclGetErrorString");
    n.GetVersion = (decltype(n.GetVersion)) sym("ncclGetVersion");
    if (!ok) {
        nb_fail(ctx, NB_E_UNSUPPORTED, "libnccl.so.2 lacks %s", n.why.c_str());
        dlclose(n.lib); n.lib = nullptr;
        return nullptr;
    }
    return &n;
}

#define NB_NCCL(ctx, n, call) do { const int r_ = (call); if (r_ != NB_NCCL_SUCCESS) \
    return nb_fail(ctx, NB_E_CUDA, "%s failed: %s", #call, (n)->GetErrorString(r_)); } while (0)

void nb_comm_free(nb_ctx *ctx) {
    if (ctx->comm && g_nccl.lib) g_nccl.CommDestroy((NbNcclComm) ctx->comm);
    ctx->comm = nullptr; ctx->comm_rank = 0; ctx->comm_size = 1;
    if (ctx->group) {
        NbGroup *g = ctx->group;
        for (size_t i = 0; i < g->members.size(); ++i) if (g->members[i] == ctx) g->members[i] = nullptr;
        bool empty = true;
        for (nb_ctx *m : g->members) if (m) empty = false;
        if (empty) delete g;
        ctx->group = nullptr;
    }
}

/* After the reduce the samples live in the root's film: the other ranks start again from zero, so that rendering more
 * subframes and reducing again (progressive display, render.cpp:725-747) counts nothing twice. Must be queued AFTER
 * ncclGroupEnd: calls inside an NCCL group reach the stream only when the group closes. */
static int nb_reduce_clear(nb_ctx *ctx) {
    if (ctx->comm_rank == 0) return NB_OK;
    float *data = nullptr, *normal = nullptr;
    uint64_t count = 0;
    const int rc = nb_render_film(ctx, &data, &normal, &count);
    if (rc != NB_OK) return rc;
    NB_CUDA(ctx, cudaMemsetAsync(data, 0, (size_t) count * sizeof(float), ctx->stream));
    if (normal) NB_CUDA(ctx, cudaMemsetAsync(normal, 0, (size_t) count * sizeof(float), ctx->stream));
    return NB_OK;
}

/* ncclReduce calls of one context: composite film, then the normal layer when the render has one */
static int nb_reduce_enqueue(nb_ctx *ctx, NbNccl *n) {
    float *data = nullptr, *normal = nullptr;
    uint64_t count = 0;
    const int rc = nb_render_film(ctx, &data, &normal, &count);
    if (rc != NB_OK) return rc;
    NB_NCCL(ctx, n, n->Reduce(data, data, (size_t) count, NB_NCCL_FLOAT, NB_NCCL_SUM, 0, (NbNcclComm) ctx->comm, ctx->stream));
    if (normal) NB_NCCL(ctx, n, n->Reduce(normal, normal, (size_t) count, NB_NCCL_FLOAT, NB_NCCL_SUM, 0, (NbNcclComm) ctx->comm, ctx->stream));
    return NB_OK;
}

extern "C" {

int nb_comm_unique_id(void *id) {
    if (!id) return NB_E_INVALID;
    NbNccl *n = nb_nccl(nullptr);
    if (!n) return NB_E_UNSUPPORTED;
    NB_NCCL(nullptr, n, n->GetUniqueId((NbNcclUniqueId *) id));
    return NB_OK;
}

int nb_comm_init_rank(nb_ctx *ctx, int n_ranks, int rank, const void *id) {
    if (!ctx || !id || n_ranks < 1 || rank < 0 || rank >= n_ranks) return nb_fail(ctx, NB_E_INVALID, "nb_comm_init_rank: bad argument");
    if (ctx->group) return nb_fail(ctx, NB_E_STATE, "nb_comm_init_rank: this context belongs to an nb_create_multi group");
    NbNccl *n = nb_nccl(ctx);
    if (!n) return NB_E_UNSUPPORTED;
    NB_CUDA(ctx, cudaSetDevice(ctx->device));
    nb_comm_free(ctx);
    NbNcclUniqueId uid;
    memcpy(&uid, id, sizeof(uid));
    NbNcclComm comm = nullptr;
    NB_NCCL(ctx, n, n->CommInitRank(&comm, n_ranks, uid, rank));
    ctx->comm = comm; ctx->comm_rank = rank; ctx->comm_size = n_ranks;
    return NB_OK;
}

int nb_comm_info(nb_ctx *ctx, int *rank, int *n_ranks, int *nccl_version) {
    if (!ctx) return NB_E_INVALID;
    if (rank) *rank = ctx->comm_rank;
    if (n_ranks) *n_ranks = ctx->comm_size;
    if (nccl_version) {
        *nccl_version = 0;
        if (ctx->comm && g_nccl.lib) g_nccl.GetVersion(nccl_version);
    }
    return NB_OK;
}

int nb_create_multi(const int *devices, int n, nb_ctx **out) {
    if (!devices || !out || n < 1) return nb_fail(nullptr, NB_E_INVALID, "nb_create_multi: bad argument");
    for (int i = 0; i < n; ++i) out[i] = nullptr;
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < i; ++j)
            if (devices[i] == devices[j]) return nb_fail(nullptr, NB_E_INVALID, "nb_create_multi: device %d listed twice", devices[i]);
    int rc = NB_OK;
    for (int i = 0; i < n && rc == NB_OK; ++i) rc = nb_create(devices[i], &out[i]);
    std::vector<NbNcclComm> comms((size_t) n, nullptr);
    NbNccl *nc = nullptr;
    if (rc == NB_OK && n > 1) {
        nc = nb_nccl(nullptr);
        if (!nc) rc = NB_E_UNSUPPORTED;
        else {
            const int r = nc->CommInitAll(comms.data(), n, devices);
            if (r != NB_NCCL_SUCCESS) rc = nb_fail(nullptr, NB_E_CUDA, "ncclCommInitAll failed: %s", nc->GetErrorString(r));
        }
    }
    if (rc != NB_OK) {
        for (int i = 0; i < n; ++i) { if (out[i]) nb_destroy(out[i]); out[i] = nullptr; }
        return rc;
    }
    NbGroup *g = new NbGroup();
    for (int i = 0; i < n; ++i) {
        out[i]->comm = comms[(size_t) i]; out[i]->comm_rank = i; out[i]->comm_size = n;
        out[i]->group = g;
        g->members.push_back(out[i]);
    }
    return NB_OK;
}

int nb_film_reduce(nb_ctx *ctx) {
    if (!ctx) return NB_E_INVALID;
    if (ctx->comm_size <= 1) return NB_OK;                   /* one GPU: the film is already the sum */
    NbNccl *n = nb_nccl(ctx);
    if (!n || !ctx->comm) return nb_fail(ctx, NB_E_STATE, "nb_film_reduce: no communicator (nb_comm_init_rank / nb_create_multi)");
    if (!ctx->group) {
        NB_CUDA(ctx, cudaSetDevice(ctx->device));
        const int rc = nb_reduce_enqueue(ctx, n);
        return rc != NB_OK ? rc : nb_reduce_clear(ctx);
    }
    /* one process, n devices: every member's reduce is issued here, inside one NCCL group */
    for (nb_ctx *m : ctx->group->members)
        if (!m) return nb_fail(ctx, NB_E_STATE, "nb_film_reduce: a context of this group was destroyed");
    NB_NCCL(ctx, n, n->GroupStart());
    int rc = NB_OK;
    for (nb_ctx *m : ctx->group->members) {
        cudaSetDevice(m->device);
        rc = nb_reduce_enqueue(m, n);
        if (rc != NB_OK) { if (m != ctx) nb_fail(ctx, rc, "nb_film_reduce: device %d: %s", m->device, m->error.c_str()); break; }
    }
    const int r = n->GroupEnd();
    if (rc == NB_OK && r == NB_NCCL_SUCCESS)
        for (nb_ctx *m : ctx->group->members) {
            cudaSetDevice(m->device);
            if ((rc = nb_reduce_clear(m)) != NB_OK) { if (m != ctx) nb_fail(ctx, rc, "nb_film_reduce: device %d: %s", m->device, m->error.c_str()); break; }
        }
    cudaSetDevice(ctx->device);
    if (rc != NB_OK) return rc;
    if (r != NB_NCCL_SUCCESS) return nb_fail(ctx, NB_E_CUDA, "ncclGroupEnd failed: %s", n->GetErrorString(r));
    return NB_OK;
}

} /* extern "C" */
