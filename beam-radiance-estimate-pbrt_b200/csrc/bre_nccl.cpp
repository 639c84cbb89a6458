// bre_nccl.cpp -- multi-GPU photon exchange over NCCL (NVLink 5 / NVSwitch).
//
// The reference is single-process; this is the one exchange step of the
// sharded path: every rank traces its slice of photon path indices, then all ranks gather
// all photon records so that every GPU builds the identical LBVH.
//
// NCCL is reached through dlopen so that libbre.so loads on machines without NCCL and shares
// whatever libnccl the host process already has (e.g. the one bundled with PyTorch).
#include "bre_internal.cuh"
#include <dlfcn.h>
#include <cstring>
#include <vector>

namespace {

typedef struct { char internal[128]; } ncclUniqueId_t;
typedef void *ncclComm_p;
typedef int (*fnGetUniqueId)(ncclUniqueId_t *);
typedef int (*fnCommInitRank)(ncclComm_p *, int, ncclUniqueId_t, int);
typedef int (*fnCommDestroy)(ncclComm_p);
typedef int (*fnAllGather)(const void *, void *, size_t, int, ncclComm_p, cudaStream_t);
typedef int (*fnBroadcast)(const void *, void *, size_t, int, int, ncclComm_p, cudaStream_t);
typedef int (*fnGroup)(void);
typedef const char *(*fnErrStr)(int);

struct Nccl {
    void *h = nullptr;
    fnGetUniqueId GetUniqueId = nullptr;
    fnCommInitRank CommInitRank = nullptr;
    fnCommDestroy CommDestroy = nullptr;
    fnAllGather AllGather = nullptr;
    fnBroadcast Broadcast = nullptr;
    fnGroup GroupStart = nullptr, GroupEnd = nullptr;
    fnErrStr GetErrorString = nullptr;
    bool ok = false;
} g;

const int kNcclUint64 = 5, kNcclFloat = 7;  // ncclDataType_t

bool load() {
    if (g.ok) return true;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *n : names) {
        g.h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (g.h) break;
    }
    if (!g.h) return false;
#define SYM(field, name) g.field = (decltype(g.field))dlsym(g.h, name); if (!g.field) return false
    SYM(GetUniqueId, "ncclGetUniqueId");
    SYM(CommInitRank, "ncclCommInitRank");
    SYM(CommDestroy, "ncclCommDestroy");
    SYM(AllGather, "ncclAllGather");
    SYM(Broadcast, "ncclBroadcast");
    SYM(GroupStart, "ncclGroupStart");
    SYM(GroupEnd, "ncclGroupEnd");
    SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
    g.ok = true;
    return true;
}

int nccl_err(bre_ctx *ctx, int r, const char *what) {
    return bre::set_err(ctx, BRE_ERR_NCCL, std::string(what) + ": " + (g.GetErrorString ? g.GetErrorString(r) : "?"));
}

}  // namespace

namespace bre {

int grow_photon_store(bre_ctx *ctx, uint64_t newCount);

int nccl_unique_id(void *out128) {
    if (!out128) return BRE_ERR_INVALID;
    if (!load()) return BRE_ERR_NCCL;
    ncclUniqueId_t id;
    if (g.GetUniqueId(&id) != 0) return BRE_ERR_NCCL;
    memcpy(out128, &id, 128);
    return BRE_OK;
}

int nccl_init(bre_ctx *ctx, const void *uid) {
    if (!load()) return set_err(ctx, BRE_ERR_NCCL, "libnccl.so.2 could not be loaded");
    ncclUniqueId_t id;
    memcpy(&id, uid, 128);
    ncclComm_p comm = nullptr;
    int r = g.CommInitRank(&comm, ctx->n_ranks, id, ctx->rank);
    if (r != 0) return nccl_err(ctx, r, "ncclCommInitRank");
    ctx->nccl_comm = comm;
    return BRE_OK;
}

void nccl_destroy(bre_ctx *ctx) {
    if (ctx->nccl_comm && g.ok) g.CommDestroy((ncclComm_p)ctx->nccl_comm);
    ctx->nccl_comm = nullptr;
}

// All-gather of variable-size photon slabs: counts first (8 bytes per rank), then one
// ncclBroadcast per (rank, array) inside a single group, each landing directly at its final
// rank-major offset -- no padding, no second compaction pass.
int nccl_allgather_photons(bre_ctx *ctx) {
    if (!ctx->nccl_comm) return set_err(ctx, BRE_ERR_INVALID, "context was created without NCCL");
    ncclComm_p comm = (ncclComm_p)ctx->nccl_comm;
    const int R = ctx->n_ranks;
    int s;
    if ((s = ensure(ctx, ctx->counters_d, 8 * (size_t)(R + 1) + 64)) != BRE_OK) return s;
    uint64_t *cnt_d = (uint64_t *)ctx->counters_d.p;
    uint64_t mine = ctx->n_photons;
    BRE_CUDA(cudaMemcpyAsync(cnt_d + R, &mine, 8, cudaMemcpyHostToDevice, ctx->stream));
    int r = g.AllGather(cnt_d + R, cnt_d, 1, kNcclUint64, comm, ctx->stream);
    if (r != 0) return nccl_err(ctx, r, "ncclAllGather(counts)");
    std::vector<uint64_t> cnt(R), off(R + 1, 0);
    BRE_CUDA(cudaMemcpyAsync(cnt.data(), cnt_d, 8 * (size_t)R, cudaMemcpyDeviceToHost, ctx->stream));
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    for (int i = 0; i < R; ++i) off[i + 1] = off[i] + cnt[i];
    const uint64_t total = off[R];
    // receive into the build's sorted-SoA buffers as staging (rebuilt afterwards anyway)
    Buf *stage[3] = {&ctx->s_pos_r, &ctx->s_pow, &ctx->s_dir};
    Buf *store[3] = {&ctx->pos_r, &ctx->pow, &ctx->dir};
    for (int a = 0; a < 3; ++a)
        if ((s = ensure(ctx, *stage[a], (size_t)total * 16)) != BRE_OK) return s;
    if ((r = g.GroupStart()) != 0) return nccl_err(ctx, r, "ncclGroupStart");
    for (int a = 0; a < 3; ++a)
        for (int root = 0; root < R; ++root) {
            if (cnt[root] == 0) continue;
            float *recv = (float *)stage[a]->p + 4 * off[root];
            const float *send = root == ctx->rank ? (const float *)store[a]->p : recv;
            r = g.Broadcast(send, recv, (size_t)cnt[root] * 4, kNcclFloat, root, comm, ctx->stream);
            if (r != 0) { g.GroupEnd(); return nccl_err(ctx, r, "ncclBroadcast"); }
        }
    if ((r = g.GroupEnd()) != 0) return nccl_err(ctx, r, "ncclGroupEnd");
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    for (int a = 0; a < 3; ++a) { Buf t = *store[a]; *store[a] = *stage[a]; *stage[a] = t; }
    ctx->n_photons = total;
    return BRE_OK;
}

}  // namespace bre
