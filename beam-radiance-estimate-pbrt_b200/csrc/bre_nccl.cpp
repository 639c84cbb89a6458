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

void p2p_destroy(bre_ctx *ctx);
void nccl_destroy(bre_ctx *ctx) {
    p2p_destroy(ctx);
    if (ctx->nccl_comm && g.ok) g.CommDestroy((ncclComm_p)ctx->nccl_comm);
    ctx->nccl_comm = nullptr;
}

// All-gather of variable-size photon slabs: counts first (8 bytes per rank), then one
// ncclBroadcast per (rank, array) inside a single group, each landing directly at its final
// rank-major offset -- no padding, no second compaction pass.
int nccl_allgather_photons(bre_ctx *ctx) {
    if (!ctx->nccl_comm) return set_err(ctx, BRE_ERR_INVALID, "context was created without NCCL");
    if (ctx->p2p_open)
        return set_err(ctx, BRE_ERR_INVALID, "this context exchanges photons over P2P (bre_trace_exchange_photons); "
                                             "set BRE_NO_P2P=1 to use the NCCL all-gather instead");
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


// ---------------------------------------------------------------- fused trace + push (NVLink P2P)
// The NCCL path above is compute (K1 + compaction) FOLLOWED by a collective.  Here the two are one
// kernel: after tracing its slice into per-path slots every rank knows its photon count; the counts
// are all-gathered (8 bytes per rank -- also the barrier that says "nobody reads its store any
// more"), and the compaction kernel then writes every photon record straight to its final,
// rank-major position in the store of EVERY rank -- the local one and the peers' through CUDA-IPC
// mappings over NVLink / NVSwitch.  No staging slab, no broadcast; the transfer overlaps the
// compaction pass record by record.  A one-word all-gather afterwards is the arrival barrier.
// Stores are exported once (and again whenever any rank had to reallocate: the decision rides on
// the count all-gather so that it is collective).
constexpr int MAX_P2P_RANKS = 8;
struct PeerStores {
    float4 *pos[MAX_P2P_RANKS], *pw[MAX_P2P_RANKS], *dr[MAX_P2P_RANKS];
};

// One block = a tile of `tilePaths` consecutive paths.  Their photons are consecutive in the output,
// so the tile is compacted into shared memory first and then written to every rank as contiguous,
// fully coalesced 16-byte-per-lane runs (NVLink moves full packets instead of scattered 16-byte stores).
__global__ void __launch_bounds__(256) k_compact_push(uint32_t nPaths, int maxDepth, uint32_t tilePaths,
                                                      const uint32_t *__restrict__ counts,
                                                      const uint64_t *__restrict__ offs,
                                                      const float4 *__restrict__ slot_pos,
                                                      const float4 *__restrict__ slot_pow,
                                                      const float4 *__restrict__ slot_dir,
                                                      PeerStores peers, int nRanks, int self,
                                                      uint64_t rankBase) {
    extern __shared__ float4 tile[];
    const uint32_t cap = tilePaths * (uint32_t)maxDepth;
    float4 *tp = tile, *tw = tile + cap, *td = tile + 2 * cap;
    const uint32_t p0 = blockIdx.x * tilePaths;
    const uint32_t pEnd = min(p0 + tilePaths, nPaths);
    const uint64_t base = __ldg(offs + p0);
    for (uint32_t pid = p0 + threadIdx.x; pid < pEnd; pid += blockDim.x) {
        const uint32_t c = __ldg(counts + pid);
        const uint32_t o = (uint32_t)(__ldg(offs + pid) - base);
        const size_t src = (size_t)pid * maxDepth;
        for (uint32_t k = 0; k < c; ++k) {
            tp[o + k] = __ldg(slot_pos + src + k);
            tw[o + k] = __ldg(slot_pow + src + k);
            td[o + k] = __ldg(slot_dir + src + k);
        }
    }
    __syncthreads();
    const uint32_t n = (uint32_t)(__ldg(offs + pEnd - 1) + __ldg(counts + pEnd - 1) - base);
    const uint64_t dst = rankBase + base;
    // start with the neighbour so that the R ranks do not all write to rank 0 at the same time
    for (int i = 0; i < nRanks; ++i) {
        const int r = (self + 1 + i) % nRanks;
        float4 *dp = peers.pos[r] + dst, *dw = peers.pw[r] + dst, *dd = peers.dr[r] + dst;
        for (uint32_t j = threadIdx.x; j < n; j += blockDim.x) { dp[j] = tp[j]; dw[j] = tw[j]; dd[j] = td[j]; }
    }
}

static void p2p_close(bre_ctx *ctx) {
    for (int r = 0; r < MAX_P2P_RANKS; ++r)
        for (int a = 0; a < 3; ++a)
            if (ctx->p2p_peer[r][a] && r != ctx->rank) { cudaIpcCloseMemHandle(ctx->p2p_peer[r][a]); ctx->p2p_peer[r][a] = nullptr; }
    ctx->p2p_open = false;
}
void p2p_destroy(bre_ctx *ctx) {
    const bool was = ctx->p2p_open;
    p2p_close(ctx);
    if (was && ctx->nccl_comm) {
        // peers must have unmapped this rank's store before bre_destroy frees it
        uint64_t *cnt_d = (uint64_t *)ctx->counters_d.p;
        const int R = ctx->n_ranks;
        if (cnt_d && g.AllGather(cnt_d + 2 * R, cnt_d, 1, kNcclUint64, (ncclComm_p)ctx->nccl_comm, ctx->stream) == 0)
            cudaStreamSynchronize(ctx->stream);
    }
    for (void *z : ctx->p2p_zombies) cudaFree(z);
    ctx->p2p_zombies.clear();
}

// all ranks: export the three store arrays, all-gather the handles, map the peers' arrays
static int p2p_exchange_handles(bre_ctx *ctx) {
    ncclComm_p comm = (ncclComm_p)ctx->nccl_comm;
    const int R = ctx->n_ranks;
    p2p_close(ctx);
    cudaIpcMemHandle_t mine[3];
    void *store[3] = {ctx->pos_r.p, ctx->pow.p, ctx->dir.p};
    for (int a = 0; a < 3; ++a) BRE_CUDA(cudaIpcGetMemHandle(&mine[a], store[a]));
    const size_t hb = sizeof(mine);   // 192 bytes per rank
    int s;
    if ((s = ensure(ctx, ctx->p2p_scratch, hb * (size_t)(R + 1))) != BRE_OK) return s;
    char *d = (char *)ctx->p2p_scratch.p;
    BRE_CUDA(cudaMemcpyAsync(d + hb * R, mine, hb, cudaMemcpyHostToDevice, ctx->stream));
    int r = g.AllGather(d + hb * R, d, hb, 0 /* ncclChar */, comm, ctx->stream);
    if (r != 0) return nccl_err(ctx, r, "ncclAllGather(ipc handles)");
    std::vector<cudaIpcMemHandle_t> all(3 * (size_t)R);
    BRE_CUDA(cudaMemcpyAsync(all.data(), d, hb * R, cudaMemcpyDeviceToHost, ctx->stream));
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    for (int q = 0; q < R; ++q)
        for (int a = 0; a < 3; ++a) {
            if (q == ctx->rank) { ctx->p2p_peer[q][a] = store[a]; continue; }
            void *ptr = nullptr;
            cudaError_t e = cudaIpcOpenMemHandle(&ptr, all[3 * q + a], cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) {
                cudaGetLastError();
                return set_err(ctx, BRE_ERR_CUDA, std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(e));
            }
            ctx->p2p_peer[q][a] = ptr;
        }
    for (int a = 0; a < 3; ++a) ctx->p2p_exported[a] = store[a];
    ctx->p2p_open = true;
    return BRE_OK;
}

// Trace this rank's slice of [0, nPathsTotal) and deliver all photons of all ranks into every
// rank's store (replaces bre_clear_photons + bre_trace_photons + bre_allgather_photons).
int trace_exchange_p2p(bre_ctx *ctx, uint64_t nPathsTotal, int maxDepth, float radius) {
    ncclComm_p comm = (ncclComm_p)ctx->nccl_comm;
    const int R = ctx->n_ranks;
    uint64_t first = 0, count = 0;
    bre_shard_range(nPathsTotal, R, ctx->rank, &first, &count);
    ctx->built = false; ctx->topo_valid = false;
    ctx->n_photons = 0;
    SlotView v;
    int s;
    {
        PhaseTimer pt(ctx, PH_TRACE);
        if ((s = trace_to_slots(ctx, first, (uint32_t)count, maxDepth, radius, &v)) != BRE_OK) return s;
    }
    // ---- counts (+ "my store moved" flag) to everyone; doubles as the pre-push barrier
    if ((s = ensure(ctx, ctx->counters_d, 16 * (size_t)(R + 1) + 64)) != BRE_OK) return s;
    uint64_t *cnt_d = (uint64_t *)ctx->counters_d.p;
    const bool moved = !ctx->p2p_open || ctx->p2p_exported[0] != ctx->pos_r.p ||
                       ctx->p2p_exported[1] != ctx->pow.p || ctx->p2p_exported[2] != ctx->dir.p;
    uint64_t mine[2] = {v.total, moved ? 1ull : 0ull};
    BRE_CUDA(cudaMemcpyAsync(cnt_d + 2 * R, mine, 16, cudaMemcpyHostToDevice, ctx->stream));
    int r = g.AllGather(cnt_d + 2 * R, cnt_d, 2, kNcclUint64, comm, ctx->stream);
    if (r != 0) return nccl_err(ctx, r, "ncclAllGather(counts)");
    std::vector<uint64_t> got(2 * (size_t)R), off(R + 1, 0);
    BRE_CUDA(cudaMemcpyAsync(got.data(), cnt_d, 16 * (size_t)R, cudaMemcpyDeviceToHost, ctx->stream));
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    bool anyMoved = false;
    for (int i = 0; i < R; ++i) { off[i + 1] = off[i] + got[2 * i]; anyMoved |= got[2 * i + 1] != 0; }
    const uint64_t total = off[R];
    if (total >= BRE_MAX_PHOTONS) return set_err(ctx, BRE_ERR_UNSUPPORTED, "photon count exceeds 2^27");
    // ---- capacity: the same decision on every rank (total is global), with 25 % head room
    const size_t need = (size_t)total * 16;
    const bool grow = ctx->pos_r.cap < need || ctx->pow.cap < need || ctx->dir.cap < need;
    if (grow || anyMoved) {
        // everyone unmaps everyone first; only then may an exported store be freed
        p2p_close(ctx);
        r = g.AllGather(cnt_d + 2 * R, cnt_d, 1, kNcclUint64, comm, ctx->stream);
        if (r != 0) return nccl_err(ctx, r, "ncclAllGather(unmap barrier)");
        BRE_CUDA(cudaStreamSynchronize(ctx->stream));
        for (void *z : ctx->p2p_zombies) cudaFree(z);
        ctx->p2p_zombies.clear();
        if (grow && (s = grow_photon_store(ctx, total + total / 4)) != BRE_OK) return s;
        if ((s = p2p_exchange_handles(ctx)) != BRE_OK) return s;
    }
    // ---- fused compaction + push
    PeerStores ps;
    for (int q = 0; q < MAX_P2P_RANKS; ++q) {
        ps.pos[q] = (float4 *)ctx->p2p_peer[q < R ? q : ctx->rank][0];
        ps.pw[q] = (float4 *)ctx->p2p_peer[q < R ? q : ctx->rank][1];
        ps.dr[q] = (float4 *)ctx->p2p_peer[q < R ? q : ctx->rank][2];
    }
    if (count > 0) {
        // tile of paths per block, bounded by 200 KB of shared memory (3 x 16 B x tilePaths x maxDepth)
        uint32_t tilePaths = 256;
        while (tilePaths > 1 && (size_t)tilePaths * maxDepth * 48 > 200 * 1024) tilePaths >>= 1;
        const size_t smem = (size_t)tilePaths * maxDepth * 48;
        if (smem > 200 * 1024) return set_err(ctx, BRE_ERR_UNSUPPORTED, "max_depth too large for the push kernel");
        BRE_CUDA(cudaFuncSetAttribute(k_compact_push, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_compact_push<<<(unsigned)((count + tilePaths - 1) / tilePaths), 256, smem, ctx->stream>>>(
            (uint32_t)count, maxDepth, tilePaths, v.counts, v.offs, v.sp, v.sw, v.sd, ps, R, ctx->rank,
            off[ctx->rank]);
    }
    BRE_CUDA(cudaGetLastError());
    // ---- arrival barrier: every rank's push kernel has completed before anyone builds
    r = g.AllGather(cnt_d + 2 * R, cnt_d, 1, kNcclUint64, comm, ctx->stream);
    if (r != 0) return nccl_err(ctx, r, "ncclAllGather(barrier)");
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->n_photons = total;
    return BRE_OK;
}

}  // namespace bre
