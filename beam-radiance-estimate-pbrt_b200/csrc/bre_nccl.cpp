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
#include <algorithm>
#include <vector>

namespace {

typedef struct { char internal[128]; } ncclUniqueId_t;
typedef void *ncclComm_p;
typedef int (*fnGetUniqueId)(ncclUniqueId_t *);
typedef int (*fnCommInitRank)(ncclComm_p *, int, ncclUniqueId_t, int);
typedef int (*fnCommDestroy)(ncclComm_p);
typedef int (*fnAllGather)(const void *, void *, size_t, int, ncclComm_p, cudaStream_t);
typedef int (*fnBroadcast)(const void *, void *, size_t, int, int, ncclComm_p, cudaStream_t);
typedef int (*fnReduce)(const void *, void *, size_t, int, int, int, ncclComm_p, cudaStream_t);
typedef int (*fnGroup)(void);
typedef const char *(*fnErrStr)(int);

struct Nccl {
    void *h = nullptr;
    fnGetUniqueId GetUniqueId = nullptr;
    fnCommInitRank CommInitRank = nullptr;
    fnCommDestroy CommDestroy = nullptr;
    fnAllGather AllGather = nullptr;
    fnBroadcast Broadcast = nullptr;
    fnReduce Reduce = nullptr;
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
    SYM(Reduce, "ncclReduce");
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
static void p2p_close(bre_ctx *ctx);

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
    ncclComm_p comm = (ncclComm_p)ctx->nccl_comm;
    const int R = ctx->n_ranks;
    int s;
    if ((s = ensure(ctx, ctx->counters_d, 32 * (size_t)(R + 1) + 64)) != BRE_OK) return s;
    uint64_t *cnt_d = (uint64_t *)ctx->counters_d.p;
    if (ctx->p2p_open) {
        // an earlier exchange went over P2P (the choice between the two paths is a function of global values,
        // so every rank is here together): unmap the peers' stores before the slabs below replace them
        p2p_close(ctx);
        int r0 = g.AllGather(cnt_d + 2 * R, cnt_d, 1, kNcclUint64, comm, ctx->stream);
        if (r0 != 0) return nccl_err(ctx, r0, "ncclAllGather(unmap barrier)");
        BRE_CUDA(cudaStreamSynchronize(ctx->stream));
        for (void *z : ctx->p2p_zombies) cudaFree(z);
        ctx->p2p_zombies.clear();
    }
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

// Single-rank use of the same kernel: the tile-in-shared-memory compaction writes full coalesced runs where the
// one-thread-per-path k_compact (bre_trace.cu) scatters 16-byte records (0.37 ms -> see profiles/ at 10 M photons).
int compact_tiled_local(bre_ctx *ctx, const SlotView &v, uint64_t storeBase) {
    if (v.nPaths == 0) return BRE_OK;
    PeerStores ps;
    for (int q = 0; q < MAX_P2P_RANKS; ++q) {
        ps.pos[q] = (float4 *)ctx->pos_r.p; ps.pw[q] = (float4 *)ctx->pow.p; ps.dr[q] = (float4 *)ctx->dir.p;
    }
    uint32_t tilePaths = 256;
    while (tilePaths > 1 && (size_t)tilePaths * v.maxDepth * 48 > 200 * 1024) tilePaths >>= 1;
    const size_t smem = (size_t)tilePaths * v.maxDepth * 48;
    if (smem > 200 * 1024) return BRE_ERR_UNSUPPORTED;      // caller falls back to k_compact
    BRE_CUDA(cudaFuncSetAttribute(k_compact_push, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_compact_push<<<(unsigned)((v.nPaths + tilePaths - 1) / tilePaths), 256, smem, ctx->stream>>>(
        v.nPaths, v.maxDepth, tilePaths, v.counts, v.offs, v.sp, v.sw, v.sd, ps, 1, 0, storeBase);
    BRE_CUDA(cudaGetLastError());
    return BRE_OK;
}

static void p2p_close(bre_ctx *ctx) {
    for (int r = 0; r < MAX_P2P_RANKS; ++r)
        for (int a = 0; a < 3; ++a)
            if (ctx->p2p_peer[r][a] && r != ctx->rank) { cudaIpcCloseMemHandle(ctx->p2p_peer[r][a]); ctx->p2p_peer[r][a] = nullptr; }
    ctx->p2p_open = false;
}
static void film_peers_close(bre_ctx *ctx) {
    for (int r = 0; r < MAX_P2P_RANKS; ++r) {
        if (ctx->film_peer[r] && r != ctx->rank) cudaIpcCloseMemHandle(ctx->film_peer[r]);
        ctx->film_peer[r] = nullptr;
    }
    ctx->film_peers_open = false;
    ctx->film_exported = nullptr;
}
void p2p_destroy(bre_ctx *ctx) {
    if (ctx->gather_peer0 && ctx->rank != 0) cudaIpcCloseMemHandle(ctx->gather_peer0);
    const bool hadGather = ctx->gather_peer0 != nullptr || ctx->film_peers_open;
    ctx->gather_peer0 = nullptr;
    film_peers_close(ctx);
    const bool was = ctx->p2p_open || hadGather;
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
    if ((s = ensure(ctx, ctx->counters_d, 32 * (size_t)(R + 1) + 64)) != BRE_OK) return s;
    uint64_t *cnt_d = (uint64_t *)ctx->counters_d.p;
    const bool moved = !ctx->p2p_open || ctx->p2p_exported[0] != ctx->pos_r.p ||
                       ctx->p2p_exported[1] != ctx->pow.p || ctx->p2p_exported[2] != ctx->dir.p;
    // {photons traced, "my store moved", my store's capacity}: everything the collective decisions below need,
    // so that every rank takes the same branches whatever its own buffers look like
    const size_t myCap = std::min(ctx->pos_r.cap, std::min(ctx->pow.cap, ctx->dir.cap));
    uint64_t mine[3] = {v.total, moved ? 1ull : 0ull, (uint64_t)myCap};
    BRE_CUDA(cudaMemcpyAsync(cnt_d + 3 * R, mine, 24, cudaMemcpyHostToDevice, ctx->stream));
    int r = g.AllGather(cnt_d + 3 * R, cnt_d, 3, kNcclUint64, comm, ctx->stream);
    if (r != 0) return nccl_err(ctx, r, "ncclAllGather(counts)");
    std::vector<uint64_t> got(3 * (size_t)R), off(R + 1, 0);
    BRE_CUDA(cudaMemcpyAsync(got.data(), cnt_d, 24 * (size_t)R, cudaMemcpyDeviceToHost, ctx->stream));
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    bool anyMoved = false;
    uint64_t minCap = ~0ull;
    for (int i = 0; i < R; ++i) {
        off[i + 1] = off[i] + got[3 * i];
        anyMoved |= got[3 * i + 1] != 0;
        minCap = std::min(minCap, got[3 * i + 2]);
    }
    const uint64_t total = off[R];
    if (total >= BRE_MAX_PHOTONS) return set_err(ctx, BRE_ERR_UNSUPPORTED, "photon count exceeds 2^27");
    // ---- capacity: ONE decision for all ranks (total and the smallest store are global), with 25 % head room
    const size_t need = (size_t)total * 16;
    const bool grow = minCap < need;
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


// ---------------------------------------------------------------- film gather (tile-sharded camera pass)
// Every rank holds the accumulators of the 16x16 tiles bre_tile_owner gives it (integrators/photonbeam.cpp:
// 345-347,444-477: the reference's tiles own their pixels too).  k_film_resolve_tiles resolves them
// (L = Ld / (iter + 1), photonbeam.cpp:578; Film::SetImage / WriteImage arithmetic, core/film.cpp:132-140,168-204)
// and writes the result straight into RANK 0's image buffer -- the local one on rank 0, a CUDA-IPC mapping over
// NVLink on the others -- so the gather IS the resolve: no staging copy, no reduction (tiles are disjoint).
// A one-word all-gather afterwards is the arrival barrier.  Rank 0's buffer is double-buffered by call parity:
// a fast rank may already push image k + 1 while rank 0 still copies image k to the host.
__device__ __forceinline__ float3 film_convert(float r, float g, float b, float scale, int meanOnly) {
    if (meanOnly) return make_float3(r, g, b);
    // RGBToXYZ spectrum.h:62-66, XYZToRGB spectrum.h:56-60, max(0, .) * scale film.cpp:178-201
    float X = fadd(fadd(fmul(0.412453f, r), fmul(0.357580f, g)), fmul(0.180423f, b));
    float Y = fadd(fadd(fmul(0.212671f, r), fmul(0.715160f, g)), fmul(0.072169f, b));
    float Z = fadd(fadd(fmul(0.019334f, r), fmul(0.119193f, g)), fmul(0.950227f, b));
    float R = fsub(fsub(fmul(3.240479f, X), fmul(1.537150f, Y)), fmul(0.498535f, Z));
    float G = fadd(fadd(fmul(-0.969256f, X), fmul(1.875991f, Y)), fmul(0.041556f, Z));
    float B = fadd(fsub(fmul(0.055648f, X), fmul(0.204043f, Y)), fmul(1.057311f, Z));
    return make_float3(fmul(fmaxf(0.f, R), scale), fmul(fmaxf(0.f, G), scale), fmul(fmaxf(0.f, B), scale));
}
// one block per tile of the image, thread = pixel; tiles of other ranks return at once
__global__ void __launch_bounds__(256) k_film_resolve_tiles(const float4 *__restrict__ film, uint32_t imgW, uint32_t imgH,
                                                            uint32_t tilesX, float div, float scale, int meanOnly,
                                                            int nRanks, int rank, float *__restrict__ dst) {
    const uint32_t tile = blockIdx.x;
    if (nRanks > 1 && (int)(tile % (uint32_t)nRanks) != rank) return;
    const uint32_t x = (tile % tilesX) * 16u + (threadIdx.x & 15u), y = (tile / tilesX) * 16u + (threadIdx.x >> 4);
    if (x >= imgW || y >= imgH) return;
    const size_t i = (size_t)y * imgW + x;
    const float4 f = __ldg(film + i);
    const float3 v = film_convert(fdiv(f.x, div), fdiv(f.y, div), fdiv(f.z, div), scale, meanOnly);
    dst[3 * i + 0] = v.x; dst[3 * i + 1] = v.y; dst[3 * i + 2] = v.z;
}


// ---------------------------------------------------------------- film gather (photon-sharded render)
// Radiance is a sum over photons (integrators/photonbeam.cpp:494-508), so R ranks that each hold 1/R of the photon
// paths produce R partial films whose sum is the film of the whole map.  The sum is a reduce-scatter written by hand
// over NVLink: rank r owns a contiguous run of pixels, PULLS that run from every rank's film through CUDA-IPC
// mappings (adds in rank order: deterministic for given partials), resolves it (photonbeam.cpp:578, core/film.cpp:
// 168-204) and writes the finished triples into rank 0's image -- 16 + 12 bytes per pixel cross the switch once,
// instead of the (R - 1) x 16 of a reduce onto rank 0.  Blocks stage their 256 pixels in shared memory so that the
// remote stores are full 16-byte vectors.
struct FilmPeers { const float4 *film[MAX_P2P_RANKS]; };
__global__ void __launch_bounds__(256) k_film_sum_resolve(FilmPeers P, int nRanks, uint32_t first, uint32_t last, float div,
                                                         float scale, int meanOnly, float *__restrict__ dst) {
    __shared__ __align__(16) float out[768];
    const uint32_t base = first + blockIdx.x * 256u, i = base + threadIdx.x;   // base is a multiple of 256
    if (i < last) {
        float4 a = P.film[0][i];
        for (int r = 1; r < nRanks; ++r) {
            const float4 b = P.film[r][i];
            a.x = fadd(a.x, b.x); a.y = fadd(a.y, b.y); a.z = fadd(a.z, b.z);
        }
        const float3 v = film_convert(fdiv(a.x, div), fdiv(a.y, div), fdiv(a.z, div), scale, meanOnly);
        out[3 * threadIdx.x] = v.x; out[3 * threadIdx.x + 1] = v.y; out[3 * threadIdx.x + 2] = v.z;
    }
    __syncthreads();
    const uint32_t nPix = min(256u, last - base), nFloats = 3u * nPix;
    float *d = dst + 3 * (size_t)base;                                         // 3072-byte aligned
    if (threadIdx.x < nFloats / 4) ((float4 *)d)[threadIdx.x] = ((const float4 *)out)[threadIdx.x];
    if (threadIdx.x < (nFloats & 3u)) d[(nFloats & ~3u) + threadIdx.x] = out[(nFloats & ~3u) + threadIdx.x];
}

static int film_gather_sum(bre_ctx *ctx, int iter, float scale, float *rgb_out_host, int meanOnly) {
    const uint32_t n = ctx->n_pixels;
    const int R = ctx->n_ranks;
    const float div = (float)(iter + 1);     // the DIVISOR: the kernels divide like the reference (spectrum.h:181-188)
    const size_t imgBytes = (size_t)n * 12, imgStride = (imgBytes + 255) & ~(size_t)255;
    ncclComm_p comm = (ncclComm_p)ctx->nccl_comm;
    int s;
    const size_t need = 2 * imgStride + (size_t)n * 16;
    if (ctx->gather.p && ctx->gather.cap < need && ctx->gather_exported == ctx->gather.p) {
        ctx->p2p_zombies.push_back(ctx->gather.p);      // peers still map it: freed after the next remap
        ctx->gather.p = nullptr; ctx->gather.cap = 0;
    }
    if ((s = ensure(ctx, ctx->gather, need)) != BRE_OK) return s;
    if ((s = ensure(ctx, ctx->counters_d, 16 * (size_t)(R + 1) + 64)) != BRE_OK) return s;
    uint64_t *cnt_d = (uint64_t *)ctx->counters_d.p;
    const size_t half = (size_t)(ctx->gather_calls & 1u) * imgStride;
    ++ctx->gather_calls;
    if (ctx->p2p_disabled || R > MAX_P2P_RANKS) {
        // NCCL path: ncclReduce(sum) of the float4 accumulators onto rank 0, which resolves
        float4 *sum = (float4 *)((char *)ctx->gather.p + 2 * imgStride);
        int r = g.Reduce(ctx->film.p, sum, (size_t)n * 4, kNcclFloat, 0 /* ncclSum */, 0, comm, ctx->stream);
        if (r != 0) return nccl_err(ctx, r, "ncclReduce(film sum)");
        if (ctx->rank == 0) {
            float *loc = (float *)((char *)ctx->gather.p + half);
            FilmPeers P{};
            P.film[0] = sum;
            k_film_sum_resolve<<<(n + 255) / 256, 256, 0, ctx->stream>>>(P, 1, 0, n, div, scale, meanOnly, loc);
            BRE_CUDA(cudaGetLastError());
            if (rgb_out_host) BRE_CUDA(cudaMemcpyAsync(rgb_out_host, loc, imgBytes, cudaMemcpyDeviceToHost, ctx->stream));
        }
        BRE_CUDA(cudaStreamSynchronize(ctx->stream));
        return BRE_OK;
    }
    // ---- arrival of every rank's camera pass + "my film / image buffer moved" flags, one all-gather
    const bool moved = !ctx->film_peers_open || ctx->film_exported != ctx->film.p || ctx->gather_mapped_pixels != n ||
                       ctx->gather_exported != ctx->gather.p;
    uint64_t mine = moved ? 1ull : 0ull;
    BRE_CUDA(cudaMemcpyAsync(cnt_d + 2 * R, &mine, 8, cudaMemcpyHostToDevice, ctx->stream));
    int r = g.AllGather(cnt_d + 2 * R, cnt_d, 1, kNcclUint64, comm, ctx->stream);
    if (r != 0) return nccl_err(ctx, r, "ncclAllGather(film flags)");
    std::vector<uint64_t> got(R);
    BRE_CUDA(cudaMemcpyAsync(got.data(), cnt_d, 8 * (size_t)R, cudaMemcpyDeviceToHost, ctx->stream));
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    bool anyMoved = false;
    for (int q = 0; q < R; ++q) anyMoved |= got[q] != 0;
    if (anyMoved) {
        // every rank: unmap, export {film, image buffer}, all-gather the handles, map
        film_peers_close(ctx);
        if (ctx->gather_peer0 && ctx->rank != 0) cudaIpcCloseMemHandle(ctx->gather_peer0);
        ctx->gather_peer0 = nullptr;
        cudaIpcMemHandle_t h[2];
        BRE_CUDA(cudaIpcGetMemHandle(&h[0], ctx->film.p));
        BRE_CUDA(cudaIpcGetMemHandle(&h[1], ctx->gather.p));
        const size_t hb = sizeof(h);
        if ((s = ensure(ctx, ctx->p2p_scratch, hb * (size_t)(R + 1))) != BRE_OK) return s;
        char *d = (char *)ctx->p2p_scratch.p;
        BRE_CUDA(cudaMemcpyAsync(d + hb * R, h, hb, cudaMemcpyHostToDevice, ctx->stream));
        r = g.AllGather(d + hb * R, d, hb, 0 /* ncclChar */, comm, ctx->stream);
        if (r != 0) return nccl_err(ctx, r, "ncclAllGather(film ipc handles)");
        std::vector<cudaIpcMemHandle_t> all(2 * (size_t)R);
        BRE_CUDA(cudaMemcpyAsync(all.data(), d, hb * R, cudaMemcpyDeviceToHost, ctx->stream));
        BRE_CUDA(cudaStreamSynchronize(ctx->stream));
        for (int q = 0; q < R; ++q) {
            if (q == ctx->rank) { ctx->film_peer[q] = ctx->film.p; continue; }
            void *ptr = nullptr;
            cudaError_t e = cudaIpcOpenMemHandle(&ptr, all[2 * q], cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) {
                cudaGetLastError();
                return set_err(ctx, BRE_ERR_CUDA, std::string("cudaIpcOpenMemHandle(peer film): ") + cudaGetErrorString(e));
            }
            ctx->film_peer[q] = ptr;
        }
        if (ctx->rank == 0) ctx->gather_peer0 = ctx->gather.p;
        else {
            void *ptr = nullptr;
            cudaError_t e = cudaIpcOpenMemHandle(&ptr, all[1], cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) {
                cudaGetLastError();
                return set_err(ctx, BRE_ERR_CUDA, std::string("cudaIpcOpenMemHandle(film image): ") + cudaGetErrorString(e));
            }
            ctx->gather_peer0 = ptr;
        }
        ctx->film_peers_open = true;
        ctx->film_exported = ctx->film.p;
        ctx->gather_mapped_pixels = n;
        ctx->gather_exported = ctx->gather.p;
        // the old films / image buffers are unmapped everywhere once every rank got here
        for (void *z : ctx->p2p_zombies) cudaFree(z);
        ctx->p2p_zombies.clear();
    }
    // ---- pixel run of this rank (multiples of 256 pixels: 16-byte aligned triples)
    const uint32_t per = (uint32_t)((((uint64_t)n + R - 1) / R + 255) / 256 * 256);
    const uint32_t first = (uint32_t)std::min<uint64_t>((uint64_t)per * ctx->rank, n);
    const uint32_t last = (uint32_t)std::min<uint64_t>((uint64_t)first + per, n);
    float *dst = (float *)((char *)ctx->gather_peer0 + half);
    if (last > first) {
        FilmPeers P{};
        for (int q = 0; q < R; ++q) P.film[q] = (const float4 *)ctx->film_peer[q];
        k_film_sum_resolve<<<(last - first + 255) / 256, 256, 0, ctx->stream>>>(P, R, first, last, div, scale, meanOnly, dst);
        BRE_CUDA(cudaGetLastError());
    }
    // ---- every run has landed in rank 0's image, and nobody writes its film again before everyone has read it
    r = g.AllGather(cnt_d + 2 * R, cnt_d, 1, kNcclUint64, comm, ctx->stream);
    if (r != 0) return nccl_err(ctx, r, "ncclAllGather(film barrier)");
    if (ctx->rank == 0 && rgb_out_host) BRE_CUDA(cudaMemcpyAsync(rgb_out_host, dst, imgBytes, cudaMemcpyDeviceToHost, ctx->stream));
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    return BRE_OK;
}

int film_gather(bre_ctx *ctx, int iter, float scale, float *rgb_out_host, int meanOnly) {
    if (!ctx->film.p) return set_err(ctx, BRE_ERR_INVALID, "bre_film_create has not been called");
    const uint32_t n = ctx->n_pixels;
    if (ctx->grid_w == 0 || (uint64_t)ctx->grid_w * ctx->grid_h != n)
        return set_err(ctx, BRE_ERR_INVALID, "bre_film_gather needs the image layout (bre_set_ray_grid width * height == film pixels)");
    const uint32_t W = ctx->grid_w, H = ctx->grid_h, tX = (W + 15) / 16, tY = (H + 15) / 16;
    const int R = ctx->n_ranks;
    const float div = (float)(iter + 1);     // the DIVISOR: the kernels divide like the reference (spectrum.h:181-188)
    const size_t imgBytes = (size_t)n * 12;
    int s;
    if ((s = ensure(ctx, ctx->gather, 2 * imgBytes)) != BRE_OK) return s;
    if (R > 1 && ctx->photon_shard) {
        if (!ctx->nccl_comm) return set_err(ctx, BRE_ERR_INVALID, "context was created without NCCL");
        return film_gather_sum(ctx, iter, scale, rgb_out_host, meanOnly);
    }
    if (R == 1 || !ctx->tile_shard) {
        k_film_resolve_tiles<<<tX * tY, 256, 0, ctx->stream>>>((const float4 *)ctx->film.p, W, H, tX, div, scale, meanOnly, 1, 0,
                                                              (float *)ctx->gather.p);
        BRE_CUDA(cudaGetLastError());
        if (rgb_out_host) BRE_CUDA(cudaMemcpyAsync(rgb_out_host, ctx->gather.p, imgBytes, cudaMemcpyDeviceToHost, ctx->stream));
        BRE_CUDA(cudaStreamSynchronize(ctx->stream));
        return BRE_OK;
    }
    if (!ctx->nccl_comm) return set_err(ctx, BRE_ERR_INVALID, "context was created without NCCL");
    // (rgb_out_host may be NULL on rank 0 too: the assembled image then stays in rank 0's device buffer)
    ncclComm_p comm = (ncclComm_p)ctx->nccl_comm;
    if ((s = ensure(ctx, ctx->counters_d, 16 * (size_t)(R + 1) + 64)) != BRE_OK) return s;
    uint64_t *cnt_d = (uint64_t *)ctx->counters_d.p;
    const size_t half = (size_t)(ctx->gather_calls & 1u) * imgBytes;
    ++ctx->gather_calls;
    if (ctx->p2p_disabled || R > MAX_P2P_RANKS) {
        // NCCL path: every rank resolves its tiles into a zeroed local image, ncclReduce(sum) to rank 0 (disjoint tiles)
        float *loc = (float *)((char *)ctx->gather.p + half);
        BRE_CUDA(cudaMemsetAsync(loc, 0, imgBytes, ctx->stream));
        k_film_resolve_tiles<<<tX * tY, 256, 0, ctx->stream>>>((const float4 *)ctx->film.p, W, H, tX, div, scale, meanOnly, R,
                                                              ctx->rank, loc);
        BRE_CUDA(cudaGetLastError());
        int r = g.Reduce(loc, loc, (size_t)n * 3, kNcclFloat, 0 /* ncclSum */, 0, comm, ctx->stream);
        if (r != 0) return nccl_err(ctx, r, "ncclReduce(film)");
        if (ctx->rank == 0 && rgb_out_host) BRE_CUDA(cudaMemcpyAsync(rgb_out_host, loc, imgBytes, cudaMemcpyDeviceToHost, ctx->stream));
        BRE_CUDA(cudaStreamSynchronize(ctx->stream));
        return BRE_OK;
    }
    // ---- rank 0's image buffer mapped on every rank (once per film size: every rank takes this branch together)
    if (ctx->gather_mapped_pixels != n || ctx->gather_exported != ctx->gather.p) {
        if (ctx->gather_peer0 && ctx->rank != 0) { cudaIpcCloseMemHandle(ctx->gather_peer0); }
        ctx->gather_peer0 = nullptr;
        cudaIpcMemHandle_t h;
        memset(&h, 0, sizeof(h));
        if (ctx->rank == 0) BRE_CUDA(cudaIpcGetMemHandle(&h, ctx->gather.p));
        if ((s = ensure(ctx, ctx->p2p_scratch, 4096)) != BRE_OK) return s;
        char *d = (char *)ctx->p2p_scratch.p;
        BRE_CUDA(cudaMemcpyAsync(d, &h, sizeof(h), cudaMemcpyHostToDevice, ctx->stream));
        int r = g.Broadcast(d, d, sizeof(h), 0 /* ncclChar */, 0, comm, ctx->stream);
        if (r != 0) return nccl_err(ctx, r, "ncclBroadcast(film ipc handle)");
        BRE_CUDA(cudaMemcpyAsync(&h, d, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
        BRE_CUDA(cudaStreamSynchronize(ctx->stream));
        if (ctx->rank == 0) ctx->gather_peer0 = ctx->gather.p;
        else {
            void *ptr = nullptr;
            cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) {
                cudaGetLastError();
                return set_err(ctx, BRE_ERR_CUDA, std::string("cudaIpcOpenMemHandle(film): ") + cudaGetErrorString(e));
            }
            ctx->gather_peer0 = ptr;
        }
        ctx->gather_mapped_pixels = n;
        ctx->gather_exported = ctx->gather.p;
    }
    float *dst = (float *)((char *)ctx->gather_peer0 + half);
    k_film_resolve_tiles<<<tX * tY, 256, 0, ctx->stream>>>((const float4 *)ctx->film.p, W, H, tX, div, scale, meanOnly, R, ctx->rank,
                                                          dst);
    BRE_CUDA(cudaGetLastError());
    // arrival barrier: every rank's tiles have landed in rank 0's buffer
    int r = g.AllGather(cnt_d + 2 * R, cnt_d, 1, kNcclUint64, comm, ctx->stream);
    if (r != 0) return nccl_err(ctx, r, "ncclAllGather(film barrier)");
    if (ctx->rank == 0 && rgb_out_host) BRE_CUDA(cudaMemcpyAsync(rgb_out_host, dst, imgBytes, cudaMemcpyDeviceToHost, ctx->stream));
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    return BRE_OK;
}

}  // namespace bre
