// bre_internal.cuh -- shared declarations of libbre.so (sm_100a only).
//
// Data layout in HBM (all float4 / 16-byte aligned, SoA):
//   photons (store order)   pos_r[n] {x,y,z,radius}  pow[n] {r,g,b,0}  dir[n] {wx,wy,wz,0}
//   photons (Morton order)  s_pos_r, s_pow, s_dir + order[n] (sorted position -> store index)
//   canonical LBVH          node_lo[n-1] {lo.xyz, left}  node_hi[n-1] {hi.xyz, right}
//                           leaf_lo[n], leaf_hi[n] float4, node_range[n-1] {first,last}
//   production tree         fast[n-1] 64-byte two-child nodes (boxes of both children + refs)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/bre.h"

namespace bre {

// ------------------------------------------------------------------ float helpers
// Everything that decides a photon hit or a Morton code uses these explicitly
// rounded, never-contracted operations so that the results are bit-identical to
// the CPU oracle (compiled -ffp-contract=off).
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }

__device__ __forceinline__ float next_up(float x) {   // nextafterf(x, +inf), finite x
    if (x == 0.f) return __int_as_float(1);
    int b = __float_as_int(x);
    return __int_as_float(x > 0.f ? b + 1 : b - 1);
}
__device__ __forceinline__ float next_down(float x) { // nextafterf(x, -inf), finite x
    if (x == 0.f) return __int_as_float(0x80000001);
    int b = __float_as_int(x);
    return __int_as_float(x > 0.f ? b - 1 : b + 1);
}

#define BRE_PI 3.14159265358979323846f
#define BRE_INV_PI 0.31830988618379067154f
#define BRE_INV4PI 0.07957747154594766788f
#define BRE_ONE_MINUS_EPS 0.99999994f
// 1 + 2*gamma(3), core/pbrt.h:263-265 with MachineEpsilon = 2^-24
#define BRE_SLAB_WIDEN 1.00000035762793f

// ------------------------------------------------------------------ device structs
struct MediumDev {
    int kind;                 // 0 homogeneous, 1 grid
    float sigma_a[3], sigma_s[3], sigma_t[3];
    float g;
    int gray;                 // sigma_t equal in all three channels
    int nx, ny, nz;
    const float *density;     // device
    float invMaxDensity;
    float w2m[12];            // affine rows
};

struct LightDev {
    int kind;
    float pos[3], I[3], axis[3], ax[3], ay[3];
    float cosTotalWidth, cosFalloffStart;
};
#define BRE_MAX_LIGHTS 64
struct LightsDev {
    int n;
    float cdf[BRE_MAX_LIGHTS + 1];
    float func[BRE_MAX_LIGHTS];
    float funcInt;
    LightDev l[BRE_MAX_LIGHTS];
};

// Node of the production tree.
// Default: the 64-byte two-child node (boxes of both children + child references).
// BRE_WIDE=1 (make NVEXTRA=-DBRE_WIDE=1): 128 bytes = one cache line, up to FOUR children -- the two
// children of the binary radix-tree node, each replaced by ITS two children when it is itself a
// kept internal node.  Box k = floats [6k, 6k+6) of q: lo.x lo.y lo.z hi.x hi.y hi.z; unused slots
// have ref FAST_NONE.  Measured on smoke10m_1080p: nodes per ray 34.0 -> 17.0, half the walk steps,
// but the SAME number of box tests (68 per ray) and the same bytes: K5a 1.31 -> 1.31 ms, k_emit_fast
// +0.10 ms.  K5a is bound by the slab-test instructions, not by step overhead; parity tests pass
// with either layout.
// ref >= 0: internal node index;  ref < 0: leaf, see leaf_ref().
#ifndef BRE_WIDE
#define BRE_WIDE 0
#endif
#define FAST_NONE 0x7fffffff
#if BRE_WIDE
struct __align__(16) FastNode {
    float4 q[6];
    int4 ref;
    int4 pad;
};
#else
struct __align__(16) FastNode {
    float4 a;  // c0.lo.x c0.lo.y c0.lo.z c0.hi.x
    float4 b;  // c0.hi.y c0.hi.z c1.lo.x c1.lo.y
    float4 c;  // c1.lo.z c1.hi.x c1.hi.y c1.hi.z
    int4 d;    // ref0 ref1 - -
};
#endif
// leaf reference: bit31 set, (count-1) in bits 27..30, first sorted photon in bits 0..26
__host__ __device__ __forceinline__ int leaf_ref(uint32_t first, uint32_t count) {
    return (int)(0x80000000u | ((count - 1u) << 27) | first);
}
#define BRE_MAX_PHOTONS (1u << 27)
// entries of the top cut of the production tree (k_top_cut in bre_build.cu, staged by TMA in k_traverse_tile)
constexpr int TOP_CUT = 512;

struct TreeDev {
    uint32_t n;               // photons
    const float4 *s_pos_r, *s_pow, *s_dir;   // s_pow / s_dir == nullptr: payload not gathered (lazy), read it through order
    const float4 *pow_u, *dir_u;             // the photon store's power / direction arrays (store order)
    const uint32_t *order;
    const float4 *node_lo, *node_hi;   // .w = child refs as int bits
    const float4 *leaf_lo, *leaf_hi;
    const FastNode *fast;
    int root_ref;             // fast-path root: internal 0, or a leaf ref when n <= max leaf
    float abs_max;            // max |coordinate| of any leaf box
};

// The eight corners of the INTERIOR cell (ix, iy, iz) of a density grid, with 32-bit index arithmetic (the grid holds
// fewer than 2^31 voxels, checked by bre_set_medium_grid): four pointers = four IMAD.WIDE, where size_t indexing of
// signed coordinates cost ~25 integer instructions per sample in K1's tracker and K6.
__device__ __forceinline__ void grid_cell_corners(const MediumDev &m, int ix, int iy, int iz, float &v000, float &v100,
                                                  float &v010, float &v110, float &v001, float &v101, float &v011,
                                                  float &v111) {
    const unsigned nx = (unsigned)m.nx, zs = nx * (unsigned)m.ny;
    const unsigned i0 = ((unsigned)iz * (unsigned)m.ny + (unsigned)iy) * nx + (unsigned)ix;
    const float *b0 = m.density + i0, *b1 = m.density + (i0 + nx), *b2 = m.density + (i0 + zs), *b3 = m.density + (i0 + zs + nx);
    v000 = __ldg(b0); v100 = __ldg(b0 + 1); v010 = __ldg(b1); v110 = __ldg(b1 + 1);
    v001 = __ldg(b2); v101 = __ldg(b2 + 1); v011 = __ldg(b3); v111 = __ldg(b3 + 1);
}

// Power / direction of the photon at sorted position s.  The image-tile path stages them once per (tile, photon)
// and reads them through the Morton permutation straight from the store (bre_build then moves only the 16-byte
// position records); the per-ray paths gather them first (ensure_sorted_payload).
__device__ __forceinline__ void photon_payload(const TreeDev &t, uint32_t s, float4 &pw, float4 &dr) {
    if (t.s_pow) { pw = __ldg(t.s_pow + s); dr = __ldg(t.s_dir + s); }
    else { const uint32_t i = __ldg(t.order + s); pw = __ldg(t.pow_u + i); dr = __ldg(t.dir_u + i); }
}

// ------------------------------------------------------------------ context
struct Buf {
    void *p = nullptr;
    size_t cap = 0;
};

}  // namespace bre

struct bre_ctx {
    int device = 0, n_ranks = 1, rank = 0;
    int max_leaf = 8, tau_steps = 64;
    bool tau_auto = true;                   // bre_config.tau_steps == 0: chosen from the grid in bre_set_medium_grid
    int refit_direct = 16;                  // nodes up to this many photons get their box straight from the photons
    int sm_count = 148;
    uint64_t pair_queue_cap_override = 0;   // BRE_PAIR_QUEUE_CAP (tests force the overflow path)
    cudaStream_t stream = nullptr;
    std::string err;

    // scene
    bre::MediumDev medium{};
    bool have_medium = false;
    bre::Buf density;
    bre::LightsDev *lights_d = nullptr;
    int n_lights = 0;
    float blo[3] = {0, 0, 0}, bhi[3] = {0, 0, 0};
    bool have_bounds = false;
    // surfaces (bre_surface.cuh) and camera (bre_camera.cu)
    uint32_t n_tri = 0;
    bre::Buf tri, tri_kdb;
    int lights_in_medium = 1;
    bool compact_tiled = true;              // BRE_COMPACT_TILED=0: one-thread-per-path k_compact (A/B)
    bool sort_onesweep = true;              // BRE_SORT_ONESWEEP=0: the three-kernel radix passes (A/B)
    bre::Buf os_scratch;
    bool lazy_payload = true;               // BRE_LAZY_PAYLOAD=0: bre_build also gathers power / direction into Morton order
    bool payload_sorted = false;            // s_pow / s_dir hold the Morton-ordered payload of the current build
    bool gather_aos = true;                 // BRE_GATHER_AOS=0: gather straight from the three store arrays
    bool trace_nested = false;              // BRE_TRACE_NESTED=1: the nested-loop K1 (A/B experiments)
    bre_camera camera{};
    bool have_camera = false;

    // photon store (store order)
    uint64_t n_photons = 0;
    bre::Buf pos_r, pow, dir;
    // build products
    bool built = false;
    bool canon_valid = false;    // canonical leaf / small-node boxes materialised (ensure_canonical)
    bool topo_valid = false;     // order/topology of the last full build still match the photon positions
    int last_build_kind = 0;     // 0 none, 1 full build, 2 refit only
    uint32_t n_built = 0;
    bre::Buf s_pos_r, s_pow, s_dir, order, codes, keys_alt, vals_alt, hist, node_lo, node_hi,
        node_range, parent, leaf_lo, leaf_hi, flags, fast, scratch, entry, aos;
    float abs_max = 0;
    int root_ref = 0;
    // query scratch
    bre::Buf packet_ctr, pair_queue, heavy, frames;
    // image-tile path of K5 (bre_query_tile.cuh): ray q = pixel (q % grid_w, q / grid_w) when grid_w * grid_h == n
    uint32_t grid_w = 0, grid_h = 0;
    bool tile_path = true;                  // BRE_TILE=0: always the packet walk
    bool tile_shard = false;                // bre_set_tile_sharding: this rank answers only the tiles bre_tile_owner gives it
    int k5t_blocks_per_sm = 0;              // BRE_K5T_BLOCKS: cap on resident k_traverse_tile blocks per SM
    cudaEvent_t ev_est = nullptr;           // recorded after the copy of the queue control words into est_host
    bre::Buf tframes, trays, tjit, tile_state, top_cut, tile_wq;
    uint64_t tile_wq_items = 0, tile_wq_cap_override = 0;   // BRE_TILE_WQ_CAP (tests force the depth-first spill)
    // heavy-first packet order fed from one radiance query to the next (bre_query.cu)
    size_t heavy_half = 0;
    uint32_t heavy_rays = 0;
    int heavy_parity = 0;
    bool heavy_tile = false;                // the heavy lists hold tile indices (tile path) instead of packet indices
    // ray batching of K5 (bre_query.cu launch_query)
    uint32_t ray_batch_override = 0;        // BRE_RAY_BATCH (experiments)
    int k5a_blocks_per_sm = 0;              // BRE_K5A_BLOCKS: cap on resident K5a blocks per SM (0 = what fits)
    int k5b_blocks_per_sm = 0;              // BRE_K5B_BLOCKS: K5b grid = SMs x this; 0 = from the ray count (bre_query.cu)
    float fat_spread = 1.f / 64.f;          // BRE_FAT_SPREAD: largest |d_i - d_0| of a packet walked as one fat ray (0 = off)
    uint32_t ctr_batches = 0, ctr_batch_rays = 0, ctr_rays = 0;
    unsigned int *est_host = nullptr;       // pinned copy of the last batch's queue control words
    uint32_t est_batches = 0, est_batch_rays = 0, est_n = 0;
    bool est_tile = false;                  // the estimate was measured on the tile path
    // asynchronous ray staging (bre_stage_rays)
    bre::Buf rays_stage;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_stage = nullptr, ev_main = nullptr;
    uint32_t staged_rays = 0;
    // optical-depth table computed ahead of the query by bre_stage_* (K6 under the photon pass + build)
    struct TauKey { const void *ro, *rd; uint32_t n; int steps; uint32_t gridW; uint64_t mediumVersion; const float *tau; const float2 *span; };
    TauKey tau_key{};
    bool tau_staged = false;
    // K6 of staged rays is launched when bre_build starts (under the latency-bound sort / gather kernels, on the
    // low-priority staging stream) instead of under the issue-bound photon tracer: BRE_TAU_DEFER=0 launches it at once
    bool tau_defer = true;
    struct TauPending { bool on = false; uint32_t n = 0; const void *ro = nullptr, *rd = nullptr; uint32_t gridW = 0, gridH = 0; bool useGrid = false; } tau_pending;
    uint64_t medium_version = 0;
    bool cam_staged = false;                // K0's rays of (cam_offx, cam_offy) are in rays_o / rays_d (bre_stage_camera)
    float cam_offx = 0, cam_offy = 0;
    long long cam_halton = -1;              // >= 0: the staged rays are the Halton camera samples of that sample number
    cudaStream_t stream2 = nullptr;
    cudaEvent_t qev[4] = {};
    bre::Buf rays_o, rays_d, out_rgb, tau, hit_counts, hit_offsets, hit_idx, counters_d, pix_idx;
    bre_counters counters{};
    // film
    bre::Buf film;
    uint32_t n_pixels = 0;
    // trace scratch
    bre::Buf path_counts, path_offsets;
    // nccl
    void *nccl_comm = nullptr;
    // CUDA-IPC mappings of every rank's photon store (fused trace + push, bre_nccl.cpp)
    void *p2p_peer[8][3] = {};
    void *p2p_exported[3] = {};
    bool p2p_open = false;
    bool p2p_disabled = false;             // BRE_NO_P2P=1, or IPC mapping failed once
    bre::Buf p2p_scratch;
    // film gather of the tile-sharded camera pass (bre_film_gather): rank 0's image buffer, mapped on every rank
    bre::Buf gather;
    void *gather_peer0 = nullptr, *gather_exported = nullptr;
    uint32_t gather_mapped_pixels = 0, gather_calls = 0;
    // photon-sharded render (bre_set_photon_sharding): every rank keeps the photons it traced, builds its own LBVH
    // and answers every pixel from it; bre_film_gather sums the partial films over CUDA-IPC mappings of the films
    bool photon_shard = false;
    void *film_peer[8] = {};
    void *film_exported = nullptr;
    bool film_peers_open = false;
    std::vector<void *> p2p_zombies;       // exported stores replaced since: freed once peers closed them
    // per-phase device timing (bre_set_timing)
    bool timing = false;
    cudaEvent_t ev[8][2] = {};
    bool ev_valid[8] = {};
};

namespace bre {

int set_err(bre_ctx *c, int code, const std::string &msg);
int ensure(bre_ctx *c, Buf &b, size_t bytes);   // grow-only device allocation
int check_cuda(bre_ctx *c, cudaError_t e, const char *what);
#define BRE_CUDA(call)                                                    \
    do {                                                                  \
        int _s = bre::check_cuda(ctx, (call), #call);                     \
        if (_s != BRE_OK) return _s;                                      \
    } while (0)

// build.cu
int build_tree(bre_ctx *ctx);
int ensure_canonical(bre_ctx *ctx);
int ensure_sorted_payload(bre_ctx *ctx);   // per-ray query paths: gather power / direction into Morton order once per build
// query.cu
int run_query(bre_ctx *ctx, uint32_t n, const float4 *ro, const float4 *rd, float nPathsNorm,
              float4 *out);
int run_count(bre_ctx *ctx, uint32_t n, const float4 *ro, const float4 *rd);
int stage_tau(bre_ctx *ctx, uint32_t n, const float4 *ro, const float4 *rd, cudaStream_t st);
int stage_tau_deferred(bre_ctx *ctx, uint32_t n, const float4 *ro, const float4 *rd, bool useGrid, uint32_t gridW, uint32_t gridH);
int launch_pending_tau(bre_ctx *ctx);      // bre_build / the query entry points: start the deferred K6 now
int run_hitdump(bre_ctx *ctx, uint32_t n, const float4 *ro, const float4 *rd, bre_hitdump *hd);
// film.cu
int film_accumulate(bre_ctx *ctx, uint32_t n, const uint32_t *pix_d, const float4 *rgb_d);
int film_resolve(bre_ctx *ctx, int iter, float scale, float *rgb_out_host, int meanOnly);
// camera.cu
int camera_rays(bre_ctx *ctx, float offx, float offy, cudaStream_t st = nullptr, long long haltonSample = -1);
// trace.cu
constexpr uint64_t TRACE_CHUNK = 1u << 22;   // paths per trace launch (bounds the slot scratch)
struct SlotView {
    float4 *sp, *sw, *sd;      // per-path slots: path * maxDepth + k
    uint32_t *counts;          // photons deposited per path
    uint64_t *offs;            // exclusive scan of counts
    uint64_t total;
    uint32_t nPaths;
    int maxDepth;
};
int trace_to_slots(bre_ctx *ctx, uint64_t first, uint32_t nPaths, int maxDepth, float radius, SlotView *out);
int trace_photons(bre_ctx *ctx, uint64_t first, uint64_t nPaths, int maxDepth, float radius);
// nccl.cpp
int nccl_init(bre_ctx *ctx, const void *uid);
int nccl_allgather_photons(bre_ctx *ctx);
int trace_exchange_p2p(bre_ctx *ctx, uint64_t nPathsTotal, int maxDepth, float radius);
int film_gather(bre_ctx *ctx, int iter, float scale, float *rgb_out_host, int meanOnly);
int compact_tiled_local(bre_ctx *ctx, const SlotView &v, uint64_t storeBase);
void nccl_destroy(bre_ctx *ctx);
int nccl_unique_id(void *out128);

TreeDev tree_view(const bre_ctx *ctx);

enum Phase { PH_TRACE = 0, PH_BUILD = 1, PH_TAU = 2, PH_QUERY = 3, PH_FILM = 4, PH_QUERY_A = 5, PH_QUERY_B = 6 };
// RAII event bracket around a phase on the context stream (no-op unless timing is enabled)
struct PhaseTimer {
    bre_ctx *c; int ph; cudaStream_t st;
    PhaseTimer(bre_ctx *ctx, int phase, cudaStream_t stream = nullptr) : c(ctx), ph(phase), st(stream ? stream : ctx->stream) {
        if (c->timing) cudaEventRecord(c->ev[ph][0], st);
    }
    ~PhaseTimer() {
        if (c->timing) { cudaEventRecord(c->ev[ph][1], st); c->ev_valid[ph] = true; }
    }
};

}  // namespace bre
