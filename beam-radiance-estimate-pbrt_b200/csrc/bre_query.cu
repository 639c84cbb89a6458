// bre_query.cu -- K5 beam query (+ K6 optical-depth table for grid media).
//
// Replaces (reference src/): PhotonBeamBVH::Intersect (core/photonbeambvh.cpp:685-723) with
// Bounds3::IntersectP (core/geometry.h:1410-1438), the per-candidate loop of the camera pass
// (integrators/photonbeam.cpp:494-508), kernel() (integrators/photonmap.cpp:170-174, normalised as
// :198), PhaseHG (core/medium.h:66-69) and Medium::Tr (media/homogeneous.cpp:44-48; for grid media a
// deterministic optical-depth table instead of the stochastic media/grid.cpp:89-121).
#include "bre_internal.cuh"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace bre {

// ---------------------------------------------------------------- per-ray state
struct RayS {
    float ox, oy, oz, tMax, dx, dy, dz;
    float ix, iy, iz;        // 1/d
    float nx, ny, nz;        // padded origin used for the near planes
    float fx, fy, fz;        // padded origin used for the far planes
    bool sx, sy, sz;         // direction negative
};

__device__ __forceinline__ RayS ray_setup(float4 o, float4 d, float absMax) {
    RayS r;
    r.ox = o.x; r.oy = o.y; r.oz = o.z; r.tMax = o.w;
    r.dx = d.x; r.dy = d.y; r.dz = d.z;
    r.ix = fdiv(1.f, d.x); r.iy = fdiv(1.f, d.y); r.iz = fdiv(1.f, d.z);
    r.sx = r.ix < 0.f; r.sy = r.iy < 0.f; r.sz = r.iz < 0.f;
    // conservative cull: pad every box by 2^-18 * (|o|_inf + max|coord|)  
    float M = fadd(fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z)), absMax);
    float pad = fmul(M, 0x1p-18f);
    float px = r.sx ? -pad : pad, py = r.sy ? -pad : pad, pz = r.sz ? -pad : pad;
    r.nx = fadd(o.x, px); r.fx = fsub(o.x, px);
    r.ny = fadd(o.y, py); r.fy = fsub(o.y, py);
    r.nz = fadd(o.z, pz); r.fz = fsub(o.z, pz);
    return r;
}

// Bounds3::IntersectP(ray, invDir, dirIsNeg), geometry.h:1410-1438, same operation order.
__device__ __forceinline__ bool slab_hit(const RayS &r, float lox, float loy, float loz, float hix,
                                         float hiy, float hiz) {
    float tMin = fmul(fsub(r.sx ? hix : lox, r.nx), r.ix);
    float tMax = fmul(fsub(r.sx ? lox : hix, r.fx), r.ix);
    float tyMin = fmul(fsub(r.sy ? hiy : loy, r.ny), r.iy);
    float tyMax = fmul(fsub(r.sy ? loy : hiy, r.fy), r.iy);
    tMax = fmul(tMax, BRE_SLAB_WIDEN);
    tyMax = fmul(tyMax, BRE_SLAB_WIDEN);
    if (tMin > tyMax || tyMin > tMax) return false;
    if (tyMin > tMin) tMin = tyMin;
    if (tyMax < tMax) tMax = tyMax;
    float tzMin = fmul(fsub(r.sz ? hiz : loz, r.nz), r.iz);
    float tzMax = fmul(fsub(r.sz ? loz : hiz, r.fz), r.iz);
    tzMax = fmul(tzMax, BRE_SLAB_WIDEN);
    if (tMin > tzMax || tzMin > tMax) return false;
    if (tzMin > tMin) tMin = tzMin;
    if (tzMax < tMax) tMax = tzMax;
    return (tMin < r.tMax) && (tMax > 0.f);
}

// Disk predicate (DESIGN.md "disk predicate"): bit-identical to oracle/bre_oracle.c orc_disk_test.
__device__ __forceinline__ bool disk_test(const RayS &r, float4 p, float &t, float &d2) {
    float vx = fsub(p.x, r.ox), vy = fsub(p.y, r.oy), vz = fsub(p.z, r.oz);
    t = fadd(fadd(fmul(vx, r.dx), fmul(vy, r.dy)), fmul(vz, r.dz));
    if (!(t > 0.f && t < r.tMax)) return false;
    float wx = fsub(vx, fmul(t, r.dx)), wy = fsub(vy, fmul(t, r.dy)), wz = fsub(vz, fmul(t, r.dz));
    d2 = fadd(fadd(fmul(wx, wx), fmul(wy, wy)), fmul(wz, wz));
    float r2 = fmul(p.w, p.w);
    return d2 < r2;
}

__device__ __forceinline__ float phase_hg(float cosTheta, float g) {  // medium.h:66-69
    float denom = fadd(fadd(1.f, fmul(g, g)), fmul(fmul(2.f, g), cosTheta));
    return fdiv(fmul(BRE_INV4PI, fsub(1.f, fmul(g, g))), fmul(denom, sqrtf(denom)));
}

struct TrState {
    const float *tau;   // this ray's table (grid) or nullptr
    float t0, t1;
    int steps;
};

__device__ __forceinline__ void eval_tr(const MediumDev &m, const TrState &ts, float t, float tr[3]) {
    if (m.kind == 0) {  // homogeneous.cpp:44-48
#pragma unroll
        for (int c = 0; c < 3; ++c) tr[c] = expf(-m.sigma_t[c] * t);
    } else {
        float od = 0.f;
        if (ts.t1 > ts.t0) {
            float dt = fdiv(fsub(ts.t1, ts.t0), (float)ts.steps);
            float s = fdiv(fsub(t, ts.t0), dt);
            if (s > 0.f) {
                if (s >= (float)ts.steps) od = ts.tau[ts.steps];
                else {
                    int k = (int)s;
                    float f = fsub(s, (float)k);
                    float a = ts.tau[k], b = ts.tau[k + 1];
                    od = fadd(a, fmul(f, fsub(b, a)));
                }
            }
        }
#pragma unroll
        for (int c = 0; c < 3; ++c) tr[c] = expf(-m.sigma_t[c] * od);
    }
}

// ---------------------------------------------------------------- K6: optical depth table
__device__ __forceinline__ float grid_D(const MediumDev &m, int x, int y, int z) {  // grid.h:78-82
    if (x < 0 || y < 0 || z < 0 || x >= m.nx || y >= m.ny || z >= m.nz) return 0.f;
    return __ldg(m.density + ((size_t)z * m.ny + y) * m.nx + x);
}
// Bounds3::IntersectP(ray,&t0,&t1) against the unit cube, geometry.h:1386-1408
__device__ __forceinline__ bool unit_box(const float o[3], const float d[3], float tMaxRay,
                                         float &t0, float &t1) {
    t0 = 0.f; t1 = tMaxRay;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float inv = fdiv(1.f, d[i]);
        float tNear = fmul(fsub(0.f, o[i]), inv);
        float tFar = fmul(fsub(1.f, o[i]), inv);
        if (tNear > tFar) { float s = tNear; tNear = tFar; tFar = s; }
        tFar = fmul(tFar, BRE_SLAB_WIDEN);
        t0 = tNear > t0 ? tNear : t0;
        t1 = tFar < t1 ? tFar : t1;
        if (t0 > t1) return false;
    }
    return true;
}
__device__ __forceinline__ void to_medium(const MediumDev &m, const float o[3], const float d[3],
                                          float om[3], float dm[3]) {
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const float *w = m.w2m + 4 * r;
        om[r] = fadd(fadd(fadd(fmul(w[0], o[0]), fmul(w[1], o[1])), fmul(w[2], o[2])), w[3]);
        dm[r] = fadd(fadd(fmul(w[0], d[0]), fmul(w[1], d[1])), fmul(w[2], d[2]));
    }
}

// K6's own density lookup, in SAMPLE space (s = p * n - 0.5): the same trilinear interpolant as
// GridDensityMedium::Density (grid.cpp:45-60) with every lerp as one fused a + t (b - a).  It differs from the
// reference's (1-t) a + t b form by rounding only (~1e-7 relative) -- the optical-depth table is
// compared with the oracle's within the radiance tolerance, never bit for bit -- and halves the
// instruction count of this ALU-bound kernel.  K1 keeps the reference-order form (grid_density_t).
__device__ __forceinline__ float lerp_f(float t, float a, float b) { return __fmaf_rn(t, b - a, a); }
__device__ __forceinline__ float grid_density_s(const MediumDev &m, float sx, float sy, float sz) {
    const float fx = floorf(sx), fy = floorf(sy), fz = floorf(sz);
    const int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    const float dx = sx - fx, dy = sy - fy, dz = sz - fz;
    float v000, v100, v010, v110, v001, v101, v011, v111;
    if ((unsigned)ix < (unsigned)(m.nx - 1) && (unsigned)iy < (unsigned)(m.ny - 1) &&
        (unsigned)iz < (unsigned)(m.nz - 1)) {
        grid_cell_corners(m, ix, iy, iz, v000, v100, v010, v110, v001, v101, v011, v111);
    } else {
        v000 = grid_D(m, ix, iy, iz); v100 = grid_D(m, ix + 1, iy, iz);
        v010 = grid_D(m, ix, iy + 1, iz); v110 = grid_D(m, ix + 1, iy + 1, iz);
        v001 = grid_D(m, ix, iy, iz + 1); v101 = grid_D(m, ix + 1, iy, iz + 1);
        v011 = grid_D(m, ix, iy + 1, iz + 1); v111 = grid_D(m, ix + 1, iy + 1, iz + 1);
    }
    const float d00 = lerp_f(dx, v000, v100), d10 = lerp_f(dx, v010, v110);
    const float d01 = lerp_f(dx, v001, v101), d11 = lerp_f(dx, v011, v111);
    return lerp_f(dz, lerp_f(dy, d00, d10), lerp_f(dy, d01, d11));
}

// One thread per ray; the block's rows are staged in shared memory (row stride steps+1, odd for
// the default 64 steps -> conflict-free both ways) and written out as one contiguous, fully
// coalesced slab instead of 65 stores per thread with a 260-byte stride.  The midpoint samples
// advance in sample space: s_k = s_0 + k ds, three FMAs per sample.
__global__ void __launch_bounds__(128) k_tau_table(MediumDev m, uint32_t n,
                                                   const float4 *__restrict__ ro,
                                                   const float4 *__restrict__ rd, int steps,
                                                   float *__restrict__ tau,
                                                   float2 *__restrict__ span) {
    extern __shared__ float rows[];
    const uint32_t q0 = blockIdx.x * blockDim.x, q = q0 + threadIdx.x;
    const int stride = steps + 1;
    float *row = rows + (size_t)threadIdx.x * stride;
    if (q < n) {
        float4 o4 = __ldg(ro + q), d4 = __ldg(rd + q);
        float o[3] = {o4.x, o4.y, o4.z}, d[3] = {d4.x, d4.y, d4.z}, om[3], dm[3];
        to_medium(m, o, d, om, dm);
        float t0, t1;
        if (!unit_box(om, dm, o4.w, t0, t1)) {
            for (int k = 0; k <= steps; ++k) row[k] = 0.f;
            span[q] = make_float2(0.f, 0.f);
        } else {
            const float dt = fdiv(fsub(t1, t0), (float)steps);
            const float th = __fmaf_rn(0.5f, dt, t0);                 // first midpoint
            const float nx = (float)m.nx, ny = (float)m.ny, nz = (float)m.nz;
            const float s0x = __fmaf_rn(__fmaf_rn(dm[0], th, om[0]), nx, -0.5f), dsx = dm[0] * dt * nx;
            const float s0y = __fmaf_rn(__fmaf_rn(dm[1], th, om[1]), ny, -0.5f), dsy = dm[1] * dt * ny;
            const float s0z = __fmaf_rn(__fmaf_rn(dm[2], th, om[2]), nz, -0.5f), dsz = dm[2] * dt * nz;
            float acc = 0.f;
            row[0] = 0.f;
            for (int k = 0; k < steps; ++k) {
                const float kf = (float)k;
                const float dens = grid_density_s(m, __fmaf_rn(kf, dsx, s0x), __fmaf_rn(kf, dsy, s0y),
                                                  __fmaf_rn(kf, dsz, s0z));
                acc = __fmaf_rn(dens, dt, acc);
                row[k + 1] = acc;
            }
            span[q] = make_float2(t0, t1);
        }
    }
    __syncthreads();
    const uint32_t rowsHere = min((uint32_t)blockDim.x, n - q0);
    const uint32_t cnt = rowsHere * (uint32_t)stride;
    float *dst = tau + (size_t)q0 * stride;
    for (uint32_t i = threadIdx.x; i < cnt; i += blockDim.x) dst[i] = rows[i];
}

// K6 for a ray grid (tile path of K5): one block per 16x16 image tile, thread = pixel of the tile.  The table is
// laid out [tile][step][pixel]: every step the block writes 256 consecutive floats (no shared-memory staging),
// and the hits of one photon -- a few neighbouring pixels at one depth -- read neighbouring words instead of one
// cache line per ray (K5b' is bound by L1 wavefronts, not by bytes).  Tiles of other ranks are skipped.
__global__ void __launch_bounds__(256) k_tau_tile(MediumDev m, uint32_t imgW, uint32_t imgH, uint32_t tilesX,
                                                  const float4 *__restrict__ ro, const float4 *__restrict__ rd,
                                                  int steps, float *__restrict__ tau, float2 *__restrict__ span,
                                                  int rankMod, int rankId) {
    const uint32_t tile = blockIdx.x;
    if (rankMod > 1 && (int)(tile % (uint32_t)rankMod) != rankId) return;
    const uint32_t x = (tile % tilesX) * 16u + (threadIdx.x & 15u), y = (tile / tilesX) * 16u + (threadIdx.x >> 4);
    if (x >= imgW || y >= imgH) return;
    const uint32_t q = y * imgW + x;
    float *col = tau + (size_t)tile * (size_t)(steps + 1) * 256u + threadIdx.x;
    const float4 o4 = __ldg(ro + q), d4 = __ldg(rd + q);
    float o[3] = {o4.x, o4.y, o4.z}, d[3] = {d4.x, d4.y, d4.z}, om[3], dm[3];
    to_medium(m, o, d, om, dm);
    float t0, t1;
    if (!unit_box(om, dm, o4.w, t0, t1)) {
        for (int k = 0; k <= steps; ++k) col[(size_t)k * 256u] = 0.f;
        span[q] = make_float2(0.f, 0.f);
        return;
    }
    const float dt = fdiv(fsub(t1, t0), (float)steps);
    const float th = __fmaf_rn(0.5f, dt, t0);                 // first midpoint
    const float nx = (float)m.nx, ny = (float)m.ny, nz = (float)m.nz;
    const float s0x = __fmaf_rn(__fmaf_rn(dm[0], th, om[0]), nx, -0.5f), dsx = dm[0] * dt * nx;
    const float s0y = __fmaf_rn(__fmaf_rn(dm[1], th, om[1]), ny, -0.5f), dsy = dm[1] * dt * ny;
    const float s0z = __fmaf_rn(__fmaf_rn(dm[2], th, om[2]), nz, -0.5f), dsz = dm[2] * dt * nz;
    float acc = 0.f;
    col[0] = 0.f;
    for (int k = 0; k < steps; ++k) {
        const float kf = (float)k;
        const float dens = grid_density_s(m, __fmaf_rn(kf, dsx, s0x), __fmaf_rn(kf, dsy, s0y), __fmaf_rn(kf, dsz, s0z));
        acc = __fmaf_rn(dens, dt, acc);
        col[(size_t)(k + 1) * 256u] = acc;
    }
    span[q] = make_float2(t0, t1);
}

}  // namespace bre
constexpr int QT = 128;
#include "bre_query_warp.cuh"
#include "bre_query_tile.cuh"
namespace bre {

// image-tile path: the caller declared the rays to be a grid_w x grid_h image (bre_set_ray_grid / K0)
static bool grid_layout(const bre_ctx *ctx, uint32_t n) {
    return !BRE_WIDE && ctx->tile_path && ctx->grid_w >= 2 && ctx->grid_h >= 2 &&
           (uint64_t)ctx->grid_w * ctx->grid_h == n;
}
static bool tile_mode(const bre_ctx *ctx, uint32_t n) { return grid_layout(ctx, n) && ctx->n_built > 0; }

// One block of launch state per ray batch, in ctx->packet_ctr (stride QB_STRIDE bytes):
//   +0   queue control words (QC_*), zeroed before every launch
//   +64  zero word + the two heavy-packet counters (persist from launch to launch)
//   +128 this batch's LeafArgs (read by the cold in-place path of K5a)
constexpr size_t QB_STRIDE = 512, QB_HEAVY = 64, QB_ARGS = 128;
constexpr uint32_t EST_MAX_BATCHES = 128;   // batches whose demand is read back for the next launch
static_assert(QB_ARGS + sizeof(LeafArgs) <= QB_STRIDE, "batch block too small");

__host__ __device__ inline LeafArgs shift_args(LeafArgs A, size_t off) {
    A.ro += off; A.rd += off;
    A.rayBase += (uint32_t)off;
    if (A.out) A.out += off;
    if (A.tau && !A.tauW) A.tau += off * (size_t)(A.tauSteps + 1);   // (the tile-blocked table is indexed by absolute pixel)
    if (A.span) A.span += off;
    if (A.hitCounts) A.hitCounts += off;
    if (A.hitOffsets) A.hitOffsets += off;
    if (A.frames) A.frames += off / FAT_P * FRAME_V4;
    return A;
}
__global__ void k_store_batch_args(LeafArgs A, char *blocks, unsigned batchRays, unsigned nBatches) {
    unsigned b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nBatches) return;
    *reinterpret_cast<LeafArgs *>(blocks + b * QB_STRIDE + QB_ARGS) = shift_args(A, (size_t)b * batchRays);
}

// Launch K5a (traversal -> pair queue) and K5b (pairs -> radiance / hit lists).
//
// Rays are processed in BATCHES so that the pair queue stays bounded whatever the ray count
// (a 4K camera emits ~5e8 pairs): batch size = queue capacity / (pairs per ray measured on the
// previous launch, +30 %).  With more than one batch the queue is split in two halves and the
// batches are software-pipelined over two streams -- K5b of batch i runs while K5a of batch i+1
// walks the tree -- so the tail of one kernel (a few heavy packets / the last waves) is filled by
// the other.  A single batch runs exactly as before on the context stream.
template <int MODE, bool COUNT>
static int launch_query(bre_ctx *ctx, uint32_t n, const float4 *ro, const float4 *rd, float nPathsNorm,
                        float4 *out, const float *tau, const float2 *span, uint32_t *hitCounts,
                        const uint64_t *hitOffsets, uint32_t *hitIdx, unsigned long long *counters) {
    int s;
    if (n >= (1u << 30)) return set_err(ctx, BRE_ERR_UNSUPPORTED, "more than 2^30 rays in one query");
    // ---- image-tile path: the caller declared the rays to be a grid_w x grid_h image (bre_set_ray_grid / K0)
    const bool tileMode = tile_mode(ctx, n);
    // the per-ray paths read a hit's power / direction per HIT: they want them in Morton order (once per build)
    if (MODE == 0 && !tileMode && (s = ensure_sorted_payload(ctx)) != BRE_OK) return s;
    const uint32_t imgW = tileMode ? ctx->grid_w : 0, imgH = tileMode ? ctx->grid_h : 0;
    const uint32_t tilesX = tileMode ? (imgW + TILE_W - 1) / TILE_W : 0, tilesY = tileMode ? (imgH + TILE_H - 1) / TILE_H : 0;
    // ---- pair queue capacity: 128 pairs per ray, at least 1M, at most 256M pairs (2 GiB)
    uint64_t want = (uint64_t)n * 128u;
    if (want < (1u << 20)) want = 1u << 20;
    if (want > (1u << 28)) want = 1u << 28;
    if (ctx->pair_queue_cap_override) want = ctx->pair_queue_cap_override;
    // ---- batch size from the last launch's measured demand (pinned host copy of its control words)
    if (!ctx->est_host) {
        BRE_CUDA(cudaHostAlloc((void **)&ctx->est_host, EST_MAX_BATCHES * QB_STRIDE, cudaHostAllocDefault));
        memset(ctx->est_host, 0, EST_MAX_BATCHES * QB_STRIDE);
        BRE_CUDA(cudaEventCreateWithFlags(&ctx->ev_est, cudaEventDisableTiming));
    } else if (ctx->est_batches) {
        // the previous launch's copy of its control words into est_host must have landed before it is read
        BRE_CUDA(cudaEventSynchronize(ctx->ev_est));
    }
    // densest batch of the previous launch (image regions differ: the last batch alone misleads)
    double est = tileMode ? 16.0 : 128.0;
    if (ctx->est_batches && ctx->est_tile == tileMode) {
        double worst = 0.0;
        for (uint32_t b = 0; b < ctx->est_batches; ++b) {
            const uint32_t r0 = b * ctx->est_batch_rays;
            const uint32_t nb = ctx->est_n - r0 < ctx->est_batch_rays ? ctx->est_n - r0 : ctx->est_batch_rays;
            const double v = (double)ctx->est_host[b * (QB_STRIDE / 4) + QC_PAIR_TAIL] / (double)(nb ? nb : 1);
            if (v > worst) worst = v;
        }
        if (worst > 0.0) est = 1.3 * worst;
        if (est < 16.0) est = 16.0;
    }
    uint32_t batchRays = n;
    if ((double)n * est > (double)want) {
        // two halves in flight: each batch must fit half the queue
        uint64_t fit = (uint64_t)((double)(want / 2) / est);
        fit &= ~(uint64_t)(WQ_P * WQ_WARPS - 1);
        if (fit < 4096) fit = 4096;
        if (fit < n) batchRays = (uint32_t)fit;
    }
    if (ctx->ray_batch_override && ctx->ray_batch_override < n)
        batchRays = (ctx->ray_batch_override + WQ_P - 1) / WQ_P * WQ_P;
    if (tileMode && batchRays < n) {
        // a batch is a band of whole tile rows
        const uint32_t band = imgW * TILE_H;
        batchRays = batchRays / band * band;
        if (batchRays == 0) batchRays = band;
    }
    const uint32_t nBatches = (n + batchRays - 1) / batchRays;
    const bool piped = nBatches > 1;
    const unsigned cap = (unsigned)(piped ? want / 2 : want);
    if ((s = ensure(ctx, ctx->pair_queue, (size_t)want * 8)) != BRE_OK) return s;
    if ((s = ensure(ctx, ctx->packet_ctr, (size_t)nBatches * QB_STRIDE)) != BRE_OK) return s;
    if (ctx->ctr_batches != nBatches || ctx->ctr_batch_rays != batchRays || ctx->ctr_rays != n) {
        // new batch layout: the persistent heavy counters of the old one mean nothing
        BRE_CUDA(cudaMemsetAsync(ctx->packet_ctr.p, 0, (size_t)nBatches * QB_STRIDE, ctx->stream));
        ctx->ctr_batches = nBatches; ctx->ctr_batch_rays = batchRays; ctx->ctr_rays = n;
        ctx->heavy_rays = 0;
    }
    char *blocks = (char *)ctx->packet_ctr.p;
    BRE_CUDA(cudaMemset2DAsync(blocks, QB_STRIDE, 0, QB_HEAVY + 4, nBatches, ctx->stream));
    if (MODE == 0) BRE_CUDA(cudaMemsetAsync(out, 0, (size_t)n * 16, ctx->stream));
    else BRE_CUDA(cudaMemsetAsync(hitCounts, 0, (size_t)n * 4, ctx->stream));
    LeafArgs A;
    A.t = tree_view(ctx); A.m = ctx->medium; A.ro = ro; A.rd = rd; A.nPathsNorm = nPathsNorm;
    A.out = out; A.tau = tau; A.span = span; A.tauSteps = ctx->tau_steps;
    A.tauW = (grid_layout(ctx, n) && tau) ? ctx->grid_w : 0u; A.tauTilesX = (ctx->grid_w + 15) / 16; A.rayBase = 0;
    A.hitCounts = hitCounts; A.hitOffsets = hitOffsets; A.hitIdx = hitIdx;
    A.frames = nullptr;
    TileArgs TA{};
    unsigned wqCap = 0;
    if (tileMode) {
        const size_t nT = (size_t)tilesX * tilesY;
        if ((s = ensure(ctx, ctx->tframes, nT * 16 * TFRAME_V4)) != BRE_OK) return s;
        if ((s = ensure(ctx, ctx->tile_state, nT)) != BRE_OK) return s;
        if ((s = ensure(ctx, ctx->trays, nT * 256 * 16)) != BRE_OK) return s;
        if ((s = ensure(ctx, ctx->tjit, nT * 256 * 8)) != BRE_OK) return s;
        TA.cut = (const float4 *)ctx->top_cut.p;
        TA.nCut = (ctx->top_cut.p && ctx->n_built > 1) ? 1u : 0u;      // the kernel reads the entry count from the cut's header
        TA.tframes = (float4 *)ctx->tframes.p;
        TA.trays = (float4 *)ctx->trays.p;
        TA.tjit = (float2 *)ctx->tjit.p;
        TA.tileState = (unsigned char *)ctx->tile_state.p;
        TA.imgW = imgW; TA.imgH = imgH; TA.tilesX = tilesX; TA.tilesY = tilesY;
        // global work queue of the tile walk: 4 items per photon (frontier of a breadth-first sweep; more is
        // walked depth-first by the producer).  Slots are EMPTY between launches: consumers reset them.
        uint64_t wc = (uint64_t)ctx->n_built * 4u;
        if (wc < (1u << 20)) wc = 1u << 20;
        if (wc > 0xfffffff0ull - 64) wc = 0xfffffff0ull - 64;
        if (ctx->tile_wq_cap_override) wc = ctx->tile_wq_cap_override;
        if (ctx->tile_wq.cap < wc * 8 || ctx->tile_wq_items != wc) {
            if ((s = ensure(ctx, ctx->tile_wq, (size_t)wc * 8)) != BRE_OK) return s;
            BRE_CUDA(cudaMemsetAsync(ctx->tile_wq.p, 0xff, ctx->tile_wq.cap, ctx->stream));
            ctx->tile_wq_items = wc;
        }
        wqCap = (unsigned)wc;
    }
    if (ctx->fat_spread > 0.f && !tileMode) {   // one 64-byte frame per packet
        if ((s = ensure(ctx, ctx->frames, ((size_t)n / FAT_P + WQ_P / FAT_P) * 16 * FRAME_V4)) != BRE_OK) return s;
        A.frames = (float4 *)ctx->frames.p;
    }
    k_store_batch_args<<<(nBatches + 63) / 64, 64, 0, ctx->stream>>>(A, blocks, batchRays, nBatches);

    // ---- heavy-first feedback between consecutive radiance queries over the same rays
    const unsigned batchPackets = (batchRays + WQ_P - 1) / WQ_P;
    const size_t flagWords = (batchPackets + 31) / 32;
    const size_t half = ((size_t)HEAVY_CAP + flagWords) * 4;
    const bool feedback = MODE == 0 && !COUNT;
    bool haveIn = false;
    int cur = 0, prev = 1;
    if (feedback) {
        if ((s = ensure(ctx, ctx->heavy, (size_t)nBatches * 2 * half)) != BRE_OK) return s;
        if (ctx->heavy_half != half || ctx->heavy_tile != tileMode) {   // buffer re-laid out / other id space
            ctx->heavy_half = half; ctx->heavy_tile = tileMode; ctx->heavy_rays = 0;
        }
        cur = ctx->heavy_parity; prev = cur ^ 1;
        haveIn = ctx->heavy_rays == n;
        ctx->heavy_rays = n;
        ctx->heavy_parity = prev;
    }
    // ---- second stream + events for the pipelined case
    if (piped && !ctx->stream2) {
        BRE_CUDA(cudaStreamCreateWithFlags(&ctx->stream2, cudaStreamNonBlocking));
        for (int i = 0; i < 4; ++i) BRE_CUDA(cudaEventCreateWithFlags(&ctx->qev[i], cudaEventDisableTiming));
    }
    int perSM = 0, perSMT = 1;
    BRE_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, k_traverse<MODE, COUNT>, WQ_WARPS * 32, 0));
    if (perSM < 1) perSM = 1;
    const int shardMod = (tileMode && ctx->tile_shard && ctx->n_ranks > 1) ? ctx->n_ranks : 1;
    if (tileMode) {
        BRE_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSMT, k_tile_walk<MODE, COUNT>, TT_THREADS,
                                                               sizeof(TileSmem)));
        if (perSMT < 1) perSMT = 1;
        if (ctx->k5t_blocks_per_sm > 0 && perSMT > ctx->k5t_blocks_per_sm) perSMT = ctx->k5t_blocks_per_sm;
    } else if (ctx->tile_shard && ctx->n_ranks > 1)
        return set_err(ctx, BRE_ERR_INVALID, "tile sharding needs a ray grid (bre_set_ray_grid) and the tile path");
    if (ctx->k5a_blocks_per_sm > 0 && perSM > ctx->k5a_blocks_per_sm) perSM = ctx->k5a_blocks_per_sm;
    // K5b grid.  The leaf kernels walk the pair queue, whose length is only known on the device, in
    // runs of 64 pairs.  With a static stride the cost of a warp's runs varies with their hits and the
    // chip waits for the unluckiest warp: blocks/SM 4 / 8 / 32 / 128 / 512 / 2048 -> K5b 2.29 / 2.23 / 2.18 /
    // 2.115 / 2.15 / 2.39 ms at 2 M rays (and 512 blocks/SM cost a 65 K-ray launch +0.07 ms in empty
    // blocks).  LP_DYNAMIC (default): resident blocks claim runs from a counter instead.
    unsigned gb;
    if (ctx->k5b_blocks_per_sm > 0) gb = (unsigned)ctx->sm_count * (unsigned)ctx->k5b_blocks_per_sm;
#if LP_DYNAMIC
    else gb = (unsigned)ctx->sm_count * 8u;
#else
    else {
        const unsigned lo = (unsigned)ctx->sm_count * 8u, hi = (unsigned)ctx->sm_count * 128u;
        const unsigned want = (unsigned)std::min<uint64_t>((uint64_t)batchRays / 32u + 1u, hi);
        gb = want < lo ? lo : want;
    }
#endif
    if (ctx->timing) cudaEventRecord(ctx->ev[PH_QUERY_A][0], ctx->stream);
    for (uint32_t b = 0; b < nBatches; ++b) {
        const uint32_t r0 = b * batchRays, nb = (n - r0 < batchRays) ? n - r0 : batchRays;
        const LeafArgs Ab = shift_args(A, r0);
        unsigned int *ctl = (unsigned int *)(blocks + b * QB_STRIDE);
        unsigned int *heavyCnt = (unsigned int *)(blocks + b * QB_STRIDE + QB_HEAVY);
        const LeafArgs *Ag = (const LeafArgs *)(blocks + b * QB_STRIDE + QB_ARGS);
        uint2 *queue = (uint2 *)ctx->pair_queue.p + (piped ? (size_t)(b & 1u) * cap : 0);
        const unsigned *heavyIn = nullptr, *heavyFlagIn = nullptr, *nHeavyIn = heavyCnt;
        unsigned *heavyOut = nullptr, *heavyFlagOut = nullptr, *nHeavyOut = heavyCnt;
        unsigned heavyCap = 0;
        if (feedback) {
            unsigned *bufCur = (unsigned *)((char *)ctx->heavy.p + ((size_t)b * 2 + cur) * half);
            unsigned *bufPrev = (unsigned *)((char *)ctx->heavy.p + ((size_t)b * 2 + prev) * half);
            heavyOut = bufCur; heavyFlagOut = bufCur + HEAVY_CAP; nHeavyOut = heavyCnt + 1 + cur;
            heavyCap = HEAVY_CAP;
            BRE_CUDA(cudaMemsetAsync(heavyFlagOut, 0, flagWords * 4, ctx->stream));
            BRE_CUDA(cudaMemsetAsync(nHeavyOut, 0, 4, ctx->stream));
            if (haveIn) { heavyIn = bufPrev; heavyFlagIn = bufPrev + HEAVY_CAP; nHeavyIn = heavyCnt + 1 + prev; }
        }
        // the queue half of batch b was last read by K5b of batch b-2
        if (piped && b >= 2) BRE_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->qev[2 + (b & 1u)], 0));
        unsigned blocksA = (unsigned)(ctx->sm_count * perSM);
        const unsigned need = ((nb + WQ_P - 1) / WQ_P + WQ_WARPS - 1) / WQ_WARPS;
        if (blocksA > need) blocksA = need;
        if (tileMode) {
            // K5a': one block per image tile of this band (absolute ray indices: the unshifted arguments of batch 0)
            const uint32_t tile0 = (r0 / imgW / TILE_H) * tilesX;
            const uint32_t nT = ((nb / imgW + TILE_H - 1) / TILE_H) * tilesX;
            const LeafArgs *Ag0 = (const LeafArgs *)(blocks + QB_ARGS);
            WalkQueue WQ{(uint2 *)ctx->tile_wq.p, wqCap, ctl + QC_WQ_HEAD};
            unsigned blocksT = (unsigned)(ctx->sm_count * perSMT);
            if (blocksT > (nT + TT_WARPS - 1) / TT_WARPS) blocksT = (nT + TT_WARPS - 1) / TT_WARPS;
            k_tile_walk<MODE, COUNT><<<blocksT, TT_THREADS, sizeof(TileSmem), ctx->stream>>>(
                A, Ag0, TA, WQ, tile0, nT, queue, cap, ctl, heavyIn, nHeavyIn, heavyFlagIn, heavyOut, heavyFlagOut,
                nHeavyOut, heavyCap, shardMod, ctx->rank, counters);
            // rays of tiles flagged incoherent: per-ray walk (returns at once when there is none)
            k_traverse<MODE, COUNT><<<blocksA, WQ_WARPS * 32, 0, ctx->stream>>>(
                Ab, Ag, nb, queue, cap, ctl, nullptr, heavyCnt, nullptr, nullptr, nullptr, heavyCnt, 0u, 0.f,
                TA.tileState, imgW, tilesX, r0, counters);
        } else
        k_traverse<MODE, COUNT><<<blocksA, WQ_WARPS * 32, 0, ctx->stream>>>(
            Ab, Ag, nb, queue, cap, ctl, heavyIn, nHeavyIn, heavyFlagIn, heavyOut, heavyFlagOut, nHeavyOut,
            heavyCap, ctx->fat_spread, nullptr, 0u, 0u, 0u, counters);
        cudaStream_t sb = ctx->stream;
        if (piped) {
            BRE_CUDA(cudaEventRecord(ctx->qev[b & 1u], ctx->stream));
            BRE_CUDA(cudaStreamWaitEvent(ctx->stream2, ctx->qev[b & 1u], 0));
            sb = ctx->stream2;
        }
        if (b + 1 == nBatches) {
            // demand of this launch (every batch's control words) -> batch size of the next one
            const uint32_t nCopy = nBatches < EST_MAX_BATCHES ? nBatches : EST_MAX_BATCHES;
            BRE_CUDA(cudaMemcpyAsync(ctx->est_host, blocks, (size_t)nCopy * QB_STRIDE, cudaMemcpyDeviceToHost,
                                     ctx->stream));
            ctx->est_batches = nCopy; ctx->est_batch_rays = batchRays; ctx->est_n = n; ctx->est_tile = tileMode;
            BRE_CUDA(cudaEventRecord(ctx->ev_est, ctx->stream));
            if (ctx->timing) {
                cudaEventRecord(ctx->ev[PH_QUERY_A][1], ctx->stream); ctx->ev_valid[PH_QUERY_A] = true;
                cudaEventRecord(ctx->ev[PH_QUERY_B][0], ctx->stream);
            }
        }
        // the leaf kernels scan the queue; each leaves at once if K5a emitted no pair of its kind
        if (tileMode) {
            if (ctx->max_leaf <= 4) {
                BRE_CUDA(cudaFuncSetAttribute(k_leaf_tile<MODE, COUNT, 4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)LEAF_TILE_SMEM));
                BRE_CUDA(cudaFuncSetAttribute(k_leaf_tile<MODE, COUNT, 4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)LEAF_TILE_SMEM));
                k_leaf_tile<MODE, COUNT, 4, false><<<gb, LP_WARPS * 32, LEAF_TILE_SMEM, sb>>>(A, TA.tframes, TA.trays, TA.tjit, imgW, queue, cap, ctl, counters);
                k_leaf_tile<MODE, COUNT, 4, true><<<gb, LP_WARPS * 32, LEAF_TILE_SMEM, sb>>>(A, TA.tframes, TA.trays, TA.tjit, imgW, queue, cap, ctl, counters);
            } else if (ctx->max_leaf <= 8) {
                BRE_CUDA(cudaFuncSetAttribute(k_leaf_tile<MODE, COUNT, 8, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)LEAF_TILE_SMEM));
                BRE_CUDA(cudaFuncSetAttribute(k_leaf_tile<MODE, COUNT, 8, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)LEAF_TILE_SMEM));
                k_leaf_tile<MODE, COUNT, 8, false><<<gb, LP_WARPS * 32, LEAF_TILE_SMEM, sb>>>(A, TA.tframes, TA.trays, TA.tjit, imgW, queue, cap, ctl, counters);
                k_leaf_tile<MODE, COUNT, 8, true><<<gb, LP_WARPS * 32, LEAF_TILE_SMEM, sb>>>(A, TA.tframes, TA.trays, TA.tjit, imgW, queue, cap, ctl, counters);
            } else {
                BRE_CUDA(cudaFuncSetAttribute(k_leaf_tile<MODE, COUNT, 16, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)LEAF_TILE_SMEM));
                BRE_CUDA(cudaFuncSetAttribute(k_leaf_tile<MODE, COUNT, 16, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)LEAF_TILE_SMEM));
                k_leaf_tile<MODE, COUNT, 16, false><<<gb, LP_WARPS * 32, LEAF_TILE_SMEM, sb>>>(A, TA.tframes, TA.trays, TA.tjit, imgW, queue, cap, ctl, counters);
                k_leaf_tile<MODE, COUNT, 16, true><<<gb, LP_WARPS * 32, LEAF_TILE_SMEM, sb>>>(A, TA.tframes, TA.trays, TA.tjit, imgW, queue, cap, ctl, counters);
            }
        }
        if (ctx->max_leaf <= 4) {
            k_leaf_fat<MODE, COUNT, 4><<<gb, LP_WARPS * 32, 0, sb>>>(Ab, queue, cap, ctl, counters);
            k_leaf_pairs<MODE, COUNT, 4><<<gb, LP_WARPS * 32, 0, sb>>>(Ab, queue, cap, ctl, counters);
        } else if (ctx->max_leaf <= 8) {
            k_leaf_fat<MODE, COUNT, 8><<<gb, LP_WARPS * 32, 0, sb>>>(Ab, queue, cap, ctl, counters);
            k_leaf_pairs<MODE, COUNT, 8><<<gb, LP_WARPS * 32, 0, sb>>>(Ab, queue, cap, ctl, counters);
        } else {
            k_leaf_fat<MODE, COUNT, 16><<<gb, LP_WARPS * 32, 0, sb>>>(Ab, queue, cap, ctl, counters);
            k_leaf_pairs<MODE, COUNT, 16><<<gb, LP_WARPS * 32, 0, sb>>>(Ab, queue, cap, ctl, counters);
        }
        if (piped) BRE_CUDA(cudaEventRecord(ctx->qev[2 + (b & 1u)], ctx->stream2));
    }
    if (piped) {   // join: everything K5b did is ordered before what follows on the context stream
        BRE_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->qev[2], 0));
        BRE_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->qev[3], 0));
    }
    if (ctx->timing) { cudaEventRecord(ctx->ev[PH_QUERY_B][1], ctx->stream); ctx->ev_valid[PH_QUERY_B] = true; }
    BRE_CUDA(cudaGetLastError());
    return BRE_OK;
}

// ---------------------------------------------------------------- canonical counting traversal
// Walks the 1-photon-per-leaf canonical LBVH exactly like oracle/bre_oracle.c orc_query_ray and
// counts V (slab tests), T (disk tests), H (hits).  counters[0..2].
__global__ void __launch_bounds__(QT) k_count_canonical(TreeDev t, uint32_t n,
                                                        const float4 *__restrict__ ro,
                                                        const float4 *__restrict__ rd,
                                                        unsigned long long *__restrict__ counters) {
    uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long V = 0, T = 0, H = 0;
    if (q < n && t.n > 0) {
        RayS r = ray_setup(__ldg(ro + q), __ldg(rd + q), t.abs_max);
        int stack[128];
        int sp = 0;
        int cur = t.n > 1 ? 0 : ~0;
        for (;;) {
            ++V;
            if (cur < 0) {
                int s = ~cur;
                float4 lo = __ldg(t.leaf_lo + s), hi = __ldg(t.leaf_hi + s);
                if (slab_hit(r, lo.x, lo.y, lo.z, hi.x, hi.y, hi.z)) {
                    ++T;
                    float th, d2;
                    if (disk_test(r, __ldg(t.s_pos_r + s), th, d2)) ++H;
                }
                if (sp == 0) break;
                cur = stack[--sp];
            } else {
                float4 lo = __ldg(t.node_lo + cur), hi = __ldg(t.node_hi + cur);
                if (slab_hit(r, lo.x, lo.y, lo.z, hi.x, hi.y, hi.z)) {
                    stack[sp++] = __float_as_int(hi.w);
                    cur = __float_as_int(lo.w);
                } else {
                    if (sp == 0) break;
                    cur = stack[--sp];
                }
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        V += __shfl_xor_sync(0xffffffffu, V, o);
        T += __shfl_xor_sync(0xffffffffu, T, o);
        H += __shfl_xor_sync(0xffffffffu, H, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(counters + 0, V);
        atomicAdd(counters + 1, T);
        atomicAdd(counters + 2, H);
    }
}

// ---------------------------------------------------------------- host drivers
// K6 on stream `st` (the context stream, or the staging stream of bre_stage_*: the table only depends on the rays and
// the medium, so a host that hands the rays over before the photon pass gets it computed UNDER the photon pass + build).
static int prepare_tau(bre_ctx *ctx, uint32_t n, const float4 *ro, const float4 *rd,
                       const float **tau, const float2 **span, cudaStream_t st) {
    *tau = nullptr; *span = nullptr;
    if (ctx->medium.kind != 1) return BRE_OK;
    int s;
    size_t rowBytes = (size_t)(ctx->tau_steps + 1) * 4;
    size_t offSpan = ((size_t)n * rowBytes + 15) & ~(size_t)15;
    float *tp = nullptr;
    float2 *sp = nullptr;
    const bool grid = grid_layout(ctx, n);
    if (grid) {
        // tile-blocked table (k_tau_tile): [tile][step][256 pixels]
        const uint32_t tX = (ctx->grid_w + 15) / 16, tY = (ctx->grid_h + 15) / 16;
        const size_t tab = (size_t)tX * tY * rowBytes * 256;
        offSpan = (tab + 15) & ~(size_t)15;
        if ((s = ensure(ctx, ctx->tau, offSpan + (size_t)n * 8)) != BRE_OK) return s;
        tp = (float *)ctx->tau.p;
        sp = (float2 *)((char *)ctx->tau.p + offSpan);
        PhaseTimer pt(ctx, PH_TAU, st);
        const int shardMod = (ctx->tile_shard && ctx->n_ranks > 1) ? ctx->n_ranks : 1;
        k_tau_tile<<<tX * tY, 256, 0, st>>>(ctx->medium, ctx->grid_w, ctx->grid_h, tX, ro, rd, ctx->tau_steps,
                                          tp, sp, shardMod, ctx->rank);
    } else {
        if ((s = ensure(ctx, ctx->tau, offSpan + (size_t)n * 8)) != BRE_OK) return s;
        tp = (float *)ctx->tau.p;
        sp = (float2 *)((char *)ctx->tau.p + offSpan);
        PhaseTimer pt(ctx, PH_TAU, st);
        const size_t smem = (size_t)128 * (ctx->tau_steps + 1) * 4;   // tau_steps <= 384 (bre_create)
        if (smem > 48 * 1024)
            BRE_CUDA(cudaFuncSetAttribute(k_tau_table, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_tau_table<<<(n + 127) / 128, 128, smem, st>>>(
            ctx->medium, n, ro, rd, ctx->tau_steps, tp, sp);
    }
    *tau = tp; *span = sp;
    ctx->tau_key = {ro, rd, n, ctx->tau_steps, grid ? ctx->grid_w : 0u, ctx->medium_version, tp, sp};
    return BRE_OK;
}

// bre_stage_*: K6 for the rays of the NEXT query, on the staging stream (after the rays themselves are there)
int stage_tau(bre_ctx *ctx, uint32_t n, const float4 *ro, const float4 *rd, cudaStream_t st) {
    const float *tau; const float2 *span;
    ctx->tau_staged = false;
    if (!ctx->have_medium || ctx->medium.kind != 1) return BRE_OK;
    int s = prepare_tau(ctx, n, ro, rd, &tau, &span, st);
    if (s == BRE_OK) ctx->tau_staged = true;
    return s;
}

// bre_stage_*: remember the rays whose table is wanted; launch_pending_tau starts K6 on the staging stream behind
// everything queued on the context stream so far (= behind the photon pass when bre_build calls it)
int stage_tau_deferred(bre_ctx *ctx, uint32_t n, const float4 *ro, const float4 *rd, bool useGrid, uint32_t gridW,
                       uint32_t gridH) {
    ctx->tau_staged = false;
    ctx->tau_pending.on = false;
    if (!ctx->have_medium || ctx->medium.kind != 1) return BRE_OK;
    // Deferring pays when K6 fits under the build (its latency-bound kernels leave issue slots free); when K6 is the
    // longer of the two -- small photon shards on several GPUs, 4K frames -- it starts at once and also runs in the
    // gaps of the photon pass.  Estimates from measured rates: 3.2e-9 ms per table sample, 1.9e-7 ms per photon of the
    // last build.
    const double k6Ms = 3.2e-9 * (double)n * (double)ctx->tau_steps, buildMs = 1.9e-7 * (double)ctx->n_built;
    if (!ctx->tau_defer || k6Ms > 0.6 * buildMs) {
        const uint32_t gw = ctx->grid_w, gh = ctx->grid_h;
        if (useGrid) { ctx->grid_w = gridW; ctx->grid_h = gridH; }
        const int s = stage_tau(ctx, n, ro, rd, ctx->copy_stream);
        ctx->grid_w = gw; ctx->grid_h = gh;
        return s;
    }
    ctx->tau_pending.on = true; ctx->tau_pending.n = n; ctx->tau_pending.ro = ro; ctx->tau_pending.rd = rd;
    ctx->tau_pending.useGrid = useGrid; ctx->tau_pending.gridW = gridW; ctx->tau_pending.gridH = gridH;
    return BRE_OK;
}
int launch_pending_tau(bre_ctx *ctx) {
    if (!ctx->tau_pending.on) return BRE_OK;
    const bre_ctx::TauPending p = ctx->tau_pending;
    ctx->tau_pending.on = false;
    if (!ctx->copy_stream || !ctx->have_medium || ctx->medium.kind != 1) return BRE_OK;
    BRE_CUDA(cudaEventRecord(ctx->ev_main, ctx->stream));
    BRE_CUDA(cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_main, 0));
    const uint32_t gw = ctx->grid_w, gh = ctx->grid_h;
    if (p.useGrid) { ctx->grid_w = p.gridW; ctx->grid_h = p.gridH; }
    const int s = stage_tau(ctx, p.n, (const float4 *)p.ro, (const float4 *)p.rd, ctx->copy_stream);
    ctx->grid_w = gw; ctx->grid_h = gh;
    if (s != BRE_OK) return s;
    BRE_CUDA(cudaEventRecord(ctx->ev_stage, ctx->copy_stream));
    return BRE_OK;
}

int run_query(bre_ctx *ctx, uint32_t n, const float4 *ro, const float4 *rd, float nPathsNorm,
              float4 *out) {
    if (n == 0) return BRE_OK;
    {
        const int sp = launch_pending_tau(ctx);      // staged, but bre_build was not called in between
        if (sp != BRE_OK) return sp;
    }
    const float *tau = nullptr; const float2 *span = nullptr;
    int s = BRE_OK;
    const bre_ctx::TauKey &k = ctx->tau_key;
    if (ctx->tau_staged && ctx->medium.kind == 1 && k.ro == ro && k.rd == rd && k.n == n && k.steps == ctx->tau_steps &&
        k.gridW == (grid_layout(ctx, n) ? ctx->grid_w : 0u) && k.mediumVersion == ctx->medium_version) {
        // the table of exactly these rays was computed by bre_stage_* under the photon pass: wait for it
        tau = k.tau; span = k.span;
        BRE_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_stage, 0));
    } else
        s = prepare_tau(ctx, n, ro, rd, &tau, &span, ctx->stream);
    ctx->tau_staged = false;
    if (s != BRE_OK) return s;
    {
        PhaseTimer pt(ctx, PH_QUERY);
        s = launch_query<0, false>(ctx, n, ro, rd, nPathsNorm, out, tau, span, nullptr, nullptr, nullptr, nullptr);
    }
    return s;
}

int run_count(bre_ctx *ctx, uint32_t n, const float4 *ro, const float4 *rd) {
    int s;
    if ((s = ensure(ctx, ctx->counters_d, 8 * 8)) != BRE_OK) return s;
    unsigned long long *cd = (unsigned long long *)ctx->counters_d.p;
    BRE_CUDA(cudaMemsetAsync(cd, 0, 64, ctx->stream));
    if (n > 0) {
        if ((s = ensure_canonical(ctx)) != BRE_OK) return s;
        if ((s = ensure(ctx, ctx->hit_counts, (size_t)n * 4)) != BRE_OK) return s;
        k_count_canonical<<<(n + QT - 1) / QT, QT, 0, ctx->stream>>>(tree_view(ctx), n, ro, rd, cd);
        if ((s = launch_query<1, true>(ctx, n, ro, rd, 1.f, nullptr, nullptr, nullptr,
                                       (uint32_t *)ctx->hit_counts.p, nullptr, nullptr, cd)) != BRE_OK) return s;
    }
    unsigned long long h[8];
    BRE_CUDA(cudaMemcpyAsync(h, cd, 64, cudaMemcpyDeviceToHost, ctx->stream));
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->counters.nodes_visited = h[0];
    ctx->counters.disk_tests = h[1];
    ctx->counters.hits = h[2];
    ctx->counters.bytes_algorithmic = 48ull * n + 32ull * h[0] + 16ull * h[1] + 32ull * h[2];
    ctx->counters.fast_nodes = h[3];
    ctx->counters.fast_disk_tests = h[4];
    uint64_t emitted = 0, inplace = 0;
    if (n > 0) {
        std::vector<unsigned int> ctl((size_t)ctx->ctr_batches * QB_STRIDE / 4);
        BRE_CUDA(cudaMemcpyAsync(ctl.data(), ctx->packet_ctr.p, ctl.size() * 4, cudaMemcpyDeviceToHost, ctx->stream));
        BRE_CUDA(cudaStreamSynchronize(ctx->stream));
        for (uint32_t b = 0; b < ctx->ctr_batches; ++b) {
            emitted += ctl[b * QB_STRIDE / 4 + QC_EMITTED];
            inplace += ctl[b * QB_STRIDE / 4 + QC_INPLACE];
        }
    }
    ctx->counters.leaf_pairs = emitted;
    ctx->counters.pair_queue_overflow = inplace;
    return BRE_OK;
}

}  // namespace bre

#include <algorithm>
#include <vector>

namespace bre {

int run_hitdump(bre_ctx *ctx, uint32_t n, const float4 *ro, const float4 *rd, bre_hitdump *hd) {
    hd->total = 0;
    if (hd->offsets) hd->offsets[0] = 0;
    if (n == 0) return BRE_OK;
    int s;
    if ((s = ensure(ctx, ctx->hit_counts, (size_t)n * 4)) != BRE_OK) return s;
    if ((s = ensure(ctx, ctx->hit_offsets, ((size_t)n + 1) * 8)) != BRE_OK) return s;
    if ((s = launch_query<1, false>(ctx, n, ro, rd, 1.f, nullptr, nullptr, nullptr,
                                    (uint32_t *)ctx->hit_counts.p, nullptr, nullptr, nullptr)) != BRE_OK) return s;
    std::vector<uint32_t> counts(n);
    BRE_CUDA(cudaMemcpyAsync(counts.data(), ctx->hit_counts.p, (size_t)n * 4, cudaMemcpyDeviceToHost,
                             ctx->stream));
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    std::vector<uint64_t> offs((size_t)n + 1);
    offs[0] = 0;
    for (uint32_t q = 0; q < n; ++q) offs[q + 1] = offs[q] + counts[q];
    hd->total = offs[n];
    if (hd->offsets) memcpy(hd->offsets, offs.data(), ((size_t)n + 1) * 8);
    if (!hd->indices || hd->total > hd->capacity || hd->total == 0) return BRE_OK;
    if ((s = ensure(ctx, ctx->hit_idx, (size_t)hd->total * 4)) != BRE_OK) return s;
    BRE_CUDA(cudaMemcpyAsync(ctx->hit_offsets.p, offs.data(), ((size_t)n + 1) * 8,
                             cudaMemcpyHostToDevice, ctx->stream));
    if ((s = launch_query<2, false>(ctx, n, ro, rd, 1.f, nullptr, nullptr, nullptr, (uint32_t *)ctx->hit_counts.p,
                                    (const uint64_t *)ctx->hit_offsets.p, (uint32_t *)ctx->hit_idx.p,
                                    nullptr)) != BRE_OK) return s;
    BRE_CUDA(cudaMemcpyAsync(hd->indices, ctx->hit_idx.p, (size_t)hd->total * 4,
                             cudaMemcpyDeviceToHost, ctx->stream));
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    for (uint32_t q = 0; q < n; ++q) std::sort(hd->indices + offs[q], hd->indices + offs[q + 1]);
    return BRE_OK;
}

}  // namespace bre
