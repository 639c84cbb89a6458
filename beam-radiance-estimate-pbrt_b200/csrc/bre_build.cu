// bre_build.cu -- K2..K4: Morton-code LBVH over photon bounding spheres.
//
// Replaces (reference src/): PhotonBeamBVH ctor + recursiveBuild / HLBVH2Build
// (core/photonbeambvh.cpp:204-248,259-425,427-495), EncodeMorton3/LeftShift3 (:109-139),
// RadixSort (:141-182), emitLBVH2 (:497-559), flattenBVH2Tree (:663-681) and
// KdTree<Photon> build (core/kdtree.h:101-149).
//
// Pipeline (all on the context stream, no host round trip except one 4-byte read of abs_max):
//   k_bounds      min/max of photon positions                  (centroid bounds, :431-436)
//   k_morton      30-bit Morton code per photon                 (:438-445)
//   radix sort    4 x 8-bit stable LSD passes of (code, index)  (integer; hist / 3-launch scan / scatter)
//   k_gather      photon SoA -> Morton order, leaf boxes, max |coord|
//   k_karras      binary radix tree over keys (code<<32 | index): one thread per internal node
//   k_refit       bottom-up AABB union with per-node arrival counters
//   k_emit_fast   64-byte two-child nodes of the production tree, leaves of <= max_leaf photons
#include "bre_internal.cuh"
#include <algorithm>
#include <cfloat>

namespace bre {

// ---------------------------------------------------------------- bounds
__device__ __forceinline__ uint32_t f2ord(float f) {
    uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

__global__ void k_bounds_init(uint32_t *b) {
    if (threadIdx.x < 3) b[threadIdx.x] = 0xffffffffu;       // min
    else if (threadIdx.x < 6) b[threadIdx.x] = 0u;            // max
    else if (threadIdx.x == 6) b[6] = 0u;                     // abs_max bits
}

__global__ void __launch_bounds__(256) k_bounds(const float4 *__restrict__ pos_r, uint32_t n,
                                                uint32_t *__restrict__ b) {
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float4 p = __ldg(pos_r + i);
        lo[0] = fminf(lo[0], p.x); hi[0] = fmaxf(hi[0], p.x);
        lo[1] = fminf(lo[1], p.y); hi[1] = fmaxf(hi[1], p.y);
        lo[2] = fminf(lo[2], p.z); hi[2] = fmaxf(hi[2], p.z);
    }
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            atomicMin(b + k, f2ord(lo[k]));
            atomicMax(b + 3 + k, f2ord(hi[k]));
        }
    }
}

// ---------------------------------------------------------------- Morton codes
__device__ __forceinline__ uint32_t left_shift3(uint32_t x) {  // photonbeambvh.cpp:109-132
    if (x == (1u << 10)) --x;
    x = (x | (x << 16)) & 0x30000ffu;
    x = (x | (x << 8)) & 0x300f00fu;
    x = (x | (x << 4)) & 0x30c30c3u;
    x = (x | (x << 2)) & 0x9249249u;
    return x;
}
__device__ __forceinline__ uint32_t morton_axis(float p, float lo, float hi) {
    // Bounds3::Offset then * mortonScale (photonbeambvh.cpp:440-444), float32, no contraction
    float o = fsub(p, lo);
    if (hi > lo) o = fdiv(o, fsub(hi, lo));
    float s = fmul(o, 1024.f);
    uint32_t v = (uint32_t)s;  // truncation toward zero, saturating like the oracle's clamp
    if (v > 1024u) v = 1024u;
    return left_shift3(v);
}

__global__ void __launch_bounds__(256) k_morton(const float4 *__restrict__ pos_r, uint32_t n,
                                                const uint32_t *__restrict__ b,
                                                uint32_t *__restrict__ codes) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float lo0 = ord2f(b[0]), lo1 = ord2f(b[1]), lo2 = ord2f(b[2]);
    float hi0 = ord2f(b[3]), hi1 = ord2f(b[4]), hi2 = ord2f(b[5]);
    float4 p = __ldg(pos_r + i);
    codes[i] = (morton_axis(p.z, lo2, hi2) << 2) | (morton_axis(p.y, lo1, hi1) << 1) |
               morton_axis(p.x, lo0, hi0);
}

// ---------------------------------------------------------------- radix sort
// Stable LSD sort of (key u32, value u32), 8 bits per pass.  Tile = 256 threads x 16 keys;
// warp w owns the contiguous chunk [w*512, (w+1)*512) of its tile, processed 32 keys at a time
// in index order, so (warp, item, lane) order == index order and the sort is stable.
constexpr int RS_THREADS = 256;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;

__global__ void __launch_bounds__(RS_THREADS) k_rs_hist(const uint32_t *__restrict__ keys,
                                                         uint32_t n, int shift,
                                                         uint32_t *__restrict__ hist,
                                                         uint32_t numBlocks) {
    __shared__ uint32_t sh[256];
    sh[threadIdx.x] = 0;
    __syncthreads();
    uint32_t base = blockIdx.x * RS_TILE;
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
        uint32_t idx = base + k * RS_THREADS + threadIdx.x;
        if (idx < n) atomicAdd(&sh[(__ldg(keys + idx) >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[threadIdx.x * numBlocks + blockIdx.x] = sh[threadIdx.x];
}

// Exclusive scan of `len` u32 counters in place, three launches: per-tile sums (tiles of 2048),
// a single-block scan of the tile sums, and the in-tile scan with the tile offset applied.
constexpr int SCN_T = 256, SCN_I = 8, SCN_TILE = SCN_T * SCN_I;

__global__ void __launch_bounds__(SCN_T) k_rs_scan_reduce(const uint32_t *__restrict__ in,
                                                          uint32_t len,
                                                          uint32_t *__restrict__ tileSums) {
    __shared__ uint32_t ws[SCN_T / 32];
    const uint32_t base = blockIdx.x * SCN_TILE;
    uint32_t sum = 0;
#pragma unroll
    for (int k = 0; k < SCN_I; ++k) {
        uint32_t i = base + k * SCN_T + threadIdx.x;
        if (i < len) sum += __ldg(in + i);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int i = 0; i < SCN_T / 32; ++i) t += ws[i];
        tileSums[blockIdx.x] = t;
    }
}

// single block of 1024 threads: exclusive scan of nTiles tile sums in place
__global__ void __launch_bounds__(1024) k_rs_scan_tiles(uint32_t *__restrict__ tileSums,
                                                        uint32_t nTiles) {
    __shared__ uint32_t warpSums[32];
    const uint32_t per = (nTiles + blockDim.x - 1) / blockDim.x;
    const uint32_t b = threadIdx.x * per, e = min(b + per, nTiles);
    uint32_t sum = 0;
    for (uint32_t i = b; i < e; ++i) sum += tileSums[i];
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) warpSums[w] = incl;
    __syncthreads();
    if (w == 0) {
        uint32_t v = warpSums[lane], iv = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, iv, o);
            if (lane >= o) iv += t;
        }
        warpSums[lane] = iv - v;
    }
    __syncthreads();
    uint32_t run = warpSums[w] + incl - sum;
    for (uint32_t i = b; i < e; ++i) {
        uint32_t v = tileSums[i];
        tileSums[i] = run;
        run += v;
    }
}

__global__ void __launch_bounds__(SCN_T) k_rs_scan_apply(uint32_t *__restrict__ data, uint32_t len,
                                                         const uint32_t *__restrict__ tileOff) {
    __shared__ uint32_t ws[SCN_T / 32];
    const uint32_t base = blockIdx.x * SCN_TILE + threadIdx.x * SCN_I;   // SCN_I consecutive elements
    uint32_t v[SCN_I], sum = 0;
#pragma unroll
    for (int k = 0; k < SCN_I; ++k) { v[k] = (base + k < len) ? data[base + k] : 0u; sum += v[k]; }
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) ws[w] = incl;
    __syncthreads();
    uint32_t woff = 0;
    for (uint32_t i = 0; i < w; ++i) woff += ws[i];
    uint32_t run = __ldg(tileOff + blockIdx.x) + woff + incl - sum;
#pragma unroll
    for (int k = 0; k < SCN_I; ++k) {
        if (base + k < len) data[base + k] = run;
        run += v[k];
    }
}

__global__ void __launch_bounds__(RS_THREADS, 4) k_rs_scatter(
    const uint32_t *__restrict__ keysIn, const uint32_t *__restrict__ valsIn,
    uint32_t *__restrict__ keysOut, uint32_t *__restrict__ valsOut, uint32_t n, int shift,
    const uint32_t *__restrict__ hist, uint32_t numBlocks, int iotaVals) {
    __shared__ uint32_t whist[RS_THREADS / 32][256];
    __shared__ uint32_t binLocal[256], binGlobal[256], wsum[RS_THREADS / 32];
    __shared__ uint32_t skey[RS_TILE], sval[RS_TILE];
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < (RS_THREADS / 32) * 256; i += RS_THREADS) (&whist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * RS_TILE + w * (RS_ITEMS * 32);
    uint32_t key[RS_ITEMS], rank[RS_ITEMS];
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
        uint32_t idx = base + k * 32 + lane;
        bool valid = idx < n;
        key[k] = valid ? __ldg(keysIn + idx) : 0xffffffffu;
        uint32_t d = (key[k] >> shift) & 255u;
        uint32_t vmask = __ballot_sync(0xffffffffu, valid);
        rank[k] = 0;
        if (valid) {
            uint32_t peers = __match_any_sync(vmask, d);
            int leader = __ffs(peers) - 1;
            uint32_t old = 0;
            if ((int)lane == leader) {
                old = whist[w][d];
                whist[w][d] = old + __popc(peers);
            }
            old = __shfl_sync(peers, old, leader);
            rank[k] = old + __popc(peers & lt);
        }
        __syncwarp();
    }
    __syncthreads();
    // digit threadIdx.x: exclusive prefix over warps (whist), the tile's exclusive prefix over
    // digits (binLocal) and the global destination of the tile's first key of that digit.
    {
        const uint32_t d = threadIdx.x;
        uint32_t run = 0;
#pragma unroll
        for (int ww = 0; ww < RS_THREADS / 32; ++ww) {
            uint32_t c = whist[ww][d];
            whist[ww][d] = run;
            run += c;
        }
        uint32_t incl = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) wsum[w] = incl;
        __syncthreads();
        uint32_t woff = 0;
        for (uint32_t i = 0; i < w; ++i) woff += wsum[i];
        const uint32_t excl = woff + incl - run;
        binLocal[d] = excl;
        binGlobal[d] = __ldg(hist + d * numBlocks + blockIdx.x) - excl;   // dst = binGlobal[d] + tile position
    }
    __syncthreads();
    // stage the tile in shared memory in its sorted (digit, original index) order ...
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
        uint32_t idx = base + k * 32 + lane;
        if (idx < n) {
            uint32_t d = (key[k] >> shift) & 255u;
            uint32_t lpos = binLocal[d] + whist[w][d] + rank[k];
            skey[lpos] = key[k];
            sval[lpos] = iotaVals ? idx : __ldg(valsIn + idx);
        }
    }
    __syncthreads();
    // ... and write it out in runs: neighbouring threads hit neighbouring addresses within a digit
    // (full sectors instead of 4096 scattered 4-byte stores per array)
    const uint32_t tile0 = blockIdx.x * RS_TILE;
    const uint32_t nValid = n - tile0 < (uint32_t)RS_TILE ? n - tile0 : (uint32_t)RS_TILE;
    for (uint32_t i = threadIdx.x; i < nValid; i += RS_THREADS) {
        uint32_t kk = skey[i];
        uint32_t dst = binGlobal[(kk >> shift) & 255u] + i;
        keysOut[dst] = kk;
        valsOut[dst] = sval[i];
    }
}

// ---------------------------------------------------------------- onesweep variant of the same sort
// The three-kernel passes above spend their time in launches and dependent latencies, not in bytes (20
// launches, 0.85 ms for 0.8 GB at 10 M keys).  Here ONE kernel histograms all four digits of every key
// (k_os_hist, global counts -> k_os_scan turns them into per-pass bin bases) and each pass is ONE scatter
// kernel whose tiles find their global offsets by decoupled look-back (Merrill & Garland; "onesweep"): a
// tile publishes its per-digit counts (AGGREGATE), then thread d walks back over the earlier tiles' words of
// digit d, adding AGGREGATEs until it meets an inclusive PREFIX, and publishes its own PREFIX.  Tile ids are
// handed out by an atomic counter, so every tile a walker waits for has already started.  Same stable order
// as the reference's RadixSort (core/photonbeambvh.cpp:141-182): ranks inside a tile are index-ordered.
constexpr uint32_t OS_AGG = 1u << 30, OS_PRE = 2u << 30, OS_MASK = (1u << 30) - 1u;

__global__ void __launch_bounds__(RS_THREADS) k_os_hist(const uint32_t *__restrict__ keys, uint32_t n,
                                                         uint32_t *__restrict__ ghist /* [4][256] */) {
    __shared__ uint32_t sh[4][256];
    for (int i = threadIdx.x; i < 1024; i += RS_THREADS) (&sh[0][0])[i] = 0;
    __syncthreads();
    for (uint32_t base = blockIdx.x * RS_TILE; base < n; base += gridDim.x * RS_TILE) {
#pragma unroll 4
        for (int k = 0; k < RS_ITEMS; ++k) {
            const uint32_t idx = base + k * RS_THREADS + threadIdx.x;
            if (idx < n) {
                const uint32_t key = __ldg(keys + idx);
                atomicAdd(&sh[0][key & 255u], 1u);
                atomicAdd(&sh[1][(key >> 8) & 255u], 1u);
                atomicAdd(&sh[2][(key >> 16) & 255u], 1u);
                atomicAdd(&sh[3][(key >> 24) & 255u], 1u);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 1024; i += RS_THREADS) {
        const uint32_t v = (&sh[0][0])[i];
        if (v) atomicAdd(ghist + i, v);
    }
}
// exclusive scan of each pass's 256 counts, in place (one block of 4 x 256 threads)
__global__ void __launch_bounds__(1024) k_os_scan(uint32_t *__restrict__ ghist) {
    __shared__ uint32_t ws[32];
    const uint32_t v = ghist[threadIdx.x];
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (uint32_t)o) incl += t;
    }
    if (lane == 31) ws[w] = incl;
    __syncthreads();
    uint32_t woff = 0;
    for (uint32_t i = (w & ~7u); i < w; ++i) woff += ws[i];      // 8 warps per pass
    ghist[threadIdx.x] = woff + incl - v;
}

__global__ void __launch_bounds__(RS_THREADS, 4) k_os_scatter(
    const uint32_t *__restrict__ keysIn, const uint32_t *__restrict__ valsIn, uint32_t *__restrict__ keysOut,
    uint32_t *__restrict__ valsOut, uint32_t n, int shift, const uint32_t *__restrict__ base /* [256] of this pass */,
    uint32_t *__restrict__ status /* [numTiles][256] of this pass, zeroed */, uint32_t *__restrict__ tileCounter,
    int iotaVals) {
    __shared__ uint32_t whist[RS_THREADS / 32][256];
    __shared__ uint32_t binLocal[256], binGlobal[256], wsum[RS_THREADS / 32];
    __shared__ uint32_t skey[RS_TILE], sval[RS_TILE];
    __shared__ uint32_t sTile;
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (threadIdx.x == 0) sTile = atomicAdd(tileCounter, 1u);
    for (int i = threadIdx.x; i < (RS_THREADS / 32) * 256; i += RS_THREADS) (&whist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = sTile;
    const uint32_t tbase = tile * RS_TILE + w * (RS_ITEMS * 32);
    uint32_t key[RS_ITEMS], rank[RS_ITEMS];
    const uint32_t lt = (1u << lane) - 1u;
    // (a) all keys of the lane in flight at once
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
        const uint32_t idx = tbase + k * 32 + lane;
        key[k] = idx < n ? __ldg(keysIn + idx) : 0xffffffffu;
    }
    // (b) the lanes of the warp that hold the same digit, for every item: RS_ITEMS independent MATCH operations
    //     (rank[k] holds the peer mask until (c))
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
        const bool valid = tbase + k * 32 + lane < n;
        const uint32_t vmask = __ballot_sync(0xffffffffu, valid);
        rank[k] = valid ? __match_any_sync(vmask, (key[k] >> shift) & 255u) : 0u;
    }
    // (c) stable rank inside the warp: the only serial chain left is one shared-memory load and one store per item
    //     (every peer reads the running count itself; the lowest peer lane adds the group)
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
        const uint32_t peers = rank[k];
        const uint32_t d = (key[k] >> shift) & 255u;
        uint32_t old = 0;
        if (peers) old = whist[w][d];
        __syncwarp();
        if (peers && (peers & lt) == 0u) whist[w][d] = old + __popc(peers);
        __syncwarp();
        rank[k] = old + __popc(peers & lt);
    }
    __syncthreads();
    {
        const uint32_t d = threadIdx.x;
        uint32_t run = 0;
#pragma unroll
        for (int ww = 0; ww < RS_THREADS / 32; ++ww) {
            const uint32_t c = whist[ww][d];
            whist[ww][d] = run;
            run += c;
        }
        // ---- decoupled look-back for digit d: exclusive count of this digit over all earlier tiles
        volatile uint32_t *st = status + (size_t)tile * 256u + d;
        if (tile > 0) *st = OS_AGG | run;
        else *st = OS_PRE | run;
        uint32_t excl = 0;
        if (tile > 0) {
            __threadfence();
            int t = (int)tile - 1;
            for (;;) {
                const uint32_t v = *(volatile uint32_t *)(status + (size_t)t * 256u + d);
                if (v & OS_PRE) { excl += v & OS_MASK; break; }
                if (v & OS_AGG) { excl += v & OS_MASK; --t; continue; }
                // not published yet: the tile has started (ids come from the counter); wait
            }
            *st = OS_PRE | (excl + run);
        }
        uint32_t incl = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t tt = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= (uint32_t)o) incl += tt;
        }
        if (lane == 31) wsum[w] = incl;
        __syncthreads();
        uint32_t woff = 0;
        for (uint32_t i = 0; i < w; ++i) woff += wsum[i];
        const uint32_t exclLocal = woff + incl - run;
        binLocal[d] = exclLocal;
        binGlobal[d] = __ldg(base + d) + excl - exclLocal;   // dst = binGlobal[d] + position inside the sorted tile
    }
    __syncthreads();
    // keys into their sorted place of the tile; then the values with all loads in flight at once (the key
    // registers are free by then)
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
        if (tbase + k * 32 + lane < n) {
            const uint32_t d = (key[k] >> shift) & 255u;
            const uint32_t lpos = binLocal[d] + whist[w][d] + rank[k];
            skey[lpos] = key[k];
            rank[k] = lpos;
        }
    }
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
        const uint32_t idx = tbase + k * 32 + lane;
        key[k] = (iotaVals || idx >= n) ? idx : __ldg(valsIn + idx);
    }
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k)
        if (tbase + k * 32 + lane < n) sval[rank[k]] = key[k];
    __syncthreads();
    const uint32_t tile0 = tile * RS_TILE;
    const uint32_t nValid = n - tile0 < (uint32_t)RS_TILE ? n - tile0 : (uint32_t)RS_TILE;
    for (uint32_t i = threadIdx.x; i < nValid; i += RS_THREADS) {
        const uint32_t kk = skey[i];
        const uint32_t dst = binGlobal[(kk >> shift) & 255u] + i;
        keysOut[dst] = kk;
        valsOut[dst] = sval[i];
    }
}

// ---------------------------------------------------------------- gather to Morton order
// leaf box: sphere bound p +- r*(1+2e-6), then one ulp outward (DESIGN.md "conservative cull")
__device__ __forceinline__ void leaf_box(const float4 p, float4 &lo, float4 &hi) {
    const float rp = fmul(p.w, 1.000002f);
    lo.x = next_down(fsub(p.x, rp)); hi.x = next_up(fadd(p.x, rp));
    lo.y = next_down(fsub(p.y, rp)); hi.y = next_up(fadd(p.y, rp));
    lo.z = next_down(fsub(p.z, rp)); hi.z = next_up(fadd(p.z, rp));
    lo.w = 0.f; hi.w = 0.f;
}
__device__ __forceinline__ float box_abs_max(const float4 lo, const float4 hi) {
    return fmaxf(fmaxf(fmaxf(fabsf(lo.x), fabsf(hi.x)), fmaxf(fabsf(lo.y), fabsf(hi.y))),
                 fmaxf(fabsf(lo.z), fabsf(hi.z)));
}

// The per-photon leaf boxes are NOT stored on the hot path: the production tree's boxes are
// derived from the Morton-ordered photons directly (k_refit_fast, k_emit_fast); the canonical
// 1-photon-per-leaf boxes are materialised on demand (ensure_canonical) for counting / parity.
__global__ void __launch_bounds__(256) k_gather(
    const float4 *__restrict__ pos_r, const float4 *__restrict__ pw, const float4 *__restrict__ dr,
    const uint32_t *__restrict__ order, uint32_t n, float4 *__restrict__ s_pos_r,
    float4 *__restrict__ s_pow, float4 *__restrict__ s_dir, uint32_t *__restrict__ bnd) {
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    float amax = 0.f;
    if (s < n) {
        uint32_t i = __ldg(order + s);
        float4 p = __ldg(pos_r + i);
        s_pos_r[s] = p;
        s_pow[s] = __ldg(pw + i);
        s_dir[s] = __ldg(dr + i);
        float4 lo, hi;
        leaf_box(p, lo, hi);
        amax = box_abs_max(lo, hi);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    if ((threadIdx.x & 31) == 0 && amax > 0.f) atomicMax(bnd + 6, __float_as_uint(amax));
}

// The random 16-byte gathers of k_gather cost a 64-byte DRAM access each (ncu: 2.65 GB moved for
// 0.48 GB of payload, 0.89 ms at 10 M photons).  Interleaving the three store arrays first (one
// streaming pass, 0.15 ms) makes a photon one 48-byte run, 1.5 accesses on average instead of 3:
// k_gather_aos 0.68 ms, build 2.47 -> 2.40 ms.  BRE_GATHER_AOS=0 restores the direct gather.
__global__ void __launch_bounds__(256) k_interleave(const float4 *__restrict__ pos_r, const float4 *__restrict__ pw,
                                                    const float4 *__restrict__ dr, uint32_t n, float4 *__restrict__ aos) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 *r = aos + (size_t)3 * i;
    r[0] = __ldg(pos_r + i); r[1] = __ldg(pw + i); r[2] = __ldg(dr + i);
}
__global__ void __launch_bounds__(256) k_gather_aos(const float4 *__restrict__ aos, const uint32_t *__restrict__ order,
                                                    uint32_t n, float4 *__restrict__ s_pos_r, float4 *__restrict__ s_pow,
                                                    float4 *__restrict__ s_dir, uint32_t *__restrict__ bnd) {
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    float amax = 0.f;
    if (s < n) {
        const float4 *r = aos + (size_t)3 * __ldg(order + s);
        const float4 p = __ldg(r), a = __ldg(r + 1), b = __ldg(r + 2);
        s_pos_r[s] = p; s_pow[s] = a; s_dir[s] = b;
        float4 lo, hi;
        leaf_box(p, lo, hi);
        amax = box_abs_max(lo, hi);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    if ((threadIdx.x & 31) == 0 && amax > 0.f) atomicMax(bnd + 6, __float_as_uint(amax));
}

// Lazy payload (default): only the 16-byte position records move into Morton order; power / direction stay in the
// store and are read through `order` where a (tile, photon) entry is staged (bre_query_tile.cuh).
__global__ void __launch_bounds__(256) k_gather_pos(const float4 *__restrict__ pos_r, const uint32_t *__restrict__ order,
                                                    uint32_t n, float4 *__restrict__ s_pos_r, uint32_t *__restrict__ bnd) {
    // four sorted positions per thread, block-strided: four random 16-byte reads in flight per thread
    const uint32_t s0 = blockIdx.x * (blockDim.x * 4u) + threadIdx.x;
    uint32_t src[4];
    float4 p[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { const uint32_t s = s0 + k * 256u; src[k] = s < n ? __ldg(order + s) : 0u; }
#pragma unroll
    for (int k = 0; k < 4; ++k) p[k] = __ldg(pos_r + src[k]);
    float amax = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t s = s0 + k * 256u;
        if (s < n) {
            s_pos_r[s] = p[k];
            float4 lo, hi;
            leaf_box(p[k], lo, hi);
            amax = fmaxf(amax, box_abs_max(lo, hi));
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    if ((threadIdx.x & 31) == 0 && amax > 0.f) atomicMax(bnd + 6, __float_as_uint(amax));
}
__global__ void __launch_bounds__(256) k_gather_payload(const float4 *__restrict__ pw, const float4 *__restrict__ dr,
                                                        const uint32_t *__restrict__ order, uint32_t n,
                                                        float4 *__restrict__ s_pow, float4 *__restrict__ s_dir) {
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const uint32_t i = __ldg(order + s);
    s_pow[s] = __ldg(pw + i);
    s_dir[s] = __ldg(dr + i);
}

__global__ void __launch_bounds__(256) k_leafboxes(const float4 *__restrict__ s_pos_r, uint32_t n,
                                                   float4 *__restrict__ leaf_lo,
                                                   float4 *__restrict__ leaf_hi) {
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    float4 lo, hi;
    leaf_box(__ldg(s_pos_r + s), lo, hi);
    leaf_lo[s] = lo;
    leaf_hi[s] = hi;
}

// Refit-only update: positions, Morton order and topology are unchanged, only the
// radii moved (progressive shrink, integrators/photonbeam.cpp:354-356,562).  Re-reads each photon's
// radius from the store and rewrites the sorted record's .w and the leaf box -- the same arithmetic
// as k_gather, so the refitted tree is bit-identical to a full rebuild.
__global__ void __launch_bounds__(256) k_releaf(const float4 *__restrict__ pos_r,
                                                const uint32_t *__restrict__ order, uint32_t n,
                                                float4 *__restrict__ s_pos_r, uint32_t *__restrict__ bnd) {
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    float amax = 0.f;
    if (s < n) {
        float4 p = s_pos_r[s];
        p.w = __ldg(&pos_r[__ldg(order + s)].w);
        s_pos_r[s] = p;
        float4 lo, hi;
        leaf_box(p, lo, hi);
        amax = box_abs_max(lo, hi);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    if ((threadIdx.x & 31) == 0 && amax > 0.f) atomicMax(bnd + 6, __float_as_uint(amax));
}

// ---------------------------------------------------------------- Karras radix tree
__device__ __forceinline__ int delta64(const uint32_t *__restrict__ codes,
                                       const uint32_t *__restrict__ order, uint64_t ki, int j,
                                       int n) {
    if (j < 0 || j >= n) return -1;
    uint64_t kj = ((uint64_t)__ldg(codes + j) << 32) | __ldg(order + j);
    return __clzll((long long)(ki ^ kj));
}

// ctx->refit_direct (default 16, env BRE_REFIT_DIRECT): nodes up to this many photons get their box
// directly from the photons in k_refit_fast; measured 8/16/32/64/128 -> build 2.71/2.65/2.67/2.78/2.97 ms

// node_lo[i].w / node_hi[i].w hold the left / right child refs (>=0 internal, <0 = ~leaf).
// parent[0..n-2] internal parents, parent[n-1+s] leaf parents.  node_range[i] = {first,last}.
// Work lists for the hot-path refit (k_refit_direct / k_refit_climb), appended block by block:
//   dlist: kept nodes of <= directMax photons -- fat-leaf roots (<= maxLeaf photons under a larger
//          node) and small internal nodes; their box comes straight from the photons
//   clist: where the atomic climb starts -- children (node, or n-1+s for a single photon s) of
//          <= directMax photons under a larger node
// lists[0] = |dlist|, lists[1] = |clist|, dlist at lists+16, clist at lists+16+n.
__global__ void __launch_bounds__(256) k_karras(const uint32_t *__restrict__ codes,
                                                const uint32_t *__restrict__ order, int n,
                                                float4 *__restrict__ node_lo,
                                                float4 *__restrict__ node_hi,
                                                int2 *__restrict__ node_range,
                                                int *__restrict__ parent, int maxLeaf, int directMax,
                                                int *__restrict__ lists) {
    __shared__ int wsum[2][8];
    __shared__ int bbase[2];
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int dItem[3] = {-1, -1, -1}, cItem[2] = {-1, -1};
    if (i < n - 1) {
        uint64_t ki = ((uint64_t)__ldg(codes + i) << 32) | __ldg(order + i);
        int d = delta64(codes, order, ki, i + 1, n) - delta64(codes, order, ki, i - 1, n) > 0 ? 1 : -1;
        int dmin = delta64(codes, order, ki, i - d, n);
        int lmax = 2;
        while (delta64(codes, order, ki, i + lmax * d, n) > dmin) lmax <<= 1;
        int l = 0;
        for (int t = lmax >> 1; t >= 1; t >>= 1)
            if (delta64(codes, order, ki, i + (l + t) * d, n) > dmin) l += t;
        int j = i + l * d;
        int dnode = delta64(codes, order, ki, j, n);
        int s = 0, t = l;
        do {
            t = (t + 1) >> 1;
            if (delta64(codes, order, ki, i + (s + t) * d, n) > dnode) s += t;
        } while (t > 1);
        int gamma = i + s * d + min(d, 0);
        int first = min(i, j), last = max(i, j);
        int left = (first == gamma) ? ~gamma : gamma;
        int right = (last == gamma + 1) ? ~(gamma + 1) : gamma + 1;
        node_lo[i].w = __int_as_float(left);
        node_hi[i].w = __int_as_float(right);
        node_range[i] = make_int2(first, last);
        if (left >= 0) parent[left] = i; else parent[n - 1 + gamma] = i;
        if (right >= 0) parent[right] = i; else parent[n - 1 + gamma + 1] = i;
        if (i == 0) parent[0] = -1;
        const int cnt = last - first + 1, cl = gamma - first + 1, cr = last - gamma;
        const bool big = cnt > maxLeaf, huge = cnt > directMax;
        if (big && left >= 0 && cl <= directMax) dItem[0] = left;
        if (big && right >= 0 && cr <= directMax) dItem[1] = right;
        if (i == 0 && !huge) dItem[2] = 0;   // the whole tree is small: root box direct
        if (huge && cl <= directMax) cItem[0] = left >= 0 ? left : n - 1 + gamma;
        if (huge && cr <= directMax) cItem[1] = right >= 0 ? right : n - 1 + gamma + 1;
    }
    // block-aggregated append: one atomicAdd per list per block
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    int nd = (dItem[0] >= 0) + (dItem[1] >= 0) + (dItem[2] >= 0), nc = (cItem[0] >= 0) + (cItem[1] >= 0);
    int pd = nd, pc = nc;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int a = __shfl_up_sync(0xffffffffu, pd, o), c = __shfl_up_sync(0xffffffffu, pc, o);
        if ((int)lane >= o) { pd += a; pc += c; }
    }
    if (lane == 31) { wsum[0][warp] = pd; wsum[1][warp] = pc; }
    __syncthreads();
    if (threadIdx.x < 2) {
        int tot = 0;
        for (int w = 0; w < 8; ++w) { int v = wsum[threadIdx.x][w]; wsum[threadIdx.x][w] = tot; tot += v; }
        bbase[threadIdx.x] = tot ? atomicAdd(lists + threadIdx.x, tot) : 0;
    }
    __syncthreads();
    int od = bbase[0] + wsum[0][warp] + pd - nd, oc = bbase[1] + wsum[1][warp] + pc - nc;
    int *dlist = lists + 16, *clist = lists + 16 + n;
    if (dItem[0] >= 0) dlist[od++] = dItem[0];
    if (dItem[1] >= 0) dlist[od++] = dItem[1];
    if (dItem[2] >= 0) dlist[od] = dItem[2];
    if (cItem[0] >= 0) clist[oc++] = cItem[0];
    if (cItem[1] >= 0) clist[oc] = cItem[1];
}

__global__ void __launch_bounds__(256) k_refit(int n, float4 *node_lo, float4 *node_hi,
                                               const float4 *__restrict__ leaf_lo,
                                               const float4 *__restrict__ leaf_hi,
                                               const int *__restrict__ parent,
                                               uint32_t *__restrict__ flags) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    int cur = __ldg(parent + n - 1 + s);
    while (cur >= 0) {
        __threadfence();
        if (atomicAdd(flags + cur, 1u) == 0u) return;  // first arrival: sibling not done yet
        __threadfence();
        float4 nlo = __ldcg(node_lo + cur), nhi = __ldcg(node_hi + cur);
        int left = __float_as_int(nlo.w), right = __float_as_int(nhi.w);
        float4 llo, lhi, rlo, rhi;
        if (left >= 0) { llo = __ldcg(node_lo + left); lhi = __ldcg(node_hi + left); }
        else { llo = __ldg(leaf_lo + ~left); lhi = __ldg(leaf_hi + ~left); }
        if (right >= 0) { rlo = __ldcg(node_lo + right); rhi = __ldcg(node_hi + right); }
        else { rlo = __ldg(leaf_lo + ~right); rhi = __ldg(leaf_hi + ~right); }
        nlo.x = fminf(llo.x, rlo.x); nlo.y = fminf(llo.y, rlo.y); nlo.z = fminf(llo.z, rlo.z);
        nhi.x = fmaxf(lhi.x, rhi.x); nhi.y = fmaxf(lhi.y, rhi.y); nhi.z = fmaxf(lhi.z, rhi.z);
        __stcg(node_lo + cur, nlo);
        __stcg(node_hi + cur, nhi);
        cur = __ldg(parent + cur);
    }
}

// Hot-path refit: only the nodes the production tree keeps.  A "fat leaf" is a subtree of
// <= maxLeaf photons whose parent is larger; its box is the union of its photons' boxes, computed
// straight from the Morton-ordered records (contiguous, <= 128 bytes), and so is the box of every
// kept node of <= directMax photons (k_refit_direct, one thread per entry of the list k_karras
// appended: full warps instead of one busy lane in six).  k_refit_climb then starts from the
// children of <= directMax photons under larger nodes with the usual second-arrival-continues
// scheme.  Nodes inside fat leaves are never touched (7/8 of the
// tree at maxLeaf 8).  min/max are exact, so every box equals the canonical bottom-up union.
__global__ void __launch_bounds__(256) k_refit_direct(const int *__restrict__ lists, float4 *node_lo, float4 *node_hi,
                                                      const int2 *__restrict__ node_range,
                                                      const float4 *__restrict__ s_pos_r) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= __ldg(lists)) return;
    const int id = __ldg(lists + 16 + k);
    // union of the node's photons' boxes (contiguous Morton range, shared with its neighbours
    // through L1/L2) -- no dependency, no atomics
    const int2 rg = __ldg(node_range + id);
    float4 lo, hi;
    leaf_box(__ldg(s_pos_r + rg.x), lo, hi);
    for (int q = rg.x + 1; q <= rg.y; ++q) {
        float4 l, h;
        leaf_box(__ldg(s_pos_r + q), l, h);
        lo.x = fminf(lo.x, l.x); lo.y = fminf(lo.y, l.y); lo.z = fminf(lo.z, l.z);
        hi.x = fmaxf(hi.x, h.x); hi.y = fmaxf(hi.y, h.y); hi.z = fmaxf(hi.z, h.z);
    }
    lo.w = __ldcg(&node_lo[id].w); hi.w = __ldcg(&node_hi[id].w);
    __stcg(node_lo + id, lo);
    __stcg(node_hi + id, hi);
}

// second-arrival climb above the direct nodes (their boxes are complete: previous launch)
__global__ void __launch_bounds__(256) k_refit_climb(int n, const int *__restrict__ lists, float4 *node_lo, float4 *node_hi,
                                                     const int *__restrict__ parent,
                                                     const float4 *__restrict__ s_pos_r,
                                                     uint32_t *__restrict__ flags) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= __ldg(lists + 1)) return;
    int cur = __ldg(parent + __ldg(lists + 16 + n + k));
    bool fence = false;     // nothing of mine to publish before the first arrival
    while (cur >= 0) {
        if (fence) __threadfence();   // my stores are visible before the arrival is counted
        if (atomicAdd(flags + cur, 1u) == 0u) return;  // first arrival: sibling not done yet
        float4 nlo = __ldcg(node_lo + cur), nhi = __ldcg(node_hi + cur);
        const int left = __float_as_int(nlo.w), right = __float_as_int(nhi.w);
        float4 llo, lhi, rlo, rhi;
        if (left >= 0) { llo = __ldcg(node_lo + left); lhi = __ldcg(node_hi + left); }
        else leaf_box(__ldg(s_pos_r + ~left), llo, lhi);
        if (right >= 0) { rlo = __ldcg(node_lo + right); rhi = __ldcg(node_hi + right); }
        else leaf_box(__ldg(s_pos_r + ~right), rlo, rhi);
        nlo.x = fminf(llo.x, rlo.x); nlo.y = fminf(llo.y, rlo.y); nlo.z = fminf(llo.z, rlo.z);
        nhi.x = fmaxf(lhi.x, rhi.x); nhi.y = fmaxf(lhi.y, rhi.y); nhi.z = fmaxf(lhi.z, rhi.z);
        __stcg(node_lo + cur, nlo);
        __stcg(node_hi + cur, nhi);
        fence = true;
        cur = __ldg(parent + cur);
    }
}

#if BRE_WIDE
// One slot of a wide node: child c of a kept node is a single photon (c < 0), a fat-leaf root
// (<= maxLeaf photons) or a kept internal node.
__device__ __forceinline__ int emit_slot(int c, int maxLeaf, const float4 *__restrict__ node_lo,
                                         const float4 *__restrict__ node_hi, const int2 *__restrict__ node_range,
                                         const float4 *__restrict__ s_pos_r, float4 &lo, float4 &hi) {
    if (c < 0) {
        leaf_box(__ldg(s_pos_r + ~c), lo, hi);
        return leaf_ref((uint32_t)~c, 1u);
    }
    lo = __ldg(node_lo + c);
    hi = __ldg(node_hi + c);
    const int2 cr = __ldg(node_range + c);
    const int cnt = cr.y - cr.x + 1;
    return cnt <= maxLeaf ? leaf_ref((uint32_t)cr.x, (uint32_t)cnt) : c;
}
// Every kept internal node gets a wide node (only those an even number of levels below the root
// are ever visited; telling them apart would need a top-down pass over the Karras tree).
__global__ void __launch_bounds__(256) k_emit_fast(int n, int maxLeaf,
                                                   const float4 *__restrict__ node_lo,
                                                   const float4 *__restrict__ node_hi,
                                                   const int2 *__restrict__ node_range,
                                                   const float4 *__restrict__ s_pos_r,
                                                   FastNode *__restrict__ fast) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int2 rg = __ldg(node_range + i);
    if (rg.y - rg.x + 1 <= maxLeaf && i != 0) return;  // inside a fat leaf: never visited
    const int ch[2] = {__float_as_int(__ldg(&node_lo[i].w)), __float_as_int(__ldg(&node_hi[i].w))};
    // slots 0,1 come from child 0 and slots 2,3 from child 1 (fixed positions: registers only)
    float bx[4][6];
    int ref[4] = {FAST_NONE, FAST_NONE, FAST_NONE, FAST_NONE};
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int j = 0; j < 6; ++j) bx[k][j] = 0.f;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        float4 lo, hi;
        const int r = emit_slot(ch[k], maxLeaf, node_lo, node_hi, node_range, s_pos_r, lo, hi);
        if (r < 0) {   // leaf: one slot
            bx[2 * k][0] = lo.x; bx[2 * k][1] = lo.y; bx[2 * k][2] = lo.z;
            bx[2 * k][3] = hi.x; bx[2 * k][4] = hi.y; bx[2 * k][5] = hi.z;
            ref[2 * k] = r;
        } else {       // kept internal child: its two children take its place
            const int g[2] = {__float_as_int(lo.w), __float_as_int(hi.w)};
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                float4 glo, ghi;
                ref[2 * k + j] = emit_slot(g[j], maxLeaf, node_lo, node_hi, node_range, s_pos_r, glo, ghi);
                bx[2 * k + j][0] = glo.x; bx[2 * k + j][1] = glo.y; bx[2 * k + j][2] = glo.z;
                bx[2 * k + j][3] = ghi.x; bx[2 * k + j][4] = ghi.y; bx[2 * k + j][5] = ghi.z;
            }
        }
    }
    FastNode *f = fast + i;
    f->q[0] = make_float4(bx[0][0], bx[0][1], bx[0][2], bx[0][3]);
    f->q[1] = make_float4(bx[0][4], bx[0][5], bx[1][0], bx[1][1]);
    f->q[2] = make_float4(bx[1][2], bx[1][3], bx[1][4], bx[1][5]);
    f->q[3] = make_float4(bx[2][0], bx[2][1], bx[2][2], bx[2][3]);
    f->q[4] = make_float4(bx[2][4], bx[2][5], bx[3][0], bx[3][1]);
    f->q[5] = make_float4(bx[3][2], bx[3][3], bx[3][4], bx[3][5]);
    f->ref = make_int4(ref[0], ref[1], ref[2], ref[3]);
}
#else
__global__ void __launch_bounds__(256) k_emit_fast(int n, int maxLeaf,
                                                   const float4 *__restrict__ node_lo,
                                                   const float4 *__restrict__ node_hi,
                                                   const int2 *__restrict__ node_range,
                                                   const float4 *__restrict__ s_pos_r,
                                                   FastNode *__restrict__ fast) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int2 rg = __ldg(node_range + i);
    if (rg.y - rg.x + 1 <= maxLeaf && i != 0) return;  // inside a fat leaf: never visited
    float4 nlo = __ldg(node_lo + i), nhi = __ldg(node_hi + i);
    int ch[2] = {__float_as_int(nlo.w), __float_as_int(nhi.w)};
    float4 lo[2], hi[2];
    int ref[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        if (ch[k] >= 0) {
            lo[k] = __ldg(node_lo + ch[k]);
            hi[k] = __ldg(node_hi + ch[k]);
            int2 cr = __ldg(node_range + ch[k]);
            int cnt = cr.y - cr.x + 1;
            ref[k] = cnt <= maxLeaf ? leaf_ref((uint32_t)cr.x, (uint32_t)cnt) : ch[k];
        } else {
            leaf_box(__ldg(s_pos_r + ~ch[k]), lo[k], hi[k]);
            ref[k] = leaf_ref((uint32_t)~ch[k], 1u);
        }
    }
    FastNode f;
    f.a = make_float4(lo[0].x, lo[0].y, lo[0].z, hi[0].x);
    f.b = make_float4(hi[0].y, hi[0].z, lo[1].x, lo[1].y);
    f.c = make_float4(lo[1].z, hi[1].x, hi[1].y, hi[1].z);
    f.d = make_int4(ref[0], ref[1], 0, 0);
    fast[i] = f;
}

#endif

// ---------------------------------------------------------------- top cut
// The tile walk of K5 (bre_query_tile.cuh) starts from a CUT through the top of the production tree
// instead of its root: up to TOP_CUT = 512 (box, ref) entries that together cover every photon,
// staged into shared memory with one TMA bulk copy per block and tested in one step by the 256
// threads of a tile -- no ramp-up through ten nearly empty levels, no L2 round trip for the top of
// the tree.  One block refines the cut breadth-first; when a whole level no longer fits, the
// entries with the largest surface area are opened first (the reference's HLBVH also treats the top
// of the tree separately: treelets on the high Morton bits, core/photonbeambvh.cpp:456-493).
// Layout: cut[2k] = {lo.xyz, ref}, cut[2k+1] = {hi.xyz, 0}; cut[2 * TOP_CUT].x = number of entries.
constexpr int TOP_CUT_N = TOP_CUT;
#if !BRE_WIDE
__global__ void __launch_bounds__(TOP_CUT_N) k_top_cut(const FastNode *__restrict__ fast, int rootRef,
                                                      float4 *__restrict__ cut) {
    __shared__ float4 slo[TOP_CUT_N], shi[TOP_CUT_N];
    __shared__ float sarea[TOP_CUT_N];
    __shared__ int wsum[TOP_CUT_N / 32];
    __shared__ int ssize;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) ssize = 0;
    if (rootRef >= 0 && tid == 0) {
        const FastNode f = fast[rootRef];
        slo[0] = make_float4(f.a.x, f.a.y, f.a.z, __int_as_float(f.d.x));
        shi[0] = make_float4(f.a.w, f.b.x, f.b.y, 0.f);
        slo[1] = make_float4(f.b.z, f.b.w, f.c.x, __int_as_float(f.d.y));
        shi[1] = make_float4(f.c.y, f.c.z, f.c.w, 0.f);
        ssize = 2;
    }
    __syncthreads();
    for (int round = 0; round < 64; ++round) {
        const int sz = ssize;
        if (sz >= TOP_CUT_N || sz == 0) break;
        const bool internal = tid < sz && __float_as_int(slo[tid].w) >= 0;
        float a = -1.f;
        if (internal) {
            const float dx = shi[tid].x - slo[tid].x, dy = shi[tid].y - slo[tid].y, dz = shi[tid].z - slo[tid].z;
            a = dx * dy + dy * dz + dz * dx;
            if (!(a >= 0.f)) a = 0.f;
        }
        sarea[tid] = a;
        const int m = __syncthreads_count(internal);
        if (m == 0) break;
        const int avail = TOP_CUT_N - sz;
        bool expand = internal;
        if (m > avail) {
            int rank = 0;
            for (int k = 0; k < sz; ++k) {
                const float b = sarea[k];
                rank += (b > a || (b == a && k < tid)) ? 1 : 0;
            }
            expand = internal && rank < avail;
        }
        // block-wide exclusive prefix of `expand`
        const unsigned mb = __ballot_sync(0xffffffffu, expand);
        if (lane == 0) wsum[warp] = __popc(mb);
        __syncthreads();
        int pos = __popc(mb & ((1u << lane) - 1u)), tot = 0;
        for (int w = 0; w < TOP_CUT_N / 32; ++w) {
            if (w < warp) pos += wsum[w];
            tot += wsum[w];
        }
        if (expand) {
            const FastNode *fn = fast + __float_as_int(slo[tid].w);
            const float4 fa = __ldg(&fn->a), fb = __ldg(&fn->b), fc = __ldg(&fn->c);
            const int4 fd = __ldg(&fn->d);
            slo[tid] = make_float4(fa.x, fa.y, fa.z, __int_as_float(fd.x));
            shi[tid] = make_float4(fa.w, fb.x, fb.y, 0.f);
            slo[sz + pos] = make_float4(fb.z, fb.w, fc.x, __int_as_float(fd.y));
            shi[sz + pos] = make_float4(fc.y, fc.z, fc.w, 0.f);
        }
        __syncthreads();
        if (tid == 0) ssize = sz + tot;
        __syncthreads();
    }
    const int sz = ssize;
    if (tid < sz) { cut[2 * tid] = slo[tid]; cut[2 * tid + 1] = shi[tid]; }
    else {
        cut[2 * tid] = make_float4(3.0e38f, 3.0e38f, 3.0e38f, __int_as_float(FAST_NONE));
        cut[2 * tid + 1] = make_float4(-3.0e38f, -3.0e38f, -3.0e38f, 0.f);
    }
    if (tid == 0) cut[2 * TOP_CUT_N] = make_float4(__uint_as_float((unsigned)sz), 0.f, 0.f, 0.f);
}
#endif

// ---------------------------------------------------------------- host driver
// production-tree boxes: fat-leaf roots and everything above them, then the 64-byte nodes
static int launch_refit_emit(bre_ctx *ctx) {
    const uint32_t n = ctx->n_built;
    cudaStream_t st = ctx->stream;
    BRE_CUDA(cudaMemsetAsync(ctx->flags.p, 0, (size_t)n * 4, st));
    // list lengths live on the device: both kernels are launched for the worst case (n items) and
    // the surplus blocks return at once
    const int gl = (int)((n + 255) / 256);
    k_refit_direct<<<gl, 256, 0, st>>>((const int *)ctx->entry.p, (float4 *)ctx->node_lo.p, (float4 *)ctx->node_hi.p,
                                       (const int2 *)ctx->node_range.p, (const float4 *)ctx->s_pos_r.p);
    k_refit_climb<<<gl, 256, 0, st>>>((int)n, (const int *)ctx->entry.p, (float4 *)ctx->node_lo.p,
                                      (float4 *)ctx->node_hi.p, (const int *)ctx->parent.p,
                                      (const float4 *)ctx->s_pos_r.p, (uint32_t *)ctx->flags.p);
    k_emit_fast<<<(int)((n + 255) / 256), 256, 0, st>>>(
        (int)n, ctx->max_leaf, (const float4 *)ctx->node_lo.p, (const float4 *)ctx->node_hi.p,
        (const int2 *)ctx->node_range.p, (const float4 *)ctx->s_pos_r.p, (FastNode *)ctx->fast.p);
#if !BRE_WIDE
    {
        int s = ensure(ctx, ctx->top_cut, (size_t)(2 * TOP_CUT_N + 1) * 16);
        if (s != BRE_OK) return s;
        const int rootRef = (n <= (uint32_t)ctx->max_leaf) ? -1 : 0;
        k_top_cut<<<1, TOP_CUT_N, 0, st>>>((const FastNode *)ctx->fast.p, rootRef, (float4 *)ctx->top_cut.p);
    }
#endif
    return BRE_OK;
}

// Canonical 1-photon-per-leaf boxes (leaf boxes + every internal node), materialised on demand for
// the counting kernel, bre_download_tree and parity tests; not part of the render path.
int ensure_sorted_payload(bre_ctx *ctx) {
    if (ctx->payload_sorted || !ctx->built || ctx->n_built == 0) return BRE_OK;
    const uint32_t n = ctx->n_built;
    int s;
    if ((s = ensure(ctx, ctx->s_pow, (size_t)n * 16)) != BRE_OK) return s;
    if ((s = ensure(ctx, ctx->s_dir, (size_t)n * 16)) != BRE_OK) return s;
    k_gather_payload<<<(n + 255) / 256, 256, 0, ctx->stream>>>((const float4 *)ctx->pow.p, (const float4 *)ctx->dir.p,
                                                              (const uint32_t *)ctx->order.p, n, (float4 *)ctx->s_pow.p,
                                                              (float4 *)ctx->s_dir.p);
    BRE_CUDA(cudaGetLastError());
    ctx->payload_sorted = true;
    return BRE_OK;
}

int ensure_canonical(bre_ctx *ctx) {
    if (ctx->canon_valid || !ctx->built || ctx->n_built == 0) return BRE_OK;
    const uint32_t n = ctx->n_built;
    cudaStream_t st = ctx->stream;
    int s;
    if ((s = ensure(ctx, ctx->leaf_lo, (size_t)n * 16)) != BRE_OK) return s;
    if ((s = ensure(ctx, ctx->leaf_hi, (size_t)n * 16)) != BRE_OK) return s;
    const int g256 = (int)((n + 255) / 256);
    k_leafboxes<<<g256, 256, 0, st>>>((const float4 *)ctx->s_pos_r.p, n, (float4 *)ctx->leaf_lo.p,
                                      (float4 *)ctx->leaf_hi.p);
    if (n > 1) {
        BRE_CUDA(cudaMemsetAsync(ctx->flags.p, 0, (size_t)n * 4, st));
        k_refit<<<g256, 256, 0, st>>>((int)n, (float4 *)ctx->node_lo.p, (float4 *)ctx->node_hi.p,
                                      (const float4 *)ctx->leaf_lo.p, (const float4 *)ctx->leaf_hi.p,
                                      (const int *)ctx->parent.p, (uint32_t *)ctx->flags.p);
    }
    BRE_CUDA(cudaGetLastError());
    ctx->canon_valid = true;
    return BRE_OK;
}

static int refit_tree(bre_ctx *ctx) {
    const uint32_t n = ctx->n_built;
    cudaStream_t st = ctx->stream;
    const int g256 = (int)((n + 255) / 256);
    uint32_t *bnd = (uint32_t *)ctx->scratch.p;
    {
        PhaseTimer pt(ctx, PH_BUILD);
        BRE_CUDA(cudaMemsetAsync(bnd + 6, 0, 4, st));
        k_releaf<<<g256, 256, 0, st>>>((const float4 *)ctx->pos_r.p, (const uint32_t *)ctx->order.p, n,
                                       (float4 *)ctx->s_pos_r.p, bnd);
        if (n > 1) {
            int s2 = launch_refit_emit(ctx);
            if (s2 != BRE_OK) return s2;
        }
    }
    BRE_CUDA(cudaGetLastError());
    uint32_t amaxBits = 0;
    BRE_CUDA(cudaMemcpyAsync(&amaxBits, bnd + 6, 4, cudaMemcpyDeviceToHost, st));
    BRE_CUDA(cudaStreamSynchronize(st));
    memcpy(&ctx->abs_max, &amaxBits, 4);
    ctx->built = true;
    ctx->canon_valid = false;
    ctx->last_build_kind = 2;
    return BRE_OK;
}

int build_tree(bre_ctx *ctx) {
    ctx->built = false;
    {
        // K6 of the rays staged by bre_stage_*: ALU-bound, it fills the issue slots the sort / gather kernels leave idle
        const int sp = launch_pending_tau(ctx);
        if (sp != BRE_OK) return sp;
    }
    // only radii changed since the last full build (bre_set_radius): keep order and topology
    if (ctx->topo_valid && ctx->n_built == ctx->n_photons && ctx->n_built > 0) return refit_tree(ctx);
    ctx->topo_valid = false;
    if (ctx->n_photons >= BRE_MAX_PHOTONS)
        return set_err(ctx, BRE_ERR_UNSUPPORTED, "photon count exceeds 2^27 (leaf reference encoding)");
    const uint32_t n = (uint32_t)ctx->n_photons;
    ctx->n_built = n;
    ctx->abs_max = 0.f;
    ctx->root_ref = 0;
    if (n == 0) { ctx->built = true; return BRE_OK; }
    cudaStream_t st = ctx->stream;
    const size_t n4 = (size_t)n * sizeof(float4);
    int s;
#define ENS(buf, bytes) if ((s = ensure(ctx, ctx->buf, (bytes))) != BRE_OK) return s
    ENS(s_pos_r, n4);
    if (!ctx->lazy_payload) { ENS(s_pow, n4); ENS(s_dir, n4); }
    ctx->payload_sorted = false;
    ENS(order, (size_t)n * 4); ENS(codes, (size_t)n * 4);
    ENS(keys_alt, (size_t)n * 4); ENS(vals_alt, (size_t)n * 4);
    ENS(node_lo, n4); ENS(node_hi, n4);
    ENS(node_range, (size_t)n * 8); ENS(parent, (size_t)2 * n * 4);
    ENS(entry, (size_t)8 * n + 64); ENS(flags, (size_t)n * 4); ENS(fast, (size_t)n * sizeof(FastNode));
    ENS(scratch, 64);
    const uint32_t numBlocks = (n + RS_TILE - 1) / RS_TILE;
    const uint32_t histLen = 256u * numBlocks;
    const uint32_t scanTiles = (histLen + SCN_TILE - 1) / SCN_TILE;
    ENS(hist, ((size_t)histLen + scanTiles) * 4);
#undef ENS
    uint32_t *bnd = (uint32_t *)ctx->scratch.p;
    const float4 *pos_r = (const float4 *)ctx->pos_r.p;
    uint32_t *codes = (uint32_t *)ctx->codes.p, *order = (uint32_t *)ctx->order.p;
    uint32_t *kalt = (uint32_t *)ctx->keys_alt.p, *valt = (uint32_t *)ctx->vals_alt.p;
    uint32_t *hist = (uint32_t *)ctx->hist.p;
    uint32_t *tileSums = hist + histLen;
    const int g256 = (int)((n + 255) / 256);

    PhaseTimer *pt = new PhaseTimer(ctx, PH_BUILD);
    k_bounds_init<<<1, 32, 0, st>>>(bnd);
    k_bounds<<<min(g256, 148 * 8), 256, 0, st>>>(pos_r, n, bnd);
    k_morton<<<g256, 256, 0, st>>>(pos_r, n, bnd, codes);
    // 4 passes: codes/order ping-pong with kalt/valt; result ends in codes/order
    uint32_t *kin = codes, *vin = order, *kout = kalt, *vout = valt;
    if (ctx->sort_onesweep) {
        // scratch: [4][256] global bin bases | 4 tile counters | [4][numBlocks][256] look-back words
        const size_t osWords = 1024 + 16 + (size_t)4 * numBlocks * 256;
        if ((s = ensure(ctx, ctx->os_scratch, osWords * 4)) != BRE_OK) { delete pt; return s; }
        uint32_t *osb = (uint32_t *)ctx->os_scratch.p;
        BRE_CUDA(cudaMemsetAsync(osb, 0, osWords * 4, st));
        k_os_hist<<<std::min<uint32_t>(numBlocks, (uint32_t)ctx->sm_count * 8u), RS_THREADS, 0, st>>>(codes, n, osb);
        k_os_scan<<<1, 1024, 0, st>>>(osb);
        for (int pass = 0; pass < 4; ++pass) {
            k_os_scatter<<<numBlocks, RS_THREADS, 0, st>>>(kin, vin, kout, vout, n, 8 * pass, osb + 256 * pass,
                                                           osb + 1040 + (size_t)pass * numBlocks * 256, osb + 1024 + pass,
                                                           pass == 0);
            uint32_t *t = kin; kin = kout; kout = t;
            t = vin; vin = vout; vout = t;
        }
    } else
    for (int pass = 0; pass < 4; ++pass) {
        int shift = 8 * pass;
        k_rs_hist<<<numBlocks, RS_THREADS, 0, st>>>(kin, n, shift, hist, numBlocks);
        k_rs_scan_reduce<<<scanTiles, SCN_T, 0, st>>>(hist, histLen, tileSums);
        k_rs_scan_tiles<<<1, 1024, 0, st>>>(tileSums, scanTiles);
        k_rs_scan_apply<<<scanTiles, SCN_T, 0, st>>>(hist, histLen, tileSums);
        k_rs_scatter<<<numBlocks, RS_THREADS, 0, st>>>(kin, vin, kout, vout, n, shift, hist,
                                                       numBlocks, pass == 0);
        uint32_t *t = kin; kin = kout; kout = t;
        t = vin; vin = vout; vout = t;
    }
    // after 4 passes kin == codes, vin == order
    if (ctx->lazy_payload) {
        k_gather_pos<<<(n + 1023) / 1024, 256, 0, st>>>(pos_r, order, n, (float4 *)ctx->s_pos_r.p, bnd);
    } else if (ctx->gather_aos) {
        if ((s = ensure(ctx, ctx->aos, (size_t)n * 48)) != BRE_OK) { delete pt; return s; }
        k_interleave<<<g256, 256, 0, st>>>(pos_r, (const float4 *)ctx->pow.p, (const float4 *)ctx->dir.p, n,
                                           (float4 *)ctx->aos.p);
        k_gather_aos<<<g256, 256, 0, st>>>((const float4 *)ctx->aos.p, order, n, (float4 *)ctx->s_pos_r.p,
                                           (float4 *)ctx->s_pow.p, (float4 *)ctx->s_dir.p, bnd);
    } else
    k_gather<<<g256, 256, 0, st>>>(pos_r, (const float4 *)ctx->pow.p, (const float4 *)ctx->dir.p,
                                   order, n, (float4 *)ctx->s_pos_r.p, (float4 *)ctx->s_pow.p,
                                   (float4 *)ctx->s_dir.p, bnd);
    if (n > 1) {
        BRE_CUDA(cudaMemsetAsync(ctx->entry.p, 0, 64, st));
        k_karras<<<g256, 256, 0, st>>>(codes, order, (int)n, (float4 *)ctx->node_lo.p,
                                       (float4 *)ctx->node_hi.p, (int2 *)ctx->node_range.p,
                                       (int *)ctx->parent.p, ctx->max_leaf, ctx->refit_direct, (int *)ctx->entry.p);
        if ((s = launch_refit_emit(ctx)) != BRE_OK) { delete pt; return s; }
    }
    delete pt;
    BRE_CUDA(cudaGetLastError());
    uint32_t amaxBits = 0;
    BRE_CUDA(cudaMemcpyAsync(&amaxBits, bnd + 6, 4, cudaMemcpyDeviceToHost, st));
    BRE_CUDA(cudaStreamSynchronize(st));
    memcpy(&ctx->abs_max, &amaxBits, 4);
    ctx->root_ref = (n <= (uint32_t)ctx->max_leaf) ? leaf_ref(0u, n) : 0;
    ctx->built = true;
    ctx->payload_sorted = !ctx->lazy_payload;
    ctx->canon_valid = false;
    ctx->topo_valid = true;
    ctx->last_build_kind = 1;
    return BRE_OK;
}

TreeDev tree_view(const bre_ctx *ctx) {
    TreeDev t;
    t.n = ctx->n_built;
    t.s_pos_r = (const float4 *)ctx->s_pos_r.p;
    t.s_pow = ctx->payload_sorted ? (const float4 *)ctx->s_pow.p : nullptr;
    t.s_dir = ctx->payload_sorted ? (const float4 *)ctx->s_dir.p : nullptr;
    t.pow_u = (const float4 *)ctx->pow.p;
    t.dir_u = (const float4 *)ctx->dir.p;
    t.order = (const uint32_t *)ctx->order.p;
    t.node_lo = (const float4 *)ctx->node_lo.p;
    t.node_hi = (const float4 *)ctx->node_hi.p;
    t.leaf_lo = (const float4 *)ctx->leaf_lo.p;
    t.leaf_hi = (const float4 *)ctx->leaf_hi.p;
    t.fast = (const FastNode *)ctx->fast.p;
    t.root_ref = ctx->root_ref;
    t.abs_max = ctx->abs_max;
    return t;
}

}  // namespace bre
