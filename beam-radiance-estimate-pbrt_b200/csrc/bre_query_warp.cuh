// bre_query_warp.cuh -- K5 production kernels (included by bre_query.cu).
//
// The beam query is split into two kernels joined by a (ray, leaf) pair queue in HBM:
//
//  K5a k_traverse   warp-cooperative collect-all walk of the photon LBVH for a PACKET of P
//                   consecutive camera rays at once.  Per warp, in shared memory: a LIFO node
//                   stack of (node ref, ray slot) entries and a table of the P precomputed rays.
//                   Every step up to 32 entries are popped -- one 64-byte two-child node per
//                   lane (4 x LDG.128) -- both child boxes are slab-tested (one FMA per plane), and
//                   the warp compacts surviving internal children back onto the stack and surviving
//                   LEAVES (<= max_leaf contiguous photons) into the global pair queue with
//                   ballot/popc.  Queue space is reserved one 256-pair chunk per warp at a time
//                   (one atomicAdd per chunk); packets are claimed dynamically from a counter,
//                   the packets that were heaviest in the previous launch first.
//                   No photon is touched here.
//  K5b k_leaf_pairs LW lanes per (ray, leaf) pair, one photon per lane: 64 pairs are fetched per
//                   warp with two coalesced loads, then four pair-groups at a time the photon
//                   {pos,radius} records (contiguous in Morton order) and rays are loaded BEFORE
//                   the first disk predicate runs, so ~12 loads per lane are in flight.  Hits are
//                   compacted (ballot/popc) into a per-warp shared-memory buffer; whenever 32 are
//                   buffered the warp shades them together -- 2D kernel x transmittance x phase x
//                   power, one hit per lane -- sums runs of equal ray index with a segmented
//                   shuffle reduction and adds one atomicAdd triple per run.
//
// Why two kernels: through the queue every leaf becomes an independent work item, so the heavy
// rays spread over the whole chip (profiles/r01_k5_warp_v1.md: 72 % of a fused kernel was tail).
// If the pair queue is full a warp shades its leaves in place: a too-small queue costs speed,
// never correctness.  (A continuation queue that split deep stacks between warps was tried and
// measured slower: DFS stacks stay shallow, see profiles/r01_k5_v4_experiments.md.)
#pragma once

namespace bre {

constexpr int WQ_P = 8;          // rays per packet
constexpr int WQ_NCAP = 512;     // node stack entries per warp
constexpr int WQ_WARPS = 4;      // warps per block
constexpr int WQ_RAYW = 12;      // floats per ray-table row (slab-test data only)
constexpr unsigned WQ_CHUNK = 256;  // pair-queue slots reserved per atomicAdd
constexpr unsigned PAIR_NONE = 0xffffffffu;  // padding pair (ray index) in a partly used chunk
constexpr unsigned HEAVY_STEPS = 256;   // a packet this long is claimed first by the next launch
constexpr unsigned HEAVY_CAP = 65536;   // heavy packets remembered per launch

// queueCtl words (zeroed before every launch)
enum { QC_PAIR_TAIL = 0, QC_PACKET = 1, QC_INPLACE = 2, QC_EMITTED = 3, QC_WORDS = 8 };

struct __align__(16) WarpSmem {
    int nref[WQ_NCAP];
    float ray[WQ_P][WQ_RAYW];
    unsigned char nslot[WQ_NCAP];
};

// Slab-test view of a ray for the PRODUCTION tree.  The walk only has to be conservative (a box
// that holds a true hit must never be culled; V/T counters and hit decisions come from elsewhere),
// so the test is restated with one FMA per plane:  t = plane * inv + c,  c = -(o +- pad) * inv.
// The rounding error of that form is <= 2^-22 |o| |inv|, 16x below what the origin padding of
// 2^-18 (|o|_inf + max|coord|) already grants (DESIGN.md "conservative cull"); the far planes carry
// the reference's 1 + 2 gamma(3) widening.  inv of a zero direction
// component is clamped to +-2^100 so that plane * inv + c never evaluates inf - inf.
struct RayF {
    float ix, iy, iz;        // 1/d (clamped)
    float cnx, cny, cnz;     // near-plane constants  -(o + pad) / d
    float cfx, cfy, cfz;     // far-plane constants   -(o - pad) / d
    float tMax;
    int sgn;                 // bit k: direction component k negative
};
__device__ __forceinline__ RayF rayf_setup(float4 o, float4 d, float absMax) {
    RayF r;
    const float BIG = 0x1p100f;
    float ix = fdiv(1.f, d.x), iy = fdiv(1.f, d.y), iz = fdiv(1.f, d.z);
    ix = fminf(fmaxf(ix, -BIG), BIG); iy = fminf(fmaxf(iy, -BIG), BIG); iz = fminf(fmaxf(iz, -BIG), BIG);
    if (d.x == 0.f) ix = BIG;  if (d.y == 0.f) iy = BIG;  if (d.z == 0.f) iz = BIG;   // also -0
    r.sgn = (ix < 0.f ? 1 : 0) | (iy < 0.f ? 2 : 0) | (iz < 0.f ? 4 : 0);
    const float M = fadd(fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z)), absMax);
    const float pad = fmul(M, 0x1p-18f);
    const float px = (r.sgn & 1) ? -pad : pad, py = (r.sgn & 2) ? -pad : pad, pz = (r.sgn & 4) ? -pad : pad;
    r.ix = ix; r.iy = iy; r.iz = iz;
    r.cnx = -fmul(fadd(o.x, px), ix); r.cny = -fmul(fadd(o.y, py), iy); r.cnz = -fmul(fadd(o.z, pz), iz);
    r.cfx = -fmul(fsub(o.x, px), ix); r.cfy = -fmul(fsub(o.y, py), iy); r.cfz = -fmul(fsub(o.z, pz), iz);
    r.tMax = o.w;
    return r;
}
__device__ __forceinline__ void rayf_store(float *row, const RayF &r) {
    float4 *v = reinterpret_cast<float4 *>(row);
    v[0] = make_float4(r.ix, r.iy, r.iz, r.tMax);
    v[1] = make_float4(r.cnx, r.cny, r.cnz, __int_as_float(r.sgn));
    v[2] = make_float4(r.cfx, r.cfy, r.cfz, 0.f);
}
__device__ __forceinline__ RayF rayf_load(const float *row) {
    const float4 *v = reinterpret_cast<const float4 *>(row);
    const float4 a = v[0], b = v[1], c = v[2];
    RayF r;
    r.ix = a.x; r.iy = a.y; r.iz = a.z; r.tMax = a.w;
    r.cnx = b.x; r.cny = b.y; r.cnz = b.z; r.sgn = __float_as_int(b.w);
    r.cfx = c.x; r.cfy = c.y; r.cfz = c.z;
    return r;
}
__device__ __forceinline__ bool slab_hit_f(const RayF &r, float lox, float loy, float loz, float hix,
                                           float hiy, float hiz) {
    const bool sx = r.sgn & 1, sy = r.sgn & 2, sz = r.sgn & 4;
    const float tx0 = __fmaf_rn(sx ? hix : lox, r.ix, r.cnx), tx1 = __fmaf_rn(sx ? lox : hix, r.ix, r.cfx);
    const float ty0 = __fmaf_rn(sy ? hiy : loy, r.iy, r.cny), ty1 = __fmaf_rn(sy ? loy : hiy, r.iy, r.cfy);
    const float tz0 = __fmaf_rn(sz ? hiz : loz, r.iz, r.cnz), tz1 = __fmaf_rn(sz ? loz : hiz, r.iz, r.cfz);
    const float tMin = fmaxf(fmaxf(tx0, ty0), tz0);
    // the reference widens every far distance by 1 + 2 gamma(3); min commutes with that scaling
    const float tMax = fmul(fminf(fminf(tx1, ty1), tz1), BRE_SLAB_WIDEN);
    // same acceptance as geometry.h:1410-1438: intervals overlap, tMin < ray.tMax, tMax > 0
    return tMin <= tMax && tMin < r.tMax && tMax > 0.f;
}

// Everything K5b needs besides the pair itself.
struct LeafArgs {
    TreeDev t;
    MediumDev m;
    const float4 *ro, *rd;
    float nPathsNorm;
    float4 *out;                 // MODE 0: radiance, accumulated with atomicAdd (zeroed by the host)
    const float *tau;            // grid media: optical-depth tables
    const float2 *span;
    int tauSteps;
    uint32_t *hitCounts;         // MODE 1: per-ray hit counts; MODE 2: per-ray write cursors
    const uint64_t *hitOffsets;  // MODE 2
    uint32_t *hitIdx;            // MODE 2
};

// Radiance of one hit (photon s, ray q, ray parameter th, squared distance d2).
// Only the hit DECISION has to be bit-exact; the weight uses the fast reciprocal / exp units
// (relative error ~1e-6, tolerance is 1e-4 mean / 1e-3 max).
__device__ __forceinline__ float3 hit_radiance(const LeafArgs &A, uint32_t s, uint32_t q, float th, float d2) {
    const float radius = __ldg(&A.t.s_pos_r[s].w);
    const float4 w = __ldg(A.t.s_dir + s);
    const float4 pw = __ldg(A.t.s_pow + s);
    const float4 d = __ldg(A.rd + q);
    const float r2 = radius * radius;
    const float invr2 = __fdividef(1.f, r2);
    const float sK = 1.f - d2 * invr2;                       // kernel(): photonmap.cpp:170-174
    const float wgt = (3.f * BRE_INV_PI) * sK * sK * invr2 * __fdividef(1.f, A.nPathsNorm);
    const float cosT = -(d.x * w.x + d.y * w.y + d.z * w.z); // Dot(wo = -d, wi)
    const float g = A.m.g;
    const float denom = 1.f + g * g + 2.f * g * cosT;        // PhaseHG medium.h:66-69
    const float rs = rsqrtf(denom);
    const float wp = wgt * BRE_INV4PI * (1.f - g * g) * rs * rs * rs;
    float od = th;                                           // homogeneous: optical depth = sigma_t * t
    if (A.m.kind == 1) {
        const float2 sp = __ldg(A.span + q);
        const float *tau = A.tau + (size_t)q * (A.tauSteps + 1);
        od = 0.f;
        if (sp.y > sp.x) {
            float sc = (th - sp.x) * __fdividef((float)A.tauSteps, sp.y - sp.x);
            if (sc > 0.f) {
                if (sc >= (float)A.tauSteps) od = __ldg(tau + A.tauSteps);
                else {
                    int k = (int)sc;
                    float f = sc - (float)k;
                    float a = __ldg(tau + k), b = __ldg(tau + k + 1);
                    od = a + f * (b - a);
                }
            }
        }
    }
    float t0, t1, t2;
    t0 = __expf(-A.m.sigma_t[0] * od);
    if (A.m.gray) { t1 = t0; t2 = t0; }
    else { t1 = __expf(-A.m.sigma_t[1] * od); t2 = __expf(-A.m.sigma_t[2] * od); }
    return make_float3(wp * t0 * pw.x, wp * t1 * pw.y, wp * t2 * pw.z);
}
__device__ __forceinline__ void shade_hit(const LeafArgs &A, uint32_t s, uint32_t q, float th, float d2) {
    const float3 v = hit_radiance(A, s, q, th, d2);
    float *dst = reinterpret_cast<float *>(A.out + q);
    atomicAdd(dst + 0, v.x);
    atomicAdd(dst + 1, v.y);
    atomicAdd(dst + 2, v.z);
}
// Whole-warp version: lanes < cnt hold one hit each.  Maximal runs of consecutive lanes with the
// same ray index are summed with a segmented shuffle reduction; the first lane of each run issues
// the three atomicAdds (pairs of one ray are mostly adjacent in the queue: ~5x fewer atomics).
__device__ __forceinline__ void shade_hits_warp(const LeafArgs &A, unsigned lane, unsigned cnt,
                                                uint32_t s, uint32_t q, float th, float d2) {
    const unsigned FULL = 0xffffffffu;
    const bool on = lane < cnt;
    float3 v = make_float3(0.f, 0.f, 0.f);
    if (on) v = hit_radiance(A, s, q, th, d2);
    const uint32_t key = on ? q : (PAIR_NONE - lane);        // inactive lanes: unique keys
    const uint32_t prev = __shfl_up_sync(FULL, key, 1);
    const bool head = lane == 0 || prev != key;
    const unsigned heads = __ballot_sync(FULL, head);
    const unsigned above = lane == 31 ? 0u : (heads & (FULL << (lane + 1)));
    const unsigned runEnd = above ? (unsigned)(__ffs(above) - 1) : 32u;   // first lane past my run
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const float x = __shfl_down_sync(FULL, v.x, off), y = __shfl_down_sync(FULL, v.y, off),
                    z = __shfl_down_sync(FULL, v.z, off);
        if (lane + off < runEnd) { v.x += x; v.y += y; v.z += z; }
    }
    if (on && head) {
        float *dst = reinterpret_cast<float *>(A.out + q);
        atomicAdd(dst + 0, v.x);
        atomicAdd(dst + 1, v.y);
        atomicAdd(dst + 2, v.z);
    }
}

__device__ __forceinline__ void record_hit(const LeafArgs &A, int mode, uint32_t s, uint32_t q) {
    uint32_t pos = atomicAdd(A.hitCounts + q, 1u);
    if (mode == 2) A.hitIdx[A.hitOffsets[q] + pos] = __ldg(A.t.order + s);
}

// Cold path of K5a: the pair queue is full, handle one (ray, leaf) pair in place.  Takes the
// arguments through a pointer to their copy in global memory: passing the kernel-parameter struct
// itself would force a 296-byte local-memory copy of it into every thread of the hot kernel.
template <int MODE>
__device__ __noinline__ void process_pair_inplace(const LeafArgs *Ag, uint32_t q, int ref) {
    const LeafArgs &A = *Ag;
    const float4 o = __ldg(A.ro + q), d = __ldg(A.rd + q);
    RayS r;
    r.ox = o.x; r.oy = o.y; r.oz = o.z; r.tMax = o.w;
    r.dx = d.x; r.dy = d.y; r.dz = d.z;
    const uint32_t first = (uint32_t)ref & 0x07ffffffu;
    const uint32_t cnt = (((uint32_t)ref >> 27) & 15u) + 1u;
    for (uint32_t j = 0; j < cnt; ++j) {
        float th, d2;
        if (disk_test(r, __ldg(A.t.s_pos_r + first + j), th, d2)) {
            if (MODE == 0) shade_hit(A, first + j, q, th, d2);
            else record_hit(A, MODE, first + j, q);
        }
    }
}

__device__ __forceinline__ unsigned ld_volatile_u32(const unsigned *p) {
    return *reinterpret_cast<const volatile unsigned *>(p);
}

// ---------------------------------------------------------------- K5a
// queueCtl: [QC_PAIR_TAIL] pair slots reserved, [QC_PACKET] claim counter, [QC_INPLACE] pairs handled
// in place, [QC_EMITTED] pairs emitted (COUNT).
// Heavy-first order: a packet that needed >= HEAVY_STEPS steps is appended to heavyOut; the next
// launch over the same ray count claims those packets FIRST (heavyIn, nHeavyIn), so the 10^5-node
// packets that look into a light source no longer finish last (measured: one such packet kept a
// single warp busy for 2.8 ms of a 4.6 ms kernel).  Progressive rendering repeats the same camera
// every iteration; with unrelated rays the list is merely a harmless permutation.
template <int MODE, bool COUNT>
__global__ void __launch_bounds__(WQ_WARPS * 32) k_traverse(
    LeafArgs A, const LeafArgs *__restrict__ Ag, uint32_t n, uint2 *__restrict__ queue,
    unsigned int queueCap, unsigned int *__restrict__ queueCtl,
    const unsigned int *__restrict__ heavyIn, const unsigned int *__restrict__ nHeavyInPtr,
    const unsigned int *__restrict__ heavyFlagIn, unsigned int *__restrict__ heavyOut,
    unsigned int *__restrict__ heavyFlagOut, unsigned int *__restrict__ nHeavyOut,
    unsigned int heavyCap,
    unsigned long long *__restrict__ counters) {
    __shared__ WarpSmem smemAll[WQ_WARPS];
    WarpSmem &sm = smemAll[threadIdx.x >> 5];
    const TreeDev &t = A.t;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt = (1u << lane) - 1u;
    const unsigned FULL = 0xffffffffu;
    const unsigned nPackets = (n + WQ_P - 1) / WQ_P;
    const unsigned nHeavyIn = min(__ldg(nHeavyInPtr), HEAVY_CAP);
    unsigned long long cNodes = 0;
    unsigned nEmitted = 0, nInplace = 0;
    unsigned chunkPos = 0, chunkEnd = 0;   // this warp's reserved queue slots [chunkPos, chunkEnd)

    for (;;) {
        // ---- claim a packet: last launch's heavy packets first, then the rest in order
        unsigned pk = 0;
        if (lane == 0) {
            for (;;) {
                const unsigned c = atomicAdd(queueCtl + QC_PACKET, 1u);
                if (c < nHeavyIn) { pk = __ldg(heavyIn + c); break; }
                pk = c - nHeavyIn;
                if (pk >= nPackets) { pk = 0xffffffffu; break; }
                if (nHeavyIn == 0 || !((__ldg(heavyFlagIn + (pk >> 5)) >> (pk & 31u)) & 1u)) break;
            }
        }
        pk = __shfl_sync(FULL, pk, 0);
        if (pk == 0xffffffffu) break;
        const unsigned base = pk * WQ_P;
        const unsigned np = min((unsigned)WQ_P, n - base);
        unsigned steps = 0;

        // ---- packet setup: lane s < np prepares ray base+s and pushes the root for it
        bool live = false;
        if (lane < np) {
            const float4 o = __ldg(A.ro + base + lane);
            RayF r = rayf_setup(o, __ldg(A.rd + base + lane), t.abs_max);
            rayf_store(sm.ray[lane], r);
            live = t.n > 0 && r.tMax > 0.f;
        }
        int nsize = 0;
        bool e0 = false, e1 = false;   // leaf children to emit this step
        int c0 = 0, c1 = 0, slot = (int)lane;
        {
            const unsigned mlive = __ballot_sync(FULL, live);
            if (t.root_ref >= 0) {
                if (live) { int pos = __popc(mlive & lt); sm.nref[pos] = t.root_ref; sm.nslot[pos] = (unsigned char)lane; }
                nsize = __popc(mlive);
            } else {            // the whole tree is one leaf
                e0 = live; c0 = t.root_ref;
            }
        }
        __syncwarp();

        for (;;) {
            // ---- emit surviving leaves as (ray, leaf) pairs
            const unsigned ml0 = __ballot_sync(FULL, e0), ml1 = __ballot_sync(FULL, e1);
            const unsigned nl = __popc(ml0) + __popc(ml1);
            if (nl) {
                if (chunkPos + nl > chunkEnd) {
                    // pad what is left of the old chunk, reserve a new one
                    for (unsigned p = chunkPos + lane; p < chunkEnd; p += 32)
                        if (p < queueCap) queue[p] = make_uint2(PAIR_NONE, 0u);
                    unsigned nb = 0;
                    if (lane == 0) nb = atomicAdd(queueCtl + QC_PAIR_TAIL, WQ_CHUNK);
                    chunkPos = __shfl_sync(FULL, nb, 0);
                    chunkEnd = chunkPos + WQ_CHUNK;
                }
                const unsigned p0 = chunkPos + __popc(ml0 & lt);
                const unsigned p1 = chunkPos + __popc(ml0) + __popc(ml1 & lt);
                chunkPos += nl;
                if (e0) {
                    if (p0 < queueCap) queue[p0] = make_uint2(base + slot, (unsigned)c0);
                    else { process_pair_inplace<MODE>(Ag, base + slot, c0); ++nInplace; }
                }
                if (e1) {
                    if (p1 < queueCap) queue[p1] = make_uint2(base + slot, (unsigned)c1);
                    else { process_pair_inplace<MODE>(Ag, base + slot, c1); ++nInplace; }
                }
                if (COUNT) nEmitted += (e0 ? 1u : 0u) + (e1 ? 1u : 0u);
            }
            if (nsize == 0) break;
            // ---- node step: one 64-byte two-child node per lane.  Popping k entries can push
            // at most 2k, so k <= room keeps the stack inside WQ_NCAP-64; with no room left the
            // walk degrades to depth-first (k = 1), which needs at most tree-depth (< 64) more.
            const int room = WQ_NCAP - 64 - nsize;
            const int k = room >= 1 ? min(min(32, nsize), room) : 1;
            const bool act = (int)lane < k;
            int ref = 0;
            if (act) { ref = sm.nref[nsize - 1 - lane]; slot = sm.nslot[nsize - 1 - lane]; }
            nsize -= k;
            ++steps;
            bool h0 = false, h1 = false;
            if (act) {
                const FastNode *fn = t.fast + ref;
                const float4 a = __ldg(&fn->a), b = __ldg(&fn->b), c = __ldg(&fn->c);
                const int4 d = __ldg(&fn->d);
                const RayF r = rayf_load(sm.ray[slot]);
                h0 = slab_hit_f(r, a.x, a.y, a.z, a.w, b.x, b.y);
                h1 = slab_hit_f(r, b.z, b.w, c.x, c.y, c.z, c.w);
                c0 = d.x; c1 = d.y;
                if (COUNT) ++cNodes;
            }
            __syncwarp();   // all pops are done before anyone pushes
            const bool n0 = h0 && c0 >= 0, n1 = h1 && c1 >= 0;
            e0 = h0 && c0 < 0; e1 = h1 && c1 < 0;
            const unsigned mn0 = __ballot_sync(FULL, n0), mn1 = __ballot_sync(FULL, n1);
            if (n0) { int p = nsize + __popc(mn0 & lt); sm.nref[p] = c0; sm.nslot[p] = (unsigned char)slot; }
            if (n1) { int p = nsize + __popc(mn0) + __popc(mn1 & lt); sm.nref[p] = c1; sm.nslot[p] = (unsigned char)slot; }
            nsize += __popc(mn0) + __popc(mn1);
            __syncwarp();
        }
        // ---- remember heavy packets for the next launch
        if (steps >= HEAVY_STEPS && lane == 0) {
            const unsigned pos = heavyCap ? atomicAdd(nHeavyOut, 1u) : 0u;
            if (pos < heavyCap) {
                heavyOut[pos] = pk;
                atomicOr(heavyFlagOut + (pk >> 5), 1u << (pk & 31u));
            }
        }
    }
    // pad the unused tail of this warp's last chunk
    for (unsigned p = chunkPos + lane; p < chunkEnd; p += 32)
        if (p < queueCap) queue[p] = make_uint2(PAIR_NONE, 0u);
    if (nInplace) atomicAdd(queueCtl + QC_INPLACE, nInplace);
    if (COUNT) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            cNodes += __shfl_xor_sync(FULL, cNodes, o);
            nEmitted += __shfl_xor_sync(FULL, nEmitted, o);
        }
        if (lane == 0) {
            atomicAdd(counters + 3, cNodes);
            atomicAdd(queueCtl + QC_EMITTED, nEmitted);
        }
    }
}

// ---------------------------------------------------------------- K5b
constexpr int LP_WARPS = 8;
constexpr unsigned LP_RUN = 64;     // consecutive pairs a warp fetches at once
struct __align__(16) HitBuf {
    uint32_t s[64], q[64];
    float th[64], d2[64];
};

// LW lanes per pair (LW >= photons per leaf), G = 32/LW pairs per group, U groups in flight.
template <int MODE, bool COUNT, int LW>
__global__ void __launch_bounds__(LP_WARPS * 32, 4) k_leaf_pairs(LeafArgs A,
                                                              const uint2 *__restrict__ queue,
                                                              unsigned int queueCap,
                                                              const unsigned int *__restrict__ queueCtl,
                                                              unsigned long long *__restrict__ counters) {
    __shared__ HitBuf bufAll[LP_WARPS];
    HitBuf &hb = bufAll[threadIdx.x >> 5];
    constexpr unsigned G = 32 / LW;
    constexpr int U = 4;             // pair groups whose photon loads are in flight together
    constexpr bool SEG = true;       // segmented reduction before the atomics
    constexpr bool RAYPF = false;    // rays are L1 hits (consecutive pairs share them): load at use
    const unsigned FULL = 0xffffffffu;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt = (1u << lane) - 1u;
    const unsigned sub = lane / LW, j = lane % LW;
    const unsigned total = min(__ldg(queueCtl + QC_PAIR_TAIL), queueCap);
    const unsigned warpsTotal = gridDim.x * LP_WARPS;
    const unsigned warpId = blockIdx.x * LP_WARPS + (threadIdx.x >> 5);
    unsigned long long cTests = 0;
    unsigned nbuf = 0;
    for (unsigned run = warpId * LP_RUN; run < total; run += warpsTotal * LP_RUN) {
        // 64 pairs, two coalesced 256-byte loads
        uint2 pr0 = make_uint2(PAIR_NONE, 0u), pr1 = make_uint2(PAIR_NONE, 0u);
        if (run + lane < total) pr0 = __ldg(queue + run + lane);
        if (run + 32 + lane < total) pr1 = __ldg(queue + run + 32 + lane);
#pragma unroll 1
        for (unsigned g0 = 0; g0 < LP_RUN / G; g0 += U) {
            // ---- issue every load of U pair-groups before the first use
            uint32_t q[U], first[U];
            bool on[U];
            float4 p[U], o[U], d[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const unsigned idx = (g0 + u) * G + sub;           // pair index inside the run
                const bool lower = (g0 + u) * G < 32u;             // warp-uniform: which register
                const unsigned qx = __shfl_sync(FULL, lower ? pr0.x : pr1.x, idx & 31u);
                const unsigned qy = __shfl_sync(FULL, lower ? pr0.y : pr1.y, idx & 31u);
                q[u] = qx;
                first[u] = qy & 0x07ffffffu;
                const uint32_t cnt = ((qy >> 27) & 15u) + 1u;
                on[u] = qx != PAIR_NONE && j < cnt;
                if (on[u]) {
                    p[u] = __ldg(A.t.s_pos_r + first[u] + j);
                    if (RAYPF) { o[u] = __ldg(A.ro + qx); d[u] = __ldg(A.rd + qx); }
                }
            }
            // ---- disk predicates, hit compaction
#pragma unroll
            for (int u = 0; u < U; ++u) {
                bool hit = false;
                float th = 0.f, d2 = 0.f;
                if (on[u]) {
                    if (!RAYPF) { o[u] = __ldg(A.ro + q[u]); d[u] = __ldg(A.rd + q[u]); }
                    RayS r;
                    r.ox = o[u].x; r.oy = o[u].y; r.oz = o[u].z; r.tMax = o[u].w;
                    r.dx = d[u].x; r.dy = d[u].y; r.dz = d[u].z;
                    hit = disk_test(r, p[u], th, d2);
                    if (COUNT) ++cTests;
                }
                if (MODE != 0) {
                    if (hit) record_hit(A, MODE, first[u] + j, q[u]);
                    continue;
                }
                const unsigned mh = __ballot_sync(FULL, hit);
                if (mh == 0u) continue;
                if (hit) {
                    const unsigned pos = nbuf + __popc(mh & lt);
                    hb.s[pos] = first[u] + j; hb.q[pos] = q[u]; hb.th[pos] = th; hb.d2[pos] = d2;
                }
                nbuf += __popc(mh);
                __syncwarp();
                if (nbuf >= 32) {   // shade the 32 newest buffered hits, one per lane
                    const unsigned e = nbuf - 32 + lane;
                    if (SEG) shade_hits_warp(A, lane, 32u, hb.s[e], hb.q[e], hb.th[e], hb.d2[e]);
                    else shade_hit(A, hb.s[e], hb.q[e], hb.th[e], hb.d2[e]);
                    nbuf -= 32;
                    __syncwarp();
                }
            }
        }
    }
    if (MODE == 0 && nbuf) {
        uint32_t s = 0, q = 0; float th = 0.f, d2 = 0.f;
        if (lane < nbuf) { s = hb.s[lane]; q = hb.q[lane]; th = hb.th[lane]; d2 = hb.d2[lane]; }
        if (SEG) shade_hits_warp(A, lane, nbuf, s, q, th, d2);
        else if (lane < nbuf) shade_hit(A, s, q, th, d2);
    }
    if (COUNT) {
#pragma unroll
        for (int o2 = 16; o2 > 0; o2 >>= 1) cTests += __shfl_xor_sync(FULL, cTests, o2);
        if (lane == 0 && cTests) atomicAdd(counters + 4, cTests);
    }
}

}  // namespace bre
