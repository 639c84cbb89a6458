// bre_query_warp.cuh -- K5 production kernels (included by bre_query.cu).
//
// The beam query is split into two kernels joined by a (ray, leaf) pair queue in HBM:
//
//  K5a k_traverse   warp-cooperative collect-all walk of the photon LBVH for a PACKET of P
//                   consecutive camera rays at once.  Per warp, in shared memory: a LIFO node
//                   stack of (node ref, ray slot) entries and a table of the P precomputed rays.
//                   Every step up to 64 entries are popped -- TWO 64-byte two-child nodes per
//                   lane (8 x LDG.128 in flight) -- all four child boxes are slab-tested, and the
//                   warp compacts surviving internal children back onto the stack and surviving
//                   LEAVES (<= max_leaf contiguous photons) into the global pair queue with
//                   ballot/popc.  Queue space is reserved one 256-pair chunk per warp at a time
//                   (one atomicAdd per chunk); packets are claimed dynamically from a counter.
//                   No photon is touched here.
//                   Load balance: a packet whose rays look into a light source walks 10^5 nodes.
//                   Once the packet counter is exhausted and some warp is idle, a warp whose stack
//                   is deep hands the BOTTOM 128 entries (the largest subtrees) to a global
//                   continuation queue as one unit {packet base, entries}; idle warps pop units,
//                   redo the 8-ray packet setup and continue the walk (and may spill again).
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
// If the pair queue is full a warp shades its leaves in place; if the continuation queue is full
// a warp keeps its stack: too-small queues cost speed, never correctness.
#pragma once

namespace bre {

constexpr int WQ_P = 8;          // rays per packet
constexpr int WQ_NCAP = 512;     // node stack entries per warp
constexpr int WQ_WARPS = 4;      // warps per block
constexpr int WQ_RAYW = 16;      // floats per ray-table row (slab-test data only)
constexpr unsigned WQ_CHUNK = 256;  // pair-queue slots reserved per atomicAdd
constexpr unsigned PAIR_NONE = 0xffffffffu;  // padding pair (ray index) in a partly used chunk
constexpr int SPILL_N = 128;     // stack entries handed over per continuation unit
constexpr int SPILL_HI = 320;    // spill only from stacks at least this deep
constexpr unsigned SPILL_CAP = 32768;   // continuation units per launch

// queueCtl words (zeroed before every launch)
enum { QC_PAIR_TAIL = 0, QC_PACKET = 1, QC_INPLACE = 2, QC_EMITTED = 3, QC_SPILL_TAIL = 4,
       QC_SPILL_HEAD = 5, QC_BUSY = 6, QC_EXITED = 7, QC_WORDS = 8 };

struct __align__(16) SpillUnit {
    unsigned base, count, pad0, pad1;
    int ref[SPILL_N];
    unsigned char slot[SPILL_N];
};

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
// the reference's 1 + 2 gamma(3) widening folded into invF / cF.  inv of a zero direction
// component is clamped to +-2^100 so that plane * inv + c never evaluates inf - inf.
struct RayF {
    float ix, iy, iz;        // 1/d (clamped)
    float fx, fy, fz;        // 1/d * (1 + 2 gamma3)
    float cnx, cny, cnz;     // near-plane constants
    float cfx, cfy, cfz;     // far-plane constants
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
    r.fx = fmul(ix, BRE_SLAB_WIDEN); r.fy = fmul(iy, BRE_SLAB_WIDEN); r.fz = fmul(iz, BRE_SLAB_WIDEN);
    r.cnx = -fmul(fadd(o.x, px), ix); r.cny = -fmul(fadd(o.y, py), iy); r.cnz = -fmul(fadd(o.z, pz), iz);
    r.cfx = -fmul(fsub(o.x, px), r.fx); r.cfy = -fmul(fsub(o.y, py), r.fy); r.cfz = -fmul(fsub(o.z, pz), r.fz);
    r.tMax = o.w;
    return r;
}
__device__ __forceinline__ void rayf_store(float *row, const RayF &r) {
    float4 *v = reinterpret_cast<float4 *>(row);
    v[0] = make_float4(r.ix, r.iy, r.iz, r.tMax);
    v[1] = make_float4(r.fx, r.fy, r.fz, __int_as_float(r.sgn));
    v[2] = make_float4(r.cnx, r.cny, r.cnz, 0.f);
    v[3] = make_float4(r.cfx, r.cfy, r.cfz, 0.f);
}
__device__ __forceinline__ RayF rayf_load(const float *row) {
    const float4 *v = reinterpret_cast<const float4 *>(row);
    const float4 a = v[0], b = v[1], c = v[2], d = v[3];
    RayF r;
    r.ix = a.x; r.iy = a.y; r.iz = a.z; r.tMax = a.w;
    r.fx = b.x; r.fy = b.y; r.fz = b.z; r.sgn = __float_as_int(b.w);
    r.cnx = c.x; r.cny = c.y; r.cnz = c.z;
    r.cfx = d.x; r.cfy = d.y; r.cfz = d.z;
    return r;
}
__device__ __forceinline__ bool slab_hit_f(const RayF &r, float lox, float loy, float loz, float hix,
                                           float hiy, float hiz) {
    const bool sx = r.sgn & 1, sy = r.sgn & 2, sz = r.sgn & 4;
    const float tx0 = __fmaf_rn(sx ? hix : lox, r.ix, r.cnx), tx1 = __fmaf_rn(sx ? lox : hix, r.fx, r.cfx);
    const float ty0 = __fmaf_rn(sy ? hiy : loy, r.iy, r.cny), ty1 = __fmaf_rn(sy ? loy : hiy, r.fy, r.cfy);
    const float tz0 = __fmaf_rn(sz ? hiz : loz, r.iz, r.cnz), tz1 = __fmaf_rn(sz ? loz : hiz, r.fz, r.cfz);
    const float tMin = fmaxf(fmaxf(tx0, ty0), tz0);
    const float tMax = fminf(fminf(tx1, ty1), tz1);
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

// Cold path of K5a: the pair queue is full, handle one (ray, leaf) pair in place.
template <int MODE>
__device__ __noinline__ void process_pair_inplace(const LeafArgs &A, uint32_t q, int ref) {
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
template <int MODE, bool COUNT>
__global__ void __launch_bounds__(WQ_WARPS * 32) k_traverse(
    LeafArgs A, uint32_t n, uint2 *__restrict__ queue, unsigned int queueCap,
    unsigned int *__restrict__ queueCtl, SpillUnit *__restrict__ spill,
    unsigned int *__restrict__ spillReady, unsigned long long *__restrict__ counters) {
    __shared__ WarpSmem smemAll[WQ_WARPS];
    WarpSmem &sm = smemAll[threadIdx.x >> 5];
    const TreeDev &t = A.t;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt = (1u << lane) - 1u;
    const unsigned FULL = 0xffffffffu;
    const unsigned totalWarps = gridDim.x * WQ_WARPS;
    unsigned long long cNodes = 0;
    unsigned nEmitted = 0, nInplace = 0;
    unsigned chunkPos = 0, chunkEnd = 0;   // this warp's reserved queue slots [chunkPos, chunkEnd)
    bool havePackets = true;
    bool counted = true;                   // this warp is included in queueCtl[QC_BUSY]
    bool idleSeen = false;                 // some warp is waiting for work (refreshed while spilling)
    bool spillOff = false;
    unsigned pollCtr = 0;
    if (lane == 0) atomicAdd(queueCtl + QC_BUSY, 1u);

    for (;;) {
        // ================= acquire work: a fresh packet, else a continuation unit
        unsigned base = 0;
        int nsize = 0;
        int unit = -1;
        if (havePackets) {
            if (lane == 0) base = atomicAdd(queueCtl + QC_PACKET, (unsigned)WQ_P);
            base = __shfl_sync(FULL, base, 0);
            if (base >= n) havePackets = false;
        }
        if (!havePackets) {
            int res = 0;   // 1 = got a unit, 2 = everything is done
            for (;;) {
                int cnt2 = counted ? 1 : 0;
                if (lane == 0) {
                    res = 0;
                    const unsigned tl = min(ld_volatile_u32(queueCtl + QC_SPILL_TAIL), SPILL_CAP);
                    const unsigned hd = ld_volatile_u32(queueCtl + QC_SPILL_HEAD);
                    if (hd < tl) {
                        // become busy BEFORE claiming, so "busy == 0" implies nobody holds work
                        if (!cnt2) { atomicAdd(queueCtl + QC_BUSY, 1u); cnt2 = 1; }
                        if (atomicCAS(queueCtl + QC_SPILL_HEAD, hd, hd + 1u) == hd) { res = 1; unit = (int)hd; }
                    } else {
                        if (cnt2) { atomicSub(queueCtl + QC_BUSY, 1u); cnt2 = 0; }
                        if (ld_volatile_u32(queueCtl + QC_BUSY) == 0u &&
                            ld_volatile_u32(queueCtl + QC_SPILL_HEAD) >=
                                min(ld_volatile_u32(queueCtl + QC_SPILL_TAIL), SPILL_CAP))
                            res = 2;
                    }
                }
                res = __shfl_sync(FULL, res, 0);
                unit = __shfl_sync(FULL, unit, 0);
                counted = __shfl_sync(FULL, cnt2, 0) != 0;
                if (res != 0) break;
                if (!counted) __nanosleep(200);
            }
            if (res == 2) break;
            // wait until the unit has been published, then read it through L2
            if (lane == 0) {
                while (ld_volatile_u32(spillReady + unit) == 0u) __nanosleep(40);
                __threadfence();
            }
            __syncwarp();
            const SpillUnit *su = spill + unit;
            base = __ldcg(&su->base);
            nsize = (int)__ldcg(&su->count);
            for (int i = lane; i < nsize; i += 32) {
                sm.nref[i] = __ldcg(&su->ref[i]);
                sm.nslot[i] = __ldcg(&su->slot[i]);
            }
        }
        const unsigned np = min((unsigned)WQ_P, n - base);

        // ================= packet setup: lane s < np prepares ray base+s
        bool live = false;
        if (lane < np) {
            const float4 o = __ldg(A.ro + base + lane);
            RayF r = rayf_setup(o, __ldg(A.rd + base + lane), t.abs_max);
            rayf_store(sm.ray[lane], r);
            live = t.n > 0 && r.tMax > 0.f;
        }
        bool rootLeaf = false;
        if (unit < 0) {       // fresh packet: push the root for every live ray
            const unsigned mlive = __ballot_sync(FULL, live);
            if (t.root_ref >= 0) {
                if (live) { int pos = __popc(mlive & lt); sm.nref[pos] = t.root_ref; sm.nslot[pos] = (unsigned char)lane; }
                nsize = __popc(mlive);
            } else {
                rootLeaf = live;   // the whole tree is one leaf
            }
        }
        __syncwarp();

        for (;;) {
            // ---- node step: up to two 64-byte two-child nodes per lane.  Popping k entries can
            // push at most 2k, so k <= room keeps the stack inside WQ_NCAP-64; with no room left
            // the walk degrades to depth-first (k = 1), which needs at most tree-depth more.
            bool e[4] = {rootLeaf, false, false, false};    // leaf children to emit
            int c[4] = {t.root_ref, 0, 0, 0};
            int slot[2] = {(int)lane, 0};
            bool push[4] = {false, false, false, false};
            const bool haveNodes = nsize > 0;
            if (haveNodes) {
                const int room = WQ_NCAP - 64 - nsize;
                const int k = room >= 1 ? min(min(64, nsize), room) : 1;
                bool act[2];
                int ref[2] = {0, 0};
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    act[u] = (int)lane + 32 * u < k;
                    if (act[u]) {
                        const int idx = nsize - 1 - (int)lane - 32 * u;
                        ref[u] = sm.nref[idx];
                        slot[u] = sm.nslot[idx];
                    }
                }
                nsize -= k;
                float4 na[2], nb[2], nc[2];
                int4 nd[2];
#pragma unroll
                for (int u = 0; u < 2; ++u)
                    if (act[u]) {
                        const FastNode *fn = t.fast + ref[u];
                        na[u] = __ldg(&fn->a); nb[u] = __ldg(&fn->b); nc[u] = __ldg(&fn->c);
                        nd[u] = __ldg(&fn->d);
                    }
#pragma unroll
                for (int u = 0; u < 2; ++u)
                    if (act[u]) {
                        const RayF r = rayf_load(sm.ray[slot[u]]);
                        const bool h0 = slab_hit_f(r, na[u].x, na[u].y, na[u].z, na[u].w, nb[u].x, nb[u].y);
                        const bool h1 = slab_hit_f(r, nb[u].z, nb[u].w, nc[u].x, nc[u].y, nc[u].z, nc[u].w);
                        c[2 * u] = nd[u].x; c[2 * u + 1] = nd[u].y;
                        push[2 * u] = h0 && nd[u].x >= 0; push[2 * u + 1] = h1 && nd[u].y >= 0;
                        e[2 * u] = h0 && nd[u].x < 0; e[2 * u + 1] = h1 && nd[u].y < 0;
                        if (COUNT) ++cNodes;
                    }
                __syncwarp();   // all pops are done before anyone pushes
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const unsigned m = __ballot_sync(FULL, push[j]);
                    if (push[j]) {
                        const int p = nsize + __popc(m & lt);
                        sm.nref[p] = c[j];
                        sm.nslot[p] = (unsigned char)slot[j >> 1];
                    }
                    nsize += __popc(m);
                }
                __syncwarp();
            }
            rootLeaf = false;
            // ---- emit surviving leaves as (ray, leaf) pairs
            unsigned me[4], nl = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) { me[j] = __ballot_sync(FULL, e[j]); nl += __popc(me[j]); }
            if (nl) {
                if (chunkPos + nl > chunkEnd) {
                    // pad what is left of the old chunk, reserve a new one
                    for (unsigned p = chunkPos + lane; p < chunkEnd; p += 32)
                        if (p < queueCap) queue[p] = make_uint2(PAIR_NONE, 0u);
                    unsigned nb2 = 0;
                    if (lane == 0) nb2 = atomicAdd(queueCtl + QC_PAIR_TAIL, WQ_CHUNK);
                    chunkPos = __shfl_sync(FULL, nb2, 0);
                    chunkEnd = chunkPos + WQ_CHUNK;
                }
                unsigned off = chunkPos;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if (e[j]) {
                        const unsigned p = off + __popc(me[j] & lt);
                        const unsigned q = base + (unsigned)slot[j >> 1];
                        if (p < queueCap) queue[p] = make_uint2(q, (unsigned)c[j]);
                        else { process_pair_inplace<MODE>(A, q, c[j]); ++nInplace; }
                        if (COUNT) ++nEmitted;
                    }
                    off += __popc(me[j]);
                }
                chunkPos += nl;
            }
            if (nsize == 0) break;
            // ---- load balance: hand the bottom of a deep stack to idle warps
            if (nsize >= SPILL_HI && !spillOff) {
                if ((pollCtr++ & 7u) == 0u) {
                    unsigned waiting = 0;
                    if (lane == 0) {
                        const unsigned busy = ld_volatile_u32(queueCtl + QC_BUSY);
                        const unsigned gone = ld_volatile_u32(queueCtl + QC_EXITED);
                        const unsigned pk = ld_volatile_u32(queueCtl + QC_PACKET);
                        waiting = (pk >= n && busy + gone < totalWarps) ? 1u : 0u;
                    }
                    idleSeen = __shfl_sync(FULL, waiting, 0) != 0u;
                }
                if (idleSeen) {
                    unsigned u = 0;
                    if (lane == 0) u = atomicAdd(queueCtl + QC_SPILL_TAIL, 1u);
                    u = __shfl_sync(FULL, u, 0);
                    if (u < SPILL_CAP) {
                        SpillUnit *su = spill + u;
                        for (int i = lane; i < SPILL_N; i += 32) {
                            su->ref[i] = sm.nref[i];
                            su->slot[i] = sm.nslot[i];
                        }
                        if (lane == 0) { su->base = base; su->count = SPILL_N; }
                        __threadfence();
                        __syncwarp();
                        if (lane == 0) atomicExch(spillReady + u, 1u);
                        const int rem = nsize - SPILL_N;
                        for (int i0 = 0; i0 < rem; i0 += 32) {
                            const int i = i0 + (int)lane;
                            int rr = 0; unsigned char ss = 0;
                            if (i < rem) { rr = sm.nref[i + SPILL_N]; ss = sm.nslot[i + SPILL_N]; }
                            __syncwarp();
                            if (i < rem) { sm.nref[i] = rr; sm.nslot[i] = ss; }
                            __syncwarp();
                        }
                        nsize = rem;
                    } else {
                        spillOff = true;    // continuation queue exhausted for this launch
                    }
                }
            }
        }
    }
    // pad the unused tail of this warp's last chunk
    for (unsigned p = chunkPos + lane; p < chunkEnd; p += 32)
        if (p < queueCap) queue[p] = make_uint2(PAIR_NONE, 0u);
    if (lane == 0) atomicAdd(queueCtl + QC_EXITED, 1u);
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
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) nInplace += __shfl_xor_sync(FULL, nInplace, o);
    if (lane == 0 && nInplace) atomicAdd(queueCtl + QC_INPLACE, nInplace);
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
__global__ void __launch_bounds__(LP_WARPS * 32) k_leaf_pairs(LeafArgs A,
                                                              const uint2 *__restrict__ queue,
                                                              unsigned int queueCap,
                                                              const unsigned int *__restrict__ queueCtl,
                                                              unsigned long long *__restrict__ counters) {
    __shared__ HitBuf bufAll[LP_WARPS];
    HitBuf &hb = bufAll[threadIdx.x >> 5];
    constexpr unsigned G = 32 / LW;
    constexpr int U = 4;
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
                    o[u] = __ldg(A.ro + qx);
                    d[u] = __ldg(A.rd + qx);
                }
            }
            // ---- disk predicates, hit compaction
#pragma unroll
            for (int u = 0; u < U; ++u) {
                bool hit = false;
                float th = 0.f, d2 = 0.f;
                if (on[u]) {
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
                    shade_hits_warp(A, lane, 32u, hb.s[e], hb.q[e], hb.th[e], hb.d2[e]);
                    nbuf -= 32;
                    __syncwarp();
                }
            }
        }
    }
    if (MODE == 0 && nbuf) {
        uint32_t s = 0, q = 0; float th = 0.f, d2 = 0.f;
        if (lane < nbuf) { s = hb.s[lane]; q = hb.q[lane]; th = hb.th[lane]; d2 = hb.d2[lane]; }
        shade_hits_warp(A, lane, nbuf, s, q, th, d2);
    }
    if (COUNT) {
#pragma unroll
        for (int o2 = 16; o2 > 0; o2 >>= 1) cTests += __shfl_xor_sync(FULL, cTests, o2);
        if (lane == 0 && cTests) atomicAdd(counters + 4, cTests);
    }
}

}  // namespace bre
