// bre_query_warp.cuh -- K5 production kernels (included by bre_query.cu).
//
// The beam query is split into two kernels joined by a (ray, leaf) pair queue in HBM:
//
//  K5a k_traverse   warp-cooperative collect-all walk of the photon LBVH for a PACKET of P
//                   consecutive camera rays at once.  Per warp, in shared memory: a LIFO node
//                   stack of (node ref, ray slot) entries and a table of the P precomputed rays.
//                   Every step up to 32 entries are popped -- one 64-byte two-child node per
//                   lane (4 x LDG.128) -- both child boxes are slab-tested, and the warp compacts
//                   surviving internal children back onto the stack and surviving LEAVES
//                   (<= max_leaf contiguous photons) into the global pair queue with ballot/popc.
//                   Queue space is reserved one 256-pair chunk per warp at a time (one atomicAdd
//                   per chunk); packets are claimed dynamically from a global counter.
//                   No photon is touched here.
//  K5b k_leaf_pairs LW lanes per (ray, leaf) pair, one photon per lane: a coalesced load of the
//                   leaf's float4 {pos,radius} records (contiguous in Morton order), the disk
//                   predicate, and the hits are compacted (ballot/popc) into a per-warp shared-
//                   memory buffer.  Whenever 32 hits are buffered the warp shades them together
//                   -- 2D kernel x transmittance x phase x power, one hit per lane, no divergence --
//                   and adds them to the rays' radiance with atomicAdd.
//
// Why two kernels: a ray that looks at a light source meets 10^4..10^5 photons; inside one
// warp that work is latency-bound and leaves the rest of the GPU idle (measured: 72 % of K5 time
// was such a tail, profiles/r01_k5_warp_v1.md).  Through the queue every leaf becomes an
// independent work item, so the heavy rays spread over the whole chip.
// If the queue is full a warp shades its leaves in place, so a too-small queue costs speed,
// never correctness.
#pragma once

namespace bre {

constexpr int WQ_P = 8;          // rays per packet
constexpr int WQ_NCAP = 384;     // node stack entries per warp
constexpr int WQ_WARPS = 4;      // warps per block
constexpr int WQ_RAYW = 12;      // floats per ray-table row (slab-test data only)
constexpr unsigned WQ_CHUNK = 256;  // pair-queue slots reserved per atomicAdd
constexpr unsigned PAIR_NONE = 0xffffffffu;  // padding pair (ray index) in a partly used chunk

struct __align__(16) WarpSmem {
    int nref[WQ_NCAP];
    float ray[WQ_P][WQ_RAYW];
    unsigned char nslot[WQ_NCAP];
};

// slab-test view of a ray
struct RaySlab {
    float ix, iy, iz, nx, ny, nz, fx, fy, fz, tMax;
    bool sx, sy, sz;
};
__device__ __forceinline__ void slab_store(float *row, const RayS &r) {
    row[0] = r.ix; row[1] = r.iy; row[2] = r.iz;
    row[3] = r.nx; row[4] = r.ny; row[5] = r.nz;
    row[6] = r.fx; row[7] = r.fy; row[8] = r.fz;
    row[9] = r.tMax;
    row[10] = __int_as_float((r.sx ? 1 : 0) | (r.sy ? 2 : 0) | (r.sz ? 4 : 0));
}
__device__ __forceinline__ RayS slab_load(const float *row) {
    RayS r;
    r.ix = row[0]; r.iy = row[1]; r.iz = row[2];
    r.nx = row[3]; r.ny = row[4]; r.nz = row[5];
    r.fx = row[6]; r.fy = row[7]; r.fz = row[8];
    r.tMax = row[9];
    int s = __float_as_int(row[10]);
    r.sx = s & 1; r.sy = s & 2; r.sz = s & 4;
    r.ox = r.oy = r.oz = r.dx = r.dy = r.dz = 0.f;
    return r;
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

// Radiance of one hit (photon s, ray q, ray parameter th, squared distance d2), added to out[q].
// Only the hit DECISION has to be bit-exact; the weight uses the fast reciprocal / exp units
// (relative error ~1e-6, tolerance is 1e-4 mean / 1e-3 max).
__device__ __forceinline__ void shade_hit(const LeafArgs &A, uint32_t s, uint32_t q, float th, float d2) {
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
    float *dst = reinterpret_cast<float *>(A.out + q);
    atomicAdd(dst + 0, wp * t0 * pw.x);
    atomicAdd(dst + 1, wp * t1 * pw.y);
    atomicAdd(dst + 2, wp * t2 * pw.z);
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

// ---------------------------------------------------------------- K5a
// queueCtl[0] = pair slots reserved (tail), [1] = packet counter, [2] = pairs handled in place,
// [3] = pairs emitted (without chunk padding)
template <int MODE, bool COUNT>
__global__ void __launch_bounds__(WQ_WARPS * 32) k_traverse(
    LeafArgs A, uint32_t n, uint2 *__restrict__ queue, unsigned int queueCap,
    unsigned int *__restrict__ queueCtl, unsigned long long *__restrict__ counters) {
    __shared__ WarpSmem smemAll[WQ_WARPS];
    WarpSmem &sm = smemAll[threadIdx.x >> 5];
    const TreeDev &t = A.t;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt = (1u << lane) - 1u;
    const unsigned FULL = 0xffffffffu;
    unsigned long long cNodes = 0;
    unsigned nEmitted = 0, nInplace = 0;
    unsigned chunkPos = 0, chunkEnd = 0;   // this warp's reserved queue slots [chunkPos, chunkEnd)

    for (;;) {
        unsigned base = 0;
        if (lane == 0) base = atomicAdd(queueCtl + 1, (unsigned)WQ_P);
        base = __shfl_sync(FULL, base, 0);
        if (base >= n) break;
        const unsigned np = min((unsigned)WQ_P, n - base);

        // ---- packet setup: lane s < np prepares ray base+s and pushes the root for it
        bool live = false;
        if (lane < np) {
            RayS r = ray_setup(__ldg(A.ro + base + lane), __ldg(A.rd + base + lane), t.abs_max);
            slab_store(sm.ray[lane], r);
            live = t.n > 0 && r.tMax > 0.f;
        }
        int nsize = 0;
        bool emit0 = false, emit1 = false;   // leaf children to emit this step
        int c0 = 0, c1 = 0, slot = (int)lane;
        {
            unsigned mlive = __ballot_sync(FULL, live);
            if (t.root_ref >= 0) {
                if (live) { int pos = __popc(mlive & lt); sm.nref[pos] = t.root_ref; sm.nslot[pos] = (unsigned char)lane; }
                nsize = __popc(mlive);
            } else {            // the whole tree is one leaf
                emit0 = live; c0 = t.root_ref;
            }
        }
        __syncwarp();

        for (;;) {
            // ---- emit surviving leaves as (ray, leaf) pairs
            const unsigned ml0 = __ballot_sync(FULL, emit0), ml1 = __ballot_sync(FULL, emit1);
            const unsigned nl = __popc(ml0) + __popc(ml1);
            if (nl) {
                if (chunkPos + nl > chunkEnd) {
                    // pad what is left of the old chunk, reserve a new one
                    for (unsigned p = chunkPos + lane; p < chunkEnd; p += 32)
                        if (p < queueCap) queue[p] = make_uint2(PAIR_NONE, 0u);
                    unsigned nb = 0;
                    if (lane == 0) nb = atomicAdd(queueCtl + 0, WQ_CHUNK);
                    chunkPos = __shfl_sync(FULL, nb, 0);
                    chunkEnd = chunkPos + WQ_CHUNK;
                }
                const unsigned p0 = chunkPos + __popc(ml0 & lt);
                const unsigned p1 = chunkPos + __popc(ml0) + __popc(ml1 & lt);
                chunkPos += nl;
                if (emit0) {
                    if (p0 < queueCap) queue[p0] = make_uint2(base + slot, (unsigned)c0);
                    else { process_pair_inplace<MODE>(A, base + slot, c0); ++nInplace; }
                }
                if (emit1) {
                    if (p1 < queueCap) queue[p1] = make_uint2(base + slot, (unsigned)c1);
                    else { process_pair_inplace<MODE>(A, base + slot, c1); ++nInplace; }
                }
                if (COUNT) nEmitted += (emit0 ? 1u : 0u) + (emit1 ? 1u : 0u);
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
            bool h0 = false, h1 = false;
            if (act) {
                const FastNode *fn = t.fast + ref;
                float4 a = __ldg(&fn->a), b = __ldg(&fn->b), c = __ldg(&fn->c);
                int4 d = __ldg(&fn->d);
                RayS r = slab_load(sm.ray[slot]);
                h0 = slab_hit(r, a.x, a.y, a.z, a.w, b.x, b.y);
                h1 = slab_hit(r, b.z, b.w, c.x, c.y, c.z, c.w);
                c0 = d.x; c1 = d.y;
                if (COUNT) ++cNodes;
            }
            __syncwarp();   // all pops are done before anyone pushes
            const bool n0 = h0 && c0 >= 0, n1 = h1 && c1 >= 0;
            emit0 = h0 && c0 < 0; emit1 = h1 && c1 < 0;
            const unsigned mn0 = __ballot_sync(FULL, n0), mn1 = __ballot_sync(FULL, n1);
            if (n0) { int p = nsize + __popc(mn0 & lt); sm.nref[p] = c0; sm.nslot[p] = (unsigned char)slot; }
            if (n1) { int p = nsize + __popc(mn0) + __popc(mn1 & lt); sm.nref[p] = c1; sm.nslot[p] = (unsigned char)slot; }
            nsize += __popc(mn0) + __popc(mn1);
            __syncwarp();
        }
    }
    // pad the unused tail of this warp's last chunk
    for (unsigned p = chunkPos + lane; p < chunkEnd; p += 32)
        if (p < queueCap) queue[p] = make_uint2(PAIR_NONE, 0u);
    if (nInplace) atomicAdd(queueCtl + 2, nInplace);
    if (COUNT) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            cNodes += __shfl_xor_sync(FULL, cNodes, o);
            nEmitted += __shfl_xor_sync(FULL, nEmitted, o);
        }
        if (lane == 0) {
            atomicAdd(counters + 3, cNodes);
            atomicAdd(queueCtl + 3, nEmitted);
        }
    }
}

// ---------------------------------------------------------------- K5b
constexpr int LP_WARPS = 8;
struct __align__(16) HitBuf {
    uint32_t s[64], q[64];
    float th[64], d2[64];
};

// LW lanes per pair (LW >= photons per leaf), 32/LW pairs per warp step.
template <int MODE, bool COUNT, int LW>
__global__ void __launch_bounds__(LP_WARPS * 32) k_leaf_pairs(LeafArgs A,
                                                              const uint2 *__restrict__ queue,
                                                              unsigned int queueCap,
                                                              const unsigned int *__restrict__ queueCtl,
                                                              unsigned long long *__restrict__ counters) {
    __shared__ HitBuf bufAll[LP_WARPS];
    HitBuf &hb = bufAll[threadIdx.x >> 5];
    constexpr unsigned G = 32 / LW;
    const unsigned FULL = 0xffffffffu;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt = (1u << lane) - 1u;
    const unsigned sub = lane / LW, j = lane % LW;
    const unsigned total = min(__ldg(queueCtl), queueCap);
    const unsigned warpsTotal = gridDim.x * LP_WARPS;
    const unsigned warpId = blockIdx.x * LP_WARPS + (threadIdx.x >> 5);
    unsigned long long cTests = 0;
    unsigned nbuf = 0;
    // each warp walks the queue in strides of 64 consecutive pairs (same-ray pairs are adjacent)
    constexpr unsigned RUN = 64;
    for (unsigned run = warpId * RUN; run < total; run += warpsTotal * RUN) {
        const unsigned runEnd = min(run + RUN, total);
        for (unsigned i0 = run; i0 < runEnd; i0 += G) {
            const unsigned i = i0 + sub;
            uint2 pr = make_uint2(PAIR_NONE, 0u);
            if (i < runEnd) pr = __ldg(queue + i);
            const uint32_t q = pr.x;
            const uint32_t first = pr.y & 0x07ffffffu;
            const uint32_t cnt = ((pr.y >> 27) & 15u) + 1u;
            bool hit = false;
            float th = 0.f, d2 = 0.f;
            if (q != PAIR_NONE && j < cnt) {
                const float4 p = __ldg(A.t.s_pos_r + first + j);
                const float4 o = __ldg(A.ro + q), d = __ldg(A.rd + q);
                RayS r;
                r.ox = o.x; r.oy = o.y; r.oz = o.z; r.tMax = o.w;
                r.dx = d.x; r.dy = d.y; r.dz = d.z;
                hit = disk_test(r, p, th, d2);
                if (COUNT) ++cTests;
            }
            if (MODE != 0) {
                if (hit) record_hit(A, MODE, first + j, q);
                continue;
            }
            const unsigned mh = __ballot_sync(FULL, hit);
            if (hit) {
                const unsigned pos = nbuf + __popc(mh & lt);
                hb.s[pos] = first + j; hb.q[pos] = q; hb.th[pos] = th; hb.d2[pos] = d2;
            }
            nbuf += __popc(mh);
            __syncwarp();
            if (nbuf >= 32) {   // shade 32 buffered hits, one per lane
                const unsigned e = nbuf - 32 + lane;
                shade_hit(A, hb.s[e], hb.q[e], hb.th[e], hb.d2[e]);
                nbuf -= 32;
                __syncwarp();
            }
        }
    }
    if (MODE == 0 && lane < nbuf) shade_hit(A, hb.s[lane], hb.q[lane], hb.th[lane], hb.d2[lane]);
    if (COUNT) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) cTests += __shfl_xor_sync(FULL, cTests, o);
        if (lane == 0 && cTests) atomicAdd(counters + 4, cTests);
    }
}

}  // namespace bre
