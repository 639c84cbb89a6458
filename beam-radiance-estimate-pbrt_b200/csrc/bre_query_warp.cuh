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
//                   (<= max_leaf contiguous photons) into the global pair queue with ballot/popc
//                   and one warp-aggregated atomicAdd.  Packets are claimed dynamically from a
//                   global counter.  No photon is touched here.
//  K5b k_leaf_pairs one thread per (ray, leaf) pair: loads the leaf's photons (float4 {pos,radius},
//                   contiguous in Morton order) four at a time, applies the disk predicate and
//                   shades the hits (2D kernel x transmittance x phase x power), then adds the
//                   pair's contribution to the ray's radiance with atomicAdd.
//
// Why two kernels: a ray that looks at a light source meets 10^4..10^5 photons; inside one
// warp that work is latency-bound and leaves the rest of the GPU idle (measured: 72 % of K5 time
// was such a tail, profiles/r01_k5_warp_v1.md).  Through the queue every leaf becomes an
// independent work item, so the heavy rays spread over the whole chip.
// If the queue is full a warp shades its leaves in place (same device function), so a too-small
// queue costs speed, never correctness.
#pragma once

namespace bre {

constexpr int WQ_P = 8;          // rays per packet
constexpr int WQ_NCAP = 512;     // node stack entries per warp
constexpr int WQ_WARPS = 4;      // warps per block
constexpr int WQ_RAYW = 20;      // floats per ray-table row

struct __align__(16) WarpSmem {
    int nref[WQ_NCAP];
    float ray[WQ_P][WQ_RAYW];
    unsigned char nslot[WQ_NCAP];
};

__device__ __forceinline__ void ray_store(float *row, const RayS &r) {
    row[0] = r.ox; row[1] = r.oy; row[2] = r.oz; row[3] = r.tMax;
    row[4] = r.dx; row[5] = r.dy; row[6] = r.dz;
    row[7] = r.ix; row[8] = r.iy; row[9] = r.iz;
    row[10] = r.nx; row[11] = r.ny; row[12] = r.nz;
    row[13] = r.fx; row[14] = r.fy; row[15] = r.fz;
    row[16] = __int_as_float((r.sx ? 1 : 0) | (r.sy ? 2 : 0) | (r.sz ? 4 : 0));
}
__device__ __forceinline__ RayS ray_load(const float *row) {
    RayS r;
    r.ox = row[0]; r.oy = row[1]; r.oz = row[2]; r.tMax = row[3];
    r.dx = row[4]; r.dy = row[5]; r.dz = row[6];
    r.ix = row[7]; r.iy = row[8]; r.iz = row[9];
    r.nx = row[10]; r.ny = row[11]; r.nz = row[12];
    r.fx = row[13]; r.fy = row[14]; r.fz = row[15];
    int s = __float_as_int(row[16]);
    r.sx = s & 1; r.sy = s & 2; r.sz = s & 4;
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

// One (ray, leaf) pair: disk tests + shading.  MODE 0 radiance, 1 count hits, 2 write hit indices.
template <int MODE>
__device__ __forceinline__ void process_pair(const LeafArgs &A, uint32_t q, int ref,
                                             unsigned long long &cTests) {
    const float4 o = __ldg(A.ro + q), d = __ldg(A.rd + q);
    RayS r;
    r.ox = o.x; r.oy = o.y; r.oz = o.z; r.tMax = o.w;
    r.dx = d.x; r.dy = d.y; r.dz = d.z;
    const uint32_t first = (uint32_t)ref & 0x07ffffffu;
    const uint32_t cnt = (((uint32_t)ref >> 27) & 15u) + 1u;
    cTests += cnt;
    float L0 = 0.f, L1 = 0.f, L2 = 0.f;
    uint32_t hitS[16];
    uint32_t nh = 0;
    TrState ts;
    ts.tau = nullptr; ts.t0 = ts.t1 = 0.f; ts.steps = A.tauSteps;
    if (MODE == 0 && A.m.kind == 1) {
        ts.tau = A.tau + (size_t)q * (A.tauSteps + 1);
        float2 sp = __ldg(A.span + q);
        ts.t0 = sp.x; ts.t1 = sp.y;
    }
    for (uint32_t j0 = 0; j0 < cnt; j0 += 4) {
        // four independent 16-byte loads in flight per thread
        float4 p[4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
            p[j] = __ldg(A.t.s_pos_r + first + min(j0 + j, cnt - 1u));
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float th, d2;
            if (j0 + j < cnt && disk_test(r, p[j], th, d2)) {
                const uint32_t s = first + j0 + j;
                if (MODE == 0) {
                    float r2 = fmul(p[j].w, p[j].w);
                    float sK = fsub(1.f, fdiv(d2, r2));
                    float kk = fmul(fmul(3.f * BRE_INV_PI, sK), sK);
                    float wgt = fdiv(kk, fmul(A.nPathsNorm, r2));
                    float4 w = __ldg(A.t.s_dir + s);
                    float4 pw = __ldg(A.t.s_pow + s);
                    float cosT = fadd(fadd(fmul(-r.dx, w.x), fmul(-r.dy, w.y)), fmul(-r.dz, w.z));
                    float wp = fmul(wgt, phase_hg(cosT, A.m.g));
                    float tr[3];
                    eval_tr(A.m, ts, th, tr);
                    L0 = fadd(L0, fmul(fmul(wp, tr[0]), pw.x));
                    L1 = fadd(L1, fmul(fmul(wp, tr[1]), pw.y));
                    L2 = fadd(L2, fmul(fmul(wp, tr[2]), pw.z));
                } else if (MODE == 2) {
                    hitS[nh] = s;
                }
                ++nh;
            }
        }
    }
    if (nh == 0) return;
    if (MODE == 0) {
        float *dst = reinterpret_cast<float *>(A.out + q);
        atomicAdd(dst + 0, L0); atomicAdd(dst + 1, L1); atomicAdd(dst + 2, L2);
    } else if (MODE == 1) {
        atomicAdd(A.hitCounts + q, nh);
    } else {
        uint32_t pos = atomicAdd(A.hitCounts + q, nh);
        uint32_t *dst = A.hitIdx + A.hitOffsets[q] + pos;
        for (uint32_t k = 0; k < nh; ++k) dst[k] = __ldg(A.t.order + hitS[k]);
    }
}

// ---------------------------------------------------------------- K5a
// queueCtl[0] = pairs written (tail), queueCtl[1] = packet counter, queueCtl[2] = overflow pairs
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
    unsigned long long cNodes = 0, cTests = 0;

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
            ray_store(sm.ray[lane], r);
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
                unsigned qb = 0;
                if (lane == 0) qb = atomicAdd(queueCtl + 0, nl);
                qb = __shfl_sync(FULL, qb, 0);
                const unsigned p0 = qb + __popc(ml0 & lt), p1 = qb + __popc(ml0) + __popc(ml1 & lt);
                if (emit0) {
                    if (p0 < queueCap) queue[p0] = make_uint2(base + slot, (unsigned)c0);
                    else { process_pair<MODE>(A, base + slot, c0, cTests); atomicAdd(queueCtl + 2, 1u); }
                }
                if (emit1) {
                    if (p1 < queueCap) queue[p1] = make_uint2(base + slot, (unsigned)c1);
                    else { process_pair<MODE>(A, base + slot, c1, cTests); atomicAdd(queueCtl + 2, 1u); }
                }
            }
            if (nsize == 0) break;
            // ---- node step: one 64-byte two-child node per lane
            const int k = (nsize > WQ_NCAP - 128) ? 1 : min(32, nsize);
            const bool act = (int)lane < k;
            int ref = 0;
            if (act) { ref = sm.nref[nsize - 1 - lane]; slot = sm.nslot[nsize - 1 - lane]; }
            nsize -= k;
            bool h0 = false, h1 = false;
            if (act) {
                const FastNode *fn = t.fast + ref;
                float4 a = __ldg(&fn->a), b = __ldg(&fn->b), c = __ldg(&fn->c);
                int4 d = __ldg(&fn->d);
                RayS r = ray_load(sm.ray[slot]);
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
    if (COUNT) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            cNodes += __shfl_xor_sync(FULL, cNodes, o);
            cTests += __shfl_xor_sync(FULL, cTests, o);
        }
        if (lane == 0) {
            atomicAdd(counters + 3, cNodes);
            atomicAdd(counters + 4, cTests);
        }
    }
}

// ---------------------------------------------------------------- K5b
template <int MODE, bool COUNT>
__global__ void __launch_bounds__(256) k_leaf_pairs(LeafArgs A, const uint2 *__restrict__ queue,
                                                    unsigned int queueCap,
                                                    const unsigned int *__restrict__ queueCtl,
                                                    unsigned long long *__restrict__ counters) {
    const unsigned total = min(__ldg(queueCtl), queueCap);
    unsigned long long cTests = 0;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        uint2 pr = __ldg(queue + i);
        process_pair<MODE>(A, pr.x, (int)pr.y, cTests);
    }
    if (COUNT) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) cTests += __shfl_xor_sync(0xffffffffu, cTests, o);
        if ((threadIdx.x & 31) == 0 && cTests) atomicAdd(counters + 4, cTests);
    }
}

}  // namespace bre
