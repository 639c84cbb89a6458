// bre_query_warp.cuh -- K5, warp-cooperative production traversal (included by bre_query.cu).
//
// One warp walks the photon LBVH for a PACKET of P consecutive camera rays at once.  The
// traversal state lives in shared memory, per warp:
//   node stack   (ref, ray slot)   LIFO, up to 32 entries popped per step -- one node per lane
//   leaf stack   (leaf ref, slot)  drained 32 leaves at a time -- one leaf per lane
//   ray table    P precomputed rays (origin, direction, 1/d, padded origins, signs)
//   accumulators [P*3][32 lanes]   per-lane partial radiance, reduced in fixed order at the end
// Every lane fetches one 64-byte two-child node per step (4 x LDG.128), tests both child boxes
// and the warp compacts the surviving children onto the stacks with ballot/popc -- so a ray
// that meets thousands of photons (a light source in view) spreads over all 32 lanes instead
// of serialising one thread, and rays with little work share the warp with their neighbours.
// Work is handed out dynamically: warps claim packets from a global counter.
#pragma once

namespace bre {

constexpr int WQ_P = 8;          // rays per packet
constexpr int WQ_NCAP = 512;     // node stack entries per warp
constexpr int WQ_LCAP = 128;     // leaf stack entries per warp
constexpr int WQ_WARPS = 4;      // warps per block
constexpr int WQ_RAYW = 20;      // floats per ray-table row
constexpr int WQ_ACCW = 33;      // accumulator row stride (32 lanes + 1 pad)

struct __align__(16) WarpSmem {
    int nref[WQ_NCAP];
    int lref[WQ_LCAP];
    float ray[WQ_P][WQ_RAYW];
    float acc[WQ_P * 3][WQ_ACCW];
    unsigned int hitc[WQ_P];
    unsigned char nslot[WQ_NCAP];
    unsigned char lslot[WQ_LCAP];
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

template <int MODE, bool COUNT>
__global__ void __launch_bounds__(WQ_WARPS * 32) k_query_warp(
    TreeDev t, MediumDev m, uint32_t n, const float4 *__restrict__ ro,
    const float4 *__restrict__ rd, float nPathsNorm, float4 *__restrict__ out,
    const float *__restrict__ tau, const float2 *__restrict__ span, int tauSteps,
    uint32_t *__restrict__ hitCounts, const uint64_t *__restrict__ hitOffsets,
    uint32_t *__restrict__ hitIdx, unsigned long long *__restrict__ counters,
    unsigned int *__restrict__ packetCounter) {
    __shared__ WarpSmem smemAll[WQ_WARPS];
    WarpSmem &sm = smemAll[threadIdx.x >> 5];
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt = (1u << lane) - 1u;
    const unsigned FULL = 0xffffffffu;
    unsigned long long cNodes = 0, cTests = 0;

    for (;;) {
        unsigned base = 0;
        if (lane == 0) base = atomicAdd(packetCounter, (unsigned)WQ_P);
        base = __shfl_sync(FULL, base, 0);
        if (base >= n) break;
        const unsigned np = min((unsigned)WQ_P, n - base);

        // ---- packet setup: lane s < np prepares ray base+s
        bool live = false;
        if (lane < np) {
            RayS r = ray_setup(__ldg(ro + base + lane), __ldg(rd + base + lane), t.abs_max);
            ray_store(sm.ray[lane], r);
            live = t.n > 0 && r.tMax > 0.f;
        }
        if (MODE == 0) {
#pragma unroll
            for (int p = 0; p < WQ_P * 3; ++p) sm.acc[p][lane] = 0.f;
        } else if (lane < WQ_P) sm.hitc[lane] = 0u;
        int nsize = 0, lsize = 0;
        {
            unsigned mlive = __ballot_sync(FULL, live);
            int pos = __popc(mlive & lt);
            if (live) {
                if (t.root_ref >= 0) { sm.nref[pos] = t.root_ref; sm.nslot[pos] = (unsigned char)lane; }
                else { sm.lref[pos] = t.root_ref; sm.lslot[pos] = (unsigned char)lane; }
            }
            if (t.root_ref >= 0) nsize = __popc(mlive); else lsize = __popc(mlive);
        }
        __syncwarp();

        // ---- traversal
        while (nsize > 0 || lsize > 0) {
            if (lsize >= 32 || nsize == 0) {
                // leaf step: one leaf (<= 16 contiguous photons) per lane
                const int k = min(32, lsize);
                const bool act = (int)lane < k;
                int ref = 0, slot = 0;
                if (act) { ref = sm.lref[lsize - 1 - lane]; slot = sm.lslot[lsize - 1 - lane]; }
                lsize -= k;
                if (act) {
                    const float *row = sm.ray[slot];
                    RayS r;
                    r.ox = row[0]; r.oy = row[1]; r.oz = row[2]; r.tMax = row[3];
                    r.dx = row[4]; r.dy = row[5]; r.dz = row[6];
                    const uint32_t first = (uint32_t)ref & 0x07ffffffu;
                    const uint32_t cnt = (((uint32_t)ref >> 27) & 15u) + 1u;
                    if (COUNT) cTests += cnt;
                    for (uint32_t j = 0; j < cnt; ++j) {
                        const uint32_t s = first + j;
                        float4 p = __ldg(t.s_pos_r + s);
                        float th, d2;
                        if (disk_test(r, p, th, d2)) {
                            if (MODE == 0) {
                                float r2 = fmul(p.w, p.w);
                                float sK = fsub(1.f, fdiv(d2, r2));
                                float kk = fmul(fmul(3.f * BRE_INV_PI, sK), sK);
                                float wgt = fdiv(kk, fmul(nPathsNorm, r2));
                                float4 w = __ldg(t.s_dir + s);
                                float cosT = fadd(fadd(fmul(-r.dx, w.x), fmul(-r.dy, w.y)), fmul(-r.dz, w.z));
                                float wp = fmul(wgt, phase_hg(cosT, m.g));
                                TrState ts;
                                ts.tau = nullptr; ts.t0 = ts.t1 = 0.f; ts.steps = tauSteps;
                                if (m.kind == 1) {
                                    const unsigned q = base + slot;
                                    ts.tau = tau + (size_t)q * (tauSteps + 1);
                                    float2 sp = __ldg(span + q);
                                    ts.t0 = sp.x; ts.t1 = sp.y;
                                }
                                float tr[3];
                                eval_tr(m, ts, th, tr);
                                float4 pw = __ldg(t.s_pow + s);
                                float *a = &sm.acc[slot * 3][lane];
                                a[0] = fadd(a[0], fmul(fmul(wp, tr[0]), pw.x));
                                a[WQ_ACCW] = fadd(a[WQ_ACCW], fmul(fmul(wp, tr[1]), pw.y));
                                a[2 * WQ_ACCW] = fadd(a[2 * WQ_ACCW], fmul(fmul(wp, tr[2]), pw.z));
                            } else {
                                unsigned pos = atomicAdd(&sm.hitc[slot], 1u);
                                if (MODE == 2) hitIdx[hitOffsets[base + slot] + pos] = __ldg(t.order + s);
                            }
                        }
                    }
                }
                __syncwarp();
            } else {
                // node step: one 64-byte two-child node per lane
                const int k = (nsize > WQ_NCAP - 128) ? 1 : min(32, nsize);
                const bool act = (int)lane < k;
                int ref = 0, slot = 0;
                if (act) { ref = sm.nref[nsize - 1 - lane]; slot = sm.nslot[nsize - 1 - lane]; }
                nsize -= k;
                bool h0 = false, h1 = false;
                int c0 = 0, c1 = 0;
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
                const bool l0 = h0 && c0 < 0, l1 = h1 && c1 < 0;
                const unsigned mn0 = __ballot_sync(FULL, n0), mn1 = __ballot_sync(FULL, n1);
                const unsigned ml0 = __ballot_sync(FULL, l0), ml1 = __ballot_sync(FULL, l1);
                if (n0) { int p = nsize + __popc(mn0 & lt); sm.nref[p] = c0; sm.nslot[p] = (unsigned char)slot; }
                if (n1) { int p = nsize + __popc(mn0) + __popc(mn1 & lt); sm.nref[p] = c1; sm.nslot[p] = (unsigned char)slot; }
                if (l0) { int p = lsize + __popc(ml0 & lt); sm.lref[p] = c0; sm.lslot[p] = (unsigned char)slot; }
                if (l1) { int p = lsize + __popc(ml0) + __popc(ml1 & lt); sm.lref[p] = c1; sm.lslot[p] = (unsigned char)slot; }
                nsize += __popc(mn0) + __popc(mn1);
                lsize += __popc(ml0) + __popc(ml1);
                __syncwarp();
            }
        }

        // ---- packet results
        if (MODE == 0) {
            float v = 0.f;
            if (lane < WQ_P * 3) {
#pragma unroll
                for (int l = 0; l < 32; ++l) v = fadd(v, sm.acc[lane][l]);
            }
            __syncwarp();
            if (lane < WQ_P * 3) sm.acc[lane][32] = v;
            __syncwarp();
            if (lane < np)
                out[base + lane] = make_float4(sm.acc[lane * 3][32], sm.acc[lane * 3 + 1][32],
                                               sm.acc[lane * 3 + 2][32], 0.f);
        } else if (MODE == 1) {
            if (lane < np) hitCounts[base + lane] = sm.hitc[lane];
        }
        __syncwarp();
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

}  // namespace bre
