// bre_query_tile.cuh -- K5 for camera images: one BVH walk per 16x16 IMAGE TILE (included by bre_query.cu).
//
// The reference's camera pass works tile by tile (16x16 pixels, integrators/photonbeam.cpp:345-347,
// 444-477).  When the caller declares that ray q is pixel (q % W, q / W) of a W x H image
// (bre_set_ray_grid, or the K0 camera path) the rays of a tile leave (nearly) one point through a
// (nearly) regular grid of directions, and the query is organised around that:
//
//  K5a' k_tile_setup      one WARP per tile fits an orthonormal frame (u, v, n) to the tile (u = central
//                        direction, v = along image rows), maps every ray to gnomonic coordinates
//                        (s, t) = (d.v, d.n)/(d.u), fits the affine model (s, t) = (s0, t0) + i (As, At)
//                        + j (Bs, Bt) through three corner rays and measures how far the real rays
//                        deviate from it (dev).  The tile's rays lie inside the pyramid s in [sLo, sHi],
//                        t in [tLo, tHi], 0 < depth <= tFar: six "max corner" plane tests per box.  The
//                        TOP CUT of the production LBVH (<= 512 boxes, staged into shared memory with
//                        one TMA bulk copy per block) is tested first.
//       k_tile_walk      then sweeps the tree below the cut breadth-first through ONE global queue of
//                        (tile, node) items and emits (tile, leaf) pairs into the pair queue.
//  K5b' k_leaf_tile      one photon per lane: the photon is projected into the tile's pixel grid;
//                        a true hit of ray (i, j) implies |(s_ij, t_ij) - (sp, tp)| <= rho with
//                        rho = (r / z) k_phi + slack (derivation below), so only the pixels of the
//                        bounding rectangle of that disk -- and of those only the ones inside the
//                        disk under the affine model -- are candidates.  Candidates are expanded
//                        across the warp (prefix sum + binary search) and each one runs the EXACT
//                        disk predicate with the ray's own origin and direction, so hit sets are
//                        bit-identical to every other path; hits are buffered and shaded 32 at a time.
//
// Bound used by K5b'.  Apex at the common origin, A = d/(d.u) the ray's point on the plane depth 1,
// B = P'/z the photon's (z = P'.u > 0).  dist(P', ray line) = z dist(B, line through A).  With e the
// perpendicular from the line to B (|e| = delta, e.A = 0):  B - A = e - (e.u) A, and e.u = -e.a for
// A = u + a, so |B - A|^2 = delta^2 + (e.u)^2 |A|^2 <= delta^2 (1 + tan^2 phi sec^2 phi), phi the angle
// between the ray and u.  A hit has delta < r / z, hence |B - A| <= (r / z) k_phi with
// k_phi = sqrt(1 + g2 (1 + g2)), g2 = max_i (s_i^2 + t_i^2).  B - A lies in the plane spanned by
// (v, n), so |B - A|^2 = (sp - s_i)^2 + (tp - t_i)^2.  Slack: r is widened by 2 x the spread of the
// origins + 2^-18 x the coordinate scale (rounding of P', the dot products and the rays' own (s, t):
// > 5x their worst case), rho by 2e-3 pixel pitches and 1e-4 relative.  Where a ray is NOT at its pixel's model
// position (sub-pixel camera jitter: the reference's Halton camera samples) the tile setup stores its offset in pixel
// units per pixel (tjit): the candidate rectangle grows by the tile's largest offset, the pretest uses the ray's own
// position, so the number of exact tests per hit stays that of a regular grid.
// Tiles whose rays do not fit (different origins, a sample more than two pixels off its pixel, a direction more than
// 26 degrees off the tile axis, NaNs, fewer than 2 columns or rows) are flagged and their rays take
// the per-ray walk of k_traverse in the same launch.
#pragma once

namespace bre {

constexpr int TILE_W = 16, TILE_H = 16;          // integrators/photonbeam.cpp:345-347
constexpr int TT_THREADS = 256;                  // threads per k_traverse_tile block
constexpr int TT_WARPS = TT_THREADS / 32;
constexpr unsigned PAIR_TILE = 0x40000000u;      // pair.x bit (with bit 31 clear): the pair stands for all rays of tile (pair.x & TILE_MASK)
constexpr unsigned TILE_MASK = 0x3fffffffu;
constexpr int TFRAME_V4 = 14;                    // float4 per tile frame: 8 for k_leaf_tile + the 6 frustum planes
// tile state bytes
enum { TS_TILE = 0, TS_INCOHERENT = 1 };

// Top cut of the production tree (bre_build.cu k_top_cut): up to TOP_CUT entries {lo.xyz, ref}{hi.xyz, -}
// covering the whole tree; staged into shared memory by cp.async.bulk (TMA) once per block.

struct TileArgs {
    const float4 *cut;       // TOP_CUT x 2 float4 (lo|ref, hi|0), padded with ref = FAST_NONE
    uint32_t nCut;           // != 0: cut[2 * TOP_CUT].x holds the number of entries (as bits); 0: start from the root
    float4 *tframes;         // TFRAME_V4 float4 per tile
    float4 *trays;           // 256 float4 per tile: {d.xyz, tMax} of pixel (i, j) at [j * 16 + i] (tile-blocked copy of the rays)
    float2 *tjit;            // 256 float2 per tile: where the ray of pixel (i, j) really is in the tile's pixel model, minus (i, j)
    unsigned char *tileState;
    uint32_t imgW, imgH, tilesX, tilesY;
};

// ---- TMA 1-D bulk copy global -> shared with an mbarrier (sm_90+; SASS: UBLKCP)
__device__ __forceinline__ void mbar_init(uint64_t *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");   // visible to the async (TMA) proxy
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dstSmem, const void *srcGlobal, unsigned bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     (unsigned)__cvta_generic_to_shared(dstSmem)),
                 "l"(srcGlobal), "r"(bytes), "r"((unsigned)__cvta_generic_to_shared(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned phase) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"((unsigned)__cvta_generic_to_shared(bar)),
        "r"(phase)
        : "memory");
}

// 16-byte asynchronous copy global -> shared (SASS: LDGSTS), bypassing L1
__device__ __forceinline__ void cp_async16(void *dstSmem, const void *srcGlobal) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(dstSmem)), "l"(srcGlobal)
                 : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// max over the box of N.X >= c   (N.xyz, c in pl); all six frustum planes have this form
__device__ __forceinline__ bool plane_in(const float4 pl, float lox, float loy, float loz, float hix, float hiy,
                                         float hiz) {
    const float x = pl.x > 0.f ? hix : lox, y = pl.y > 0.f ? hiy : loy, z = pl.z > 0.f ? hiz : loz;
    const float m = __fmaf_rn(pl.z, z, __fmaf_rn(pl.y, y, pl.x * x));
    return !(m < pl.w);       // NaN keeps the box
}
__device__ __forceinline__ bool frustum_hit(const float4 *pl, float lox, float loy, float loz, float hix, float hiy,
                                            float hiz) {
    bool in = plane_in(pl[0], lox, loy, loz, hix, hiy, hiz);
    in = in && plane_in(pl[1], lox, loy, loz, hix, hiy, hiz);
    in = in && plane_in(pl[2], lox, loy, loz, hix, hiy, hiz);
    in = in && plane_in(pl[3], lox, loy, loz, hix, hiy, hiz);
    in = in && plane_in(pl[4], lox, loy, loz, hix, hiy, hiz);
    in = in && plane_in(pl[5], lox, loy, loz, hix, hiy, hiz);
    return in;
}

// Cold path: the pair queue is full, one (tile, leaf) pair handled in place (every ray x every photon).
template <int MODE>
__device__ __noinline__ void process_tile_pair_inplace(const LeafArgs *Ag, uint32_t tile, int ref, uint32_t imgW,
                                                       uint32_t imgH, uint32_t tilesX) {
    const LeafArgs &A = *Ag;
    const uint32_t first = (uint32_t)ref & 0x07ffffffu;
    const uint32_t cnt = (((uint32_t)ref >> 27) & 15u) + 1u;
    const uint32_t x0 = (tile % tilesX) * TILE_W, y0 = (tile / tilesX) * TILE_H;
    for (uint32_t j = 0; j < (uint32_t)TILE_H && y0 + j < imgH; ++j)
        for (uint32_t i = 0; i < (uint32_t)TILE_W && x0 + i < imgW; ++i) {
            const uint32_t q = (y0 + j) * imgW + x0 + i;
            const float4 o = __ldg(A.ro + q), d = __ldg(A.rd + q);
            RayS r;
            r.ox = o.x; r.oy = o.y; r.oz = o.z; r.tMax = o.w;
            r.dx = d.x; r.dy = d.y; r.dz = d.z;
            for (uint32_t k = 0; k < cnt; ++k) {
                float th, d2;
                if (disk_test(r, __ldg(A.t.s_pos_r + first + k), th, d2)) {
                    if (MODE == 0) shade_hit(A, first + k, q, th, d2);
                    else record_hit(A, MODE, first + k, q);
                }
            }
        }
}

// ---------------------------------------------------------------- K5a'
// Photon maps are extremely uneven on the screen (smoke10m_1080p: the tile that looks into the point
// light holds 284 000 photons, 230x the mean; 8 of 8 160 tiles hold 12 % of the map), so a per-tile
// unit of work alone leaves the chip waiting for a handful of walks:
//   a block per tile, 256 nodes per step                                      K5a' 0.85 ms
//   a warp per tile, wide walks give work to the other warps of their block        0.73 ms (SMs busy 23 % of it)
//   one global breadth-first queue of (tile, node) items, 32 per claim             1.17 ms (3-4 same-address
//                                                     atomics per 32 nodes: the counters serialise the chip)
// Kept: a WARP per tile walk (deque of node refs in shared memory, up to 32 nodes per step, the TOP CUT
// of the tree tested first from shared memory, where one TMA bulk copy per block put it), and a walk
// whose deque grows past DONATE_MIN appends its 32 OLDEST entries -- the largest subtrees -- as one
// batch to a GLOBAL queue (one atomicAdd per batch).  A warp that finds no tile left claims the next
// batch position and waits until it is written (or until `pending` -- walks in progress + batches not
// yet finished -- reads 0).  Consumers reset the slots to EMPTY, so the queue is clean for the next launch.
constexpr int TW_STACK = 512;            // node refs per warp (a deque: donated from the bottom, popped from the top)
constexpr int DONATE_MIN = 96;           // a walk with this many pending nodes gives 32 away
constexpr unsigned HEAVY_STEPS_TILE = 24; // a tile walk this long (or one that gave work away) is claimed first by the next launch
struct __align__(16) TileSmem {
    float4 cut[2 * TOP_CUT];
    float4 planes[TT_WARPS][6];
    int stack[TT_WARPS][TW_STACK];
    int tilesDone;
    uint64_t bar;
};
constexpr unsigned WQ_EMPTY = 0xffffffffu;
struct WalkQueue {
    uint2 *items;            // {tile, node ref}, batches of 32
    unsigned cap;
    unsigned *ctl;           // [0] head (claimed), [1] tail (reserved), [2] pending
};

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ int ld_vol_s(const int *p) { return *reinterpret_cast<const volatile int *>(p); }

template <int MODE, bool COUNT>
#ifndef TW_MINB
#define TW_MINB 3
#endif
__global__ void __launch_bounds__(TT_THREADS, TW_MINB) k_tile_walk(
    LeafArgs A, const LeafArgs *__restrict__ Ag, TileArgs T, WalkQueue W, uint32_t tile0, uint32_t nTiles,
    uint2 *__restrict__ queue, unsigned int queueCap, unsigned int *__restrict__ queueCtl,
    const unsigned int *__restrict__ heavyIn, const unsigned int *__restrict__ nHeavyInPtr,
    const unsigned int *__restrict__ heavyFlagIn, unsigned int *__restrict__ heavyOut,
    unsigned int *__restrict__ heavyFlagOut, unsigned int *__restrict__ nHeavyOut, unsigned int heavyCap,
    int rankMod, int rankId, unsigned long long *__restrict__ counters) {
    extern __shared__ __align__(16) unsigned char tt_smem_raw[];
    TileSmem &sm = *reinterpret_cast<TileSmem *>(tt_smem_raw);
    const TreeDev &t = A.t;
    const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const unsigned lt = (1u << lane) - 1u;
    const unsigned FULL = 0xffffffffu;
    const unsigned nHeavyIn = min(__ldg(nHeavyInPtr), HEAVY_CAP);
    int *stack = sm.stack[warp];
    float4 *planes = sm.planes[warp];
    unsigned long long cNodes = 0;
    unsigned nEmitted = 0, nInplace = 0;
    bool anyPairs = false;
    unsigned chunkPos = 0, chunkEnd = 0;   // this warp's reserved queue slots

    // ---- stage the top cut of the tree into shared memory: one TMA bulk copy per block
    const unsigned nCut = T.nCut ? min(__float_as_uint(__ldg(&T.cut[2 * TOP_CUT].x)), (unsigned)TOP_CUT) : 0u;
    if (tid == 0) {
        sm.tilesDone = 0;
        if (nCut) mbar_init(&sm.bar, 1);
    }
    __syncthreads();
    if (nCut && tid == 0) {
        const unsigned bytes = nCut * 32u;
        mbar_expect_tx(&sm.bar, bytes);
        bulk_g2s(sm.cut, T.cut, bytes, &sm.bar);
    }
    bool cutReady = nCut == 0;

    // emit up to two leaves per lane as (tile, leaf) pairs into this warp's queue chunk
    auto emit = [&](unsigned tile, bool e0, int c0, bool e1, int c1) {
        const unsigned ml0 = __ballot_sync(FULL, e0), ml1 = __ballot_sync(FULL, e1);
        const unsigned nl = __popc(ml0) + __popc(ml1);
        if (!nl) return;
        if (chunkPos + nl > chunkEnd) {
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
        anyPairs = true;
        const unsigned px = tile | PAIR_TILE;
        if (e0) {
            if (p0 < queueCap) queue[p0] = make_uint2(px, (unsigned)c0);
            else { process_tile_pair_inplace<MODE>(Ag, tile, c0, T.imgW, T.imgH, T.tilesX); ++nInplace; }
        }
        if (e1) {
            if (p1 < queueCap) queue[p1] = make_uint2(px, (unsigned)c1);
            else { process_tile_pair_inplace<MODE>(Ag, tile, c1, T.imgW, T.imgH, T.tilesX); ++nInplace; }
        }
        if (COUNT) nEmitted += (e0 ? 1u : 0u) + (e1 ? 1u : 0u);
    };

    for (;;) {
        // ================= get work: a new tile while there are any, else a donated batch from the global queue
        unsigned tile = 0xffffffffu;
        unsigned nbot = 0, ntop = 0;       // deque [nbot, ntop) of node refs, indices mod TW_STACK
        bool stolen = false;
        unsigned pkClaimed = 0xffffffffu;
        {
            unsigned pk = 0xffffffffu;
            if (lane == 0 && ld_vol_s(&sm.tilesDone) == 0) {
                for (;;) {
                    const unsigned c = atomicAdd(queueCtl + QC_TILE, 1u);
                    if (c < nHeavyIn) { pk = __ldg(heavyIn + c); break; }
                    pk = c - nHeavyIn;
                    if (pk >= nTiles) { pk = 0xffffffffu; *reinterpret_cast<volatile int *>(&sm.tilesDone) = 1; break; }
                    if (nHeavyIn == 0 || !((__ldg(heavyFlagIn + (pk >> 5)) >> (pk & 31u)) & 1u)) break;
                }
                if (pk != 0xffffffffu) atomicAdd(W.ctl + 2, 1u);      // a walk in progress
            }
            pk = __shfl_sync(FULL, pk, 0);
            if (pk != 0xffffffffu) { pkClaimed = pk; tile = tile0 + pk; }
        }
        if (tile == 0xffffffffu) {
            // no tile left: take the next batch of 32 (tile, node) items some wide walk gives away -- it may not
            // have been written yet: wait for it, or for the end (no walk in progress, nothing pending)
            unsigned h = 0;
            if (lane == 0) h = atomicAdd(W.ctl + 0, 32u);
            h = __shfl_sync(FULL, h, 0);
            if (h + 32u <= W.cap) {
                uint2 it = make_uint2(WQ_EMPTY, WQ_EMPTY);
                for (;;) {
                    if (it.x == WQ_EMPTY) it = __ldcv(W.items + h + lane);
                    if (__all_sync(FULL, it.x != WQ_EMPTY)) { stolen = true; break; }
                    if (__shfl_sync(FULL, *reinterpret_cast<volatile unsigned *>(W.ctl + 2), 0) == 0u) break;
                    __nanosleep(200);
                }
                if (stolen) {
                    __stcg(W.items + h + lane, make_uint2(WQ_EMPTY, WQ_EMPTY));   // clean for the next launch
                    tile = __shfl_sync(FULL, it.x, 0);
                    stack[lane] = (int)it.y;
                    ntop = 32;
                }
            }
        }
        if (tile == 0xffffffffu) break;

        unsigned steps = 0, donated = 0;
        bool walk = true;
        // one walk step: pop up to 32 nodes from the top of the deque, test both children of each
        auto step = [&]() {
            const unsigned nsize = ntop - nbot;
            // popping k pushes at most 2k; with no room left the walk degrades to depth-first (k = 1), which needs
            // at most the depth of the tree (< 64) more
            const unsigned room = TW_STACK - 64u - nsize;
            const unsigned k = (nsize + 64u < (unsigned)TW_STACK) ? min(min(32u, nsize), max(room, 1u)) : 1u;
            const bool act = lane < k;
            int ref = 0;
            if (act) ref = stack[(ntop - 1u - lane) & (TW_STACK - 1)];
            ntop -= k;
            ++steps;
            bool n0 = false, n1 = false, e0 = false, e1 = false;
            int c0 = 0, c1 = 0;
            if (act) {
                const FastNode *fn = t.fast + ref;
                const float4 a = __ldg(&fn->a), b = __ldg(&fn->b), c = __ldg(&fn->c);
                const int4 dd = __ldg(&fn->d);
                const bool h0 = frustum_hit(planes, a.x, a.y, a.z, a.w, b.x, b.y);
                const bool h1 = frustum_hit(planes, b.z, b.w, c.x, c.y, c.z, c.w);
                c0 = dd.x; c1 = dd.y;
                n0 = h0 && c0 >= 0; n1 = h1 && c1 >= 0;
                e0 = h0 && c0 < 0; e1 = h1 && c1 < 0;
                if (COUNT) ++cNodes;
            }
            __syncwarp();   // all pops are done before anyone pushes
            emit(tile, e0, c0, e1, c1);
            const unsigned mn0 = __ballot_sync(FULL, n0), mn1 = __ballot_sync(FULL, n1);
            if (n0) stack[(ntop + __popc(mn0 & lt)) & (TW_STACK - 1)] = c0;
            if (n1) stack[(ntop + __popc(mn0) + __popc(mn1 & lt)) & (TW_STACK - 1)] = c1;
            ntop += __popc(mn0) + __popc(mn1);
            __syncwarp();
            // ---- a wide walk gives its 32 OLDEST entries (the largest subtrees) to the global queue
            if (ntop - nbot >= (unsigned)DONATE_MIN) {
                unsigned base = 0;
                if (lane == 0) {
                    atomicAdd(W.ctl + 2, 1u);                 // pending first: the count never reads 0 while a batch is unprocessed
                    base = atomicAdd(W.ctl + 1, 32u);
                }
                base = __shfl_sync(FULL, base, 0);
                if (base + 32u <= W.cap) {
                    __stcg(W.items + base + lane, make_uint2(tile, (unsigned)stack[(nbot + lane) & (TW_STACK - 1)]));
                    nbot += 32;
                    ++donated;
                } else if (lane == 0) atomicSub(W.ctl + 2, 1u);   // queue full: keep walking it here
            }
        };
        auto drain = [&](unsigned limit) { while (ntop - nbot > limit) step(); };
        if (stolen) {
            // the donor's tile: planes from its frame
            if (lane < 6) planes[lane] = __ldcg(T.tframes + (size_t)tile * TFRAME_V4 + 8 + lane);
            __syncwarp();
        } else {
            // ================= tile setup: frame, model, frustum
            const unsigned tx = tile % T.tilesX, ty = tile / T.tilesX;
            walk = false;
            // multi-GPU: a rank walks only the tiles it owns (bre_tile_owner: round-robin)
            const bool mineTile = !(rankMod > 1 && (int)(tile % (unsigned)rankMod) != rankId);
            const unsigned x0 = tx * TILE_W, y0 = ty * TILE_H;
            const unsigned twv = min((unsigned)TILE_W, T.imgW - x0), thv = min((unsigned)TILE_H, T.imgH - y0);
            unsigned state = TS_TILE;
            if (mineTile && t.n > 0) {
                const unsigned q00 = y0 * T.imgW + x0;
                const float4 o0 = __ldg(A.ro + q00), d00 = __ldg(A.rd + q00);
                const float4 d10 = __ldg(A.rd + q00 + twv - 1), d01 = __ldg(A.rd + q00 + (thv - 1) * T.imgW);
                const float4 d11 = __ldg(A.rd + q00 + (thv - 1) * T.imgW + twv - 1);
                // frame: u = bisector of the diagonal corners, v = along the row, n = u x v
                float ux = d00.x + d11.x, uy = d00.y + d11.y, uz = d00.z + d11.z;
                const float ul = rsqrtf(ux * ux + uy * uy + uz * uz);
                ux *= ul; uy *= ul; uz *= ul;
                float wx = d10.x - d00.x, wy = d10.y - d00.y, wz = d10.z - d00.z;
                const float wu = wx * ux + wy * uy + wz * uz;
                wx -= wu * ux; wy -= wu * uy; wz -= wu * uz;
                const float wl2 = wx * wx + wy * wy + wz * wz;
                const float wi = rsqrtf(wl2);
                const float vx = wx * wi, vy = wy * wi, vz = wz * wi;
                const float nx = uy * vz - uz * vy, ny = uz * vx - ux * vz, nz = ux * vy - uy * vx;
                auto gno = [&](const float4 dd, float &s, float &tt, float &da) {
                    da = dd.x * ux + dd.y * uy + dd.z * uz;
                    const float ia = __fdividef(1.f, da);
                    s = (dd.x * vx + dd.y * vy + dd.z * vz) * ia;
                    tt = (dd.x * nx + dd.y * ny + dd.z * nz) * ia;
                };
                float s00, t00, s10, t10, s01, t01, daC;
                gno(d00, s00, t00, daC);
                gno(d10, s10, t10, daC);
                gno(d01, s01, t01, daC);
                const float fi = 1.f / (float)max(twv - 1u, 1u), fj = 1.f / (float)max(thv - 1u, 1u);
                float As = (s10 - s00) * fi, At = (t10 - t00) * fi, Bs = (s01 - s00) * fj, Bt = (t01 - t00) * fj;
                // ---- the tile's rays, 8 per lane: bounds, deviation from the model, coherence
                const float BIGF = 3.0e38f;
                float sLo = BIGF, sHi = -BIGF, tLo = BIGF, tHi = -BIGF, dev = 0.f, g2 = 0.f, tFar = 0.f, oSp = 0.f, oAbs = 0.f;
                float bad = 0.f, oDiff = 0.f, jI = 0.f, jJ = 0.f, nLive = 0.f;
                // inverse of the affine pixel model: the sample position of a ray inside its pixel (camera jitter) is taken
                // out of the model error and stored per pixel, so that the candidate pretest of k_leaf_tile stays as sharp
                // for jittered samples (the reference's Halton camera samples) as for a regular grid
                const float idet0 = 1.f / (As * Bt - Bs * At);
                const float Ni00 = Bt * idet0, Ni01 = -Bs * idet0, Ni10 = -At * idet0, Ni11 = As * idet0;
                const unsigned ti = lane & (TILE_W - 1);
#pragma unroll 4
                for (unsigned k = 0; k < TILE_H / 2; ++k) {
                    const unsigned tj = (lane >> 4) + 2 * k;
                    if (ti < twv && tj < thv) {
                        const unsigned q = q00 + tj * T.imgW + ti;
                        const float4 o = __ldg(A.ro + q), d = __ldg(A.rd + q);
                        // tile-blocked copy for k_leaf_tile: a candidate reads 16 bytes next to its row neighbours'
                        // instead of two 32-byte sectors of the image-ordered arrays
                        __stcg(T.trays + (size_t)tile * 256u + tj * 16u + ti, make_float4(d.x, d.y, d.z, o.w));
                        if (o.w > 0.f) {
                            nLive += 1.f;
                            if (__float_as_uint(o.x) != __float_as_uint(o0.x) || __float_as_uint(o.y) != __float_as_uint(o0.y) ||
                                __float_as_uint(o.z) != __float_as_uint(o0.z)) oDiff = 1.f;
                            float myS, myT, myDa;
                            gno(d, myS, myT, myDa);
                            const float ps = s00 + (float)ti * As + (float)tj * Bs, pt = t00 + (float)ti * At + (float)tj * Bt;
                            const float es = myS - ps, et = myT - pt;
                            const float2 jit = make_float2(Ni00 * es + Ni01 * et, Ni10 * es + Ni11 * et);
                            jI = fmaxf(jI, fabsf(jit.x)); jJ = fmaxf(jJ, fabsf(jit.y));
                            const float dv = fmaxf(fabsf(jit.x), fabsf(jit.y));      // in pixels
                            const float gg = myS * myS + myT * myT;
                            const float os = fmaxf(fmaxf(fabsf(o.x - o0.x), fabsf(o.y - o0.y)), fabsf(o.z - o0.z));
                            sLo = fminf(sLo, myS); sHi = fmaxf(sHi, myS); tLo = fminf(tLo, myT); tHi = fmaxf(tHi, myT);
                            dev = fmaxf(dev, dv); g2 = fmaxf(g2, gg); tFar = fmaxf(tFar, o.w); oSp = fmaxf(oSp, os);
                            oAbs = fmaxf(oAbs, fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z)));
                            // a live ray that breaks the model (NaN anywhere, far off axis) makes the tile incoherent
                            if (!(myDa > 0.85f && dv == dv && gg == gg && os == os)) bad = 1.f;
                        } else if (!(o.w <= 0.f)) bad = 1.f;      // NaN tMax
                    }
                }
                sLo = warp_min(sLo); sHi = warp_max(sHi); tLo = warp_min(tLo); tHi = warp_max(tHi);
                dev = warp_max(dev); g2 = warp_max(g2); tFar = warp_max(tFar); oSp = warp_max(oSp);
                oAbs = warp_max(oAbs); bad = warp_max(bad); oDiff = warp_max(oDiff); jI = warp_max(jI); jJ = warp_max(jJ);
#pragma unroll
                for (int o2 = 16; o2 > 0; o2 >>= 1) nLive += __shfl_xor_sync(FULL, nLive, o2);
                // ---- jittered samples: the three corner rays carry their own jitter, so the model through them is skewed by
                // up to a pixel across the tile.  With every ray alive the least-squares affine fit over the full rectangle
                // is closed-form (i and j are orthogonal there); the per-pixel offsets are then re-measured against it and
                // shrink to the jitter itself (|.| <= ~0.5 pixels for samples inside their pixels).
                // (the offset table is written for jittered tiles only -- 17 MB at 1080p that a regular grid never reads and
                // that would otherwise sit dirty in L2 while k_leaf_tile runs)
                const bool jittered = jI > 1e-3f || jJ > 1e-3f;
                const bool refit = jittered && bad == 0.f && twv >= 2 && thv >= 2 && nLive == (float)(twv * thv);
                if (jittered && !refit) {
                    // rays missing or a ragged tile: the offsets against the model through the corner rays
                    for (unsigned k = 0; k < TILE_H / 2; ++k) {
                        const unsigned tj = (lane >> 4) + 2 * k;
                        if (ti < twv && tj < thv) {
                            const float4 d = __ldcg(T.trays + (size_t)tile * 256u + tj * 16u + ti);
                            float2 jit = make_float2(0.f, 0.f);
                            if (d.w > 0.f) {
                                float myS, myT, myDa;
                                gno(d, myS, myT, myDa);
                                const float es = myS - (s00 + (float)ti * As + (float)tj * Bs), et = myT - (t00 + (float)ti * At + (float)tj * Bt);
                                jit = make_float2(Ni00 * es + Ni01 * et, Ni10 * es + Ni11 * et);
                            }
                            __stcg(T.tjit + (size_t)tile * 256u + tj * 16u + ti, jit);
                        }
                    }
                }
                if (refit) {
                    const float ib = 0.5f * (float)(twv - 1u), jb = 0.5f * (float)(thv - 1u);
                    float S = 0.f, Tt = 0.f, Sis = 0.f, Sit = 0.f, Sjs = 0.f, Sjt = 0.f;
                    for (unsigned k = 0; k < TILE_H / 2; ++k) {
                        const unsigned tj = (lane >> 4) + 2 * k;
                        if (ti < twv && tj < thv) {
                            const float4 d = __ldcg(T.trays + (size_t)tile * 256u + tj * 16u + ti);
                            float myS, myT, myDa;
                            gno(d, myS, myT, myDa);
                            const float wi2 = (float)ti - ib, wj2 = (float)tj - jb;
                            S += myS; Tt += myT; Sis += wi2 * myS; Sit += wi2 * myT; Sjs += wj2 * myS; Sjt += wj2 * myT;
                        }
                    }
#pragma unroll
                    for (int o2 = 16; o2 > 0; o2 >>= 1) {
                        S += __shfl_xor_sync(FULL, S, o2); Tt += __shfl_xor_sync(FULL, Tt, o2);
                        Sis += __shfl_xor_sync(FULL, Sis, o2); Sit += __shfl_xor_sync(FULL, Sit, o2);
                        Sjs += __shfl_xor_sync(FULL, Sjs, o2); Sjt += __shfl_xor_sync(FULL, Sjt, o2);
                    }
                    const float tw = (float)twv, th = (float)thv;
                    const float Sii = th * tw * (tw * tw - 1.f) * (1.f / 12.f), Sjj = tw * th * (th * th - 1.f) * (1.f / 12.f);
                    As = Sis / Sii; At = Sit / Sii; Bs = Sjs / Sjj; Bt = Sjt / Sjj;
                    s00 = S / nLive - ib * As - jb * Bs;
                    t00 = Tt / nLive - ib * At - jb * Bt;
                    const float idet1 = 1.f / (As * Bt - Bs * At);
                    const float R00 = Bt * idet1, R01 = -Bs * idet1, R10 = -At * idet1, R11 = As * idet1;
                    jI = 0.f; jJ = 0.f;
                    for (unsigned k = 0; k < TILE_H / 2; ++k) {
                        const unsigned tj = (lane >> 4) + 2 * k;
                        if (ti < twv && tj < thv) {
                            const float4 d = __ldcg(T.trays + (size_t)tile * 256u + tj * 16u + ti);
                            float myS, myT, myDa;
                            gno(d, myS, myT, myDa);
                            const float es = myS - (s00 + (float)ti * As + (float)tj * Bs), et = myT - (t00 + (float)ti * At + (float)tj * Bt);
                            const float2 jit = make_float2(R00 * es + R01 * et, R10 * es + R11 * et);
                            jI = fmaxf(jI, fabsf(jit.x)); jJ = fmaxf(jJ, fabsf(jit.y));
                            __stcg(T.tjit + (size_t)tile * 256u + tj * 16u + ti, jit);
                        }
                    }
                    jI = warp_max(jI); jJ = warp_max(jJ);
                    dev = fmaxf(jI, jJ);       // (NaN -> the comparisons below fail -> incoherent)
                    if (!(jI == jI && jJ == jJ)) bad = 1.f;
                }
                const bool dead = !(tFar > 0.f) && bad == 0.f;
                const float M = oAbs + t.abs_max;
                const float det = As * Bt - Bs * At;
                const float pitch = sqrtf(fminf(As * As + At * At, Bs * Bs + Bt * Bt));
                // (comparisons written so that a NaN anywhere makes the tile incoherent)
                const bool coherent = !dead && bad == 0.f && twv >= 2 && thv >= 2 && wl2 > 1e-30f && g2 <= 0.25f &&
                                      oSp <= M * 0x1p-10f && pitch > 0.f && dev <= 2.f && jI <= 2.f && jJ <= 2.f &&
                                      fabsf(det) > 1e-3f * pitch * pitch;     // (tFar may be +inf: no far cull)
                if (!dead && !coherent) state = TS_INCOHERENT;
                if (coherent) {
                    walk = true;
                    const float oSlack = 2.f * oSp + M * 0x1p-18f;
                    // frustum planes: max corner of N.X >= c.  The rays' points have s in [sLo, sHi], t in [tLo, tHi],
                    // 0 < depth <= tFar relative to their own origin; oSlack covers the origins' spread and rounding.
                    float4 pl = make_float4(0.f, 0.f, 0.f, 0.f);
                    {
                        const float e = oSlack + M * 0x1p-18f;
                        float Nx = 0.f, Ny = 0.f, Nz = 0.f, extra = 0.f;
                        switch (lane) {
                            case 0: Nx = vx - sLo * ux; Ny = vy - sLo * uy; Nz = vz - sLo * uz; break;
                            case 1: Nx = sHi * ux - vx; Ny = sHi * uy - vy; Nz = sHi * uz - vz; break;
                            case 2: Nx = nx - tLo * ux; Ny = ny - tLo * uy; Nz = nz - tLo * uz; break;
                            case 3: Nx = tHi * ux - nx; Ny = tHi * uy - ny; Nz = tHi * uz - nz; break;
                            case 4: Nx = ux; Ny = uy; Nz = uz; break;                                     // some depth > 0
                            default: Nx = -ux; Ny = -uy; Nz = -uz; extra = tFar * 1.000001f; break;       // some depth <= tFar
                        }
                        const float l1 = fabsf(Nx) + fabsf(Ny) + fabsf(Nz);
                        pl = make_float4(Nx, Ny, Nz, Nx * o0.x + Ny * o0.y + Nz * o0.z - e * l1 - extra);
                    }
                    if (lane < 6) planes[lane] = pl;
                    // frame for k_leaf_tile (+ the planes, for warps that take over part of this walk)
                    const float idet = 1.f / det;
                    const float Mi00 = Bt * idet, Mi01 = -Bs * idet, Mi10 = -At * idet, Mi11 = As * idet;
                    const float kphi = sqrtf(1.f + g2 * (1.f + g2)) * 1.0001f;
                    float4 fv;
                    switch (lane) {
                        case 0: fv = make_float4(o0.x, o0.y, o0.z, tFar); break;
                        case 1: fv = make_float4(ux, uy, uz, 2e-3f * sqrtf(fmaxf(As * As + At * At, Bs * Bs + Bt * Bt))); break;   // .w: slack of rho (2e-3 pixels: rounding of the stored jitter, jitter below 1e-3 pixels is ignored)
                        case 2: fv = make_float4(vx, vy, vz, s00); break;
                        case 3: fv = make_float4(nx, ny, nz, t00); break;
                        case 4: fv = make_float4(Mi00, Mi01, Mi10, Mi11); break;
                        case 5: fv = make_float4(sqrtf(Mi00 * Mi00 + Mi01 * Mi01), sqrtf(Mi10 * Mi10 + Mi11 * Mi11), oSlack, kphi); break;
                        case 6: fv = make_float4(As * As + At * At, As * Bs + At * Bt, Bs * Bs + Bt * Bt,
                                                 __uint_as_float(twv | (thv << 8))); break;
                        // .y: every live ray starts at exactly o0 (a pinhole camera): candidates take the origin from the frame
                        default: fv = make_float4(__uint_as_float(q00), oDiff == 0.f ? 1.f : 0.f, jI, jJ); break;   // .zw: largest |jitter| in pixels
                    }
                    float4 *fr = T.tframes + (size_t)tile * TFRAME_V4;
                    if (lane < 8) __stcg(fr + lane, fv);
                    if (lane == 0 && (jI > 1e-3f || jJ > 1e-3f)) atomicOr(queueCtl + QC_KINDS, (unsigned)KIND_TILE_JIT);
                }
            }
            if (lane == 0) {
                T.tileState[tile] = (unsigned char)state;
                if (state == TS_INCOHERENT) atomicAdd(queueCtl + QC_NINCOH, 1u);
            }
            __syncwarp();
            if (walk) {
                // the planes again, behind the frame, for thieves
                if (lane < 6) __stcg(T.tframes + (size_t)tile * TFRAME_V4 + 8 + lane, planes[lane]);
                if (!cutReady) { mbar_wait(&sm.bar, 0); cutReady = true; }
                // ---- step 0: the top cut, 16 entries per lane
                if (nCut) {
                    for (unsigned k0 = 0; k0 < nCut; k0 += 32) {
                        const unsigned k = k0 + lane;
                        bool n0 = false, e0 = false;
                        int c0 = 0;
                        if (k < nCut) {
                            const float4 lo = sm.cut[2 * k], hi = sm.cut[2 * k + 1];
                            c0 = __float_as_int(lo.w);
                            const bool h = frustum_hit(planes, lo.x, lo.y, lo.z, hi.x, hi.y, hi.z);
                            n0 = h && c0 >= 0; e0 = h && c0 < 0;
                        }
                        emit(tile, e0, c0, false, 0);
                        const unsigned mn = __ballot_sync(FULL, n0);
                        if (n0) stack[(ntop + __popc(mn & lt)) & (TW_STACK - 1)] = c0;
                        ntop += __popc(mn);
                        __syncwarp();
                        drain((unsigned)TW_STACK - 160u);      // a wide tile passes most of the cut: keep the deque short
                    }
                } else {
                    const int c0 = t.root_ref;
                    if (c0 >= 0) { if (lane == 0) stack[0] = c0; ntop = 1; }
                    else emit(tile, lane == 0, c0, false, 0);
                }
                ++steps;
                __syncwarp();
            }
        }

        // ================= walk until the deque is empty
        if (walk) drain(0u);
        if (lane == 0) atomicSub(W.ctl + 2, 1u);
        if (!stolen && lane == 0 && (steps >= HEAVY_STEPS_TILE || donated)) {
            const unsigned pos = heavyCap ? atomicAdd(nHeavyOut, 1u) : 0u;
            if (pos < heavyCap) {
                heavyOut[pos] = pkClaimed;
                atomicOr(heavyFlagOut + (pkClaimed >> 5), 1u << (pkClaimed & 31u));
            }
        }
    }
    if (!cutReady) mbar_wait(&sm.bar, 0);   // never leave with the bulk copy in flight
    for (unsigned p = chunkPos + lane; p < chunkEnd; p += 32)
        if (p < queueCap) queue[p] = make_uint2(PAIR_NONE, 0u);
    if (nInplace) atomicAdd(queueCtl + QC_INPLACE, nInplace);
    if (lane == 0 && anyPairs) atomicOr(queueCtl + QC_KINDS, (unsigned)KIND_TILE);
    if (COUNT) {
#pragma unroll
        for (int o2 = 16; o2 > 0; o2 >>= 1) {
            cNodes += __shfl_xor_sync(FULL, cNodes, o2);
            nEmitted += __shfl_xor_sync(FULL, nEmitted, o2);
        }
        if (lane == 0) {
            atomicAdd(counters + 3, cNodes);
            atomicAdd(queueCtl + QC_EMITTED, nEmitted);
        }
    }
}

// ---------------------------------------------------------------- K5b'
// Per warp and pass: TE_SUB x 32 photons (stage 1) -> their candidate pixels (stage 2) -> hits, shaded
// before the next pass overwrites the entries, so a hit only carries {entry, ray, t, d2} and takes the
// photon's radius / power / direction from the entry in shared memory (staged ONCE per (tile, photon)
// instead of three L2 sectors per hit), and adds its radiance with ONE vector atomic (RED.128).
#ifndef BRE_TE_SUB
#define BRE_TE_SUB 1
#endif
constexpr int TE_SUB = BRE_TE_SUB;        // stage-1 sub-passes per stage 2
constexpr int TE_N = 32 * TE_SUB;
struct __align__(16) TileEnt {      // stage-1 results, one entry per photon that has candidate pixels
    float4 p[TE_N];        // photon {pos, radius}
    float4 pw[TE_N];       // power
    float4 dr[TE_N];       // direction
    float4 c[TE_N];        // ci, cj, q11, q22   (centre in tile pixel coordinates, quadratic form / rho^2)
    float q12[TE_N];
    uint32_t s[TE_N];      // sorted photon index
    uint32_t rb[TE_N];     // ray index of the tile's pixel (0, 0)
    uint32_t tl[TE_N];     // tile index
    uint32_t rect[TE_N];   // i0 | j0 << 8 | w << 16 | h << 24
    float4 o[TE_N];        // the tile's ray origin, .w != 0: every ray of the tile starts exactly there
    uint32_t cs[TE_N];     // first candidate of the entry (entries are compacted: only photons with candidates, ascending cs)
};
struct __align__(16) TileHits {
    uint32_t e[64], q[64];   // e = entry | pixel-in-tile << 8
    float th[64], d2[64], ct[64];
};
#ifndef TILE_MINB
#define TILE_MINB 5
#endif
constexpr size_t LEAF_TILE_SMEM = LP_WARPS * (sizeof(TileEnt) + sizeof(TileHits));

// radiance of one hit of the tile path: photon data from the entry, cos(theta) from the candidate pass,
// optical depth from the tile-blocked table (see hit_radiance for the formula)
__device__ __forceinline__ float3 hit_radiance_p(const LeafArgs &A, float radius, const float4 pw, float cosT,
                                                 uint32_t q, uint32_t tile, uint32_t pp, float th, float d2) {
    const float r2 = radius * radius;
    const float invr2 = __fdividef(1.f, r2);
    const float sK = __fmaf_rn(-d2, invr2, 1.f);             // kernel(): photonmap.cpp:170-174
    const float wgt = (3.f * BRE_INV_PI) * sK * sK * invr2 * __fdividef(1.f, A.nPathsNorm);
    const float g = A.m.g;
    const float denom = __fmaf_rn(2.f * g, cosT, __fmaf_rn(g, g, 1.f));   // PhaseHG medium.h:66-69
    const float rs = rsqrtf(denom);
    const float wp = wgt * BRE_INV4PI * __fmaf_rn(-g, g, 1.f) * rs * rs * rs;
    float od = th;
    if (A.m.kind == 1) {
        const float2 sp = __ldg(A.span + q);
        if (A.tauW) od = optical_depth_at(A, sp, A.tau + (size_t)tile * (size_t)(A.tauSteps + 1) * 256u + pp, 256, th);
        else od = optical_depth_at(A, sp, A.tau + (size_t)q * (A.tauSteps + 1), 1, th);
    }
    float t0, t1, t2;
    t0 = __expf(-A.m.sigma_t[0] * od);
    if (A.m.gray) { t1 = t0; t2 = t0; }
    else { t1 = __expf(-A.m.sigma_t[1] * od); t2 = __expf(-A.m.sigma_t[2] * od); }
    return make_float3(wp * t0 * pw.x, wp * t1 * pw.y, wp * t2 * pw.z);
}

template <int MODE, bool COUNT, int LW, bool JIT>
__global__ void __launch_bounds__(LP_WARPS * 32, TILE_MINB) k_leaf_tile(LeafArgs A, const float4 *__restrict__ tframes,
                                                                    const float4 *__restrict__ trays, const float2 *__restrict__ tjit, uint32_t imgW,
                                                                    const uint2 *__restrict__ queue,
                                                                    unsigned int queueCap,
                                                                    unsigned int *__restrict__ queueCtl,
                                                                    unsigned long long *__restrict__ counters) {
    extern __shared__ __align__(16) unsigned char lt_smem_raw[];     // LP_WARPS x (TileEnt + TileHits): > 48 KB
    TileEnt &te = reinterpret_cast<TileEnt *>(lt_smem_raw)[threadIdx.x >> 5];
    TileHits &hb = reinterpret_cast<TileHits *>(lt_smem_raw + LP_WARPS * sizeof(TileEnt))[threadIdx.x >> 5];
    constexpr unsigned G = 32 / LW;
    const unsigned FULL = 0xffffffffu;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt = (1u << lane) - 1u;
    const unsigned sub = lane / LW, j = lane % LW;
    {
        // JIT: some tile's rays sit off their pixels' model positions (camera jitter): this variant widens the candidate
        // rectangles and reads the per-pixel offsets; exactly one of the two variants runs
        const unsigned kinds = __ldg(queueCtl + QC_KINDS);
        if (!(kinds & KIND_TILE) || ((kinds & KIND_TILE_JIT) != 0u) != JIT) return;
    }
    const unsigned total = min(__ldg(queueCtl + QC_PAIR_TAIL), queueCap);
    const unsigned warpsTotal = gridDim.x * LP_WARPS;
    unsigned long long cTests = 0;
    unsigned nbuf = 0;

    // shade `cnt` buffered hits starting at hb[from], one per lane
    auto shade = [&](unsigned from, unsigned cnt) {
        if (lane < cnt) {
            const unsigned ep = hb.e[from + lane], q = hb.q[from + lane];
            const unsigned e = ep & 255u;
            const float3 v = hit_radiance_p(A, te.p[e].w, te.pw[e], hb.ct[from + lane], q, te.tl[e] & TILE_MASK, ep >> 8,
                                            hb.th[from + lane], hb.d2[from + lane]);
            atomicAdd(A.out + q, make_float4(v.x, v.y, v.z, 0.f));   // RED.E.ADD.F32x4
        }
    };

    const unsigned totalRuns = (total + LP_RUN - 1) / LP_RUN;
    const unsigned rpc = max(1u, min(64u, totalRuns / (warpsTotal * 16u)));
    unsigned cur = 0, curEnd = 0;
    for (;;) {
        if (cur == curEnd) {
            unsigned c = 0;
            if (lane == 0) c = atomicAdd(queueCtl + QC_RUN_TILE, rpc);
            cur = __shfl_sync(FULL, c, 0);
            if (cur >= totalRuns) break;
            curEnd = min(cur + rpc, totalRuns);
        }
        const unsigned run = (cur++) * LP_RUN;
        uint2 pr0 = make_uint2(PAIR_NONE, 0u), pr1 = make_uint2(PAIR_NONE, 0u);
        if (run + lane < total) pr0 = __ldg(queue + run + lane);
        if (run + 32 + lane < total) pr1 = __ldg(queue + run + 32 + lane);
        const bool t0 = pr0.x != PAIR_NONE && (pr0.x & 0xc0000000u) == PAIR_TILE;
        const bool t1 = pr1.x != PAIR_NONE && (pr1.x & 0xc0000000u) == PAIR_TILE;
        if (!__any_sync(FULL, t0 || t1)) continue;
#pragma unroll 1
        for (unsigned g0 = 0; g0 < LP_RUN / G; g0 += TE_SUB) {
            // ---- stage 1: TE_SUB x one photon per lane against its tile's pixel grid
            // (a pass without hits never waited for its payload copies: they must have landed before the entries
            // are written again -- two copies in flight to one address may land in either order)
            if (MODE == 0) cp_async_wait_all();
            unsigned totalC = 0, nEnt = 0;
#pragma unroll
            for (int u = 0; u < TE_SUB; ++u) {
                const unsigned idx = (g0 + u) * G + sub;
                const bool lower = (g0 + u) * G < 32u;              // warp-uniform
                const unsigned qx = __shfl_sync(FULL, lower ? pr0.x : pr1.x, idx & 31u);
                const unsigned qy = __shfl_sync(FULL, lower ? pr0.y : pr1.y, idx & 31u);
                const uint32_t firstP = qy & 0x07ffffffu;
                const uint32_t cntP = ((qy >> 27) & 15u) + 1u;
                const bool on = qx != PAIR_NONE && (qx & 0xc0000000u) == PAIR_TILE && j < cntP;
                unsigned count = 0, ent_rect = 0;
                float4 ent_p = make_float4(0.f, 0.f, 0.f, 0.f), ent_c = ent_p;
                float ent_q12 = 0.f;
                const float4 *ent_fr = tframes;
                bool ent_jit = false;
                if (on) {
                    const float4 p = __ldg(A.t.s_pos_r + firstP + j);
                    const float4 *fr = tframes + (size_t)(qx & TILE_MASK) * TFRAME_V4;
                    const float4 f0 = __ldg(fr), f1 = __ldg(fr + 1), f2 = __ldg(fr + 2), f3 = __ldg(fr + 3);
                    const float4 f5 = __ldg(fr + 5), f7 = __ldg(fr + 7);
                    const float x = p.x - f0.x, y = p.y - f0.y, z = p.z - f0.z;
                    const float dz = __fmaf_rn(z, f1.z, __fmaf_rn(y, f1.y, x * f1.x));
                    const float da = __fmaf_rn(z, f2.z, __fmaf_rn(y, f2.y, x * f2.x));
                    const float db = __fmaf_rn(z, f3.z, __fmaf_rn(y, f3.y, x * f3.x));
                    const float len2 = __fmaf_rn(z, z, __fmaf_rn(y, y, x * x));
                    const float Reff = __fmaf_rn(p.w, 1.00001f, f5.z);
                    const float4 f6 = __ldg(fr + 6);
                    const unsigned dims = __float_as_uint(f6.w);
                    const int twv = (int)(dims & 255u), thv = (int)((dims >> 8) & 255u);
                    int i0 = 0, i1 = -1, j0 = 0, j1 = -1;
                    float ci = 0.f, cj = 0.f, q11 = 0.f, q12 = 0.f, q22 = 0.f;
                    if (!(dz * dz >= 0.0025f * len2 && dz > 0.f)) {
                        // more than ~87 degrees off the tile axis (or behind the apex): every tile ray is at least
                        // 61 degrees away from P', so a hit needs |P'| sin(61 deg) < r
                        if (!(len2 > 1.44f * Reff * Reff)) { i0 = 0; i1 = twv - 1; j0 = 0; j1 = thv - 1; }   // all pixels, no pretest
                    } else if (!(dz - Reff > f0.w * 1.000001f)) {        // depth <= t + r <= tFar + r
                        const float iz = __fdividef(1.f, dz);
                        const float sp = da * iz - f2.w, tp = db * iz - f3.w;
                        const float rho = __fmaf_rn(Reff * iz, f5.w, f1.w);
                        const float4 f4 = __ldg(fr + 4);
                        ci = __fmaf_rn(f4.y, tp, f4.x * sp);
                        cj = __fmaf_rn(f4.w, tp, f4.z * sp);
                        const float ei = __fmaf_rn(rho, f5.x, 1e-3f) + (JIT ? f7.z : 0.f), ej = __fmaf_rn(rho, f5.y, 1e-3f) + (JIT ? f7.w : 0.f);   // + the tile's largest jitter
                        const float lo_i = fmaxf(ceilf(ci - ei), 0.f), hi_i = fminf(floorf(ci + ei), (float)(twv - 1));
                        const float lo_j = fmaxf(ceilf(cj - ej), 0.f), hi_j = fminf(floorf(cj + ej), (float)(thv - 1));
                        if (lo_i <= hi_i && lo_j <= hi_j) { i0 = (int)lo_i; i1 = (int)hi_i; j0 = (int)lo_j; j1 = (int)hi_j; }
                        else if (!(ci == ci && cj == cj && rho == rho)) { i0 = 0; i1 = twv - 1; j0 = 0; j1 = thv - 1; }
                        const float ir2 = __fdividef(1.f, rho * rho * 1.002f);
                        q11 = f6.x * ir2; q12 = f6.y * ir2; q22 = f6.z * ir2;
                        if (!(q11 == q11 && q12 == q12 && q22 == q22) || !(ci == ci && cj == cj)) { q11 = q12 = q22 = 0.f; ci = cj = 0.f; }
                    }
                    if (i1 >= i0 && j1 >= j0) {
                        const unsigned w = (unsigned)(i1 - i0 + 1), h = (unsigned)(j1 - j0 + 1);
                        count = w * h;
                    }
                    if (COUNT) ++cTests;
                    ent_p = p; ent_c = make_float4(ci, cj, q11, q22); ent_q12 = q12;
                    ent_rect = (unsigned)i0 | ((unsigned)j0 << 8) | ((unsigned)(i1 - i0 + 1) << 16) | ((unsigned)(j1 - j0 + 1) << 24);
                    ent_fr = fr;
                    ent_jit = JIT && (f7.z > 1e-3f || f7.w > 1e-3f);
                }
                // exclusive prefix of the candidate counts, continued over the sub-passes; entries are compacted
                unsigned incl = count;
#pragma unroll
                for (int o2 = 1; o2 < 32; o2 <<= 1) {
                    const unsigned v = __shfl_up_sync(FULL, incl, o2);
                    if ((int)lane >= o2) incl += v;
                }
                const unsigned nz = __ballot_sync(FULL, count != 0u);
                if (count) {
                    const unsigned e = nEnt + __popc(nz & lt);
                    {
                        const float4 p = ent_p;
                        const float4 *fr = ent_fr;
                        const float ci = ent_c.x, cj = ent_c.y, q11 = ent_c.z, q22 = ent_c.w, q12 = ent_q12;
                        te.cs[e] = totalC + incl - count;
                        te.p[e] = p;
                        if (MODE == 0) {
                            // power / direction are not needed before the first hit is shaded: asynchronous copies
                            // (LDGSTS) hide the latency of the store-order reads behind the candidate passes
                            const float4 *gpw, *gdr;
                            if (A.t.s_pow) { gpw = A.t.s_pow + firstP + j; gdr = A.t.s_dir + firstP + j; }
                            else { const uint32_t oi = __ldg(A.t.order + firstP + j); gpw = A.t.pow_u + oi; gdr = A.t.dir_u + oi; }
                            cp_async16(&te.pw[e], gpw);
                            cp_async16(&te.dr[e], gdr);
                        }
                        te.c[e] = make_float4(ci, cj, q11, q22);
                        te.q12[e] = q12;
                        te.s[e] = firstP + j;
                        const float4 f7 = __ldg(fr + 7), f0o = __ldg(fr);
                        te.rb[e] = __float_as_uint(f7.x);
                        te.o[e] = make_float4(f0o.x, f0o.y, f0o.z, f7.y);
                        te.tl[e] = (qx & TILE_MASK) | (ent_jit ? 0x80000000u : 0u);      // bit 31: the tile's rays carry per-pixel jitter
                        te.rect[e] = ent_rect;
                    }
                }
                nEnt += __popc(nz);
                totalC += __shfl_sync(FULL, incl, 31);
            }
            if (totalC == 0u) continue;
            __syncwarp();
            // ---- stage 2: one candidate (photon, pixel) per lane
            const unsigned cs0 = lane < nEnt ? te.cs[lane] : 0u, cs1 = lane + 32u < nEnt ? te.cs[lane + 32u] : 0u;
            unsigned nBefore = 0;
            bool payloadIn = MODE != 0;
            for (unsigned cb = 0; cb < totalC; cb += 32) {
                const unsigned c = cb + lane;
                bool hit = false;
                float th = 0.f, d2 = 0.f, ct = 0.f, hdx = 0.f, hdy = 0.f, hdz = 0.f;
                uint32_t e = 0, q = 0, pp = 0;
                // entry with cs <= c < cs + count.  Every entry has candidates, so the cs are distinct: the entries that START
                // inside this window of 32 candidates set one bit each (a warp OR-reduction), and candidate lane c
                // belongs to entry (entries started before the window) - 1 + (heads at or below lane c)
                unsigned h = 0;
                {
                    const unsigned d0 = cs0 - cb, d1 = cs1 - cb;
                    if (lane < nEnt && d0 < 32u) h = 1u << d0;
                    if (lane + 32u < nEnt && d1 < 32u) h |= 1u << d1;
                }
                const unsigned heads = __reduce_or_sync(FULL, h);
                const unsigned eSel = nBefore - 1u + __popc(heads & (lt | (1u << lane)));
                nBefore += __popc(heads);
                if (c < totalC) {
                    e = eSel;
                    const unsigned k = c - te.cs[e];
                    const unsigned rect = te.rect[e];
                    const unsigned w = (rect >> 16) & 255u;
                    const unsigned jj = (k * ((65536u + w - 1u) / w)) >> 16;   // k / w for k < 256, w <= 16
                    const unsigned ii = k - jj * w;
                    const unsigned pi = (rect & 255u) + ii, pj = ((rect >> 8) & 255u) + jj;
                    const float4 cc = te.c[e];
                    // (without jitter the tile index is read by the survivors of the pretest only: the kernel is bound by
                    // shared-memory / L1 wavefronts)
                    unsigned tlv = JIT ? te.tl[e] : 0u;
                    float di = (float)pi - cc.x, dj = (float)pj - cc.y;
                    if (JIT && (tlv & 0x80000000u)) {      // where this pixel's ray really is in the pixel model
                        const float2 jt = __ldg(tjit + (size_t)(tlv & TILE_MASK) * 256u + pj * 16u + pi);
                        di += jt.x; dj += jt.y;
                    }
                    const float qf = __fmaf_rn(cc.z * di, di, __fmaf_rn(2.f * te.q12[e] * di, dj, cc.w * dj * dj));
                    if (!(qf > 1.f)) {
                        if (!JIT) tlv = te.tl[e];
                        q = te.rb[e] + pj * imgW + pi;
                        const float4 p = te.p[e];
                        const float4 d = __ldg(trays + (size_t)(tlv & TILE_MASK) * 256u + pj * 16u + pi);     // {d.xyz, tMax}
                        float4 o = te.o[e];
                        if (o.w == 0.f) o = __ldg(A.ro + q);
                        RayS r;
                        r.ox = o.x; r.oy = o.y; r.oz = o.z; r.tMax = d.w;
                        r.dx = d.x; r.dy = d.y; r.dz = d.z;
                        hit = disk_test(r, p, th, d2);
                        if (COUNT) ++cTests;
                        if (MODE == 0 && hit) { hdx = d.x; hdy = d.y; hdz = d.z; pp = pj * 16u + pi; }
                    }
                }
                if (MODE != 0) {
                    if (hit) record_hit(A, MODE, te.s[e], q);
                    continue;
                }
                const unsigned mh = __ballot_sync(FULL, hit);
                if (mh == 0u) continue;
                // the first hit of the pass needs the photon directions: their asynchronous copies must have landed
                if (!payloadIn) { cp_async_wait_all(); __syncwarp(); payloadIn = true; }
                if (hit) {
                    const float4 w = te.dr[e];
                    ct = -__fmaf_rn(hdz, w.z, __fmaf_rn(hdy, w.y, hdx * w.x));   // Dot(wo = -d, wi)
                    const unsigned pos = nbuf + __popc(mh & lt);
                    hb.e[pos] = e | (pp << 8); hb.q[pos] = q; hb.th[pos] = th; hb.d2[pos] = d2; hb.ct[pos] = ct;
                }
                nbuf += __popc(mh);
                __syncwarp();
                if (nbuf >= 32) {
                    shade(nbuf - 32, 32u);
                    nbuf -= 32;
                    __syncwarp();
                }
            }
            if (MODE == 0 && nbuf) {     // the entries are overwritten by the next pass: shade what is left
                shade(0u, nbuf);
                nbuf = 0;
            }
            __syncwarp();
        }
    }
    if (COUNT) {
#pragma unroll
        for (int o2 = 16; o2 > 0; o2 >>= 1) cTests += __shfl_xor_sync(FULL, cTests, o2);
        if (lane == 0 && cTests) atomicAdd(counters + 4, cTests);
    }
}

}  // namespace bre
