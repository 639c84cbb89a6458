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
#include <type_traits>

namespace bre {

// rays per packet: 8 / 16 / 32 measured on smoke10m_1080p -> K5a 1.40 / 1.39 / 1.65 ms (coherent),
// 3.45 / 4.78 / 6.35 ms (per-ray walk): more rays per packet means more distinct ray rows per
// shared-memory load (bank conflicts instead of broadcasts) and heavier tail packets
#ifndef BRE_WQ_P
#define BRE_WQ_P 8
#endif
#ifndef BRE_RED_V4
#define BRE_RED_V4 0      // 1: shade_hit adds a hit with one vector atomic (RED.128) instead of three scalar ones
#endif
#ifndef BRE_WQ_NCAP
#define BRE_WQ_NCAP 512
#endif
constexpr int WQ_P = BRE_WQ_P;   // rays per packet (lane s < WQ_P prepares ray s)
constexpr int FAT_P = 8;         // rays per coherent group ("fat ray"): WQ_P / FAT_P groups per packet
constexpr int WQ_NCAP = BRE_WQ_NCAP;   // node stack entries per warp
constexpr int WQ_WARPS = 4;      // warps per block
constexpr int WQ_RAYW = 12;      // floats per ray-table row (slab-test data only)
#ifndef BRE_WQ_CHUNK
#define BRE_WQ_CHUNK 256
#endif
constexpr unsigned WQ_CHUNK = BRE_WQ_CHUNK;  // pair-queue slots reserved per atomicAdd (>= 64: a step emits up to 64 pairs);
                                             // 64 / 128 / 256 / 512 -> K5a 1.50 / 1.36 / 1.31 / 1.30, K5b 2.02 / 1.96 / 1.95 / 1.97 ms
constexpr unsigned PAIR_NONE = 0xffffffffu;  // padding pair (ray index) in a partly used chunk
constexpr unsigned PAIR_FAT = 0x80000000u;   // pair.x bit: the pair stands for the FAT_P rays of a coherent group
constexpr unsigned HEAVY_STEPS = BRE_WIDE ? 256 : 512;   // a per-ray packet this long is claimed first by the next launch
#ifndef BRE_HEAVY_FAT
#define BRE_HEAVY_FAT (BRE_WIDE ? 128 : 256)
#endif
constexpr unsigned HEAVY_STEPS_FAT = BRE_HEAVY_FAT;   // same for a coherent packet (measured: 256 / 512 -> K5a 1.31 / 1.40 ms
                                            // coherent, 3.78 / 3.45 ms per-ray)
constexpr unsigned HEAVY_CAP = 65536;   // heavy packets remembered per launch

// queueCtl words (zeroed before every launch)
enum { QC_PAIR_TAIL = 0, QC_PACKET = 1, QC_INPLACE = 2, QC_EMITTED = 3, QC_KINDS = 4, QC_RUN_THIN = 5, QC_RUN_FAT = 6,
       QC_RUN_TILE = 7, QC_NINCOH = 8, QC_WQ_HEAD = 9, QC_WQ_TAIL = 10, QC_WQ_INFLIGHT = 11, QC_TILE = 12, QC_WORDS = 16 };
enum { KIND_THIN = 1u, KIND_FAT = 2u, KIND_TILE = 4u, KIND_TILE_JIT = 8u };   // QC_KINDS bits: which kinds of pair K5a put into the queue

struct __align__(16) WarpSmem {
    int nref[WQ_NCAP];
    float ray[WQ_P][WQ_RAYW];
    float fat[WQ_P / FAT_P][24]; // FatRay of each coherent group
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

// Coherent packets ("fat ray").  When the WQ_P rays of a packet leave (nearly) one point in (nearly)
// one direction -- camera rays of neighbouring pixels -- the walk is done ONCE for the packet with an
// interval version of the slab test and emits (packet, leaf) pairs; K5b then runs the exact disk
// predicate for every ray of the packet, so hit sets are unchanged (the walk is only a cull).
// Per axis, with all inv_i of one sign and x the plane offset from the padded common origin,
//   near_i = x_i * inv_i >= min(x * invLo, x * invHi),   far_i <= max(x' * invLo, x' * invHi)
// because the pad (the per-ray pad plus the spread of the origins) moves x below every x_i (above,
// for the far plane).  If any ray passes its own slab test then every near_k <= far_l of that ray,
// hence of these bounds: the fat test never culls a box a member ray would enter.
struct FatRay {
    float ax, ay, az;        // smallest 1/d per axis over the packet
    float bx, by, bz;        // largest
    float nax, nay, naz, nbx, nby, nbz;   // near constants for a / b
    float fax, fay, faz, fbx, fby, fbz;   // far constants
    float tMax;              // largest tMax of the packet
    int sgn;
};
__device__ __forceinline__ void fat_store(float *row, const FatRay &r) {
    float4 *v = reinterpret_cast<float4 *>(row);
    v[0] = make_float4(r.ax, r.ay, r.az, r.tMax);
    v[1] = make_float4(r.bx, r.by, r.bz, __int_as_float(r.sgn));
    v[2] = make_float4(r.nax, r.nay, r.naz, r.nbx);
    v[3] = make_float4(r.nby, r.nbz, r.fax, r.fay);
    v[4] = make_float4(r.faz, r.fbx, r.fby, r.fbz);
}
__device__ __forceinline__ FatRay fat_load(const float *row) {
    const float4 *v = reinterpret_cast<const float4 *>(row);
    const float4 a = v[0], b = v[1], c = v[2], d = v[3], e = v[4];
    FatRay r;
    r.ax = a.x; r.ay = a.y; r.az = a.z; r.tMax = a.w;
    r.bx = b.x; r.by = b.y; r.bz = b.z; r.sgn = __float_as_int(b.w);
    r.nax = c.x; r.nay = c.y; r.naz = c.z; r.nbx = c.w;
    r.nby = d.x; r.nbz = d.y; r.fax = d.z; r.fay = d.w;
    r.faz = e.x; r.fbx = e.y; r.fby = e.z; r.fbz = e.w;
    return r;
}
__device__ __forceinline__ bool slab_hit_fat(const FatRay &r, float lox, float loy, float loz, float hix,
                                             float hiy, float hiz) {
    const bool sx = r.sgn & 1, sy = r.sgn & 2, sz = r.sgn & 4;
    const float nx = sx ? hix : lox, fx = sx ? lox : hix;
    const float ny = sy ? hiy : loy, fy = sy ? loy : hiy;
    const float nz = sz ? hiz : loz, fz = sz ? loz : hiz;
    const float tx0 = fminf(__fmaf_rn(nx, r.ax, r.nax), __fmaf_rn(nx, r.bx, r.nbx));
    const float ty0 = fminf(__fmaf_rn(ny, r.ay, r.nay), __fmaf_rn(ny, r.by, r.nby));
    const float tz0 = fminf(__fmaf_rn(nz, r.az, r.naz), __fmaf_rn(nz, r.bz, r.nbz));
    const float tx1 = fmaxf(__fmaf_rn(fx, r.ax, r.fax), __fmaf_rn(fx, r.bx, r.fbx));
    const float ty1 = fmaxf(__fmaf_rn(fy, r.ay, r.fay), __fmaf_rn(fy, r.by, r.fby));
    const float tz1 = fmaxf(__fmaf_rn(fz, r.az, r.faz), __fmaf_rn(fz, r.bz, r.fbz));
    const float tMin = fmaxf(fmaxf(tx0, ty0), tz0);
    const float tMax = fmul(fminf(fminf(tx1, ty1), tz1), BRE_SLAB_WIDEN);
    return tMin <= tMax && tMin < r.tMax && tMax > 0.f;
}
// min / max over the FAT_P lanes of a group (lane = ray of the packet)
__device__ __forceinline__ float pk_min(float v) {
#pragma unroll
    for (int o = 1; o < FAT_P; o <<= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float pk_max(float v) {
#pragma unroll
    for (int o = 1; o < FAT_P; o <<= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Photon pretest of a coherent packet (k_leaf_fat).  K5a fits an orthonormal frame (u, v, n) to the
// packet -- u the bisector of its first and last direction, v the direction the rays fan out in, n
// their common normal -- and stores, in this frame, the range [mLo, mHi] of the slopes m_i =
// (d_i.v)/(d_i.u), 1/kappa with kappa = min_i (d_i.u)/|(d_i.u, d_i.v)|, and hT = max_i|d_i.n| * tau
// + slack (tau bounds the ray parameter of any hit).  For P' = P - o and ANY orthonormal frame
//   dist(P, ray_i) >= |P'.n| - |d_i.n| tau_i          and   dist(P, ray_i) >= |P'.v - m_i P'.u| kappa_i,
// so a photon with |P'.n| > r + hT, or with P'.v outside [min(mLo t, mHi t) - R, max(..) + R],
// R = r / kappa + hT, t = P'.u, is missed by every ray of the packet.  The slack (4 x origin spread
// + 2^-16 x coordinate scale) covers the origins' differences and every rounding error of this test
// and of the exact predicate by more than 10x.  Only photons that pass are tested exactly, per ray.
constexpr int FRAME_V4 = 5;      // float4 per packet frame: a, b, c, d below + the slope ranges of the two half packets
struct PacketFrame {
    float4 a;   // o.xyz, hT
    float4 b;   // u.xyz, 1/kappa
    float4 c;   // v.xyz, mLo
    float4 d;   // n.xyz, mHi
};
__device__ __forceinline__ bool frame_may_hit(const float4 fa, const float4 fb, const float4 fc, const float4 fd,
                                              const float4 p) {
    const float x = p.x - fa.x, y = p.y - fa.y, z = p.z - fa.z;
    const float h = __fmaf_rn(z, fd.z, __fmaf_rn(y, fd.y, x * fd.x));
    const float t = __fmaf_rn(z, fb.z, __fmaf_rn(y, fb.y, x * fb.x));
    const float s = __fmaf_rn(z, fc.z, __fmaf_rn(y, fc.y, x * fc.x));
    const float R = __fmaf_rn(p.w, fb.w, fa.w);
    const float m0 = t * fc.w, m1 = t * fd.w;
    // written so that a NaN anywhere keeps the photon
    return !(fabsf(h) > p.w + fa.w) && !(s > fmaxf(m0, m1) + R) && !(s < fminf(m0, m1) - R);
}

// Same test, then per half packet (rays 0-3: slopes [fe.x, fe.y], rays 4-7: [fe.z, fe.w]): bit 0 / bit 1 set
// when the photon may reach that half.  0 = no ray of the packet can hit.  NaNs keep both halves.
__device__ __forceinline__ unsigned frame_halves(const float4 fa, const float4 fb, const float4 fc, const float4 fd,
                                                 const float4 fe, const float4 p) {
    const float x = p.x - fa.x, y = p.y - fa.y, z = p.z - fa.z;
    const float h = __fmaf_rn(z, fd.z, __fmaf_rn(y, fd.y, x * fd.x));
    const float t = __fmaf_rn(z, fb.z, __fmaf_rn(y, fb.y, x * fb.x));
    const float s = __fmaf_rn(z, fc.z, __fmaf_rn(y, fc.y, x * fc.x));
    const float R = __fmaf_rn(p.w, fb.w, fa.w);
    if (fabsf(h) > p.w + fa.w) return 0u;
    const float a0 = t * fe.x, a1 = t * fe.y, b0 = t * fe.z, b1 = t * fe.w;
    const bool inA = !(s > fmaxf(a0, a1) + R) && !(s < fminf(a0, a1) - R);
    const bool inB = !(s > fmaxf(b0, b1) + R) && !(s < fminf(b0, b1) - R);
    return (inA ? 1u : 0u) | (inB ? 2u : 0u);
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
    uint32_t tauW, tauTilesX;    // tauW != 0: the table is tile-blocked [tile][step][256 px] for an image tauW wide (k_tau_tile)
    uint32_t rayBase;            // index of ro[0] in the whole launch (ray batches): the tile-blocked table wants absolute pixels
    uint32_t *hitCounts;         // MODE 1: per-ray hit counts; MODE 2: per-ray write cursors
    const uint64_t *hitOffsets;  // MODE 2
    uint32_t *hitIdx;            // MODE 2
    float4 *frames;              // FRAME_V4 float4 per packet: PacketFrame of coherent packets (K5a -> k_leaf_fat)
};

// Optical depth of a grid medium at ray parameter th: linear interpolation in the ray's table.
// base/stride address the table entries of this ray (ray layout: stride 1; tile-blocked: stride 256).
__device__ __forceinline__ float optical_depth_at(const LeafArgs &A, const float2 sp, const float *base, size_t stride,
                                                  float th) {
    float od = 0.f;
    if (sp.y > sp.x) {
        const float sc = (th - sp.x) * __fdividef((float)A.tauSteps, sp.y - sp.x);
        if (sc > 0.f) {
            if (sc >= (float)A.tauSteps) od = __ldg(base + (size_t)A.tauSteps * stride);
            else {
                const int k = (int)sc;
                const float f = sc - (float)k;
                const float a = __ldg(base + (size_t)k * stride), b = __ldg(base + (size_t)(k + 1) * stride);
                od = __fmaf_rn(f, b - a, a);
            }
        }
    }
    return od;
}
__device__ __forceinline__ float optical_depth(const LeafArgs &A, uint32_t q, float th) {
    const float2 sp = __ldg(A.span + q);
    if (A.tauW) {
        const uint32_t qa = A.rayBase + q, x = qa % A.tauW, y = qa / A.tauW;
        const size_t tile = (size_t)(y >> 4) * A.tauTilesX + (x >> 4);
        return optical_depth_at(A, sp, A.tau + tile * (size_t)(A.tauSteps + 1) * 256u + ((y & 15u) * 16u + (x & 15u)), 256, th);
    }
    return optical_depth_at(A, sp, A.tau + (size_t)q * (A.tauSteps + 1), 1, th);
}

// Radiance of one hit (photon s, ray q, ray parameter th, squared distance d2).
// Only the hit DECISION has to be bit-exact; the weight uses the fast reciprocal / exp units
// (relative error ~1e-6, tolerance is 1e-4 mean / 1e-3 max).
__device__ __forceinline__ float3 hit_radiance(const LeafArgs &A, uint32_t s, uint32_t q, float th, float d2) {
    const float radius = __ldg(&A.t.s_pos_r[s].w);
    float4 w, pw;
    photon_payload(A.t, s, pw, w);
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
    if (A.m.kind == 1) od = optical_depth(A, q, th);
    float t0, t1, t2;
    t0 = __expf(-A.m.sigma_t[0] * od);
    if (A.m.gray) { t1 = t0; t2 = t0; }
    else { t1 = __expf(-A.m.sigma_t[1] * od); t2 = __expf(-A.m.sigma_t[2] * od); }
    return make_float3(wp * t0 * pw.x, wp * t1 * pw.y, wp * t2 * pw.z);
}
__device__ __forceinline__ void shade_hit(const LeafArgs &A, uint32_t s, uint32_t q, float th, float d2) {
    const float3 v = hit_radiance(A, s, q, th, d2);
#if BRE_RED_V4
    atomicAdd(A.out + q, make_float4(v.x, v.y, v.z, 0.f));   // one RED.128 instead of three REDs
#else
    float *dst = reinterpret_cast<float *>(A.out + q);
    atomicAdd(dst + 0, v.x);
    atomicAdd(dst + 1, v.y);
    atomicAdd(dst + 2, v.z);
#endif
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
__device__ __noinline__ void process_pair_inplace(const LeafArgs *Ag, uint32_t px, int ref) {
    const LeafArgs &A = *Ag;
    const uint32_t q0 = px & ~PAIR_FAT, nq = (px & PAIR_FAT) ? (uint32_t)FAT_P : 1u;
    const uint32_t first = (uint32_t)ref & 0x07ffffffu;
    const uint32_t cnt = (((uint32_t)ref >> 27) & 15u) + 1u;
    for (uint32_t q = q0; q < q0 + nq; ++q) {
        const float4 o = __ldg(A.ro + q), d = __ldg(A.rd + q);
        RayS r;
        r.ox = o.x; r.oy = o.y; r.oz = o.z; r.tMax = o.w;
        r.dx = d.x; r.dy = d.y; r.dz = d.z;
        for (uint32_t j = 0; j < cnt; ++j) {
            float th, d2;
            if (disk_test(r, __ldg(A.t.s_pos_r + first + j), th, d2)) {
                if (MODE == 0) shade_hit(A, first + j, q, th, d2);
                else record_hit(A, MODE, first + j, q);
            }
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
// WQ_MINB (make NVEXTRA=-DWQ_MINB=n) forces n blocks per SM: 10 / 12 (48 / 40 registers) spill and measured
// 1.79 / 2.19 ms against 1.31 ms for the compiler's own 56 registers (9 blocks of 4 warps per SM)
template <int MODE, bool COUNT>
#ifdef WQ_MINB
__global__ void __launch_bounds__(WQ_WARPS * 32, WQ_MINB) k_traverse(
#else
__global__ void __launch_bounds__(WQ_WARPS * 32) k_traverse(
#endif
    LeafArgs A, const LeafArgs *__restrict__ Ag, uint32_t n, uint2 *__restrict__ queue,
    unsigned int queueCap, unsigned int *__restrict__ queueCtl,
    const unsigned int *__restrict__ heavyIn, const unsigned int *__restrict__ nHeavyInPtr,
    const unsigned int *__restrict__ heavyFlagIn, unsigned int *__restrict__ heavyOut,
    unsigned int *__restrict__ heavyFlagOut, unsigned int *__restrict__ nHeavyOut,
    unsigned int heavyCap, float fatSpread,
    const unsigned char *__restrict__ tileState, uint32_t imgW, uint32_t tilesX, uint32_t rayOffset,
    unsigned long long *__restrict__ counters) {
    __shared__ WarpSmem smemAll[WQ_WARPS];
    // tile mode (bre_query_tile.cuh): only the rays of tiles k_traverse_tile flagged incoherent are walked here
    if (tileState && ld_volatile_u32(queueCtl + QC_NINCOH) == 0u) return;
    WarpSmem &sm = smemAll[threadIdx.x >> 5];
    const TreeDev &t = A.t;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt = (1u << lane) - 1u;
    const unsigned FULL = 0xffffffffu;
    const unsigned nPackets = (n + WQ_P - 1) / WQ_P;
    const unsigned nHeavyIn = min(__ldg(nHeavyInPtr), HEAVY_CAP);
    unsigned long long cNodes = 0;
    unsigned nEmitted = 0, nInplace = 0, kinds = 0;
    unsigned chunkPos = 0, chunkEnd = 0;   // this warp's reserved queue slots [chunkPos, chunkEnd)

    for (;;) {
        // ---- claim a packet: last launch's heavy packets first, then the rest in order
        // (Other claim orders were measured and lost, K5a ms against 1.31: 2 / 4 / 8 consecutive packets per
        // claim 1.35 / 1.54 / 1.84; the image walked in vertical strips of 8..120 packets 1.38-1.43; the
        // next claim issued one packet ahead 1.39.  Row order with single-packet claims keeps the
        // packets in flight closest together in the tree.)
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

        // ---- packet setup: lane s < np prepares ray base+s.  A full packet whose groups of FAT_P
        // rays (one group with the default WQ_P = 8) are each coherent or dead is walked with one
        // fat ray per group; any other packet with one stack entry stream per ray.
        bool live = false;
        float4 o = make_float4(0.f, 0.f, 0.f, 0.f), d = make_float4(0.f, 0.f, 1.f, 0.f);
        RayF r;
        r.sgn = 0; r.tMax = 0.f; r.ix = r.iy = r.iz = 1.f;
        if (lane < np) {
            o = __ldg(A.ro + base + lane);
            d = __ldg(A.rd + base + lane);
            r = rayf_setup(o, d, t.abs_max);
            rayf_store(sm.ray[lane], r);
            live = t.n > 0 && r.tMax > 0.f;
            if (tileState) {
                const uint32_t qa = rayOffset + base + lane;
                live = live && tileState[(qa / imgW / 16u) * tilesX + (qa % imgW) / 16u] != 0;
            }
        }
        if (tileState && !__any_sync(FULL, live)) continue;
        bool fat = false;
        bool grpLive = false;      // my group takes part in the fat walk
        if (np == WQ_P && fatSpread > 0.f && t.n > 0) {
            const unsigned g0 = lane & ~(unsigned)(FAT_P - 1), gl = g0 + FAT_P - 1;   // first / last lane of my group
            const float ox0 = __shfl_sync(FULL, o.x, g0), oy0 = __shfl_sync(FULL, o.y, g0), oz0 = __shfl_sync(FULL, o.z, g0);
            const float dx0 = __shfl_sync(FULL, d.x, g0), dy0 = __shfl_sync(FULL, d.y, g0), dz0 = __shfl_sync(FULL, d.z, g0);
            const int sgn0 = __shfl_sync(FULL, r.sgn, g0);
            const float oSpread = pk_max(fmaxf(fmaxf(fabsf(o.x - ox0), fabsf(o.y - oy0)), fabsf(o.z - oz0)));
            const float dSpread = pk_max(fmaxf(fmaxf(fabsf(d.x - dx0), fabsf(d.y - dy0)), fabsf(d.z - dz0)));
            const float M = pk_max(fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z))) + t.abs_max;
            const float tMaxAll = pk_max(r.tMax);
            const unsigned sameSgn = __ballot_sync(FULL, r.sgn == sgn0);
            const unsigned grpBits = ((1u << FAT_P) - 1u) << g0;
            // the group's frame for k_leaf_fat (not bit-critical: any orthonormal frame is valid)
            const float d7x = __shfl_sync(FULL, d.x, gl), d7y = __shfl_sync(FULL, d.y, gl), d7z = __shfl_sync(FULL, d.z, gl);
            float ux = dx0 + d7x, uy = dy0 + d7y, uz = dz0 + d7z;
            const float ul = rsqrtf(ux * ux + uy * uy + uz * uz);
            ux *= ul; uy *= ul; uz *= ul;
            float wx = d7x - dx0, wy = d7y - dy0, wz = d7z - dz0;
            const float wu = wx * ux + wy * uy + wz * uz;
            wx -= wu * ux; wy -= wu * uy; wz -= wu * uz;
            float wl = wx * wx + wy * wy + wz * wz;
            if (!(wl > 1e-24f)) {        // all directions equal: any perpendicular will do
                const float ax = fabsf(ux), ay = fabsf(uy), az = fabsf(uz);
                if (ax <= ay && ax <= az) { wx = 0.f; wy = -uz; wz = uy; }
                else if (ay <= az) { wx = uz; wy = 0.f; wz = -ux; }
                else { wx = -uy; wy = ux; wz = 0.f; }
                wl = wx * wx + wy * wy + wz * wz;
            }
            const float wi = rsqrtf(wl);
            const float vx = wx * wi, vy = wy * wi, vz = wz * wi;
            const float nx2 = uy * vz - uz * vy, ny2 = uz * vx - ux * vz, nz2 = ux * vy - uy * vx;
            const float da = d.x * ux + d.y * uy + d.z * uz, db = d.x * vx + d.y * vy + d.z * vz,
                        dc = d.x * nx2 + d.y * ny2 + d.z * nz2;
            const float m = __fdividef(db, da);
            const float kap = da * rsqrtf(da * da + db * db);
            const float aMin = pk_min(da), kMin = pk_min(kap), cMax = pk_max(fabsf(dc));
            // slope range of the packet and of its two halves (rays 0-3, 4-7): k_leaf_fat runs the exact
            // predicate only for the half packet(s) a photon can reach.  The butterfly's first two steps
            // are the half-group reductions, its last step joins the halves.
            static_assert(FAT_P == 8, "half-packet slope ranges assume 8 rays per group");
            float hLo = m, hHi = m;
#pragma unroll
            for (int o2 = 1; o2 < FAT_P / 2; o2 <<= 1) {
                hLo = fminf(hLo, __shfl_xor_sync(FULL, hLo, o2));
                hHi = fmaxf(hHi, __shfl_xor_sync(FULL, hHi, o2));
            }
            const float loB = __shfl_sync(FULL, hLo, g0 + FAT_P / 2), hiB = __shfl_sync(FULL, hHi, g0 + FAT_P / 2);
            const float mLo = fminf(hLo, __shfl_xor_sync(FULL, hLo, FAT_P / 2));
            const float mHi = fmaxf(hHi, __shfl_xor_sync(FULL, hHi, FAT_P / 2));
            // NaNs fail the comparisons and fall back to the per-ray walk
            const bool dead = !(tMaxAll > 0.f) && tMaxAll == tMaxAll;
            const bool coh = (sameSgn & grpBits) == grpBits && dSpread <= fatSpread && oSpread <= M * 0x1p-10f &&
                             aMin > 0.5f && tMaxAll > 0.f;
            fat = __all_sync(FULL, dead || coh);
            grpLive = fat && coh;
            // (group reductions use full-mask shuffles: every lane must execute them)
            FatRay f;
            f.ax = pk_min(r.ix); f.ay = pk_min(r.iy); f.az = pk_min(r.iz);
            f.bx = pk_max(r.ix); f.by = pk_max(r.iy); f.bz = pk_max(r.iz);
            if (grpLive) {
                const float pad = 2.f * (M * 0x1p-18f) + 2.f * oSpread;
                const float px = (sgn0 & 1) ? -pad : pad, py = (sgn0 & 2) ? -pad : pad, pz = (sgn0 & 4) ? -pad : pad;
                const float nx = fadd(ox0, px), ny = fadd(oy0, py), nz = fadd(oz0, pz);
                const float fx = fsub(ox0, px), fy = fsub(oy0, py), fz = fsub(oz0, pz);
                f.nax = -fmul(nx, f.ax); f.nay = -fmul(ny, f.ay); f.naz = -fmul(nz, f.az);
                f.nbx = -fmul(nx, f.bx); f.nby = -fmul(ny, f.by); f.nbz = -fmul(nz, f.bz);
                f.fax = -fmul(fx, f.ax); f.fay = -fmul(fy, f.ay); f.faz = -fmul(fz, f.az);
                f.fbx = -fmul(fx, f.bx); f.fby = -fmul(fy, f.by); f.fbz = -fmul(fz, f.bz);
                f.tMax = tMaxAll; f.sgn = sgn0;
                if (lane == g0) {
                    fat_store(sm.fat[lane / FAT_P], f);
                    const float tau = fminf(tMaxAll, 4.f * M);
                    const float hT = 1.01f * cMax * tau + 4.f * oSpread + M * 0x1p-16f;
                    float4 *fr = A.frames + ((size_t)pk * (WQ_P / FAT_P) + lane / FAT_P) * FRAME_V4;
                    fr[0] = make_float4(ox0, oy0, oz0, hT);
                    fr[1] = make_float4(ux, uy, uz, 1.001f * __fdividef(1.f, kMin));
                    fr[2] = make_float4(vx, vy, vz, mLo);
                    fr[3] = make_float4(nx2, ny2, nz2, mHi);
                    fr[4] = make_float4(hLo, hHi, loB, hiB);     // (lane g0 holds the first half's range)
                }
            }
        }
#if BRE_WIDE
        int nsize = 0;
        unsigned eMask = 0;            // bit k: child k of my node is a leaf to emit this step
        int c[4] = {0, 0, 0, 0};
        int slot = (int)lane;
        {
            // fat walk: one root entry per live group (slot = group); per-ray walk: one per live ray
            const bool push = fat ? (grpLive && (lane & (FAT_P - 1)) == 0) : live;
            if (fat) slot = (int)(lane / FAT_P);
            const unsigned mpush = __ballot_sync(FULL, push);
            if (t.root_ref >= 0) {
                if (push) { int pos = __popc(mpush & lt); sm.nref[pos] = t.root_ref; sm.nslot[pos] = (unsigned char)slot; }
                nsize = __popc(mpush);
            } else {            // the whole tree is one leaf
                eMask = push ? 1u : 0u; c[0] = t.root_ref;
            }
        }
        __syncwarp();

        auto walk = [&](auto fatTag) {
        constexpr bool FAT = decltype(fatTag)::value;
        for (;;) {
            // ---- emit surviving leaves as (ray, leaf) pairs
            unsigned ml[4], nl = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) { ml[k] = __ballot_sync(FULL, (eMask >> k) & 1u); nl += __popc(ml[k]); }
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
                const unsigned px = FAT ? ((base + FAT_P * slot) | PAIR_FAT) : base + slot;
                kinds |= FAT ? KIND_FAT : KIND_THIN;
                unsigned off = chunkPos;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if ((eMask >> k) & 1u) {
                        const unsigned p = off + __popc(ml[k] & lt);
                        if (p < queueCap) queue[p] = make_uint2(px, (unsigned)c[k]);
                        else { process_pair_inplace<MODE>(Ag, px, c[k]); ++nInplace; }
                        if (COUNT) ++nEmitted;
                    }
                    off += __popc(ml[k]);
                }
                chunkPos += nl;
            }
            if (nsize == 0) break;
            // ---- node step: one 128-byte four-child node per lane.  Popping k entries can push at
            // most 4k, so 3k <= room keeps the stack inside WQ_NCAP-192; with no room left the walk
            // degrades to depth-first (k = 1), which needs at most 3 x wide-tree depth (< 64) more.
            const int room = (WQ_NCAP - 192 - nsize) / 3;
            const int k = room >= 1 ? min(min(32, nsize), room) : 1;
            const bool act = (int)lane < k;
            int ref = 0;
            if (act) { ref = sm.nref[nsize - 1 - lane]; slot = sm.nslot[nsize - 1 - lane]; }
            nsize -= k;
            ++steps;
            unsigned hit = 0;
            if (act) {
                const FastNode *fn = t.fast + ref;
                const float4 q0 = __ldg(&fn->q[0]), q1 = __ldg(&fn->q[1]), q2 = __ldg(&fn->q[2]);
                const float4 q3 = __ldg(&fn->q[3]), q4 = __ldg(&fn->q[4]), q5 = __ldg(&fn->q[5]);
                const int4 r4 = __ldg(&fn->ref);
                c[0] = r4.x; c[1] = r4.y; c[2] = r4.z; c[3] = r4.w;
                if (FAT) {
                    const FatRay f = fat_load(sm.fat[slot]);
                    hit = (slab_hit_fat(f, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y) ? 1u : 0u) |
                          (slab_hit_fat(f, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w) ? 2u : 0u) |
                          (slab_hit_fat(f, q3.x, q3.y, q3.z, q3.w, q4.x, q4.y) ? 4u : 0u) |
                          (slab_hit_fat(f, q4.z, q4.w, q5.x, q5.y, q5.z, q5.w) ? 8u : 0u);
                } else {
                    const RayF r = rayf_load(sm.ray[slot]);
                    hit = (slab_hit_f(r, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y) ? 1u : 0u) |
                          (slab_hit_f(r, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w) ? 2u : 0u) |
                          (slab_hit_f(r, q3.x, q3.y, q3.z, q3.w, q4.x, q4.y) ? 4u : 0u) |
                          (slab_hit_f(r, q4.z, q4.w, q5.x, q5.y, q5.z, q5.w) ? 8u : 0u);
                }
                hit &= (c[0] != FAST_NONE ? 1u : 0u) | (c[1] != FAST_NONE ? 2u : 0u) |
                       (c[2] != FAST_NONE ? 4u : 0u) | (c[3] != FAST_NONE ? 8u : 0u);
                if (COUNT) ++cNodes;
            }
            __syncwarp();   // all pops are done before anyone pushes
            const unsigned leafBits = (c[0] < 0 ? 1u : 0u) | (c[1] < 0 ? 2u : 0u) | (c[2] < 0 ? 4u : 0u) | (c[3] < 0 ? 8u : 0u);
            eMask = hit & leafBits;
            const unsigned nMask = hit & ~leafBits;
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                const bool on = (nMask >> kk) & 1u;
                const unsigned mn = __ballot_sync(FULL, on);
                if (on) { int p = nsize + __popc(mn & lt); sm.nref[p] = c[kk]; sm.nslot[p] = (unsigned char)slot; }
                nsize += __popc(mn);
            }
            __syncwarp();
        }
        };
#else
        int nsize = 0;
        bool e0 = false, e1 = false;   // leaf children to emit this step
        int c0 = 0, c1 = 0, slot = (int)lane;
        {
            // fat walk: one root entry per live group (slot = group); per-ray walk: one per live ray
            const bool push = fat ? (grpLive && (lane & (FAT_P - 1)) == 0) : live;
            if (fat) slot = (int)(lane / FAT_P);
            const unsigned mpush = __ballot_sync(FULL, push);
            if (t.root_ref >= 0) {
                if (push) { int pos = __popc(mpush & lt); sm.nref[pos] = t.root_ref; sm.nslot[pos] = (unsigned char)slot; }
                nsize = __popc(mpush);
            } else {            // the whole tree is one leaf
                e0 = push; c0 = t.root_ref;
            }
        }
        __syncwarp();

        auto walk = [&](auto fatTag) {
        constexpr bool FAT = decltype(fatTag)::value;
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
                const unsigned px = FAT ? ((base + FAT_P * slot) | PAIR_FAT) : base + slot;
                kinds |= FAT ? KIND_FAT : KIND_THIN;
                if (e0) {
                    if (p0 < queueCap) queue[p0] = make_uint2(px, (unsigned)c0);
                    else { process_pair_inplace<MODE>(Ag, px, c0); ++nInplace; }
                }
                if (e1) {
                    if (p1 < queueCap) queue[p1] = make_uint2(px, (unsigned)c1);
                    else { process_pair_inplace<MODE>(Ag, px, c1); ++nInplace; }
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
                if (FAT) {
                    const FatRay f = fat_load(sm.fat[slot]);
                    h0 = slab_hit_fat(f, a.x, a.y, a.z, a.w, b.x, b.y);
                    h1 = slab_hit_fat(f, b.z, b.w, c.x, c.y, c.z, c.w);
                } else {
                    const RayF r = rayf_load(sm.ray[slot]);
                    h0 = slab_hit_f(r, a.x, a.y, a.z, a.w, b.x, b.y);
                    h1 = slab_hit_f(r, b.z, b.w, c.x, c.y, c.z, c.w);
                }
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
        };
#endif
        if (fat) walk(std::true_type{}); else walk(std::false_type{});
        // ---- remember heavy packets for the next launch
        if (steps >= (fat ? HEAVY_STEPS_FAT : HEAVY_STEPS) && lane == 0) {
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
    if (lane == 0 && kinds) atomicOr(queueCtl + QC_KINDS, kinds);
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
#ifndef LP_DYNAMIC
#define LP_DYNAMIC 1
#endif
struct __align__(16) HitBuf {
    uint32_t s[64], q[64];
    float th[64], d2[64];
};

// LW lanes per pair (LW >= photons per leaf), G = 32/LW pairs per group, U groups in flight.
template <int MODE, bool COUNT, int LW>
__global__ void __launch_bounds__(LP_WARPS * 32, 4) k_leaf_pairs(LeafArgs A,
                                                              const uint2 *__restrict__ queue,
                                                              unsigned int queueCap,
                                                              unsigned int *__restrict__ queueCtl,
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
    if (!(__ldg(queueCtl + QC_KINDS) & KIND_THIN)) return;
    const unsigned total = min(__ldg(queueCtl + QC_PAIR_TAIL), queueCap);
    const unsigned warpsTotal = gridDim.x * LP_WARPS;
    const unsigned warpId = blockIdx.x * LP_WARPS + (threadIdx.x >> 5);
    (void)warpId;
    unsigned long long cTests = 0;
    unsigned nbuf = 0;
#if LP_DYNAMIC
    // Runs of LP_RUN pairs are claimed from a counter: the cost of a run varies with its hits, a
    // static stride leaves the chip waiting for the unluckiest warp.  A claim takes `rpc` runs so
    // that a launch issues ~16 claims per warp however long the queue is (one claim per run made
    // the counter itself the bottleneck of a 180 M-pair queue: K5b 6.4 -> 8.7 ms).
    const unsigned totalRuns = (total + LP_RUN - 1) / LP_RUN;
    const unsigned rpc = max(1u, min(64u, totalRuns / (warpsTotal * 16u)));
    unsigned cur = 0, curEnd = 0;      // runs [cur, curEnd) belong to this warp
    for (;;) {
        if (cur == curEnd) {
            // (claiming the NEXT batch of runs early, to hide the counter's round trip, measured
            // slower: K5b 1.90 -> 2.00 ms)
            unsigned c = 0;
            if (lane == 0) c = atomicAdd(queueCtl + QC_RUN_THIN, rpc);
            cur = __shfl_sync(FULL, c, 0);
            if (cur >= totalRuns) break;
            curEnd = min(cur + rpc, totalRuns);
        }
        const unsigned run = (cur++) * LP_RUN;
#else
    for (unsigned run = warpId * LP_RUN; run < total; run += warpsTotal * LP_RUN) {
#endif
        // 64 pairs, two coalesced 256-byte loads
        uint2 pr0 = make_uint2(PAIR_NONE, 0u), pr1 = make_uint2(PAIR_NONE, 0u);
        if (run + lane < total) pr0 = __ldg(queue + run + lane);
        if (run + 32 + lane < total) pr1 = __ldg(queue + run + 32 + lane);
        // pairs of coherent packets (PAIR_FAT) belong to k_leaf_fat; a run is almost always all of
        // one kind (a K5a warp fills its chunks packet by packet)
        if (!__any_sync(FULL, !(pr0.x & 0xc0000000u) || !(pr1.x & 0xc0000000u))) continue;   // (bit 30: tile pairs, k_leaf_tile)
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
                on[u] = !(qx & 0xc0000000u) && j < cnt;   // PAIR_NONE has the bits set too
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

// ---------------------------------------------------------------- K5b for coherent packets
// (packet, leaf) pairs, two stages per warp:
//  1. one photon per lane (LW lanes per pair, like k_leaf_pairs): the photon is tested ONCE against
//     the packet's frame and the slope ranges of its two half packets (frame_halves); it is entered
//     into a shared-memory buffer once per half packet (rays 0-3, 4-7) it may reach.
//  2. whenever 8 entries are buffered: lane = (entry, ray of its half packet) runs the exact disk
//     predicate, 32 per pass with every lane busy.  Hits go through the same buffer / warp shading
//     as in k_leaf_pairs.
// On the bench scene 28 % of the photons of a packet's leaves survive stage 1 and reach 1.2 half
// packets on average, so a photon costs ~1 + 0.28 x 1.2 x 4 predicates instead of 8.
// Work distribution: resident warps claim runs of LP_RUN pairs from a counter, in queue order.
struct __align__(16) FatBuf {      // one entry per (surviving photon, half packet it may reach)
    float4 p[96];        // photon {pos, radius}
    uint32_t s[96];      // its index
    uint32_t base[96];   // first ray of the HALF packet (packet base, +4 for the second half)
};
// Tuning knobs (make NVEXTRA=-D...), measured on smoke10m_1080p, K5b ms:
//   U=2,MINB=3,SEG=1 2.88 | U=4 2.75 | MINB=4 2.39 | SEG=0 2.35 | SEG=0,MINB=5 2.36 | SEG=0,MINB=6 2.65
//   | SEG=0,U=4,MINB=4 2.24 (default).  Occupancy is what pays: stage 2 hands same-ray hits to
//   non-adjacent lanes, so the segmented reduction merges little and only costs registers.
#ifndef FAT_U
#define FAT_U 4
#endif
#ifndef FAT_MINB
#define FAT_MINB 4
#endif
#ifndef FAT_SEG
#define FAT_SEG 0
#endif
#ifndef FAT_PREF
#define FAT_PREF 0
#endif

template <int MODE, bool COUNT, int LW>
__global__ void __launch_bounds__(LP_WARPS * 32, FAT_MINB) k_leaf_fat(LeafArgs A, const uint2 *__restrict__ queue,
                                                            unsigned int queueCap,
                                                            unsigned int *__restrict__ queueCtl,
                                                            unsigned long long *__restrict__ counters) {
    __shared__ HitBuf bufAll[LP_WARPS];
    __shared__ FatBuf fatAll[LP_WARPS];
    HitBuf &hb = bufAll[threadIdx.x >> 5];
    FatBuf &fb = fatAll[threadIdx.x >> 5];
    constexpr unsigned G = 32 / LW;
    constexpr int U = FAT_U;
    constexpr unsigned HALF = FAT_P / 2;               // rays of a half packet
    constexpr unsigned SLOTS = 32 / HALF;              // (survivor, half packet) entries per stage-2 pass
    const unsigned FULL = 0xffffffffu;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt = (1u << lane) - 1u;
    const unsigned sub = lane / LW, j = lane % LW;
    const unsigned slot = lane / HALF, ray = lane % HALF;   // stage 2: entry `slot` of the pass, ray `ray` of its half packet
    if (!(__ldg(queueCtl + QC_KINDS) & KIND_FAT)) return;
    const unsigned total = min(__ldg(queueCtl + QC_PAIR_TAIL), queueCap);
    const unsigned warpsTotal = gridDim.x * LP_WARPS;
    const unsigned warpId = blockIdx.x * LP_WARPS + (threadIdx.x >> 5);
    (void)warpId;
    unsigned long long cTests = 0;
    unsigned nbuf = 0, npb = 0;

    // stage 2 over the newest `take` entries (take <= SLOTS): HALF rays each
    auto exact_pass = [&](unsigned take) {
        bool hit = false;
        float th = 0.f, d2 = 0.f;
        uint32_t s = 0, q = 0;
        if (slot < take) {
            const unsigned e = npb - take + slot;
            const float4 p = fb.p[e];
            s = fb.s[e];
            q = fb.base[e] + ray;
            const float4 o = __ldg(A.ro + q), d = __ldg(A.rd + q);
            RayS r;
            r.ox = o.x; r.oy = o.y; r.oz = o.z; r.tMax = o.w;
            r.dx = d.x; r.dy = d.y; r.dz = d.z;
            hit = disk_test(r, p, th, d2);
            if (COUNT) ++cTests;
        }
        npb -= take;
        if (MODE != 0) {
            if (hit) record_hit(A, MODE, s, q);
            return;
        }
        const unsigned mh = __ballot_sync(FULL, hit);
        if (mh == 0u) return;
        if (hit) {
            const unsigned pos = nbuf + __popc(mh & lt);
            hb.s[pos] = s; hb.q[pos] = q; hb.th[pos] = th; hb.d2[pos] = d2;
        }
        nbuf += __popc(mh);
        __syncwarp();
        if (nbuf >= 32) {
            const unsigned e = nbuf - 32 + lane;
            if (FAT_SEG) shade_hits_warp(A, lane, 32u, hb.s[e], hb.q[e], hb.th[e], hb.d2[e]);
            else shade_hit(A, hb.s[e], hb.q[e], hb.th[e], hb.d2[e]);
            nbuf -= 32;
            __syncwarp();
        }
    };

#if LP_DYNAMIC
    // Runs of LP_RUN pairs are claimed from a counter: the cost of a run varies with its hits, a
    // static stride leaves the chip waiting for the unluckiest warp.  A claim takes `rpc` runs so
    // that a launch issues ~16 claims per warp however long the queue is (one claim per run made
    // the counter itself the bottleneck of a 180 M-pair queue: K5b 6.4 -> 8.7 ms).
    const unsigned totalRuns = (total + LP_RUN - 1) / LP_RUN;
    const unsigned rpc = max(1u, min(64u, totalRuns / (warpsTotal * 16u)));
    unsigned cur = 0, curEnd = 0;      // runs [cur, curEnd) belong to this warp
    for (;;) {
        if (cur == curEnd) {
            // (claiming the NEXT batch of runs early, to hide the counter's round trip, measured
            // slower: K5b 1.90 -> 2.00 ms)
            unsigned c = 0;
            if (lane == 0) c = atomicAdd(queueCtl + QC_RUN_FAT, rpc);
            cur = __shfl_sync(FULL, c, 0);
            if (cur >= totalRuns) break;
            curEnd = min(cur + rpc, totalRuns);
        }
        const unsigned run = (cur++) * LP_RUN;
#else
    for (unsigned run = warpId * LP_RUN; run < total; run += warpsTotal * LP_RUN) {
#endif
        uint2 pr0 = make_uint2(PAIR_NONE, 0u), pr1 = make_uint2(PAIR_NONE, 0u);
        if (run + lane < total) pr0 = __ldg(queue + run + lane);
        if (run + 32 + lane < total) pr1 = __ldg(queue + run + 32 + lane);
        if (!__any_sync(FULL, (pr0.x != PAIR_NONE && (pr0.x & PAIR_FAT)) || (pr1.x != PAIR_NONE && (pr1.x & PAIR_FAT))))
            continue;
#pragma unroll 1
        for (unsigned g0 = 0; g0 < LP_RUN / G; g0 += U) {
            uint32_t base[U], first[U];
            bool on[U];
            float4 p[U];
#if FAT_PREF
            float4 f0[U], f1[U], f2[U], f3[U];
#endif
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const unsigned idx = (g0 + u) * G + sub;
                const bool lower = (g0 + u) * G < 32u;             // warp-uniform
                const unsigned qx = __shfl_sync(FULL, lower ? pr0.x : pr1.x, idx & 31u);
                const unsigned qy = __shfl_sync(FULL, lower ? pr0.y : pr1.y, idx & 31u);
                base[u] = qx & ~PAIR_FAT;
                first[u] = qy & 0x07ffffffu;
                const uint32_t cnt = ((qy >> 27) & 15u) + 1u;
                on[u] = qx != PAIR_NONE && (qx & PAIR_FAT) && j < cnt;
                if (on[u]) {
                    p[u] = __ldg(A.t.s_pos_r + first[u] + j);
#if FAT_PREF
                    const float4 *fr = A.frames + (size_t)(base[u] / FAT_P) * FRAME_V4;
                    f0[u] = __ldg(fr); f1[u] = __ldg(fr + 1); f2[u] = __ldg(fr + 2); f3[u] = __ldg(fr + 3);
#endif
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                unsigned halves = 0;       // bit 0: rays 0-3 of the packet may hit this photon, bit 1: rays 4-7
                if (on[u]) {
#if FAT_PREF
                    const float4 f4 = __ldg(A.frames + (size_t)(base[u] / FAT_P) * FRAME_V4 + 4);
                    halves = frame_halves(f0[u], f1[u], f2[u], f3[u], f4, p[u]);
#else
                    const float4 *fr = A.frames + (size_t)(base[u] / FAT_P) * FRAME_V4;
                    halves = frame_halves(__ldg(fr), __ldg(fr + 1), __ldg(fr + 2), __ldg(fr + 3), __ldg(fr + 4), p[u]);
#endif
                    if (COUNT) ++cTests;
                }
                const unsigned mkA = __ballot_sync(FULL, halves & 1u), mkB = __ballot_sync(FULL, halves & 2u);
                if ((mkA | mkB) == 0u) continue;
                if (halves & 1u) {
                    const unsigned pos = npb + __popc(mkA & lt);
                    fb.p[pos] = p[u]; fb.s[pos] = first[u] + j; fb.base[pos] = base[u];
                }
                if (halves & 2u) {
                    const unsigned pos = npb + __popc(mkA) + __popc(mkB & lt);
                    fb.p[pos] = p[u]; fb.s[pos] = first[u] + j; fb.base[pos] = base[u] + HALF;
                }
                npb += __popc(mkA) + __popc(mkB);
                __syncwarp();
                while (npb >= SLOTS) exact_pass(SLOTS);
                __syncwarp();
            }
        }
    }
    if (npb) exact_pass(npb);
    if (MODE == 0 && nbuf) {
        uint32_t s = 0, q = 0; float th = 0.f, d2 = 0.f;
        if (lane < nbuf) { s = hb.s[lane]; q = hb.q[lane]; th = hb.th[lane]; d2 = hb.d2[lane]; }
        if (FAT_SEG) shade_hits_warp(A, lane, nbuf, s, q, th, d2);
        else if (lane < nbuf) shade_hit(A, s, q, th, d2);
    }
    if (COUNT) {
#pragma unroll
        for (int o2 = 16; o2 > 0; o2 >>= 1) cTests += __shfl_xor_sync(FULL, cTests, o2);
        if (lane == 0 && cTests) atomicAdd(counters + 4, cTests);
    }
}

}  // namespace bre
