/* oracle/bre_oracle.c -- TEST INFRASTRUCTURE ONLY (see bre_oracle.h header comment).
 *
 * CPU restatement of the beam-radiance-estimate hot path.  float32 everywhere
 * (core/pbrt.h:142), compile with -ffp-contract=off so every operation rounds
 * once, in the order written.  `file:line` citations are relative to
 * /root/reference/src/.
 */
#include "bre_oracle.h"
#define ORC_TAU_STEPS_MAX 384
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <stdatomic.h>
#include <unistd.h>

#define ORC_PI 3.14159265358979323846f      /* core/pbrt.h:150-156 */
#define ORC_INV_PI 0.31830988618379067154f
#define ORC_INV4PI 0.07957747154594766788f
#define ORC_MACHINE_EPS 5.9604644775390625e-8f /* numeric_limits<float>::epsilon()*0.5, pbrt.h:145 */
#define ORC_ONE_MINUS_EPS 0.99999994f         /* FloatOneMinusEpsilon, rng.h:50 */

static inline float orc_gamma(int n) { /* core/pbrt.h:263-265 */
    return (n * ORC_MACHINE_EPS) / (1 - n * ORC_MACHINE_EPS);
}
static inline float fmaxf_(float a, float b) { return a > b ? a : b; } /* std::max */
static inline float fminf_(float a, float b) { return b < a ? b : a; } /* std::min */

/* ------------------------------------------------------------------ PCG32 */
#define PCG32_DEFAULT_STATE 0x853c49e6748fea9bULL
#define PCG32_MULT 0x5851f42d4c957f2dULL

uint32_t orc_rng_u32(orc_rng *r) { /* rng.h:138-144 */
    uint64_t oldstate = r->state;
    r->state = oldstate * PCG32_MULT + r->inc;
    uint32_t xorshifted = (uint32_t)(((oldstate >> 18u) ^ oldstate) >> 27u);
    uint32_t rot = (uint32_t)(oldstate >> 59u);
    return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
}
void orc_rng_seed(orc_rng *r, uint64_t initseq) { /* rng.h:130-136 */
    r->state = 0u;
    r->inc = (initseq << 1u) | 1u;
    orc_rng_u32(r);
    r->state += PCG32_DEFAULT_STATE;
    orc_rng_u32(r);
}
float orc_rng_float(orc_rng *r) { /* rng.h:78-85 */
    return fminf_(ORC_ONE_MINUS_EPS, (float)(orc_rng_u32(r) * 0x1p-32f));
}

/* --------------------------------------------------------------- sampling */
void orc_uniform_sample_sphere(float u0, float u1, float out[3]) { /* sampling.cpp:98-103 */
    float z = 1 - 2 * u0;
    float r = sqrtf(fmaxf_(0.f, 1.f - z * z));
    float phi = 2 * ORC_PI * u1;
    out[0] = r * cosf(phi);
    out[1] = r * sinf(phi);
    out[2] = z;
}
void orc_uniform_sample_cone(float u0, float u1, float cosThetaMax, float out[3]) {
    /* sampling.cpp:136-142 */
    float cosTheta = (1.f - u0) + u0 * cosThetaMax;
    float sinTheta = sqrtf(1.f - cosTheta * cosTheta);
    float phi = u1 * 2 * ORC_PI;
    out[0] = cosf(phi) * sinTheta;
    out[1] = sinf(phi) * sinTheta;
    out[2] = cosTheta;
}

/* ---------------------------------------------------------------- phase fn */
float orc_phase_hg(float cosTheta, float g) { /* medium.h:66-69 */
    float denom = 1 + g * g + 2 * g * cosTheta;
    return ORC_INV4PI * (1 - g * g) / (denom * sqrtf(denom));
}

/* CoordinateSystem geometry.h:1020-1027; Vector3::operator/ multiplies by the
 * reciprocal (geometry.h:Vector3 operator/); Cross is evaluated in double
 * (geometry.h Cross()). */
static void coordinate_system(const float v1[3], float v2[3], float v3[3]) {
    if (fabsf(v1[0]) > fabsf(v1[1])) {
        float inv = 1.f / sqrtf(v1[0] * v1[0] + v1[2] * v1[2]);
        v2[0] = -v1[2] * inv; v2[1] = 0 * inv; v2[2] = v1[0] * inv;
    } else {
        float inv = 1.f / sqrtf(v1[1] * v1[1] + v1[2] * v1[2]);
        v2[0] = 0 * inv; v2[1] = v1[2] * inv; v2[2] = -v1[1] * inv;
    }
    double v1x = v1[0], v1y = v1[1], v1z = v1[2];
    double v2x = v2[0], v2y = v2[1], v2z = v2[2];
    v3[0] = (float)((v1y * v2z) - (v1z * v2y));
    v3[1] = (float)((v1z * v2x) - (v1x * v2z));
    v3[2] = (float)((v1x * v2y) - (v1y * v2x));
}

float orc_hg_sample_p(const float wo[3], float g, float u0, float u1, float wi[3]) {
    /* medium.cpp:186-205 */
    float cosTheta;
    if (fabsf(g) < 1e-3)
        cosTheta = 1 - 2 * u0;
    else {
        float sqrTerm = (1 - g * g) / (1 - g + 2 * g * u0);
        cosTheta = (1 + g * g - sqrTerm * sqrTerm) / (2 * g);
    }
    float sinTheta = sqrtf(fmaxf_(0.f, 1 - cosTheta * cosTheta));
    float phi = 2 * ORC_PI * u1;
    float v1[3], v2[3];
    coordinate_system(wo, v1, v2);
    /* SphericalDirection(sinTheta, cosTheta, phi, v1, v2, -wo) geometry.h:1465-1470 */
    float a = sinTheta * cosf(phi), b = sinTheta * sinf(phi);
    for (int k = 0; k < 3; ++k) wi[k] = a * v1[k] + b * v2[k] + cosTheta * (-wo[k]);
    return orc_phase_hg(-cosTheta, g);
}

/* ------------------------------------------------------------------ media */
void orc_medium_homogeneous(orc_medium *m, const float sa[3], const float ss[3], float g) {
    memset(m, 0, sizeof(*m));
    m->kind = 0;
    for (int k = 0; k < 3; ++k) {
        m->sigma_a[k] = sa[k]; m->sigma_s[k] = ss[k]; m->sigma_t[k] = sa[k] + ss[k];
    }
    m->g = g;
}
void orc_medium_grid(orc_medium *m, const float sa[3], const float ss[3], float g, int nx, int ny,
                     int nz, const float *density, const float w2m[16]) {
    orc_medium_homogeneous(m, sa, ss, g);
    m->kind = 1;
    m->nx = nx; m->ny = ny; m->nz = nz;
    m->density = density;
    float maxDensity = 0; /* grid.h:69-73 */
    for (long i = 0; i < (long)nx * ny * nz; ++i) maxDensity = fmaxf_(maxDensity, density[i]);
    m->invMaxDensity = 1 / maxDensity;
    memcpy(m->w2m, w2m, sizeof(float) * 16);
    m->tau_steps = orc_tau_steps_auto(nx, ny, nz);
}
/* Steps of the per-ray optical-depth table (the rule libbre.so applies when bre_config.tau_steps == 0):
 * 1.5 table steps per voxel of the longest grid axis, rounded up to a multiple of 32, within [64, 128].
 * A step that EQUALS the voxel pitch aliases with the lattice of the trilinear interpolant (64 steps
 * through a 64^3 grid: 1.3e-3 relative error of Tr; 96 steps: 1.5e-4, tests/test_oracle_golden.py). */
int orc_tau_steps_auto(int nx, int ny, int nz) {
    int m = nx > ny ? nx : ny;
    if (nz > m) m = nz;
    int s = (3 * m / 2 + 31) / 32 * 32;
    if (s < 64) s = 64;
    if (s > 128) s = 128;
    return s;
}
static inline float grid_D(const orc_medium *m, int x, int y, int z) { /* grid.h:78-82 */
    if (x < 0 || y < 0 || z < 0 || x >= m->nx || y >= m->ny || z >= m->nz) return 0;
    return m->density[((long)z * m->ny + y) * m->nx + x];
}
static inline float lerp_(float t, float a, float b) { return (1 - t) * a + t * b; } /* pbrt.h Lerp */
float orc_grid_density(const orc_medium *m, const float p[3]) { /* grid.cpp:45-60 */
    float sx = p[0] * m->nx - .5f, sy = p[1] * m->ny - .5f, sz = p[2] * m->nz - .5f;
    int ix = (int)floorf(sx), iy = (int)floorf(sy), iz = (int)floorf(sz);
    float dx = sx - (float)ix, dy = sy - (float)iy, dz = sz - (float)iz;
    float d00 = lerp_(dx, grid_D(m, ix, iy, iz), grid_D(m, ix + 1, iy, iz));
    float d10 = lerp_(dx, grid_D(m, ix, iy + 1, iz), grid_D(m, ix + 1, iy + 1, iz));
    float d01 = lerp_(dx, grid_D(m, ix, iy, iz + 1), grid_D(m, ix + 1, iy, iz + 1));
    float d11 = lerp_(dx, grid_D(m, ix, iy + 1, iz + 1), grid_D(m, ix + 1, iy + 1, iz + 1));
    float d0 = lerp_(dy, d00, d10);
    float d1 = lerp_(dy, d01, d11);
    return lerp_(dz, d0, d1);
}
void orc_homogeneous_tr(const orc_medium *m, float t, float tr[3]) { /* homogeneous.cpp:44-48 */
    for (int k = 0; k < 3; ++k) tr[k] = expf(-m->sigma_t[k] * fminf_(t, 3.402823466e+38f));
}

/* Bounds3::IntersectP(ray, &t0, &t1), geometry.h:1386-1408 */
int orc_ray_box(const float o[3], const float d[3], float tMaxRay, const float lo[3],
                const float hi[3], float *out0, float *out1) {
    float t0 = 0, t1 = tMaxRay;
    for (int i = 0; i < 3; ++i) {
        float invRayDir = 1 / d[i];
        float tNear = (lo[i] - o[i]) * invRayDir;
        float tFar = (hi[i] - o[i]) * invRayDir;
        if (tNear > tFar) { float s = tNear; tNear = tFar; tFar = s; }
        tFar *= 1 + 2 * orc_gamma(3);
        t0 = tNear > t0 ? tNear : t0;
        t1 = tFar < t1 ? tFar : t1;
        if (t0 > t1) return 0;
    }
    if (out0) *out0 = t0;
    if (out1) *out1 = t1;
    return 1;
}

/* world -> medium for a point / a vector: plain affine matrix product
 * (transform.h:278-300 without the error-bound origin offset of :251-264). */
static void xf_point(const float m[16], const float p[3], float out[3]) {
    for (int r = 0; r < 3; ++r)
        out[r] = m[4 * r + 0] * p[0] + m[4 * r + 1] * p[1] + m[4 * r + 2] * p[2] + m[4 * r + 3];
}
static void xf_vector(const float m[16], const float v[3], float out[3]) {
    for (int r = 0; r < 3; ++r) out[r] = m[4 * r + 0] * v[0] + m[4 * r + 1] * v[1] + m[4 * r + 2] * v[2];
}

int orc_grid_tau_table(const orc_medium *m, const float o[3], const float d[3], float tMax,
                       int nSteps, float *tau, float *out0, float *out1) {
    float om[3], dm[3];
    xf_point(m->w2m, o, om);
    xf_vector(m->w2m, d, dm);
    const float lo[3] = {0, 0, 0}, hi[3] = {1, 1, 1};
    float t0, t1;
    for (int k = 0; k <= nSteps; ++k) tau[k] = 0;
    if (!orc_ray_box(om, dm, tMax, lo, hi, &t0, &t1)) { *out0 = 0; *out1 = 0; return 0; }
    float dt = (t1 - t0) / (float)nSteps;
    float acc = 0;
    for (int k = 0; k < nSteps; ++k) {
        float t = t0 + ((float)k + 0.5f) * dt;
        float p[3] = {om[0] + dm[0] * t, om[1] + dm[1] * t, om[2] + dm[2] * t};
        acc = acc + orc_grid_density(m, p) * dt;
        tau[k + 1] = acc;
    }
    *out0 = t0; *out1 = t1;
    return 1;
}
/* transmittance at ray parameter t from a tau table */
static void grid_tr_from_table(const orc_medium *m, const float *tau, int nSteps, float t0, float t1,
                               float t, float tr[3]) {
    float od = 0;
    if (t1 > t0) {
        float dt = (t1 - t0) / (float)nSteps;
        float s = (t - t0) / dt;
        if (s > 0) {
            if (s >= (float)nSteps) od = tau[nSteps];
            else {
                int k = (int)s;
                float f = s - (float)k;
                od = tau[k] + f * (tau[k + 1] - tau[k]);
            }
        }
    }
    for (int k = 0; k < 3; ++k) tr[k] = expf(-m->sigma_t[k] * od);
}

/* Tr(o -> o + t d) of a grid medium exactly as the query evaluates it: nSteps-step table over the
 * ray's overlap with the medium box, linear interpolation at t. */
void orc_grid_tr_table(const orc_medium *m, const float o[3], const float d[3], float tMax, float t,
                       int nSteps, float tr[3]) {
    float tau[ORC_TAU_STEPS_MAX + 1], t0, t1;
    if (nSteps < 1 || nSteps > ORC_TAU_STEPS_MAX) nSteps = 64;
    orc_grid_tau_table(m, o, d, tMax, nSteps, tau, &t0, &t1);
    grid_tr_from_table(m, tau, nSteps, t0, t1, t, tr);
}

/* ------------------------------------------------------------------ Morton */
static inline uint32_t left_shift3(uint32_t x) { /* photonbeambvh.cpp:109-132 */
    if (x == (1 << 10)) --x;
    x = (x | (x << 16)) & 0x30000ff;
    x = (x | (x << 8)) & 0x300f00f;
    x = (x | (x << 4)) & 0x30c30c3;
    x = (x | (x << 2)) & 0x9249249;
    return x;
}
uint32_t orc_morton30(const float p[3], const float lo[3], const float hi[3]) {
    /* Bounds3::Offset (geometry.h) then *1024, EncodeMorton3 photonbeambvh.cpp:134-139,440-444 */
    uint32_t q[3];
    for (int k = 0; k < 3; ++k) {
        float o = p[k] - lo[k];
        if (hi[k] > lo[k]) o /= hi[k] - lo[k];
        float s = o * 1024.f;
        uint32_t v = (uint32_t)s;
        if (v > 1024u) v = 1024u;
        q[k] = left_shift3(v);
    }
    return (q[2] << 2) | (q[1] << 1) | q[0];
}

/* ------------------------------------------------------------ canonical LBVH */
static int cmp_u64(const void *a, const void *b) {
    uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
    return x < y ? -1 : (x > y ? 1 : 0);
}
static inline int clz64(uint64_t x) { return x ? __builtin_clzll(x) : 64; }

typedef struct { const uint64_t *keys; orc_lbvh *t; } build_ctx;

static void leaf_box(const orc_lbvh *t, int i, float lo[3], float hi[3]) {
    memcpy(lo, t->leaf_lo + 3 * (size_t)i, 12);
    memcpy(hi, t->leaf_hi + 3 * (size_t)i, 12);
}

/* internal node `idx` covers sorted range [f,l], f<l (numbering as in Karras 2012:
 * left child = split, right child = split+1, root = 0) */
static void build_rec(build_ctx *c, int idx, int f, int l) {
    const uint64_t *keys = c->keys;
    orc_node *nd = &c->t->nodes[idx];
    int delta = clz64(keys[f] ^ keys[l]);
    /* largest g in [f,l) sharing more than delta leading bits with keys[f] */
    int a = f, b = l; /* invariant: a shares > delta, b does not */
    while (b - a > 1) {
        int mid = a + (b - a) / 2;
        if (clz64(keys[f] ^ keys[mid]) > delta) a = mid; else b = mid;
    }
    int g = a;
    int bit = 63 - delta;
    nd->axis = bit >= 32 ? (bit - 32) % 3 : 0;
    nd->first = f; nd->count = l - f + 1;
    float llo[3], lhi[3], rlo[3], rhi[3];
    if (g == f) { nd->left = ~f; leaf_box(c->t, f, llo, lhi); }
    else {
        nd->left = g; build_rec(c, g, f, g);
        memcpy(llo, c->t->nodes[g].lo, 12); memcpy(lhi, c->t->nodes[g].hi, 12);
    }
    if (g + 1 == l) { nd->right = ~l; leaf_box(c->t, l, rlo, rhi); }
    else {
        nd->right = g + 1; build_rec(c, g + 1, g + 1, l);
        memcpy(rlo, c->t->nodes[g + 1].lo, 12); memcpy(rhi, c->t->nodes[g + 1].hi, 12);
    }
    for (int k = 0; k < 3; ++k) {
        nd->lo[k] = fminf_(llo[k], rlo[k]);
        nd->hi[k] = fmaxf_(lhi[k], rhi[k]);
    }
}

orc_lbvh *orc_lbvh_build(const orc_photons *ph) {
    uint32_t n = ph->n;
    orc_lbvh *t = (orc_lbvh *)calloc(1, sizeof(orc_lbvh));
    t->n = n;
    t->order = (uint32_t *)malloc(sizeof(uint32_t) * (n ? n : 1));
    t->codes = (uint32_t *)malloc(sizeof(uint32_t) * (n ? n : 1));
    t->leaf_lo = (float *)malloc(sizeof(float) * 3 * (n ? n : 1));
    t->leaf_hi = (float *)malloc(sizeof(float) * 3 * (n ? n : 1));
    t->nodes = (orc_node *)calloc(n > 1 ? n - 1 : 1, sizeof(orc_node));
    if (n == 0) return t;
    for (int k = 0; k < 3; ++k) { t->cb_lo[k] = INFINITY; t->cb_hi[k] = -INFINITY; }
    for (uint32_t i = 0; i < n; ++i)
        for (int k = 0; k < 3; ++k) {
            t->cb_lo[k] = fminf_(t->cb_lo[k], ph->pos[3 * (size_t)i + k]);
            t->cb_hi[k] = fmaxf_(t->cb_hi[k], ph->pos[3 * (size_t)i + k]);
        }
    uint64_t *keys = (uint64_t *)malloc(sizeof(uint64_t) * n);
    for (uint32_t i = 0; i < n; ++i)
        keys[i] = ((uint64_t)orc_morton30(ph->pos + 3 * (size_t)i, t->cb_lo, t->cb_hi) << 32) | i;
    qsort(keys, n, sizeof(uint64_t), cmp_u64);
    float amax = 0;
    for (uint32_t s = 0; s < n; ++s) {
        uint32_t i = (uint32_t)(keys[s] & 0xffffffffu);
        t->order[s] = i;
        t->codes[s] = (uint32_t)(keys[s] >> 32);
        /* leaf box: sphere bound p +- r, r inflated by 2^-19 relative, then one ulp outward */
        float rp = ph->radius[i] * 1.000002f;
        for (int k = 0; k < 3; ++k) {
            float lo = nextafterf(ph->pos[3 * (size_t)i + k] - rp, -INFINITY);
            float hi = nextafterf(ph->pos[3 * (size_t)i + k] + rp, INFINITY);
            t->leaf_lo[3 * (size_t)s + k] = lo;
            t->leaf_hi[3 * (size_t)s + k] = hi;
            amax = fmaxf_(amax, fmaxf_(fabsf(lo), fabsf(hi)));
        }
    }
    t->abs_max = amax;
    if (n > 1) {
        build_ctx c = {keys, t};
        build_rec(&c, 0, 0, (int)n - 1);
    }
    free(keys);
    return t;
}
void orc_lbvh_free(orc_lbvh *t) {
    if (!t) return;
    free(t->order); free(t->codes); free(t->leaf_lo); free(t->leaf_hi); free(t->nodes); free(t);
}

/* ------------------------------------------------------------- disk predicate */
int orc_disk_test(const float o[3], const float d[3], float tMax, const float p[3], float r,
                  float *t_out, float *d2_out) {
    float vx = p[0] - o[0], vy = p[1] - o[1], vz = p[2] - o[2];
    float t = vx * d[0] + vy * d[1] + vz * d[2];
    if (!(t > 0 && t < tMax)) return 0;
    float wx = vx - t * d[0], wy = vy - t * d[1], wz = vz - t * d[2];
    float d2 = wx * wx + wy * wy + wz * wz;
    float r2 = r * r;
    if (!(d2 < r2)) return 0;   /* strict <, as kdtree.h:182-184 / photonbeam.cpp:502 */
    *t_out = t; *d2_out = d2;
    return 1;
}

/* contribution of a hit photon: K * Tr * phase * alpha / nPaths
 * kernel(): photonmap.cpp:170-174; normalisation photonmap.cpp:198 */
static inline void accumulate_hit(const orc_photons *ph, uint32_t i, const float d[3], float t,
                                  float d2, const float tr[3], float g, float nPathsNorm,
                                  float rgb[3]) {
    float r = ph->radius[i];
    float r2 = r * r;
    float s = 1.f - d2 / r2;
    float k = 3.f * ORC_INV_PI * s * s;
    float wgt = k / (nPathsNorm * r2);
    const float *wi = ph->dir + 3 * (size_t)i;
    /* phase->p(wo = -d, wi) = PhaseHG(Dot(wo, wi), g)  medium.cpp:207-210, vsppm.cpp:488-489 */
    float cosTheta = (-d[0]) * wi[0] + (-d[1]) * wi[1] + (-d[2]) * wi[2];
    float wp = wgt * orc_phase_hg(cosTheta, g);
    const float *a = ph->power + 3 * (size_t)i;
    for (int c = 0; c < 3; ++c) rgb[c] += (wp * tr[c]) * a[c];
    (void)t;
}


typedef struct {
    const orc_medium *m;
    float tau[ORC_TAU_STEPS_MAX + 1];
    int steps;
    float t0, t1;
} ray_tr_state;
static void ray_tr_init(ray_tr_state *s, const orc_medium *m, const float o[3], const float d[3],
                        float tMax) {
    s->m = m;
    s->steps = (m->tau_steps >= 1 && m->tau_steps <= ORC_TAU_STEPS_MAX) ? m->tau_steps : 64;
    if (m->kind == 1) orc_grid_tau_table(m, o, d, tMax, s->steps, s->tau, &s->t0, &s->t1);
}
static void ray_tr_eval(const ray_tr_state *s, float t, float tr[3]) {
    if (s->m->kind == 0) orc_homogeneous_tr(s->m, t, tr);
    else grid_tr_from_table(s->m, s->tau, s->steps, s->t0, s->t1, t, tr);
}

/* slab test of geometry.h:1410-1438 with per-ray padded origins (DESIGN.md "conservative cull") */
typedef struct {
    float oNear[3], oFar[3], invDir[3], tMax;
    int neg[3];
} ray_slab;
static void ray_slab_init_pad(ray_slab *rs, const float o[3], const float d[3], float tMax,
                              float pad) {
    for (int k = 0; k < 3; ++k) {
        rs->invDir[k] = 1 / d[k];
        rs->neg[k] = rs->invDir[k] < 0;
        float sp = rs->neg[k] ? -pad : pad;
        rs->oNear[k] = o[k] + sp;
        rs->oFar[k] = o[k] - sp;
    }
    rs->tMax = tMax;
}
static void ray_slab_init(ray_slab *rs, const float o[3], const float d[3], float tMax,
                          float abs_max) {
    float M = fmaxf_(fmaxf_(fabsf(o[0]), fabsf(o[1])), fabsf(o[2])) + abs_max;
    ray_slab_init_pad(rs, o, d, tMax, M * 0x1p-18f);
}
static inline int slab_hit(const ray_slab *rs, const float lo[3], const float hi[3]) {
    const float g3 = 1 + 2 * orc_gamma(3);
    float tMin = ((rs->neg[0] ? hi[0] : lo[0]) - rs->oNear[0]) * rs->invDir[0];
    float tMax = ((rs->neg[0] ? lo[0] : hi[0]) - rs->oFar[0]) * rs->invDir[0];
    float tyMin = ((rs->neg[1] ? hi[1] : lo[1]) - rs->oNear[1]) * rs->invDir[1];
    float tyMax = ((rs->neg[1] ? lo[1] : hi[1]) - rs->oFar[1]) * rs->invDir[1];
    tMax *= g3;
    tyMax *= g3;
    if (tMin > tyMax || tyMin > tMax) return 0;
    if (tyMin > tMin) tMin = tyMin;
    if (tyMax < tMax) tMax = tyMax;
    float tzMin = ((rs->neg[2] ? hi[2] : lo[2]) - rs->oNear[2]) * rs->invDir[2];
    float tzMax = ((rs->neg[2] ? lo[2] : hi[2]) - rs->oFar[2]) * rs->invDir[2];
    tzMax *= g3;
    if (tMin > tzMax || tzMin > tMax) return 0;
    if (tzMin > tMin) tMin = tzMin;
    if (tzMax < tMax) tMax = tzMax;
    return (tMin < rs->tMax) && (tMax > 0);
}

int orc_slab_test(const float o[3], const float d[3], float tMax, const float lo[3],
                  const float hi[3], float pad) {
    ray_slab rs;
    ray_slab_init_pad(&rs, o, d, tMax, pad);
    return slab_hit(&rs, lo, hi);
}

uint32_t orc_query_ray(const orc_lbvh *t, const orc_photons *ph, const orc_medium *m,
                       const float o[3], const float d[3], float tMax, float nPathsNorm,
                       float rgb[3], uint32_t *hits, uint32_t hits_cap, orc_counters *c) {
    rgb[0] = rgb[1] = rgb[2] = 0;
    uint32_t nh = 0;
    if (t->n == 0) return 0;
    ray_slab rs;
    ray_slab_init(&rs, o, d, tMax, t->abs_max);
    ray_tr_state trs;
    ray_tr_init(&trs, m, o, d, tMax);
    uint64_t V = 0, T = 0;
    int32_t stack[128];
    int sp = 0;
    int32_t cur = t->n > 1 ? 0 : ~0;
    for (;;) {
        ++V;
        if (cur < 0) { /* leaf: one photon */
            int s = ~cur;
            if (slab_hit(&rs, t->leaf_lo + 3 * (size_t)s, t->leaf_hi + 3 * (size_t)s)) {
                ++T;
                uint32_t i = t->order[s];
                float th, d2;
                if (orc_disk_test(o, d, tMax, ph->pos + 3 * (size_t)i, ph->radius[i], &th, &d2)) {
                    float tr[3];
                    ray_tr_eval(&trs, th, tr);
                    accumulate_hit(ph, i, d, th, d2, tr, m->g, nPathsNorm, rgb);
                    if (hits && nh < hits_cap) hits[nh] = i;
                    ++nh;
                }
            }
            if (sp == 0) break;
            cur = stack[--sp];
        } else {
            const orc_node *nd = &t->nodes[cur];
            if (slab_hit(&rs, nd->lo, nd->hi)) {
                /* near child first, photonbeambvh.cpp:709-715 */
                if (rs.neg[nd->axis]) { stack[sp++] = nd->left; cur = nd->right; }
                else { stack[sp++] = nd->right; cur = nd->left; }
            } else {
                if (sp == 0) break;
                cur = stack[--sp];
            }
        }
    }
    if (c) { c->nodes_visited += V; c->disk_tests += T; c->hits += nh; }
    return nh;
}

uint32_t orc_query_ray_brute(const orc_photons *ph, const orc_medium *m, const float o[3],
                             const float d[3], float tMax, float nPathsNorm, float rgb[3],
                             uint32_t *hits, uint32_t hits_cap) {
    rgb[0] = rgb[1] = rgb[2] = 0;
    uint32_t nh = 0;
    ray_tr_state trs;
    ray_tr_init(&trs, m, o, d, tMax);
    for (uint32_t i = 0; i < ph->n; ++i) {
        float th, d2;
        if (orc_disk_test(o, d, tMax, ph->pos + 3 * (size_t)i, ph->radius[i], &th, &d2)) {
            float tr[3];
            ray_tr_eval(&trs, th, tr);
            accumulate_hit(ph, i, d, th, d2, tr, m->g, nPathsNorm, rgb);
            if (hits && nh < hits_cap) hits[nh] = i;
            ++nh;
        }
    }
    return nh;
}

static int cmp_u32(const void *a, const void *b) {
    uint32_t x = *(const uint32_t *)a, y = *(const uint32_t *)b;
    return x < y ? -1 : (x > y ? 1 : 0);
}

/* Parallel-for over rays with plain pthreads (chunks of 64 rays claimed from an atomic counter). */
typedef struct {
    const orc_lbvh *t; const orc_photons *ph; const orc_medium *m;
    uint32_t nRays; const float *o_tmax; const float *dirs; float nPathsNorm; float *rgb_out;
    uint64_t *hit_offsets; uint32_t *hit_indices; uint32_t *counts;
    int pass;
    atomic_long next;
    atomic_ullong V, T, H;
} batch_job;

static void *batch_worker(void *arg) {
    batch_job *j = (batch_job *)arg;
    uint64_t V = 0, T = 0, H = 0;
    for (;;) {
        long q0 = atomic_fetch_add(&j->next, 64);
        if (q0 >= (long)j->nRays) break;
        long q1 = q0 + 64 < (long)j->nRays ? q0 + 64 : (long)j->nRays;
        for (long q = q0; q < q1; ++q) {
            const float *o = j->o_tmax + 4 * q, *d = j->dirs + 3 * q;
            float tMax = j->o_tmax[4 * q + 3], rgb[3];
            if (j->pass == 1) {
                orc_counters lc = {0, 0, 0};
                uint32_t nh;
                if (j->t) nh = orc_query_ray(j->t, j->ph, j->m, o, d, tMax, j->nPathsNorm, rgb, NULL, 0, &lc);
                else nh = orc_query_ray_brute(j->ph, j->m, o, d, tMax, j->nPathsNorm, rgb, NULL, 0);
                if (j->rgb_out) memcpy(j->rgb_out + 3 * q, rgb, 12);
                if (j->counts) j->counts[q] = nh;
                V += lc.nodes_visited; T += lc.disk_tests; H += nh;
            } else {
                uint32_t cap = (uint32_t)(j->hit_offsets[q + 1] - j->hit_offsets[q]);
                uint32_t *dst = j->hit_indices + j->hit_offsets[q];
                if (j->t) orc_query_ray(j->t, j->ph, j->m, o, d, tMax, j->nPathsNorm, rgb, dst, cap, NULL);
                else orc_query_ray_brute(j->ph, j->m, o, d, tMax, j->nPathsNorm, rgb, dst, cap);
                qsort(dst, cap, sizeof(uint32_t), cmp_u32);
            }
        }
    }
    atomic_fetch_add(&j->V, V); atomic_fetch_add(&j->T, T); atomic_fetch_add(&j->H, H);
    return NULL;
}

static void batch_run(batch_job *j, int nThreads) {
    if (nThreads <= 0) nThreads = (int)sysconf(_SC_NPROCESSORS_ONLN);
    if (nThreads < 1) nThreads = 1;
    if (nThreads > 256) nThreads = 256;
    atomic_store(&j->next, 0);
    pthread_t th[256];
    for (int i = 1; i < nThreads; ++i) pthread_create(&th[i], NULL, batch_worker, j);
    batch_worker(j);
    for (int i = 1; i < nThreads; ++i) pthread_join(th[i], NULL);
}

static int query_batch_impl(const orc_lbvh *t, const orc_photons *ph, const orc_medium *m,
                            uint32_t nRays, const float *o_tmax, const float *dirs,
                            float nPathsNorm, float *rgb_out, uint64_t *hit_offsets,
                            uint32_t *hit_indices, uint64_t hit_cap, orc_counters *c,
                            int nThreads) {
    batch_job j;
    memset(&j, 0, sizeof(j));
    j.t = t; j.ph = ph; j.m = m; j.nRays = nRays; j.o_tmax = o_tmax; j.dirs = dirs;
    j.nPathsNorm = nPathsNorm; j.rgb_out = rgb_out; j.hit_offsets = hit_offsets;
    j.hit_indices = hit_indices;
    int want_hits = hit_offsets != NULL;
    if (want_hits) j.counts = (uint32_t *)calloc(nRays ? nRays : 1, sizeof(uint32_t));
    j.pass = 1; /* radiance + counts */
    batch_run(&j, nThreads);
    if (c) { c->nodes_visited += j.V; c->disk_tests += j.T; c->hits += j.H; }
    if (!want_hits) return 0;
    hit_offsets[0] = 0;
    for (uint32_t q = 0; q < nRays; ++q) hit_offsets[q + 1] = hit_offsets[q] + j.counts[q];
    free(j.counts); j.counts = NULL;
    if (hit_offsets[nRays] > hit_cap) return -1;
    if (!hit_indices) return 0;
    j.pass = 2; /* indices, sorted per ray */
    batch_run(&j, nThreads);
    return 0;
}

int orc_query_batch(const orc_lbvh *t, const orc_photons *ph, const orc_medium *m, uint32_t nRays,
                    const float *o_tmax, const float *dirs, float nPathsNorm, float *rgb_out,
                    uint64_t *hit_offsets, uint32_t *hit_indices, uint64_t hit_cap,
                    orc_counters *c, int nThreads) {
    return query_batch_impl(t, ph, m, nRays, o_tmax, dirs, nPathsNorm, rgb_out, hit_offsets,
                            hit_indices, hit_cap, c, nThreads);
}
int orc_query_batch_brute(const orc_photons *ph, const orc_medium *m, uint32_t nRays,
                          const float *o_tmax, const float *dirs, float nPathsNorm, float *rgb_out,
                          uint64_t *hit_offsets, uint32_t *hit_indices, uint64_t hit_cap,
                          int nThreads) {
    return query_batch_impl(NULL, ph, m, nRays, o_tmax, dirs, nPathsNorm, rgb_out, hit_offsets,
                            hit_indices, hit_cap, NULL, nThreads);
}

/* -------------------------------------------------------------------- film */
void orc_film_accumulate(float *ld, uint32_t nRays, const uint32_t *pixel, const float *rgb) {
    /* pixel.Ld += ..., photonbeam.cpp:504 */
    for (uint32_t q = 0; q < nRays; ++q) {
        size_t p = pixel ? pixel[q] : q;
        for (int c = 0; c < 3; ++c) ld[3 * p + c] += rgb[3 * (size_t)q + c];
    }
}
void orc_film_resolve(const float *ld, uint32_t nPixels, int iter, float scale, float *out) {
    for (uint32_t i = 0; i < nPixels; ++i) {
        /* L = Ld / (iter+1)  photonbeam.cpp:578; CoefficientSpectrum::operator/(Float) divides every channel
         * (spectrum.h:181-188: ret.c[i] /= a), it does not multiply by a reciprocal */
        const float div = (float)(iter + 1);
        float rgb[3] = {ld[3 * (size_t)i] / div, ld[3 * (size_t)i + 1] / div, ld[3 * (size_t)i + 2] / div};
        /* Film::SetImage: RGB -> XYZ, weight 1  film.cpp:132-140, spectrum.h:62-66 */
        float xyz[3];
        xyz[0] = 0.412453f * rgb[0] + 0.357580f * rgb[1] + 0.180423f * rgb[2];
        xyz[1] = 0.212671f * rgb[0] + 0.715160f * rgb[1] + 0.072169f * rgb[2];
        xyz[2] = 0.019334f * rgb[0] + 0.119193f * rgb[1] + 0.950227f * rgb[2];
        /* Film::WriteImage: XYZ -> RGB, /weight (=1), max(0,.), splat (=0), *scale  film.cpp:168-204 */
        float o[3];
        o[0] = 3.240479f * xyz[0] - 1.537150f * xyz[1] - 0.498535f * xyz[2];
        o[1] = -0.969256f * xyz[0] + 1.875991f * xyz[1] + 0.041556f * xyz[2];
        o[2] = 0.055648f * xyz[0] - 0.204043f * xyz[1] + 1.057311f * xyz[2];
        for (int c = 0; c < 3; ++c) {
            float v = fmaxf_(0.f, o[c] * 1.f);
            v += 1.f * 0.f;
            out[3 * (size_t)i + c] = v * scale;
        }
    }
}

/* -------------------------------------------------------- photon tracing (K1) */
static inline float lum_y(const float c[3]) { /* RGBSpectrum::y spectrum.h */
    return 0.212671f * c[0] + 0.715160f * c[1] + 0.072169f * c[2];
}
static inline int is_black(const float c[3]) { return c[0] == 0 && c[1] == 0 && c[2] == 0; }

static float spot_falloff(const orc_light *L, const float w[3]) { /* spot.cpp:64-73 */
    float len = sqrtf(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
    float cosTheta = (w[0] * L->axis[0] + w[1] * L->axis[1] + w[2] * L->axis[2]) / len;
    if (cosTheta < L->cosTotalWidth) return 0;
    if (cosTheta > L->cosFalloffStart) return 1;
    float delta = (cosTheta - L->cosTotalWidth) / (L->cosFalloffStart - L->cosTotalWidth);
    return (delta * delta) * (delta * delta);
}
static float light_power_y(const orc_light *L) {
    float p[3];
    for (int c = 0; c < 3; ++c) {
        if (L->kind == 0) p[c] = 4 * ORC_PI * L->I[c]; /* point.cpp:55 */
        else p[c] = L->I[c] * 2 * ORC_PI * (1 - .5f * (L->cosFalloffStart + L->cosTotalWidth)); /* spot.cpp:75-77 */
    }
    return lum_y(p);
}

/* Medium::Sample restated for a segment [0,tMax) of a unit-direction ray.
 * Returns 1 when a scattering event was sampled at *t_out; weight multiplies alpha. */
static int medium_sample(const orc_medium *m, orc_rng *rng, const float o[3], const float d[3],
                         float tMax, float *t_out, float weight[3]) {
    if (m->kind == 0) { /* homogeneous.cpp:50-77 */
        int channel = (int)(orc_rng_float(rng) * 3);
        if (channel > 2) channel = 2;
        float dist = -logf(1 - orc_rng_float(rng)) / m->sigma_t[channel];
        float len = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        float t = fminf_(dist * len, tMax);
        int sampled = t < tMax;
        float Tr[3], density[3], pdf = 0;
        for (int c = 0; c < 3; ++c) {
            Tr[c] = expf(-m->sigma_t[c] * fminf_(t, 3.402823466e+38f) * len);
            density[c] = sampled ? m->sigma_t[c] * Tr[c] : Tr[c];
            pdf += density[c];
        }
        pdf *= 1 / 3.f;
        if (pdf == 0) pdf = 1;
        for (int c = 0; c < 3; ++c) weight[c] = sampled ? (Tr[c] * m->sigma_s[c] / pdf) : (Tr[c] / pdf);
        *t_out = t;
        return sampled;
    }
    /* grid.cpp:62-87 (delta tracking) */
    float om[3], dm[3];
    xf_point(m->w2m, o, om);
    xf_vector(m->w2m, d, dm);
    const float lo[3] = {0, 0, 0}, hi[3] = {1, 1, 1};
    float tMin, tEnd;
    weight[0] = weight[1] = weight[2] = 1.f;
    if (!orc_ray_box(om, dm, tMax, lo, hi, &tMin, &tEnd)) return 0;
    float t = tMin;
    float st = m->sigma_t[0];
    for (;;) {
        t -= logf(1 - orc_rng_float(rng)) * m->invMaxDensity / st;
        if (t >= tEnd) break;
        float p[3] = {om[0] + dm[0] * t, om[1] + dm[1] * t, om[2] + dm[2] * t};
        if (orc_grid_density(m, p) * m->invMaxDensity > orc_rng_float(rng)) {
            for (int c = 0; c < 3; ++c) weight[c] = m->sigma_s[c] / st;
            *t_out = t;
            return 1;
        }
    }
    return 0;
}

uint64_t orc_trace_photons(const orc_medium *m, const orc_light *lights, int nLights,
                           const float blo[3], const float bhi[3], uint64_t first, uint64_t nPaths,
                           int maxDepth, float *pos, float *power, float *dir, uint32_t *path_id,
                           uint64_t cap) {
    /* light power distribution: integrator.cpp:217-225, Distribution1D sampling.h:55-100 */
    float cdf[65];
    if (nLights > 64) nLights = 64;
    cdf[0] = 0;
    for (int i = 0; i < nLights; ++i) cdf[i + 1] = cdf[i] + light_power_y(&lights[i]) / nLights;
    float funcInt = cdf[nLights];
    if (funcInt == 0) for (int i = 1; i <= nLights; ++i) cdf[i] = (float)i / (float)nLights;
    else for (int i = 1; i <= nLights; ++i) cdf[i] /= funcInt;
    uint64_t stored = 0;
    for (uint64_t pi = first; pi < first + nPaths; ++pi) {
        orc_rng rng;
        orc_rng_seed(&rng, pi + 1); /* photonbeam.cpp:386-391 */
        float uPick = orc_rng_float(&rng);
        /* SampleDiscrete sampling.h:90-100: FindInterval -> last i with cdf[i] <= u */
        int li = 0;
        for (int i = 0; i < nLights; ++i) if (cdf[i] <= uPick) li = i;
        float func = light_power_y(&lights[li]);
        float lightPdf = funcInt > 0 ? func / (funcInt * nLights) : 0;
        if (funcInt == 0) lightPdf = 1.f / nLights;
        float u0 = orc_rng_float(&rng), u1 = orc_rng_float(&rng);
        (void)orc_rng_float(&rng); (void)orc_rng_float(&rng); /* uLight1 */
        (void)orc_rng_float(&rng);                            /* uLightTime */
        const orc_light *L = &lights[li];
        float o[3] = {L->pos[0], L->pos[1], L->pos[2]}, d[3], Le[3], pdfDir;
        if (L->kind == 0) { /* point.cpp:61-71 */
            orc_uniform_sample_sphere(u0, u1, d);
            pdfDir = ORC_INV4PI;
            for (int c = 0; c < 3; ++c) Le[c] = L->I[c];
        } else { /* spot.cpp:83-93 */
            float w[3], ax[3], ay[3];
            orc_uniform_sample_cone(u0, u1, L->cosTotalWidth, w);
            coordinate_system(L->axis, ax, ay);
            for (int k = 0; k < 3; ++k) d[k] = w[0] * ax[k] + w[1] * ay[k] + w[2] * L->axis[k];
            pdfDir = 1 / (2 * ORC_PI * (1 - L->cosTotalWidth));
            float f = spot_falloff(L, d);
            for (int c = 0; c < 3; ++c) Le[c] = L->I[c] * f;
        }
        if (pdfDir == 0 || is_black(Le)) continue;
        float nd = fabsf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]); /* AbsDot(nLight, d), nLight = d */
        float alpha[3];
        float denom = lightPdf * 1.f * pdfDir; /* photonbeam.cpp:413-414 */
        for (int c = 0; c < 3; ++c) alpha[c] = (nd * Le[c]) / denom;
        if (is_black(alpha)) continue;
        for (int nInt = 1;; ++nInt) {
            /* scene.Intersect stand-in: distance to the absorbing box wall */
            float t0, t1;
            if (!orc_ray_box(o, d, INFINITY, blo, bhi, &t0, &t1) || t0 > 0) break;
            float tMax = t1, ts = 0, w[3];
            int scattered = medium_sample(m, &rng, o, d, tMax, &ts, w);
            for (int c = 0; c < 3; ++c) alpha[c] *= w[c];
            if (is_black(alpha)) break;
            if (!scattered) break; /* reached the wall: absorbed */
            float p[3] = {o[0] + d[0] * ts, o[1] + d[1] * ts, o[2] + d[2] * ts};
            float wo[3] = {-d[0], -d[1], -d[2]}, wi[3];
            /* Photon(misect.p, alpha, wo)  photonmap.cpp:720-723 */
            if (stored < cap) {
                memcpy(pos + 3 * stored, p, 12);
                memcpy(power + 3 * stored, alpha, 12);
                memcpy(dir + 3 * stored, wo, 12);
                if (path_id) path_id[stored] = (uint32_t)pi;
            }
            ++stored;
            float ua = orc_rng_float(&rng), ub = orc_rng_float(&rng);
            orc_hg_sample_p(wo, m->g, ua, ub, wi);
            memcpy(o, p, 12); memcpy(d, wi, 12); /* mi.SpawnRay(wi): origin exactly p */
            if (nInt >= maxDepth) break; /* photonmap.cpp:794 */
            /* Russian roulette photonmap.cpp:799-804: alphaNew == alpha => continueProb == 1 */
            float sample = orc_rng_float(&rng);
            if (sample > 1.f) break;
        }
    }
    return stored;
}

/* ------------------------------------------------------------ synthetic data */
void orc_synth_photons_uniform(uint32_t n, float *pos, float *power, float *dir) {
    orc_rng r1, r2, r3;
    orc_rng_seed(&r1, 1); orc_rng_seed(&r2, 2); orc_rng_seed(&r3, 3);
    float invN = 1.f / (float)n;
    for (uint32_t i = 0; i < n; ++i) {
        for (int k = 0; k < 3; ++k) pos[3 * (size_t)i + k] = orc_rng_float(&r1);
        float u0 = orc_rng_float(&r2), u1 = orc_rng_float(&r2);
        orc_uniform_sample_sphere(u0, u1, dir + 3 * (size_t)i);
        for (int k = 0; k < 3; ++k) power[3 * (size_t)i + k] = orc_rng_float(&r3) * invN;
    }
}
void orc_synth_rays_pinhole(int w, int h, float *o_tmax, float *dirs) {
    /* pinhole at (0.5,0.5,-1.5) looking +z, vertical fov 40 deg, pixel centres;
     * tMax = exit of the unit cube (0 if missed) */
    const float eye[3] = {0.5f, 0.5f, -1.5f};
    const float lo[3] = {0, 0, 0}, hi[3] = {1, 1, 1};
    float tanHalf = tanf(0.5f * 40.f * ORC_PI / 180.f);
    float aspect = (float)w / (float)h;
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            size_t q = (size_t)y * w + x;
            float sx = (2.f * (((float)x + 0.5f) / (float)w) - 1.f) * tanHalf * aspect;
            float sy = (1.f - 2.f * (((float)y + 0.5f) / (float)h)) * tanHalf;
            float d[3] = {sx, sy, 1.f};
            float inv = 1.f / sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
            for (int k = 0; k < 3; ++k) d[k] *= inv;
            float t0, t1;
            float tMax = 0;
            if (orc_ray_box(eye, d, INFINITY, lo, hi, &t0, &t1)) tMax = t1;
            memcpy(o_tmax + 4 * q, eye, 12);
            o_tmax[4 * q + 3] = tMax;
            memcpy(dirs + 3 * q, d, 12);
        }
}
void orc_synth_rays_incoherent(uint32_t n, float *o_tmax, float *dirs) {
    orc_rng r;
    orc_rng_seed(&r, 4);
    for (uint32_t q = 0; q < n; ++q) {
        for (int k = 0; k < 3; ++k) o_tmax[4 * (size_t)q + k] = orc_rng_float(&r);
        float u0 = orc_rng_float(&r), u1 = orc_rng_float(&r);
        orc_uniform_sample_sphere(u0, u1, dirs + 3 * (size_t)q);
        const float lo[3] = {0, 0, 0}, hi[3] = {1, 1, 1};
        float t0, t1, tMax = 0;
        if (orc_ray_box(o_tmax + 4 * (size_t)q, dirs + 3 * (size_t)q, INFINITY, lo, hi, &t0, &t1)) tMax = t1;
        o_tmax[4 * (size_t)q + 3] = tMax;
    }
}

/* ===================================================================================
 * Surfaces: triangle intersector, Lambertian bounce, camera rays.
 * Restates, in float32 with the reference's operation order:
 *   Triangle::Intersect            shapes/triangle.cpp:177-392 (watertight test, hit point, error bounds)
 *   OffsetRayOrigin / SpawnRay     core/geometry.h:1438-1458, core/interaction.h:64-67
 *   GeometricPrimitive::Intersect  core/primitive.cpp:97-109 (medium interface of the hit)
 *   MatteMaterial (sigma == 0)     materials/matte.cpp:55-60 -> LambertianReflection
 *   BSDF::Sample_f / BxDF::Sample_f core/reflection.cpp:703-760, :378-389
 *   ConcentricSampleDisk / CosineSampleHemisphere  core/sampling.cpp:113-130, sampling.h:159-163
 *   PerspectiveCamera::GenerateRayDifferential     cameras/perspective.cpp:95-103,
 *   Transform::operator()(Ray)     core/transform.h:251-264 (+ :278-300 error bound of the origin)
 * =================================================================================== */
static inline float dot3(const float a[3], const float b[3]) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
static inline void cross3(const float a[3], const float b[3], float out[3]) { /* geometry.h Cross: in double */
    double ax = a[0], ay = a[1], az = a[2], bx = b[0], by = b[1], bz = b[2];
    out[0] = (float)((ay * bz) - (az * by));
    out[1] = (float)((az * bx) - (ax * bz));
    out[2] = (float)((ax * by) - (ay * bx));
}
static inline void normalize3(float v[3]) { /* v / Length(): multiplies by the reciprocal */
    float inv = 1.f / sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    v[0] *= inv; v[1] *= inv; v[2] *= inv;
}
static inline float next_up_(float x) { return nextafterf(x, INFINITY); }
static inline float next_down_(float x) { return nextafterf(x, -INFINITY); }
static inline float max3abs(float a, float b, float c) {
    a = fabsf(a); b = fabsf(b); c = fabsf(c);
    return a > b ? (a > c ? a : c) : (b > c ? b : c);
}

int orc_intersect_triangle(const float o[3], const float d[3], float tMax, const float p0[3],
                           const float p1[3], const float p2[3], float *tHit, float bary[3]) {
    /* translate to the ray origin, permute so that |d| is largest along z, shear (triangle.cpp:191-215) */
    float ad[3] = {fabsf(d[0]), fabsf(d[1]), fabsf(d[2])};
    int kz = ad[0] > ad[1] ? (ad[0] > ad[2] ? 0 : 2) : (ad[1] > ad[2] ? 1 : 2); /* MaxDimension */
    int kx = kz + 1 == 3 ? 0 : kz + 1, ky = kx + 1 == 3 ? 0 : kx + 1;
    float dx = d[kx], dy = d[ky], dz = d[kz];
    float a[3] = {p0[kx] - o[kx], p0[ky] - o[ky], p0[kz] - o[kz]};
    float b[3] = {p1[kx] - o[kx], p1[ky] - o[ky], p1[kz] - o[kz]};
    float c[3] = {p2[kx] - o[kx], p2[ky] - o[ky], p2[kz] - o[kz]};
    float Sx = -dx / dz, Sy = -dy / dz, Sz = 1.f / dz;
    a[0] += Sx * a[2]; a[1] += Sy * a[2];
    b[0] += Sx * b[2]; b[1] += Sy * b[2];
    c[0] += Sx * c[2]; c[1] += Sy * c[2];
    /* edge functions, double fallback on exact zeros (:218-234) */
    float e0 = b[0] * c[1] - b[1] * c[0];
    float e1 = c[0] * a[1] - c[1] * a[0];
    float e2 = a[0] * b[1] - a[1] * b[0];
    if (e0 == 0.f || e1 == 0.f || e2 == 0.f) {
        e0 = (float)((double)c[1] * (double)b[0] - (double)c[0] * (double)b[1]);
        e1 = (float)((double)a[1] * (double)c[0] - (double)a[0] * (double)c[1]);
        e2 = (float)((double)b[1] * (double)a[0] - (double)b[0] * (double)a[1]);
    }
    if ((e0 < 0 || e1 < 0 || e2 < 0) && (e0 > 0 || e1 > 0 || e2 > 0)) return 0;
    float det = e0 + e1 + e2;
    if (det == 0) return 0;
    a[2] *= Sz; b[2] *= Sz; c[2] *= Sz;
    float tScaled = e0 * a[2] + e1 * b[2] + e2 * c[2];
    if (det < 0 && (tScaled >= 0 || tScaled < tMax * det)) return 0;
    else if (det > 0 && (tScaled <= 0 || tScaled > tMax * det)) return 0;
    float invDet = 1 / det;
    float t = tScaled * invDet;
    /* conservative t > 0 check (:262-280) */
    float maxZt = max3abs(a[2], b[2], c[2]);
    float deltaZ = orc_gamma(3) * maxZt;
    float maxXt = max3abs(a[0], b[0], c[0]);
    float maxYt = max3abs(a[1], b[1], c[1]);
    float deltaX = orc_gamma(5) * (maxXt + maxZt);
    float deltaY = orc_gamma(5) * (maxYt + maxZt);
    float deltaE = 2 * (orc_gamma(2) * maxXt * maxYt + deltaY * maxXt + deltaX * maxYt);
    float maxE = max3abs(e0, e1, e2);
    float deltaT = 3 * (orc_gamma(3) * maxE * maxZt + deltaE * maxZt + deltaZ * maxE) * fabsf(invDet);
    if (t <= deltaT) return 0;
    bary[0] = e0 * invDet; bary[1] = e1 * invDet; bary[2] = e2 * invDet;
    *tHit = t;
    return 1;
}

static void mesh_tri(const orc_mesh *mesh, uint32_t i, const float **p0, const float **p1, const float **p2) {
    *p0 = mesh->P + 3 * (size_t)mesh->idx[3 * (size_t)i];
    *p1 = mesh->P + 3 * (size_t)mesh->idx[3 * (size_t)i + 1];
    *p2 = mesh->P + 3 * (size_t)mesh->idx[3 * (size_t)i + 2];
}

int orc_mesh_intersect(const orc_mesh *mesh, const float o[3], const float d[3], float tMax,
                       orc_surface_hit *hit) {
    /* Scene::Intersect (core/scene.cpp:50-54): closest hit; every primitive is tested against the
     * ray's CURRENT tMax (core/primitive.cpp:97-101), here in index order instead of BVH order. */
    int found = -1;
    float bestB[3] = {0, 0, 0};
    for (uint32_t i = 0; i < mesh->n_tri; ++i) {
        const float *p0, *p1, *p2;
        float t, b[3];
        mesh_tri(mesh, i, &p0, &p1, &p2);
        if (orc_intersect_triangle(o, d, tMax, p0, p1, p2, &t, b)) {
            tMax = t; found = (int)i; memcpy(bestB, b, 12);
        }
    }
    if (found < 0) return 0;
    if (!hit) return 1;
    const float *p0, *p1, *p2;
    mesh_tri(mesh, (uint32_t)found, &p0, &p1, &p2);
    hit->t = tMax;
    hit->tri = (uint32_t)found;
    for (int k = 0; k < 3; ++k) {
        /* pHit = b0*p0 + b1*p1 + b2*p2, pError = gamma(7) * sum |b_i p_i| (:302-311) */
        hit->p[k] = bestB[0] * p0[k] + bestB[1] * p1[k] + bestB[2] * p2[k];
        hit->pError[k] = orc_gamma(7) * (fabsf(bestB[0] * p0[k]) + fabsf(bestB[1] * p1[k]) + fabsf(bestB[2] * p2[k]));
    }
    float dp02[3], dp12[3], dpdu[3];
    for (int k = 0; k < 3; ++k) { dp02[k] = p0[k] - p2[k]; dp12[k] = p1[k] - p2[k]; }
    /* default uv (0,0),(1,0),(1,1): dpdu = (duv12[1]*dp02 - duv02[1]*dp12) * 1 = -dp02 + dp12 (:288-295) */
    for (int k = 0; k < 3; ++k) dpdu[k] = (-1.f * dp02[k] - -1.f * dp12[k]) * 1.f;
    cross3(dp02, dp12, hit->n);
    normalize3(hit->n); /* :328 */
    uint8_t attr = mesh->attr ? mesh->attr[found] : 0;
    if (attr & ORC_TRI_FLIP_NORMAL) for (int k = 0; k < 3; ++k) hit->n[k] = -hit->n[k]; /* :387-388 */
    memcpy(hit->ss, dpdu, 12);
    normalize3(hit->ss); /* BSDF::ss = Normalize(shading.dpdu), reflection.h BSDF ctor */
    return 1;
}

void orc_offset_ray_origin(const float p[3], const float pError[3], const float n[3], const float w[3],
                           float out[3]) {
    /* geometry.h:1438-1458 */
    float dd = fabsf(n[0]) * pError[0] + fabsf(n[1]) * pError[1] + fabsf(n[2]) * pError[2];
    float off[3] = {dd * n[0], dd * n[1], dd * n[2]};
    if (dot3(w, n) < 0) for (int k = 0; k < 3; ++k) off[k] = -off[k];
    for (int k = 0; k < 3; ++k) {
        float po = p[k] + off[k];
        if (off[k] > 0) po = next_up_(po);
        else if (off[k] < 0) po = next_down_(po);
        out[k] = po;
    }
}

/* medium state of a ray leaving triangle `attr` towards w: Interaction::GetMedium (interaction.h:86-88)
 * over the interface chosen by GeometricPrimitive::Intersect (primitive.cpp:104-109) */
static int medium_after(uint8_t attr, const float w[3], const float n[3], int current) {
    int side = attr & 3;
    if (side == ORC_TRI_MEDIUM_INHERIT) return current;
    int outside = dot3(w, n) > 0;
    return side == ORC_TRI_MEDIUM_OUTSIDE ? outside : !outside;
}

void orc_concentric_sample_disk(float u0, float u1, float out[2]) { /* sampling.cpp:113-130 */
    float ux = 2.f * u0 - 1, uy = 2.f * u1 - 1;
    if (ux == 0 && uy == 0) { out[0] = out[1] = 0; return; }
    float theta, r;
    if (fabsf(ux) > fabsf(uy)) { r = ux; theta = (ORC_PI / 4) * (uy / ux); }
    else { r = uy; theta = (ORC_PI / 2) - (ORC_PI / 4) * (ux / uy); }
    out[0] = r * cosf(theta); out[1] = r * sinf(theta);
}

/* BSDF::Sample_f for a single LambertianReflection lobe of reflectance R (reflection.cpp:703-760).
 * Returns 0 when the sample carries nothing (f black or pdf 0); else fills wi (world), f, pdf. */
int orc_lambert_sample(const orc_surface_hit *h, const float R[3], const float woW[3], float u0,
                       float u1, float wiW[3], float f[3], float *pdf) {
    if (R[0] == 0 && R[1] == 0 && R[2] == 0) return 0; /* matte.cpp:57: no lobe added */
    const float *ns = h->n, *ss = h->ss;
    float ts[3];
    cross3(ns, ss, ts);
    float wo[3] = {dot3(woW, ss), dot3(woW, ts), dot3(woW, ns)};
    if (wo[2] == 0) return 0;
    float dsk[2];
    orc_concentric_sample_disk(fminf_(u0 * 1 - 0, ORC_ONE_MINUS_EPS), u1, dsk);
    float z = sqrtf(fmaxf_(0.f, 1 - dsk[0] * dsk[0] - dsk[1] * dsk[1]));
    float wi[3] = {dsk[0], dsk[1], z};
    if (wo[2] < 0) wi[2] *= -1;
    *pdf = (wo[2] * wi[2] > 0) ? fabsf(wi[2]) * ORC_INV_PI : 0;
    if (*pdf == 0) return 0;
    for (int k = 0; k < 3; ++k) wiW[k] = ss[k] * wi[0] + ts[k] * wi[1] + ns[k] * wi[2];
    /* one matching lobe: the geometric-normal reflect/transmit filter of :756-765 is skipped */
    for (int c = 0; c < 3; ++c) f[c] = R[c] * ORC_INV_PI;
    return 1;
}

uint64_t orc_trace_photons_mesh(const orc_medium *m, const orc_light *lights, int nLights,
                                const orc_mesh *mesh, int lightsInMedium, uint64_t first,
                                uint64_t nPaths, int maxDepth, float *pos, float *power, float *dir,
                                uint32_t *path_id, uint64_t cap) {
    /* Path loop of TracePhotonBeamRecursive (integrators/photonbeam.cpp:258-325) unrolled into the
     * iteration of the pbrt-side adapter's host tracer, with the volume-photon deposit rule of
     * photonmap.cpp:696-724: emission as orc_trace_photons; then per segment Scene::Intersect,
     * Medium::Sample, deposit + phase scatter, or surface bounce + Russian roulette (:296-323). */
    float cdf[65];
    if (nLights > 64) nLights = 64;
    cdf[0] = 0;
    for (int i = 0; i < nLights; ++i) cdf[i + 1] = cdf[i] + light_power_y(&lights[i]) / nLights;
    float funcInt = cdf[nLights];
    if (funcInt == 0) for (int i = 1; i <= nLights; ++i) cdf[i] = (float)i / (float)nLights;
    else for (int i = 1; i <= nLights; ++i) cdf[i] /= funcInt;
    uint64_t stored = 0;
    for (uint64_t pi = first; pi < first + nPaths; ++pi) {
        orc_rng rng;
        orc_rng_seed(&rng, pi + 1);
        float uPick = orc_rng_float(&rng);
        int li = 0;
        for (int i = 0; i < nLights; ++i) if (cdf[i] <= uPick) li = i;
        float func = light_power_y(&lights[li]);
        float lightPdf = funcInt > 0 ? func / (funcInt * nLights) : 1.f / nLights;
        float u0 = orc_rng_float(&rng), u1 = orc_rng_float(&rng);
        (void)orc_rng_float(&rng); (void)orc_rng_float(&rng);
        (void)orc_rng_float(&rng);
        const orc_light *L = &lights[li];
        float o[3] = {L->pos[0], L->pos[1], L->pos[2]}, d[3], Le[3], pdfDir;
        if (L->kind == 0) {
            orc_uniform_sample_sphere(u0, u1, d);
            pdfDir = ORC_INV4PI;
            for (int c = 0; c < 3; ++c) Le[c] = L->I[c];
        } else {
            float w[3], ax[3], ay[3];
            orc_uniform_sample_cone(u0, u1, L->cosTotalWidth, w);
            coordinate_system(L->axis, ax, ay);
            for (int k = 0; k < 3; ++k) d[k] = w[0] * ax[k] + w[1] * ay[k] + w[2] * L->axis[k];
            pdfDir = 1 / (2 * ORC_PI * (1 - L->cosTotalWidth));
            float f = spot_falloff(L, d);
            for (int c = 0; c < 3; ++c) Le[c] = L->I[c] * f;
        }
        if (pdfDir == 0 || is_black(Le)) continue;
        float nd = fabsf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        float beta[3];
        float denom = lightPdf * 1.f * pdfDir;
        for (int c = 0; c < 3; ++c) beta[c] = (nd * Le[c]) / denom;
        if (is_black(beta)) continue;
        int inMedium = lightsInMedium;
        for (int depth = 0; depth < maxDepth; ++depth) {
            orc_surface_hit hit;
            if (!orc_mesh_intersect(mesh, o, d, INFINITY, &hit)) break; /* escaped the scene */
            int scattered = 0;
            float ts = 0, w[3];
            if (inMedium) {
                scattered = medium_sample(m, &rng, o, d, hit.t, &ts, w);
                for (int c = 0; c < 3; ++c) beta[c] *= w[c];
            }
            if (is_black(beta)) break;
            if (scattered) {
                float p[3] = {o[0] + d[0] * ts, o[1] + d[1] * ts, o[2] + d[2] * ts};
                float wo[3] = {-d[0], -d[1], -d[2]}, wi[3];
                if (stored < cap) {
                    memcpy(pos + 3 * stored, p, 12);
                    memcpy(power + 3 * stored, beta, 12);
                    memcpy(dir + 3 * stored, wo, 12);
                    if (path_id) path_id[stored] = (uint32_t)pi;
                }
                ++stored;
                float ua = orc_rng_float(&rng), ub = orc_rng_float(&rng);
                orc_hg_sample_p(wo, m->g, ua, ub, wi);
                memcpy(o, p, 12); memcpy(d, wi, 12);
                if (depth + 1 >= maxDepth) break;
                (void)orc_rng_float(&rng); /* roulette draw, continue probability 1 */
                continue;
            }
            uint8_t attr = mesh->attr ? mesh->attr[hit.tri] : 0;
            const float *Kd = mesh->Kd + 3 * (size_t)hit.tri;
            if ((attr & ORC_TRI_NO_MATERIAL) || Kd[0] < 0) { /* photonbeam.cpp:300-304 */
                --depth;
                float no[3];
                orc_offset_ray_origin(hit.p, hit.pError, hit.n, d, no);
                inMedium = medium_after(attr, d, hit.n, inMedium);
                memcpy(o, no, 12);
                continue;
            }
            float R[3], woW[3] = {-d[0], -d[1], -d[2]}, wiW[3], f[3], pdf;
            for (int c = 0; c < 3; ++c) R[c] = Kd[c] < 0 ? 0 : (Kd[c] > INFINITY ? INFINITY : Kd[c]); /* Clamp() */
            float ub0 = orc_rng_float(&rng), ub1 = orc_rng_float(&rng);
            if (!orc_lambert_sample(&hit, R, woW, ub0, ub1, wiW, f, &pdf)) break;
            float betaNew[3], absdot = fabsf(dot3(wiW, hit.n));
            for (int c = 0; c < 3; ++c) betaNew[c] = beta[c] * f[c] * absdot / pdf;
            float no[3];
            orc_offset_ray_origin(hit.p, hit.pError, hit.n, wiW, no);
            inMedium = medium_after(attr, wiW, hit.n, inMedium);
            memcpy(o, no, 12); memcpy(d, wiW, 12);
            float q = fmaxf_(0.f, 1 - lum_y(betaNew) / lum_y(beta));
            if (orc_rng_float(&rng) < q) break;
            for (int c = 0; c < 3; ++c) beta[c] = betaNew[c] / (1 - q);
        }
    }
    return stored;
}

/* ---- camera samples of the reference's HaltonSampler (samplers/halton.cpp) ----------------------------------
 * The camera pass draws the film position of pixel p in iteration `iter` as p + (dimension 0, dimension 1) of a
 * HaltonSampler(nIterations, pixelBounds) after StartPixel(p), SetSampleNumber(iter) (integrators/photonbeam.cpp:341,
 * 459-466; GetCameraSample core/sampler.cpp:46-52; the AwesomeSampler wrapper passes the first 1000 draws through,
 * photonbeam.cpp:199-213).  Restated: constructor halton.cpp:64-91, GetIndexForSample :94-114, SampleDimension
 * :116-125 (dimensions 0 and 1 are not scrambled), InverseRadicalInverse core/lowdiscrepancy.h:82-91,
 * RadicalInverse core/lowdiscrepancy.cpp:437-447 with RadicalInverseSpecialized<3> :395-408. */
static int64_t orc_mod64(int64_t a, int64_t b) { int64_t r = a - (a / b) * b; return r < 0 ? r + b : r; } /* pbrt.h Mod */
static void orc_ext_gcd(uint64_t a, uint64_t b, int64_t *x, int64_t *y) {                                 /* halton.cpp:51-61 */
    if (b == 0) { *x = 1; *y = 0; return; }
    int64_t d = (int64_t)(a / b), xp, yp;
    orc_ext_gcd(b, a % b, &xp, &yp);
    *x = yp;
    *y = xp - (d * yp);
}
static uint64_t orc_mult_inverse(int64_t a, int64_t n) { int64_t x, y; orc_ext_gcd((uint64_t)a, (uint64_t)n, &x, &y); return (uint64_t)orc_mod64(x, n); }
static uint64_t orc_inv_radical_inverse(uint64_t inverse, int base, int nDigits) {
    uint64_t index = 0;
    for (int i = 0; i < nDigits; ++i) { uint64_t digit = inverse % (uint64_t)base; inverse /= (uint64_t)base; index = index * (uint64_t)base + digit; }
    return index;
}
static uint64_t orc_reverse_bits64(uint64_t n) {
    uint64_t r = 0;
    for (int i = 0; i < 64; ++i) { r = (r << 1) | (n & 1u); n >>= 1; }
    return r;
}
static float orc_radical_inverse2(uint64_t a) { return (float)((double)orc_reverse_bits64(a) * 0x1p-64); } /* uint64 * double, narrowed */
static float orc_radical_inverse3(uint64_t a) {
    const float invBase = (float)1 / (float)3;
    uint64_t reversedDigits = 0;
    float invBaseN = 1;
    while (a) {
        uint64_t next = a / 3;
        uint64_t digit = a - next * 3;
        reversedDigits = reversedDigits * 3 + digit;
        invBaseN *= invBase;
        a = next;
    }
    float v = (float)reversedDigits * invBaseN;
    return v < 0x1.fffffep-1f ? v : 0x1.fffffep-1f;       /* min(., OneMinusEpsilon) */
}
void orc_halton_setup(int resX, int resY, orc_halton *hs) {
    const int kMaxResolution = 128;
    const int res[2] = {resX, resY};
    for (int i = 0; i < 2; ++i) {
        int base = (i == 0) ? 2 : 3, scale = 1, e = 0;
        const int lim = res[i] < kMaxResolution ? res[i] : kMaxResolution;
        while (scale < lim) { scale *= base; ++e; }
        hs->baseScales[i] = scale;
        hs->baseExponents[i] = e;
    }
    hs->sampleStride = (int64_t)hs->baseScales[0] * hs->baseScales[1];
    hs->multInverse[0] = orc_mult_inverse(hs->baseScales[1], hs->baseScales[0]);
    hs->multInverse[1] = orc_mult_inverse(hs->baseScales[0], hs->baseScales[1]);
}
void orc_halton_camera_samples(int x0, int y0, int x1, int y1, int64_t sampleNum, float *out_xy) {
    orc_halton hs;
    orc_halton_setup(x1 - x0, y1 - y0, &hs);
    for (int y = y0; y < y1; ++y)
        for (int x = x0; x < x1; ++x) {
            int64_t offset = 0;
            if (hs.sampleStride > 1) {
                const int64_t pm[2] = {orc_mod64(x, 128), orc_mod64(y, 128)};
                uint64_t acc = 0;
                for (int i = 0; i < 2; ++i) {
                    const uint64_t dimOffset = orc_inv_radical_inverse((uint64_t)pm[i], i == 0 ? 2 : 3, hs.baseExponents[i]);
                    acc += dimOffset * (uint64_t)(hs.sampleStride / hs.baseScales[i]) * hs.multInverse[i];
                }
                offset = (int64_t)(acc % (uint64_t)hs.sampleStride);
            }
            const int64_t index = offset + sampleNum * hs.sampleStride;
            const size_t q = (size_t)(y - y0) * (size_t)(x1 - x0) + (size_t)(x - x0);
            out_xy[2 * q] = orc_radical_inverse2((uint64_t)(index >> hs.baseExponents[0]));
            out_xy[2 * q + 1] = orc_radical_inverse3((uint64_t)(index / hs.baseScales[1]));
        }
}

void orc_camera_rays(const float r2c[16], const float c2w[16], int x0, int y0, int w, int h,
                     float offx, float offy, const orc_mesh *mesh, int cameraInMedium,
                     float *o_tmax, float *dirs) {
    orc_camera_rays_px(r2c, c2w, x0, y0, w, h, offx, offy, NULL, mesh, cameraInMedium, o_tmax, dirs);
}
/* offs_xy != NULL: the sample position of pixel q inside the pixel is (offs_xy[2q], offs_xy[2q+1]) */
void orc_camera_rays_px(const float r2c[16], const float c2w[16], int x0, int y0, int w, int h,
                        float offx, float offy, const float *offs_xy, const orc_mesh *mesh, int cameraInMedium,
                        float *o_tmax, float *dirs) {
    /* pinhole PerspectiveCamera (lensRadius 0), one ray per pixel of [x0,x0+w) x [y0,y0+h);
     * then what the adapter's camera pass does (integrators/photonbeam.cpp:465-481): Scene::Intersect
     * shortens tMax, direction normalised, tMax scaled by its length; tMax = 0 outside the medium. */
    float oErr[3];
    for (int r = 0; r < 3; ++r) /* transform.h:288-294 for p = (0,0,0) */
        oErr[r] = orc_gamma(3) * (fabsf(c2w[4 * r] * 0.f) + fabsf(c2w[4 * r + 1] * 0.f) + fabsf(c2w[4 * r + 2] * 0.f) + fabsf(c2w[4 * r + 3]));
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            size_t q = (size_t)y * w + x;
            if (offs_xy) { offx = offs_xy[2 * q]; offy = offs_xy[2 * q + 1]; }
            float fx = (float)(x0 + x) + offx, fy = (float)(y0 + y) + offy, fz = 0.f;
            float pc[3];
            for (int r = 0; r < 3; ++r) pc[r] = r2c[4 * r] * fx + r2c[4 * r + 1] * fy + r2c[4 * r + 2] * fz + r2c[4 * r + 3];
            float wp = r2c[12] * fx + r2c[13] * fy + r2c[14] * fz + r2c[15];
            if (wp != 1) { float inv = 1 / wp; pc[0] *= inv; pc[1] *= inv; pc[2] *= inv; }
            normalize3(pc);
            float o[3], d[3];
            for (int r = 0; r < 3; ++r) {
                o[r] = c2w[4 * r] * 0.f + c2w[4 * r + 1] * 0.f + c2w[4 * r + 2] * 0.f + c2w[4 * r + 3];
                d[r] = c2w[4 * r] * pc[0] + c2w[4 * r + 1] * pc[1] + c2w[4 * r + 2] * pc[2];
            }
            float l2 = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
            float tMax = INFINITY;
            if (l2 > 0) { /* transform.h:256-262 */
                float dt = (fabsf(d[0]) * oErr[0] + fabsf(d[1]) * oErr[1] + fabsf(d[2]) * oErr[2]) / l2;
                for (int r = 0; r < 3; ++r) o[r] += d[r] * dt;
                tMax -= dt;
            }
            orc_surface_hit hit;
            if (mesh && mesh->n_tri && orc_mesh_intersect(mesh, o, d, tMax, &hit)) tMax = hit.t;
            float len = sqrtf(l2), inv = 1 / len;
            o_tmax[4 * q] = o[0]; o_tmax[4 * q + 1] = o[1]; o_tmax[4 * q + 2] = o[2];
            o_tmax[4 * q + 3] = cameraInMedium ? fminf_(tMax * len, 3.402823466e+38f) : 0.f;
            dirs[3 * q] = d[0] * inv; dirs[3 * q + 1] = d[1] * inv; dirs[3 * q + 2] = d[2] * inv;
        }
}
