/* oracle/bre_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (plain C11, float32, compiled -ffp-contract=off) of the beam
 * radiance estimate hot path of bwiberg/beam-radiance-estimate-pbrt.  It is the
 * CHECKER for the CUDA library (libbre.so); nothing in the product path may call
 * it.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs load this.
 *
 * Parity status: the primitives (PCG32, sphere/cone sampling, HG phase function,
 * homogeneous/grid media, slab test, Morton code, 2D kernel, film conversion) are
 * pinned against the reference's own code by tests/golden/ref_primitives.json,
 * produced by oracle/ref_harness.cpp linked against the reference sources
 * (oracle/Makefile `ref`).  The photon-DISK beam query itself has NO
 * implementation in the reference: for that function
 * "parity unpinned" -- this file is the definition, built from the pinned pieces.
 *
 * All `file:line` citations are relative to /root/reference/src/.
 */
#ifndef BRE_ORACLE_H
#define BRE_ORACLE_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- random numbers: core/rng.h:60-145 ---- */
typedef struct { uint64_t state, inc; } orc_rng;
void orc_rng_seed(orc_rng *r, uint64_t sequence);     /* RNG::SetSequence rng.h:130-136 */
uint32_t orc_rng_u32(orc_rng *r);                     /* RNG::UniformUInt32 rng.h:138-144 */
float orc_rng_float(orc_rng *r);                      /* RNG::UniformFloat rng.h:78-85 */

/* ---- sampling: core/sampling.cpp:98-103,136-142 ---- */
void orc_uniform_sample_sphere(float u0, float u1, float out[3]);
void orc_uniform_sample_cone(float u0, float u1, float cosThetaMax, float out[3]);

/* ---- phase function: core/medium.h:66-69, core/medium.cpp:186-205 ---- */
float orc_phase_hg(float cosTheta, float g);
float orc_hg_sample_p(const float wo[3], float g, float u0, float u1, float wi[3]);

/* ---- media ---- */
typedef struct {
    int kind;              /* 0 = homogeneous, 1 = grid */
    float sigma_a[3], sigma_s[3], sigma_t[3];
    float g;
    /* grid only (media/grid.h:51-82) */
    int nx, ny, nz;
    const float *density;  /* nx*ny*nz, x fastest */
    float invMaxDensity;
    float w2m[16];         /* world -> medium, row major 4x4 (affine) */
    int tau_steps;         /* steps of the per-ray optical-depth table (orc_medium_grid sets orc_tau_steps_auto) */
} orc_medium;
void orc_medium_homogeneous(orc_medium *m, const float sa[3], const float ss[3], float g);
void orc_medium_grid(orc_medium *m, const float sa[3], const float ss[3], float g, int nx, int ny,
                     int nz, const float *density, const float w2m[16]);
float orc_grid_density(const orc_medium *m, const float p[3]);  /* media/grid.cpp:45-60 */
/* HomogeneousMedium::Tr, media/homogeneous.cpp:44-48, unit-length direction */
void orc_homogeneous_tr(const orc_medium *m, float t, float tr[3]);
/* Deterministic optical depth of the grid medium along o + s*d, s in [0,t], by the
 * midpoint rule with nSteps uniform steps over the ray's overlap with the medium
 * box (the reference's GridDensityMedium::Tr, media/grid.cpp:89-121, is stochastic
 * ratio tracking; see DESIGN.md "parity mode"). Fills tau[0..nSteps] (prefix sums,
 * tau[0]=0) and the overlap interval. Returns 0 if the ray misses the box. */
int orc_grid_tau_table(const orc_medium *m, const float o[3], const float d[3], float tMax,
                       int nSteps, float *tau, float *t0, float *t1);
int orc_tau_steps_auto(int nx, int ny, int nz);
void orc_grid_tr_table(const orc_medium *m, const float o[3], const float d[3], float tMax, float t,
                       int nSteps, float tr[3]);

/* ---- photon map container ---- */
typedef struct {
    uint32_t n;
    const float *pos;     /* n*3 */
    const float *power;   /* n*3  (Photon::alpha, integrators/photonmap.cpp:61-70) */
    const float *dir;     /* n*3  (Photon::wi) */
    const float *radius;  /* n */
} orc_photons;

/* ---- canonical LBVH (DESIGN.md section "canonical tree") ----
 * Morton code: core/photonbeambvh.cpp:109-139 + :431-445 (10 bits/axis of the
 * centroid offset in the centroid bounds).  Sort key = (code << 32) | index.
 * Topology = binary radix tree over the sorted keys, 1 photon per leaf.
 * Node layout mirrors LinearPBBVHNode (core/photonbeambvh.cpp:97-106). */
typedef struct {
    float lo[3], hi[3];
    int32_t left, right;    /* children: >=0 internal node index, <0 leaf ~sortedIndex */
    int32_t first, count;   /* covered range of sorted photons */
    int32_t axis;
} orc_node;
typedef struct {
    uint32_t n;
    uint32_t *order;        /* sorted position -> original photon index */
    uint32_t *codes;        /* sorted morton codes */
    orc_node *nodes;        /* n-1 internal nodes, root = 0 (n==1: none) */
    float *leaf_lo, *leaf_hi; /* n*3 leaf boxes in sorted order */
    float cb_lo[3], cb_hi[3]; /* bounds of photon positions */
    float abs_max;          /* max |coordinate| over leaf boxes */
} orc_lbvh;
uint32_t orc_morton30(const float p[3], const float cb_lo[3], const float cb_hi[3]);
orc_lbvh *orc_lbvh_build(const orc_photons *ph);
void orc_lbvh_free(orc_lbvh *t);

/* ---- the beam query ---- */
typedef struct {
    uint64_t nodes_visited;  /* V: slab tests performed (internal nodes + leaves) */
    uint64_t disk_tests;     /* T: photons whose leaf box passed the slab test */
    uint64_t hits;           /* H */
} orc_counters;

/* Disk predicate + contribution of one photon to one ray (DESIGN.md section
 * "disk predicate"): returns 1 on hit and sets *t_out. */
int orc_disk_test(const float o[3], const float d[3], float tMax, const float p[3], float r,
                  float *t_out, float *d2_out);

/* The traversal's slab test: Bounds3::IntersectP(ray, invDir, dirIsNeg), core/geometry.h:1410-1438,
 * with the ray origin padded by `pad` (conservative cull).  pad = 0 is exactly the reference test. */
int orc_slab_test(const float o[3], const float d[3], float tMax, const float lo[3],
                  const float hi[3], float pad);

/* One ray against the canonical LBVH (collect-all traversal, core/photonbeambvh.cpp:685-723,
 * slab test core/geometry.h:1410-1438).  hits (optional) receives ORIGINAL photon indices
 * in traversal order, at most hits_cap; returns the number of hits. */
uint32_t orc_query_ray(const orc_lbvh *t, const orc_photons *ph, const orc_medium *m,
                       const float o[3], const float d[3], float tMax, float nPathsNorm,
                       float rgb[3], uint32_t *hits, uint32_t hits_cap, orc_counters *c);
/* Brute force over all photons (no tree): the ground truth for hit sets. */
uint32_t orc_query_ray_brute(const orc_photons *ph, const orc_medium *m, const float o[3],
                             const float d[3], float tMax, float nPathsNorm, float rgb[3],
                             uint32_t *hits, uint32_t hits_cap);
/* Batch versions (OpenMP over rays).  o_tmax: nRays*4 {ox,oy,oz,tMax}; dirs: nRays*3.
 * hit_offsets (nRays+1) / hit_indices: optional CSR output, each ray's list sorted ascending;
 * hit_cap = capacity of hit_indices.  Returns 0 or -1 on overflow. */
int orc_query_batch(const orc_lbvh *t, const orc_photons *ph, const orc_medium *m, uint32_t nRays,
                    const float *o_tmax, const float *dirs, float nPathsNorm, float *rgb_out,
                    uint64_t *hit_offsets, uint32_t *hit_indices, uint64_t hit_cap,
                    orc_counters *c, int nThreads);
int orc_query_batch_brute(const orc_photons *ph, const orc_medium *m, uint32_t nRays,
                          const float *o_tmax, const float *dirs, float nPathsNorm, float *rgb_out,
                          uint64_t *hit_offsets, uint32_t *hit_indices, uint64_t hit_cap,
                          int nThreads);

/* ---- film: integrators/photonbeam.cpp:504,578; core/film.cpp:132-140,168-204 ---- */
void orc_film_accumulate(float *ld, uint32_t nRays, const uint32_t *pixel, const float *rgb);
void orc_film_resolve(const float *ld, uint32_t nPixels, int iter, float scale, float *rgb_out);

/* ---- volume photon tracing (K1 restatement; DESIGN.md section "photon tracing") ---- */
typedef struct {
    int kind;              /* 0 = point (lights/point.cpp:61-71), 1 = spot (lights/spot.cpp:64-93) */
    float pos[3];
    float I[3];
    float axis[3];         /* spot: unit direction */
    float cosTotalWidth, cosFalloffStart;
} orc_light;
/* Traces paths [first, first+nPaths) inside the axis-aligned absorbing box [blo,bhi];
 * appends photons to out arrays (capacity cap); returns the number stored. Deterministic
 * per path index (PCG32 stream index+1, integrators/photonbeam.cpp:386-391). */
uint64_t orc_trace_photons(const orc_medium *m, const orc_light *lights, int nLights,
                           const float blo[3], const float bhi[3], uint64_t first, uint64_t nPaths,
                           int maxDepth, float *pos, float *power, float *dir, uint32_t *path_id,
                           uint64_t cap);

/* ---- surfaces: triangle intersector, Lambertian bounce, camera rays ---- */
#define ORC_TRI_MEDIUM_INHERIT 0  /* no medium transition: the ray keeps its medium (primitive.cpp:104-109) */
#define ORC_TRI_MEDIUM_INSIDE 1   /* MediumInterface "M" "": medium on the side opposite the normal */
#define ORC_TRI_MEDIUM_OUTSIDE 2  /* MediumInterface "" "M": medium on the normal's side */
#define ORC_TRI_FLIP_NORMAL 4     /* reverseOrientation ^ transformSwapsHandedness (triangle.cpp:387-388) */
#define ORC_TRI_NO_MATERIAL 8     /* interface-only surface, photons pass through (photonbeam.cpp:300-304) */
typedef struct {
    uint32_t n_tri;
    const float *P;        /* vertex positions xyz */
    const uint32_t *idx;   /* 3 per triangle */
    const float *Kd;       /* 3 per triangle, Lambertian reflectance; Kd[0] < 0: no material */
    const uint8_t *attr;   /* ORC_TRI_* bits per triangle, may be NULL (all 0) */
} orc_mesh;
typedef struct {
    float t;
    uint32_t tri;
    float p[3], pError[3];
    float n[3];            /* geometric = shading normal, after the orientation flip */
    float ss[3];           /* Normalize(dpdu) */
} orc_surface_hit;
int orc_intersect_triangle(const float o[3], const float d[3], float tMax, const float p0[3],
                           const float p1[3], const float p2[3], float *tHit, float bary[3]);
int orc_mesh_intersect(const orc_mesh *mesh, const float o[3], const float d[3], float tMax,
                       orc_surface_hit *hit);
void orc_offset_ray_origin(const float p[3], const float pError[3], const float n[3], const float w[3],
                           float out[3]);
void orc_concentric_sample_disk(float u0, float u1, float out[2]);
int orc_lambert_sample(const orc_surface_hit *h, const float R[3], const float woW[3], float u0,
                       float u1, float wiW[3], float f[3], float *pdf);
/* As orc_trace_photons, with Scene::Intersect over a triangle mesh instead of the absorbing box. */
uint64_t orc_trace_photons_mesh(const orc_medium *m, const orc_light *lights, int nLights,
                                const orc_mesh *mesh, int lightsInMedium, uint64_t first,
                                uint64_t nPaths, int maxDepth, float *pos, float *power, float *dir,
                                uint32_t *path_id, uint64_t cap);
/* Pinhole camera rays + first-surface tMax; o_tmax 4 floats per pixel, dirs 3 (unit length). */
/* HaltonSampler camera samples (samplers/halton.cpp): see bre_oracle.c */
typedef struct { int baseScales[2], baseExponents[2]; int64_t sampleStride; uint64_t multInverse[2]; } orc_halton;
void orc_halton_setup(int resX, int resY, orc_halton *hs);
/* out_xy: 2 floats per pixel of [x0,x1) x [y0,y1) (row major): dimensions 0 and 1 of sample `sampleNum` of that pixel */
void orc_halton_camera_samples(int x0, int y0, int x1, int y1, int64_t sampleNum, float *out_xy);
void orc_camera_rays_px(const float raster_to_camera[16], const float camera_to_world[16], int x0, int y0,
                        int w, int h, float offx, float offy, const float *offs_xy, const orc_mesh *mesh,
                        int cameraInMedium, float *o_tmax, float *dirs);
void orc_camera_rays(const float raster_to_camera[16], const float camera_to_world[16], int x0, int y0,
                     int w, int h, float offx, float offy, const orc_mesh *mesh, int cameraInMedium,
                     float *o_tmax, float *dirs);

/* ---- synthetic workloads ---- */
void orc_synth_photons_uniform(uint32_t n, float *pos, float *power, float *dir);
void orc_synth_rays_pinhole(int w, int h, float *o_tmax, float *dirs);
void orc_synth_rays_incoherent(uint32_t n, float *o_tmax, float *dirs);
/* intersection of a ray with the box [lo,hi]: core/geometry.h:1386-1408 */
int orc_ray_box(const float o[3], const float d[3], float tMax, const float lo[3],
                const float hi[3], float *t0, float *t1);

#ifdef __cplusplus
}
#endif
#endif
