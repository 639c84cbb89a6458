/* include/bre.h -- C ABI of libbre.so, the B200-native (sm_100a) beam radiance estimate.
 *
 * This is the drop-in boundary for the volumetric photon-mapping Integrator of the
 * pbrt-v3 fork bwiberg/beam-radiance-estimate-pbrt.  A pbrt-side adapter
 * (beam-radiance-estimate-pbrt_b200/host/bre_integrator.{h,cpp})
 * implements `class Integrator { virtual void Render(const Scene&) }`
 * (reference src/core/integrator.h:53-58) and calls these entry points instead of
 * the CPU code cited beside each of them.  `file:line` citations are relative to
 * the reference's src/ directory.
 *
 * Conventions: plain C, opaque handle, every call returns BRE_OK (0) or a negative
 * bre_status; bre_last_error(ctx) gives a message.  All pointers are HOST pointers
 * unless the parameter name ends in `_d` (device pointer on the context's GPU).
 * The caller owns every host buffer; the library owns device memory until
 * bre_destroy.  One context per host thread, or serialise calls externally.
 * Work is issued on one CUDA stream per context; calls that return host data
 * synchronise before returning, `_d` calls are asynchronous on bre_get_stream().
 * There is NO CPU fallback: without a CUDA device bre_create fails with
 * BRE_ERR_NO_DEVICE.
 */
#ifndef BRE_H
#define BRE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BRE_ABI_VERSION 3

typedef enum {
    BRE_OK = 0,
    BRE_ERR_INVALID = -1,     /* bad argument / call order */
    BRE_ERR_NO_DEVICE = -2,   /* no usable CUDA device (no CPU fallback exists) */
    BRE_ERR_CUDA = -3,        /* CUDA runtime error, see bre_last_error */
    BRE_ERR_NOMEM = -4,
    BRE_ERR_UNSUPPORTED = -5,
    BRE_ERR_NCCL = -6
} bre_status;

typedef struct bre_ctx bre_ctx;

typedef struct {
    int device;                /* CUDA device ordinal for this rank */
    int n_ranks;               /* 1 = single GPU */
    int rank;
    const void *nccl_unique_id; /* 128-byte ncclUniqueId shared by all ranks, or NULL when n_ranks==1 */
    int max_leaf_photons;      /* photons per LBVH leaf on the fast path, 1..16; 0 = default (8) */
    int tau_steps;             /* optical-depth table steps per ray for grid media, <= 384; 0 = from the grid:
                                  1.5 steps per voxel of the longest axis, a multiple of 32 in [64, 128] */
} bre_config;

typedef struct {
    int kind;                  /* 0 = point (lights/point.cpp:61-71), 1 = spot (lights/spot.cpp:64-93) */
    float pos[3];
    float I[3];
    float axis[3];             /* spot only: unit direction the cone points along */
    float cos_total_width;     /* spot only */
    float cos_falloff_start;   /* spot only */
} bre_light;

typedef struct {
    uint64_t nodes_visited;    /* V: slab tests on the canonical 1-photon/leaf LBVH */
    uint64_t disk_tests;       /* T: photons whose canonical leaf box passed the slab test */
    uint64_t hits;             /* H */
    uint64_t bytes_algorithmic;/* sum over rays of 32 + 32 V + 16 T + 32 H + 16 */
    uint64_t fast_nodes;       /* 64-byte two-child nodes fetched by the production kernel */
    uint64_t fast_disk_tests;  /* disk tests done by the production kernel (leaf size > 1) */
    uint64_t leaf_pairs;       /* (ray, leaf) pairs the traversal kernel produced */
    uint64_t pair_queue_overflow; /* pairs shaded in place because the pair queue was full */
} bre_counters;

typedef struct {
    uint64_t *offsets;         /* nRays+1 entries, filled by the library */
    uint32_t *indices;         /* caller buffer for ORIGINAL photon indices, ascending per ray */
    uint64_t capacity;         /* entries available in `indices` */
    uint64_t total;            /* out: total hits (if > capacity nothing is written to indices) */
} bre_hitdump;

/* ---- lifetime ------------------------------------------------------------------ */
int bre_abi_version(void);
/* Replaces the construction done in CreatePhotonBeamIntegrator (integrators/photonbeam.cpp:589-611). */
int bre_create(bre_ctx **ctx, const bre_config *cfg);
int bre_destroy(bre_ctx *ctx);
const char *bre_last_error(const bre_ctx *ctx);
/* cudaStream_t the context issues work on (as void*). */
void *bre_get_stream(bre_ctx *ctx);
int bre_synchronize(bre_ctx *ctx);
/* rank 0 creates the id every rank passes in bre_config.nccl_unique_id (128 bytes). */
int bre_nccl_unique_id(void *out128);
/* Page-locked host memory for arrays handed to bre_stage_rays / bre_film_resolve (truly
 * asynchronous, full-rate PCIe copies).  NULL on failure.  Needs no context. */
void *bre_alloc_pinned(size_t bytes);
void bre_free_pinned(void *p);

/* ---- scene inputs --------------------------------------------------------------- */
/* HomogeneousMedium (media/homogeneous.h, factory core/api.cpp:547-596). */
int bre_set_medium_homogeneous(bre_ctx *ctx, const float sigma_a[3], const float sigma_s[3], float g);
/* GridDensityMedium (media/grid.h:51-75): density nx*ny*nz floats, x fastest; world_to_medium row-major 4x4. */
int bre_set_medium_grid(bre_ctx *ctx, const float sigma_a[3], const float sigma_s[3], float g,
                        int nx, int ny, int nz, const float *density, const float world_to_medium[16]);
/* scene.lights + ComputeLightPowerDistribution (core/integrator.cpp:217-225). */
int bre_set_lights(bre_ctx *ctx, int n, const bre_light *lights);
/* Absorbing axis-aligned scene box that ends photon paths (stand-in for Scene::Intersect,
 * core/scene.cpp:50-54) for scenes given without triangles: it ends photon paths (K1) and, for the device
 * camera (K0), the camera rays -- tMax = where the ray leaves the box. */
int bre_set_bounds(bre_ctx *ctx, const float lo[3], const float hi[3]);
/* Surfaces: Scene::Intersect (core/scene.cpp:50-54) over triangle shapes (shapes/triangle.cpp:177-392)
 * for device photon tracing (K1) and device camera rays (K0).  P = vertex positions xyz (as many as
 * idx references), idx = 3 vertex indices per triangle, Kd = 3 floats per triangle: the reflectance
 * of MatteMaterial with sigma 0 (materials/matte.cpp:55-60); Kd[0] < 0 = no material.  n_tri = 0
 * removes the mesh (photon tracing then falls back to the absorbing box of bre_set_bounds).
 * Triangles are tested by brute force: intended for box-like scenes of tens of triangles. */
int bre_set_triangles(bre_ctx *ctx, uint32_t n_tri, const float *P, const uint32_t *idx, const float *Kd);
/* Per-triangle attribute bits (optional, default 0), after bre_set_triangles with the same n_tri. */
#define BRE_TRI_MEDIUM_INHERIT 0  /* not a medium transition: a ray keeps its medium (core/primitive.cpp:104-109) */
#define BRE_TRI_MEDIUM_INSIDE 1   /* MediumInterface "M" "": medium on the side opposite the normal */
#define BRE_TRI_MEDIUM_OUTSIDE 2  /* MediumInterface "" "M": medium on the normal's side */
#define BRE_TRI_FLIP_NORMAL 4     /* reverseOrientation ^ transformSwapsHandedness (shapes/triangle.cpp:387-388) */
#define BRE_TRI_NO_MATERIAL 8     /* interface-only surface: photons pass through (integrators/photonbeam.cpp:300-304) */
int bre_set_triangle_attrs(bre_ctx *ctx, uint32_t n_tri, const uint8_t *attrs);
/* Whether light sources sit inside the medium (Light::mediumInterface.inside); default 1. */
int bre_set_lights_in_medium(bre_ctx *ctx, int in_medium);

/* ---- K0: camera rays ---------------------------------------------------------------
 * Replaces the camera feed of the camera pass (integrators/photonbeam.cpp:444-481):
 * PerspectiveCamera::GenerateRayDifferential for a pinhole (cameras/perspective.cpp:95-103),
 * Transform::operator()(Ray) (core/transform.h:251-264), Scene::Intersect shortening tMax
 * (core/primitive.cpp:97-101).  One ray per pixel of the window, pixel (x,y) -> ray y*width + x. */
typedef struct {
    float raster_to_camera[16];  /* ProjectiveCamera::RasterToCamera, row major */
    float camera_to_world[16];   /* Camera::CameraToWorld (not animated), row major */
    int32_t x0, y0, width, height; /* Film::croppedPixelBounds */
    int32_t in_medium;           /* camera->medium != nullptr; 0 -> every tMax = 0 (nothing gathered) */
} bre_camera;
int bre_set_camera(bre_ctx *ctx, const bre_camera *cam);
/* Generate the rays of sample position (x + offset_x, y + offset_y) on the device; any output
 * pointer may be NULL (the rays stay resident for bre_query_camera_to_film either way).
 * o_xyz_tmax: width*height*4, d_xyz: width*height*3. */
int bre_camera_rays(bre_ctx *ctx, float offset_x, float offset_y, float *o_xyz_tmax, float *d_xyz);
/* K0 + K5/K6 + K7 without any host<->device ray traffic: generate, query, accumulate into the film
 * (created with width*height pixels). */
int bre_query_camera_to_film(bre_ctx *ctx, float offset_x, float offset_y, float n_paths_norm);
/* The same three calls with the sample position of every pixel drawn as the reference's camera pass draws it:
 * pixel p of iteration `sample_num` looks through p + (dimension 0, dimension 1) of HaltonSampler(nIterations,
 * pixelBounds) after StartPixel(p), SetSampleNumber(sample_num) (integrators/photonbeam.cpp:341,459-466;
 * samplers/halton.cpp:64-125; GetCameraSample core/sampler.cpp:46-52), computed per pixel on the device.
 * pixelBounds = the window given to bre_set_camera.  bre_stage_camera_halton -> bre_query_camera_halton_to_film
 * with the same sample number picks up the staged rays. */
int bre_camera_rays_halton(bre_ctx *ctx, int64_t sample_num, float *o_xyz_tmax, float *d_xyz);
int bre_stage_camera_halton(bre_ctx *ctx, int64_t sample_num);
int bre_query_camera_halton_to_film(bre_ctx *ctx, int64_t sample_num, float n_paths_norm);

/* ---- K1: volume photon tracing ---------------------------------------------------
 * Replaces the photon pass of PhotonBeamIntegrator::Render (integrators/photonbeam.cpp:365-437,
 * TracePhotonBeamRecursive :258-325) with the volume-photon deposit rule of
 * PhotonIntegrator::ShootPhotons (integrators/photonmap.cpp:696-724).
 * Traces path indices [first_index, first_index+n_paths); photons are appended to the
 * context's photon store in path-index order (deterministic).  `radius` is the gather
 * radius given to every new photon.  *n_stored (optional) = photons in the store afterwards. */
int bre_trace_photons(bre_ctx *ctx, uint64_t first_index, uint64_t n_paths, int max_depth,
                      float radius, uint64_t *n_stored);
/* Parity path: load photons dumped from the reference.  AoS float triples as in
 * struct Photon (integrators/photonmap.cpp:61-70).  radii==NULL -> every photon gets `radius`.
 * append=0 replaces the store. */
int bre_upload_photons(bre_ctx *ctx, uint64_t n, const float *p_xyz, const float *alpha_rgb,
                       const float *wi_xyz, float radius, const float *radii, int append);
int bre_clear_photons(bre_ctx *ctx);
int bre_photon_count(bre_ctx *ctx, uint64_t *n);
/* Copy the photon store back (any pointer may be NULL). */
int bre_download_photons(bre_ctx *ctx, uint64_t cap, float *p_xyz, float *alpha_rgb, float *wi_xyz,
                         float *radii);
/* Set the radius of every stored photon (progressive shrink, integrators/photonbeam.cpp:354-356,562). */
int bre_set_radius(bre_ctx *ctx, float radius);
/* Multi-GPU, fused: trace this rank's slice (bre_shard_range) of path indices [0, n_paths_total) and
 * deliver the photons of ALL ranks into every rank's store in rank-major (= path index) order.
 * One kernel compacts the traced records and writes them straight into the local store and, through
 * CUDA-IPC mappings over NVLink / NVSwitch, into the peers' stores; NCCL carries only the counts
 * and the barriers.  Collective: every rank of the communicator must call it with the same
 * arguments.  Replaces bre_clear_photons + bre_trace_photons + bre_allgather_photons; with one rank
 * it is clear + trace.  BRE_NO_P2P=1 in the environment selects the NCCL all-gather instead. */
int bre_trace_exchange_photons(bre_ctx *ctx, uint64_t n_paths_total, int max_depth, float radius,
                               uint64_t *n_stored);
/* Multi-GPU: all-gather every rank's photon store in rank-major order (NCCL over NVLink). */
int bre_allgather_photons(bre_ctx *ctx);

/* ---- K2-K4: LBVH build -----------------------------------------------------------
 * Replaces PhotonBeamBVH::PhotonBeamBVH (core/photonbeambvh.cpp:204-248: recursiveBuild :259-425
 * / HLBVH2Build :427-495 with EncodeMorton3 :134-139, RadixSort :141-182, emitLBVH2 :497-559,
 * flattenBVH2Tree :663-681) and KdTree<Photon> (core/kdtree.h:101-149). */
int bre_build(bre_ctx *ctx);
/* When only bre_set_radius was called since the last full build (progressive radius shrink,
 * integrators/photonbeam.cpp:354-356,562) bre_build keeps the Morton order and the topology and only
 * refits the boxes; the result is bit-identical to a full rebuild.  Returns 0 = never built,
 * 1 = the last bre_build was a full build, 2 = it was a refit. */
int bre_last_build_kind(bre_ctx *ctx);
/* Debug/parity: copy the canonical tree back.  order: n u32 (sorted pos -> photon index);
 * codes: n u32; node_lo/node_hi: (n-1)*3 floats; node_children: (n-1)*2 int32 (>=0 internal,
 * <0 = ~sorted leaf position).  Any pointer may be NULL. */
int bre_download_tree(bre_ctx *ctx, uint32_t *order, uint32_t *codes, float *node_lo, float *node_hi,
                      int32_t *node_children);

/* ---- K5/K6: beam query -----------------------------------------------------------
 * Replaces PhotonBeamBVH::Intersect (core/photonbeambvh.cpp:685-723) + the candidate loop of
 * the camera pass (integrators/photonbeam.cpp:494-508) + Medium::Tr (media/homogeneous.cpp:44-48,
 * media/grid.cpp:89-121).  o_xyz_tmax: n_rays*4 {ox,oy,oz,tMax}; d_xyz: n_rays*3 unit directions.
 * n_paths_norm = number of emitted photon paths (integrators/photonmap.cpp:198).
 * out_rgb: n_rays*3.  hits: optional per-ray hit-index dump (parity). */
int bre_query(bre_ctx *ctx, uint32_t n_rays, const float *o_xyz_tmax, const float *d_xyz,
              float n_paths_norm, float *out_rgb, bre_hitdump *hits);
/* Device-resident variant: rays_o_d / rays_d_d are float4 arrays {ox,oy,oz,tMax} / {dx,dy,dz,0};
 * out_rgb_d float4 {r,g,b,0}.  Asynchronous on the context stream. */
int bre_query_d(bre_ctx *ctx, uint32_t n_rays, const void *rays_o_d, const void *rays_d_d,
                float n_paths_norm, void *out_rgb_d);
/* Host rays -> device query -> film accumulation on the device, nothing copied back: the call
 * the pbrt adapter makes once per camera pass (integrators/photonbeam.cpp:444-557).
 * pixel_idx==NULL -> ray q belongs to pixel q. */
int bre_query_to_film(bre_ctx *ctx, uint32_t n_rays, const float *o_xyz_tmax, const float *d_xyz,
                      const uint32_t *pixel_idx, float n_paths_norm);
/* The same camera pass in two calls, so that the ray upload overlaps the photon pass: the rays of a
 * pass do not depend on the photons, so a host generates them first, calls bre_stage_rays (returns
 * at once; the copy runs on its own stream -- use page-locked host arrays and keep them untouched
 * until bre_query_staged_to_film has returned), then traces / builds, then calls
 * bre_query_staged_to_film. */
int bre_stage_rays(bre_ctx *ctx, uint32_t n_rays, const float *o_xyz_tmax, const float *d_xyz);
/* The same head start for the other two ray sources.  A camera pass needs nothing from the photons until
 * the walk: the rays (K0) and, for grid media, their optical-depth tables (K6) can be computed on the
 * staging stream while the photon pass and the LBVH build run (the reference generates its camera rays
 * inside the camera pass, integrators/photonbeam.cpp:444-481, after the photon pass).  Call right after
 * bre_set_camera / before the photon pass; the matching query call picks the staged results up when its
 * arguments are the same, and computes them itself otherwise.
 *   bre_stage_camera   -> bre_query_camera_to_film(same offsets)
 *   bre_stage_rays_d   -> bre_query_d(same n and device pointers; the arrays must not change in between) */
int bre_stage_camera(bre_ctx *ctx, float offset_x, float offset_y);
int bre_stage_rays_d(bre_ctx *ctx, uint32_t n_rays, const void *rays_o_d, const void *rays_d_d);
int bre_query_staged_to_film(bre_ctx *ctx, const uint32_t *pixel_idx, float n_paths_norm);
/* Ray layout hint for every query entry point: when width * height == n_rays, ray q is pixel
 * (q % width, q / width) of an image, the way the reference's camera pass visits them tile by tile
 * (16x16 pixels, integrators/photonbeam.cpp:345-347,444-477).  The query then walks the photon LBVH
 * once per 16x16 tile with the tile's frustum and tests each photon only against the pixels its disk
 * projects onto (exact predicate per candidate: hit sets do not change).  Tiles whose rays are not a
 * near-regular fan from one point fall back to the per-ray walk.  0, 0 removes the hint.
 * bre_query_camera_to_film always uses its camera's own width x height. */
int bre_set_ray_grid(bre_ctx *ctx, uint32_t width, uint32_t height);
/* Multi-GPU camera pass: when enabled on a context with n_ranks > 1, a query over a ray grid answers
 * only the 16x16 tiles bre_tile_owner() gives to this rank (the other pixels' radiance stays 0);
 * bre_film_gather() then assembles the image on rank 0. */
int bre_set_tile_sharding(bre_ctx *ctx, int enabled);
/* Multi-GPU render sharded by PHOTON PATH instead (the estimate is a sum over photons, integrators/
 * photonbeam.cpp:494-508): when enabled on a context with n_ranks > 1, bre_trace_exchange_photons traces this
 * rank's slice of the paths (bre_shard_range) and keeps it -- no exchange --, bre_build builds the LBVH of that
 * slice only, every query answers all its rays from it, and bre_film_gather SUMS the partial films (a
 * reduce-scatter over CUDA-IPC mappings of the films, the owner of a pixel run resolving it into rank 0's image;
 * BRE_NO_P2P=1: ncclReduce).  Trace, build and the photon tests of the query all shrink with the rank count;
 * only K6 (the rays' optical-depth tables) is replicated.  Enabling one sharding mode disables the other. */
int bre_set_photon_sharding(bre_ctx *ctx, int enabled);
/* Counting pass on the canonical tree for the same rays (fills V,T,H,bytes) plus the
 * production kernel's own counters. */
int bre_count(bre_ctx *ctx, uint32_t n_rays, const float *o_xyz_tmax, const float *d_xyz);
int bre_count_d(bre_ctx *ctx, uint32_t n_rays, const void *rays_o_d, const void *rays_d_d);
int bre_get_counters(bre_ctx *ctx, bre_counters *out);

/* ---- K7: film ---------------------------------------------------------------------
 * Replaces PhotonBeamPixel::Ld accumulation (integrators/photonbeam.cpp:504), L = Ld/(iter+1)
 * (:578), Film::SetImage (core/film.cpp:132-140) and the arithmetic of Film::WriteImage
 * (core/film.cpp:168-204). */
int bre_film_create(bre_ctx *ctx, uint32_t n_pixels);
int bre_film_clear(bre_ctx *ctx);
/* pixel_idx==NULL -> ray q belongs to pixel q.  rgb = n_rays*3 host floats. */
int bre_film_accumulate(bre_ctx *ctx, uint32_t n_rays, const uint32_t *pixel_idx, const float *rgb);
/* Device variant: rgb_d float4 per ray; pixel_idx_d may be NULL. */
int bre_film_accumulate_d(bre_ctx *ctx, uint32_t n_rays, const uint32_t *pixel_idx_d, const void *rgb_d);
/* rgb_out: n_pixels*3 floats = what Film::WriteImage hands to pbrt::WriteImage. */
int bre_film_resolve(bre_ctx *ctx, int iter, float scale, float *rgb_out);
/* rgb_out: n_pixels*3 floats = L = Ld/(iter+1) (integrators/photonbeam.cpp:578), no colour conversion:
 * the array a pbrt host passes to Film::SetImage (core/film.cpp:132-140) before Film::WriteImage. */
int bre_film_mean(bre_ctx *ctx, int iter, float *rgb_out);
/* bre_film_resolve / bre_film_mean (mean_only = 1) for callers that must not allocate -- code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc.  NO counterpart in the reference.
 * The device staging buffer of the resolved triples is a file-scope static device
 * array of BRE_STATIC_RESOLVE_MAX_PIXELS pixels (24 MiB, held by every GPU that loads the library until exit); the
 * call makes no cudaMalloc / cudaFree / allocating library call.  Films with more pixels are rejected with
 * BRE_ERR_UNSUPPORTED and a message (bre_last_error).  Every call on one GPU shares the array: calls that overlap
 * on the GPU (other contexts, streams or threads) race -- serialise them.  Results are bit-identical to
 * bre_film_resolve / bre_film_mean. */
#define BRE_STATIC_RESOLVE_MAX_PIXELS (1u << 21)   /* 2 097 152 >= 1920 x 1080 */
int bre_film_resolve_static(bre_ctx *ctx, int iter, float scale, float *rgb_out, int mean_only);
/* Multi-GPU camera pass.  bre_set_photon_sharding: the partial films of all ranks are summed pixel by pixel (rank
 * order), resolved and written into rank 0's image.  bre_set_tile_sharding: every rank resolves the 16x16 tiles it owns (bre_tile_owner,
 * integrators/photonbeam.cpp:345-347) and writes them straight into rank 0's image over NVLink (CUDA-IPC
 * mapping; BRE_NO_P2P=1: ncclReduce of disjoint tiles).  Collective: every rank calls it with the same iter /
 * scale / mean_only; rgb_out (n_pixels*3 floats) is filled on rank 0 only and may be NULL on the others (NULL on
 * rank 0 too: the assembled image stays in rank 0's device buffer).
 * The image layout comes from bre_set_ray_grid (width * height == film pixels).  mean_only = 1 gives
 * L = Ld/(iter+1) like bre_film_mean, 0 the Film::WriteImage arithmetic like bre_film_resolve.  With one rank
 * (or sharding off) it is a plain resolve. */
int bre_film_gather(bre_ctx *ctx, int iter, float scale, float *rgb_out, int mean_only);

/* ---- on-disk interchange -----------------------------------------------
 * Photon file: 64-byte header {"BREPHOT\0", u32 version = 1, u32 0, u64 n, u64 n_paths, zeros} then
 * the three float4 arrays pos_r {x,y,z,radius}, pow {r,g,b,0}, dir {wx,wy,wz,0}, little endian.
 * n_paths = emitted photon paths (the estimate's normaliser), carried for the consumer. */
int bre_save_photons(bre_ctx *ctx, const char *path, uint64_t n_paths);
int bre_load_photons(bre_ctx *ctx, const char *path, int append, uint64_t *n_loaded, uint64_t *n_paths);
/* Film checkpoint: 64-byte header {"BREFILM\0", u32 version = 1, u32 n_pixels, u64 iterations_done,
 * u64 0, f32 radius, zeros} then the float4 accumulators.  Makes a multi-iteration render resumable
 * (the reference's "startiteration", integrators/photonbeam.cpp:354-356,594-595, only restores the
 * radius).  A failed load leaves the film untouched. */
int bre_film_save(bre_ctx *ctx, const char *path, uint64_t iterations_done, float radius);
int bre_film_load(bre_ctx *ctx, const char *path, uint64_t *iterations_done, float *radius);

/* ---- device timing ------------------------------------------------------------------
 * With timing enabled every phase is bracketed by CUDA events on the context stream.
 * bre_get_timings synchronises and returns the LAST duration of each phase in milliseconds:
 * ms[0]=K1 trace, ms[1]=K2-K4 build, ms[2]=K6 optical-depth table, ms[3]=K5 beam query (K5a+K5b),
 * ms[4]=K7 film accumulate, ms[5]=K5a traversal kernel, ms[6]=K5b leaf-pair kernel, ms[7]=0. */
int bre_set_timing(bre_ctx *ctx, int enabled);
int bre_get_timings(bre_ctx *ctx, float ms[8]);

/* ---- host-side sharding helpers (pure functions, no GPU needed) -------------------- */
/* Photon path indices [first, first+count) owned by `rank` out of n_ranks. */
void bre_shard_range(uint64_t n, int n_ranks, int rank, uint64_t *first, uint64_t *count);
/* Owner rank of a 16x16 image tile (integrators/photonbeam.cpp:345-347) dealt round-robin. */
int bre_tile_owner(uint32_t tile_x, uint32_t tile_y, uint32_t tiles_x, int n_ranks);

#ifdef __cplusplus
}
#endif
#endif /* BRE_H */
