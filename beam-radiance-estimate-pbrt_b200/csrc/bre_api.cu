// bre_api.cu -- the extern "C" boundary of libbre.so (see include/bre.h).
#include "bre_internal.cuh"
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace bre {

int set_err(bre_ctx *c, int code, const std::string &msg) {
    if (c) c->err = msg;
    return code;
}
int check_cuda(bre_ctx *c, cudaError_t e, const char *what) {
    if (e == cudaSuccess) return BRE_OK;
    return set_err(c, BRE_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}
int ensure(bre_ctx *c, Buf &b, size_t bytes) {
    if (bytes == 0) bytes = 16;
    if (b.cap >= bytes) return BRE_OK;
    if (b.p) { cudaFree(b.p); b.p = nullptr; b.cap = 0; }
    size_t cap = bytes + bytes / 8;
    cudaError_t e = cudaMalloc(&b.p, cap);
    if (e != cudaSuccess) {
        b.p = nullptr;
        return set_err(c, BRE_ERR_NOMEM, std::string("cudaMalloc(") + std::to_string(cap) + "): " + cudaGetErrorString(e));
    }
    b.cap = cap;
    return BRE_OK;
}
int grow_photon_store(bre_ctx *ctx, uint64_t newCount);

// pack AoS float triples into float4 (w from a scalar or a per-element array)
__global__ void __launch_bounds__(256) k_pack3(const float *__restrict__ in3, uint64_t n,
                                               const float *__restrict__ w, float wconst,
                                               float4 *__restrict__ out) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = make_float4(in3[3 * i], in3[3 * i + 1], in3[3 * i + 2], w ? w[i] : wconst);
}
__global__ void __launch_bounds__(256) k_unpack3(const float4 *__restrict__ in, uint64_t n,
                                                 float *__restrict__ out3, float *__restrict__ w) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 v = in[i];
    if (out3) { out3[3 * i] = v.x; out3[3 * i + 1] = v.y; out3[3 * i + 2] = v.z; }
    if (w) w[i] = v.w;
}
__global__ void __launch_bounds__(256) k_set_w(float4 *__restrict__ a, uint64_t n, float w) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) a[i].w = w;
}

static void host_coordinate_system(const float v1[3], float v2[3], float v3[3]) {
    // same arithmetic as oracle coordinate_system / reference geometry.h:1020-1027
    if (std::fabs(v1[0]) > std::fabs(v1[1])) {
        float inv = 1.f / std::sqrt(v1[0] * v1[0] + v1[2] * v1[2]);
        v2[0] = -v1[2] * inv; v2[1] = 0.f; v2[2] = v1[0] * inv;
    } else {
        float inv = 1.f / std::sqrt(v1[1] * v1[1] + v1[2] * v1[2]);
        v2[0] = 0.f; v2[1] = v1[2] * inv; v2[2] = -v1[1] * inv;
    }
    double ax = v1[0], ay = v1[1], az = v1[2], bx = v2[0], by = v2[1], bz = v2[2];
    v3[0] = (float)((ay * bz) - (az * by));
    v3[1] = (float)((az * bx) - (ax * bz));
    v3[2] = (float)((ax * by) - (ay * bx));
}

// upload n*3 host floats and pack to float4 at dst[dstOff...]
static int upload_pack(bre_ctx *ctx, const float *h3, uint64_t n, const float *hw, float wconst,
                       float4 *dst) {
    int s;
    size_t bytes3 = (size_t)n * 12, bytesw = hw ? (size_t)n * 4 : 0;
    size_t offw = (bytes3 + 15) & ~(size_t)15;
    if ((s = ensure(ctx, ctx->scratch, offw + bytesw + 64)) != BRE_OK) return s;
    float *d3 = (float *)ctx->scratch.p;
    float *dw = hw ? (float *)((char *)ctx->scratch.p + offw) : nullptr;
    BRE_CUDA(cudaMemcpyAsync(d3, h3, bytes3, cudaMemcpyHostToDevice, ctx->stream));
    if (hw) BRE_CUDA(cudaMemcpyAsync(dw, hw, bytesw, cudaMemcpyHostToDevice, ctx->stream));
    k_pack3<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(d3, n, dw, wconst, dst);
    BRE_CUDA(cudaGetLastError());
    // scratch is reused by the next call: finish before returning
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    return BRE_OK;
}

}  // namespace bre

using namespace bre;

#define CTX_CHECK()                          \
    if (!ctx) return BRE_ERR_INVALID;        \
    cudaSetDevice(ctx->device)

extern "C" {

int bre_abi_version(void) { return BRE_ABI_VERSION; }

int bre_create(bre_ctx **out, const bre_config *cfg) {
    if (!out) return BRE_ERR_INVALID;
    *out = nullptr;
    int nDev = 0;
    cudaError_t e = cudaGetDeviceCount(&nDev);
    if (e != cudaSuccess || nDev == 0) return BRE_ERR_NO_DEVICE;  // no CPU fallback
    bre_ctx *ctx = new bre_ctx();
    if (cfg) {
        ctx->device = cfg->device;
        ctx->n_ranks = cfg->n_ranks > 0 ? cfg->n_ranks : 1;
        ctx->rank = cfg->rank;
        if (cfg->max_leaf_photons > 0) ctx->max_leaf = cfg->max_leaf_photons;
        if (cfg->tau_steps > 0) { ctx->tau_steps = cfg->tau_steps; ctx->tau_auto = false; }
    }
    if (ctx->device < 0 || ctx->device >= nDev || ctx->max_leaf > 16 || ctx->tau_steps > 384 || ctx->rank < 0 ||
        ctx->rank >= ctx->n_ranks) {
        delete ctx;
        return BRE_ERR_INVALID;
    }
    int prLeast = 0, prGreatest = 0;
    // the context stream outranks the staging stream (bre_stage_*: K0 / K6 fill the slots its kernels leave free)
    if (cudaSetDevice(ctx->device) != cudaSuccess ||
        cudaDeviceGetStreamPriorityRange(&prLeast, &prGreatest) != cudaSuccess ||
        cudaStreamCreateWithPriority(&ctx->stream, cudaStreamNonBlocking, prGreatest) != cudaSuccess) {
        delete ctx;
        return BRE_ERR_CUDA;
    }
    {
        const char *qc = getenv("BRE_PAIR_QUEUE_CAP");
        if (qc && atoll(qc) > 0) ctx->pair_queue_cap_override = (uint64_t)atoll(qc);
        const char *lf = getenv("BRE_L2_FETCH");   // experiment: 32 / 64 / 128 byte L2 fetch granularity
        if (lf && atoi(lf) > 0) cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)atoi(lf));
        const char *fs = getenv("BRE_FAT_SPREAD");
        if (fs) ctx->fat_spread = (float)atof(fs);
        const char *ka = getenv("BRE_K5A_BLOCKS");
        if (ka && atoi(ka) > 0) ctx->k5a_blocks_per_sm = atoi(ka);
        const char *kb = getenv("BRE_K5B_BLOCKS");
        if (kb && atoi(kb) > 0) ctx->k5b_blocks_per_sm = atoi(kb);
        const char *rd = getenv("BRE_REFIT_DIRECT");
        if (rd && atoi(rd) > 0) ctx->refit_direct = atoi(rd) < ctx->max_leaf ? ctx->max_leaf : atoi(rd);
        const char *np = getenv("BRE_NO_P2P");
        ctx->p2p_disabled = np && atoi(np) != 0;
        const char *ct = getenv("BRE_COMPACT_TILED");
        ctx->compact_tiled = !(ct && atoi(ct) == 0);
        const char *so = getenv("BRE_SORT_ONESWEEP");
        ctx->sort_onesweep = !(so && atoi(so) == 0);
        const char *ga = getenv("BRE_GATHER_AOS");
        ctx->gather_aos = !(ga && atoi(ga) == 0);
        const char *td = getenv("BRE_TAU_DEFER");
        ctx->tau_defer = !(td && atoi(td) == 0);
        const char *lp = getenv("BRE_LAZY_PAYLOAD");
        ctx->lazy_payload = !(lp && atoi(lp) == 0);
        const char *tn = getenv("BRE_TRACE_NESTED");
        ctx->trace_nested = tn && atoi(tn) != 0;
        const char *tp = getenv("BRE_TILE");
        ctx->tile_path = !(tp && atoi(tp) == 0);
        const char *wq = getenv("BRE_TILE_WQ_CAP");
        if (wq && atoll(wq) > 0) ctx->tile_wq_cap_override = (uint64_t)atoll(wq);
        const char *kt = getenv("BRE_K5T_BLOCKS");
        if (kt && atoi(kt) > 0) ctx->k5t_blocks_per_sm = atoi(kt);
        const char *rb = getenv("BRE_RAY_BATCH");
        if (rb && atoll(rb) > 0) ctx->ray_batch_override = (uint32_t)atoll(rb);
        cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, ctx->device);
    }
    if (ctx->n_ranks > 1) {
        if (!cfg->nccl_unique_id) { cudaStreamDestroy(ctx->stream); delete ctx; return BRE_ERR_INVALID; }
        int s = nccl_init(ctx, cfg->nccl_unique_id);
        if (s != BRE_OK) { cudaStreamDestroy(ctx->stream); delete ctx; return s; }
    }
    *out = ctx;
    return BRE_OK;
}

int bre_destroy(bre_ctx *ctx) {
    if (!ctx) return BRE_ERR_INVALID;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    nccl_destroy(ctx);
    Buf *bufs[] = {&ctx->density, &ctx->pos_r, &ctx->pow, &ctx->dir, &ctx->s_pos_r, &ctx->s_pow,
                   &ctx->s_dir, &ctx->order, &ctx->codes, &ctx->keys_alt, &ctx->vals_alt, &ctx->hist,
                   &ctx->node_lo, &ctx->node_hi, &ctx->node_range, &ctx->parent, &ctx->leaf_lo,
                   &ctx->leaf_hi, &ctx->flags, &ctx->fast, &ctx->scratch, &ctx->rays_o, &ctx->rays_d,
                   &ctx->out_rgb, &ctx->tau, &ctx->hit_counts, &ctx->hit_offsets, &ctx->hit_idx,
                   &ctx->counters_d, &ctx->pix_idx, &ctx->packet_ctr, &ctx->pair_queue, &ctx->heavy, &ctx->frames, &ctx->film, &ctx->path_counts, &ctx->path_offsets, &ctx->tri, &ctx->tri_kdb, &ctx->entry, &ctx->rays_stage, &ctx->p2p_scratch, &ctx->aos, &ctx->tframes, &ctx->trays, &ctx->tjit, &ctx->tile_state, &ctx->top_cut, &ctx->tile_wq, &ctx->gather, &ctx->os_scratch};
    for (Buf *b : bufs) if (b->p) cudaFree(b->p);
    if (ctx->lights_d) cudaFree(ctx->lights_d);
    if (ctx->est_host) cudaFreeHost(ctx->est_host);
    if (ctx->ev_est) cudaEventDestroy(ctx->ev_est);
    if (ctx->copy_stream) {
        cudaStreamSynchronize(ctx->copy_stream);
        cudaEventDestroy(ctx->ev_stage); cudaEventDestroy(ctx->ev_main);
        cudaStreamDestroy(ctx->copy_stream);
    }
    if (ctx->stream2) {
        cudaStreamSynchronize(ctx->stream2);
        for (int i = 0; i < 4; ++i) if (ctx->qev[i]) cudaEventDestroy(ctx->qev[i]);
        cudaStreamDestroy(ctx->stream2);
    }
    for (int i = 0; i < 8; ++i)
        for (int j = 0; j < 2; ++j) if (ctx->ev[i][j]) cudaEventDestroy(ctx->ev[i][j]);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
    return BRE_OK;
}

const char *bre_last_error(const bre_ctx *ctx) { return ctx ? ctx->err.c_str() : "null context"; }
void *bre_get_stream(bre_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }
int bre_synchronize(bre_ctx *ctx) {
    CTX_CHECK();
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    return BRE_OK;
}
int bre_nccl_unique_id(void *out128) { return nccl_unique_id(out128); }

void *bre_alloc_pinned(size_t bytes) {
    void *p = nullptr;
    return cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) == cudaSuccess ? p : nullptr;
}
void bre_free_pinned(void *p) { if (p) cudaFreeHost(p); }

// ---------------------------------------------------------------- scene
int bre_set_medium_homogeneous(bre_ctx *ctx, const float sa[3], const float ss[3], float g) {
    CTX_CHECK();
    if (!sa || !ss) return set_err(ctx, BRE_ERR_INVALID, "null sigma");
    MediumDev &m = ctx->medium;
    memset(&m, 0, sizeof(m));
    m.kind = 0;
    for (int k = 0; k < 3; ++k) { m.sigma_a[k] = sa[k]; m.sigma_s[k] = ss[k]; m.sigma_t[k] = sa[k] + ss[k]; }
    m.g = g;
    m.gray = (m.sigma_t[0] == m.sigma_t[1] && m.sigma_t[1] == m.sigma_t[2]) ? 1 : 0;
    ctx->have_medium = true;
    ++ctx->medium_version;
    ctx->tau_staged = false;
    return BRE_OK;
}

int bre_set_medium_grid(bre_ctx *ctx, const float sa[3], const float ss[3], float g, int nx, int ny,
                        int nz, const float *density, const float w2m[16]) {
    CTX_CHECK();
    if (!sa || !ss || !density || !w2m || nx < 1 || ny < 1 || nz < 1)
        return set_err(ctx, BRE_ERR_INVALID, "bad grid medium arguments");
    if ((uint64_t)nx * (uint64_t)ny * (uint64_t)nz >= (1ull << 31))
        return set_err(ctx, BRE_ERR_UNSUPPORTED, "density grid of 2^31 voxels or more (32-bit voxel indices on the device)");
    // GridDensityMedium requires a spectrally uniform sigma_t (media/grid.h:63-67)
    float st0 = sa[0] + ss[0];
    if (sa[1] + ss[1] != st0 || sa[2] + ss[2] != st0)
        return set_err(ctx, BRE_ERR_INVALID, "GridDensityMedium requires a spectrally uniform attenuation coefficient");
    int s = bre_set_medium_homogeneous(ctx, sa, ss, g);
    if (s != BRE_OK) return s;
    MediumDev &m = ctx->medium;
    size_t cnt = (size_t)nx * ny * nz;
    if ((s = ensure(ctx, ctx->density, cnt * 4)) != BRE_OK) { ctx->have_medium = false; return s; }
    BRE_CUDA(cudaMemcpyAsync(ctx->density.p, density, cnt * 4, cudaMemcpyHostToDevice, ctx->stream));
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    float maxDensity = 0;  // grid.h:69-73
    for (size_t i = 0; i < cnt; ++i) maxDensity = std::max(maxDensity, density[i]);
    m.kind = 1;
    m.nx = nx; m.ny = ny; m.nz = nz;
    m.density = (const float *)ctx->density.p;
    m.invMaxDensity = 1 / maxDensity;
    memcpy(m.w2m, w2m, 12 * sizeof(float));
    ++ctx->medium_version;
    if (ctx->tau_auto) {
        // 1.5 table steps per voxel of the longest axis, a multiple of 32 within [64, 128]: a step equal to the
        // voxel pitch aliases with the trilinear lattice (64 steps through 64^3: 1.3e-3 error of Tr, 96: 1.5e-4;
        // pinned against the reference's GridDensityMedium::Tr by tests/golden/ref_gridtr.npz)
        int mx = nx > ny ? nx : ny;
        if (nz > mx) mx = nz;
        int st = (3 * mx / 2 + 31) / 32 * 32;
        ctx->tau_steps = st < 64 ? 64 : (st > 128 ? 128 : st);
    }
    return BRE_OK;
}

int bre_set_lights(bre_ctx *ctx, int n, const bre_light *lights) {
    CTX_CHECK();
    if (n < 1 || n > BRE_MAX_LIGHTS || !lights) return set_err(ctx, BRE_ERR_INVALID, "1..64 lights required");
    std::vector<LightsDev> hv(1);
    LightsDev &h = hv[0];
    memset(&h, 0, sizeof(h));
    h.n = n;
    const float Pi = BRE_PI;
    for (int i = 0; i < n; ++i) {
        LightDev &L = h.l[i];
        L.kind = lights[i].kind;
        if (L.kind != 0 && L.kind != 1) return set_err(ctx, BRE_ERR_UNSUPPORTED, "only point and spot lights");
        memcpy(L.pos, lights[i].pos, 12); memcpy(L.I, lights[i].I, 12); memcpy(L.axis, lights[i].axis, 12);
        L.cosTotalWidth = lights[i].cos_total_width;
        L.cosFalloffStart = lights[i].cos_falloff_start;
        if (L.kind == 1) host_coordinate_system(L.axis, L.ax, L.ay);
        float p[3];
        for (int c = 0; c < 3; ++c) {
            if (L.kind == 0) p[c] = 4 * Pi * L.I[c];                                                // point.cpp:55
            else p[c] = L.I[c] * 2 * Pi * (1 - .5f * (L.cosFalloffStart + L.cosTotalWidth));        // spot.cpp:75-77
        }
        h.func[i] = 0.212671f * p[0] + 0.715160f * p[1] + 0.072169f * p[2];
    }
    // Distribution1D, sampling.h:55-75
    h.cdf[0] = 0;
    for (int i = 0; i < n; ++i) h.cdf[i + 1] = h.cdf[i] + h.func[i] / n;
    h.funcInt = h.cdf[n];
    if (h.funcInt == 0) for (int i = 1; i <= n; ++i) h.cdf[i] = (float)i / (float)n;
    else for (int i = 1; i <= n; ++i) h.cdf[i] /= h.funcInt;
    if (!ctx->lights_d) BRE_CUDA(cudaMalloc(&ctx->lights_d, sizeof(LightsDev)));
    BRE_CUDA(cudaMemcpyAsync(ctx->lights_d, &h, sizeof(LightsDev), cudaMemcpyHostToDevice, ctx->stream));
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->n_lights = n;
    return BRE_OK;
}

int bre_set_bounds(bre_ctx *ctx, const float lo[3], const float hi[3]) {
    CTX_CHECK();
    if (!lo || !hi) return set_err(ctx, BRE_ERR_INVALID, "null bounds");
    for (int k = 0; k < 3; ++k) {
        if (!(lo[k] < hi[k])) return set_err(ctx, BRE_ERR_INVALID, "empty bounds");
        ctx->blo[k] = lo[k]; ctx->bhi[k] = hi[k];
    }
    ctx->have_bounds = true;
    return BRE_OK;
}

// ---------------------------------------------------------------- photons
int bre_trace_photons(bre_ctx *ctx, uint64_t first, uint64_t nPaths, int maxDepth, float radius,
                      uint64_t *nStored) {
    CTX_CHECK();
    if (!(radius > 0.f)) return set_err(ctx, BRE_ERR_INVALID, "radius must be > 0");
    int s = trace_photons(ctx, first, nPaths, maxDepth, radius);
    if (s != BRE_OK) return s;
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    if (nStored) *nStored = ctx->n_photons;
    return BRE_OK;
}

int bre_upload_photons(bre_ctx *ctx, uint64_t n, const float *p, const float *alpha, const float *wi,
                       float radius, const float *radii, int append) {
    CTX_CHECK();
    if (n > 0 && (!p || !alpha || !wi)) return set_err(ctx, BRE_ERR_INVALID, "null photon arrays");
    if (!radii && n > 0 && !(radius > 0.f)) return set_err(ctx, BRE_ERR_INVALID, "radius must be > 0");
    if (!append) ctx->n_photons = 0;
    ctx->built = false; ctx->topo_valid = false;
    if (n == 0) return BRE_OK;
    int s = grow_photon_store(ctx, ctx->n_photons + n);
    if (s != BRE_OK) return s;
    uint64_t off = ctx->n_photons;
    if ((s = upload_pack(ctx, p, n, radii, radius, (float4 *)ctx->pos_r.p + off)) != BRE_OK) return s;
    if ((s = upload_pack(ctx, alpha, n, nullptr, 0.f, (float4 *)ctx->pow.p + off)) != BRE_OK) return s;
    if ((s = upload_pack(ctx, wi, n, nullptr, 0.f, (float4 *)ctx->dir.p + off)) != BRE_OK) return s;
    ctx->n_photons += n;
    return BRE_OK;
}

int bre_clear_photons(bre_ctx *ctx) {
    CTX_CHECK();
    ctx->n_photons = 0;
    ctx->built = false; ctx->topo_valid = false;
    return BRE_OK;
}

int bre_photon_count(bre_ctx *ctx, uint64_t *n) {
    if (!ctx || !n) return BRE_ERR_INVALID;
    *n = ctx->n_photons;
    return BRE_OK;
}

int bre_download_photons(bre_ctx *ctx, uint64_t cap, float *p, float *alpha, float *wi, float *radii) {
    CTX_CHECK();
    uint64_t n = ctx->n_photons < cap ? ctx->n_photons : cap;
    if (n == 0) return BRE_OK;
    int s;
    if ((s = ensure(ctx, ctx->scratch, (size_t)n * 16 + 64)) != BRE_OK) return s;
    float *d3 = (float *)ctx->scratch.p, *dw = d3 + 3 * n;
    const float4 *src[3] = {(const float4 *)ctx->pos_r.p, (const float4 *)ctx->pow.p, (const float4 *)ctx->dir.p};
    float *dst[3] = {p, alpha, wi};
    for (int a = 0; a < 3; ++a) {
        if (!dst[a] && !(a == 0 && radii)) continue;
        k_unpack3<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(src[a], n, d3, a == 0 ? dw : nullptr);
        BRE_CUDA(cudaGetLastError());
        if (dst[a]) BRE_CUDA(cudaMemcpyAsync(dst[a], d3, (size_t)n * 12, cudaMemcpyDeviceToHost, ctx->stream));
        if (a == 0 && radii) BRE_CUDA(cudaMemcpyAsync(radii, dw, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
        BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return BRE_OK;
}

int bre_set_radius(bre_ctx *ctx, float radius) {
    CTX_CHECK();
    if (!(radius > 0.f)) return set_err(ctx, BRE_ERR_INVALID, "radius must be > 0");
    ctx->built = false;   // topology stays valid: the next bre_build only refits
    if (ctx->n_photons)
        k_set_w<<<(unsigned)((ctx->n_photons + 255) / 256), 256, 0, ctx->stream>>>((float4 *)ctx->pos_r.p, ctx->n_photons, radius);
    BRE_CUDA(cudaGetLastError());
    return BRE_OK;
}

int bre_trace_exchange_photons(bre_ctx *ctx, uint64_t nPathsTotal, int maxDepth, float radius,
                               uint64_t *nStored) {
    CTX_CHECK();
    if (!(radius > 0.f)) return set_err(ctx, BRE_ERR_INVALID, "radius must be > 0");
    uint64_t first = 0, count = nPathsTotal;
    bre_shard_range(nPathsTotal, ctx->n_ranks, ctx->rank, &first, &count);
    int s;
    // the path must be the same on every rank (both are collectives): decide from GLOBAL values only --
    // the largest shard any rank gets, not this rank's own
    const uint64_t maxShard = (nPathsTotal + (uint64_t)ctx->n_ranks - 1) / (uint64_t)ctx->n_ranks;
    if (ctx->photon_shard && ctx->n_ranks > 1) {
        // photon-sharded render: the slice stays here, nothing is exchanged (the films are summed instead)
        ctx->n_photons = 0;
        ctx->built = false; ctx->topo_valid = false;
        s = trace_photons(ctx, first, count, maxDepth, radius);
    } else if (ctx->n_ranks > 1 && !ctx->p2p_disabled && ctx->n_ranks <= 8 && maxShard <= TRACE_CHUNK) {
        s = trace_exchange_p2p(ctx, nPathsTotal, maxDepth, radius);
    } else {
        ctx->n_photons = 0;
        s = trace_photons(ctx, first, count, maxDepth, radius);
        if (s == BRE_OK && ctx->n_ranks > 1) s = nccl_allgather_photons(ctx);
    }
    if (s != BRE_OK) return s;
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    if (nStored) *nStored = ctx->n_photons;
    return BRE_OK;
}

int bre_allgather_photons(bre_ctx *ctx) {
    CTX_CHECK();
    if (ctx->n_ranks == 1) return BRE_OK;
    ctx->built = false; ctx->topo_valid = false;
    return nccl_allgather_photons(ctx);
}

// ---------------------------------------------------------------- build
int bre_build(bre_ctx *ctx) {
    CTX_CHECK();
    return build_tree(ctx);
}

int bre_last_build_kind(bre_ctx *ctx) { return ctx ? ctx->last_build_kind : BRE_ERR_INVALID; }

int bre_download_tree(bre_ctx *ctx, uint32_t *order, uint32_t *codes, float *node_lo, float *node_hi,
                      int32_t *node_children) {
    CTX_CHECK();
    if (!ctx->built) return set_err(ctx, BRE_ERR_INVALID, "bre_build has not been called");
    uint32_t n = ctx->n_built;
    if (n == 0) return BRE_OK;
    {
        int s = ensure_canonical(ctx);
        if (s != BRE_OK) return s;
    }
    if (order) BRE_CUDA(cudaMemcpyAsync(order, ctx->order.p, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (codes) BRE_CUDA(cudaMemcpyAsync(codes, ctx->codes.p, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    if (n > 1 && (node_lo || node_hi || node_children)) {
        std::vector<float4> lo(n - 1), hi(n - 1);
        BRE_CUDA(cudaMemcpyAsync(lo.data(), ctx->node_lo.p, (size_t)(n - 1) * 16, cudaMemcpyDeviceToHost, ctx->stream));
        BRE_CUDA(cudaMemcpyAsync(hi.data(), ctx->node_hi.p, (size_t)(n - 1) * 16, cudaMemcpyDeviceToHost, ctx->stream));
        BRE_CUDA(cudaStreamSynchronize(ctx->stream));
        for (uint32_t i = 0; i + 1 < n; ++i) {
            if (node_lo) { node_lo[3 * i] = lo[i].x; node_lo[3 * i + 1] = lo[i].y; node_lo[3 * i + 2] = lo[i].z; }
            if (node_hi) { node_hi[3 * i] = hi[i].x; node_hi[3 * i + 1] = hi[i].y; node_hi[3 * i + 2] = hi[i].z; }
            if (node_children) {
                memcpy(&node_children[2 * i], &lo[i].w, 4);
                memcpy(&node_children[2 * i + 1], &hi[i].w, 4);
            }
        }
    }
    return BRE_OK;
}

// ---------------------------------------------------------------- query
static int upload_rays(bre_ctx *ctx, uint32_t n, const float *o4, const float *d3) {
    int s;
    if ((s = ensure(ctx, ctx->rays_o, (size_t)n * 16)) != BRE_OK) return s;
    if ((s = ensure(ctx, ctx->rays_d, (size_t)n * 16)) != BRE_OK) return s;
    BRE_CUDA(cudaMemcpyAsync(ctx->rays_o.p, o4, (size_t)n * 16, cudaMemcpyHostToDevice, ctx->stream));
    return upload_pack(ctx, d3, n, nullptr, 0.f, (float4 *)ctx->rays_d.p);
}

static int query_ready(bre_ctx *ctx) {
    if (!ctx->built) return set_err(ctx, BRE_ERR_INVALID, "bre_build has not been called since the photons changed");
    if (!ctx->have_medium) return set_err(ctx, BRE_ERR_INVALID, "no medium set");
    return BRE_OK;
}

int bre_query(bre_ctx *ctx, uint32_t n, const float *o4, const float *d3, float nPathsNorm,
              float *out_rgb, bre_hitdump *hits) {
    CTX_CHECK();
    int s = query_ready(ctx);
    if (s != BRE_OK) return s;
    if (n == 0) { if (hits) { hits->total = 0; if (hits->offsets) hits->offsets[0] = 0; } return BRE_OK; }
    if (!o4 || !d3) return set_err(ctx, BRE_ERR_INVALID, "null ray arrays");
    if ((s = upload_rays(ctx, n, o4, d3)) != BRE_OK) return s;
    if (out_rgb) {
        if ((s = ensure(ctx, ctx->out_rgb, (size_t)n * 16 + (size_t)n * 12)) != BRE_OK) return s;
        float4 *o4d = (float4 *)ctx->out_rgb.p;
        float *o3d = (float *)(o4d + n);
        if ((s = run_query(ctx, n, (const float4 *)ctx->rays_o.p, (const float4 *)ctx->rays_d.p, nPathsNorm, o4d)) != BRE_OK) return s;
        k_unpack3<<<(n + 255) / 256, 256, 0, ctx->stream>>>(o4d, n, o3d, nullptr);
        BRE_CUDA(cudaGetLastError());
        BRE_CUDA(cudaMemcpyAsync(out_rgb, o3d, (size_t)n * 12, cudaMemcpyDeviceToHost, ctx->stream));
        BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    if (hits) return run_hitdump(ctx, n, (const float4 *)ctx->rays_o.p, (const float4 *)ctx->rays_d.p, hits);
    return BRE_OK;
}

int bre_query_d(bre_ctx *ctx, uint32_t n, const void *ro, const void *rd, float nPathsNorm, void *out) {
    CTX_CHECK();
    int s = query_ready(ctx);
    if (s != BRE_OK) return s;
    if (n > 0 && (!ro || !rd || !out)) return set_err(ctx, BRE_ERR_INVALID, "null device pointer");
    return run_query(ctx, n, (const float4 *)ro, (const float4 *)rd, nPathsNorm, (float4 *)out);
}

int bre_set_ray_grid(bre_ctx *ctx, uint32_t width, uint32_t height) {
    CTX_CHECK();
    if ((width == 0) != (height == 0)) return set_err(ctx, BRE_ERR_INVALID, "width and height must both be 0 or both be set");
    ctx->grid_w = width; ctx->grid_h = height;
    return BRE_OK;
}

int bre_set_tile_sharding(bre_ctx *ctx, int enabled) {
    CTX_CHECK();
    ctx->tile_shard = enabled != 0;
    if (ctx->tile_shard) ctx->photon_shard = false;
    return BRE_OK;
}

int bre_set_photon_sharding(bre_ctx *ctx, int enabled) {
    CTX_CHECK();
    ctx->photon_shard = enabled != 0;
    if (ctx->photon_shard) ctx->tile_shard = false;
    return BRE_OK;
}

int bre_query_to_film(bre_ctx *ctx, uint32_t n, const float *o4, const float *d3, const uint32_t *pix,
                      float nPathsNorm) {
    CTX_CHECK();
    int s = query_ready(ctx);
    if (s != BRE_OK) return s;
    if (!ctx->film.p) return set_err(ctx, BRE_ERR_INVALID, "bre_film_create has not been called");
    if (n == 0) return BRE_OK;
    if (!o4 || !d3) return set_err(ctx, BRE_ERR_INVALID, "null ray arrays");
    if ((s = upload_rays(ctx, n, o4, d3)) != BRE_OK) return s;
    if ((s = ensure(ctx, ctx->out_rgb, (size_t)n * 16)) != BRE_OK) return s;
    const uint32_t *pd = nullptr;
    if (pix) {
        if ((s = ensure(ctx, ctx->pix_idx, (size_t)n * 4)) != BRE_OK) return s;
        BRE_CUDA(cudaMemcpyAsync(ctx->pix_idx.p, pix, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
        pd = (const uint32_t *)ctx->pix_idx.p;
    }
    if ((s = run_query(ctx, n, (const float4 *)ctx->rays_o.p, (const float4 *)ctx->rays_d.p, nPathsNorm,
                       (float4 *)ctx->out_rgb.p)) != BRE_OK) return s;
    return film_accumulate(ctx, n, pd, (const float4 *)ctx->out_rgb.p);
}

// Asynchronous ray upload on a copy stream: the camera rays of a pass do not depend on the photon
// pass, so a host can hand them over first and let the PCIe transfer overlap K1 and the LBVH build.
// staging stream of bre_stage_*: starts after everything queued on the context stream so far (the ray / table
// buffers may still be read by the previous camera pass)
static int staging_begin(bre_ctx *ctx) {
    if (!ctx->copy_stream) {
        int prLeast = 0, prGreatest = 0;
        BRE_CUDA(cudaDeviceGetStreamPriorityRange(&prLeast, &prGreatest));
        // lowest priority: its kernels take the SM slots the context stream's kernels leave free
        BRE_CUDA(cudaStreamCreateWithPriority(&ctx->copy_stream, cudaStreamNonBlocking, prLeast));
        BRE_CUDA(cudaEventCreateWithFlags(&ctx->ev_stage, cudaEventDisableTiming));
        BRE_CUDA(cudaEventCreateWithFlags(&ctx->ev_main, cudaEventDisableTiming));
    }
    BRE_CUDA(cudaEventRecord(ctx->ev_main, ctx->stream));
    BRE_CUDA(cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_main, 0));
    ctx->tau_staged = false;
    ctx->tau_pending.on = false;
    ctx->cam_staged = false;
    return BRE_OK;
}

int bre_stage_rays(bre_ctx *ctx, uint32_t n, const float *o4, const float *d3) {
    CTX_CHECK();
    if (n == 0 || !o4 || !d3) return set_err(ctx, BRE_ERR_INVALID, "bad ray arrays");
    int s;
    if ((s = ensure(ctx, ctx->rays_o, (size_t)n * 16)) != BRE_OK) return s;
    if ((s = ensure(ctx, ctx->rays_d, (size_t)n * 16)) != BRE_OK) return s;
    if ((s = ensure(ctx, ctx->rays_stage, (size_t)n * 12)) != BRE_OK) return s;
    if ((s = staging_begin(ctx)) != BRE_OK) return s;
    BRE_CUDA(cudaMemcpyAsync(ctx->rays_o.p, o4, (size_t)n * 16, cudaMemcpyHostToDevice, ctx->copy_stream));
    BRE_CUDA(cudaMemcpyAsync(ctx->rays_stage.p, d3, (size_t)n * 12, cudaMemcpyHostToDevice, ctx->copy_stream));
    k_pack3<<<(n + 255) / 256, 256, 0, ctx->copy_stream>>>((const float *)ctx->rays_stage.p, n, nullptr, 0.f,
                                                           (float4 *)ctx->rays_d.p);
    BRE_CUDA(cudaGetLastError());
    // grid media: the optical-depth table of these rays (K6) right behind the upload, under the photon pass
    if ((s = stage_tau_deferred(ctx, n, (const float4 *)ctx->rays_o.p, (const float4 *)ctx->rays_d.p, false, 0, 0)) != BRE_OK) return s;
    BRE_CUDA(cudaEventRecord(ctx->ev_stage, ctx->copy_stream));
    ctx->staged_rays = n;
    return BRE_OK;
}

int bre_stage_rays_d(bre_ctx *ctx, uint32_t n, const void *ro, const void *rd) {
    CTX_CHECK();
    if (n == 0 || !ro || !rd) return set_err(ctx, BRE_ERR_INVALID, "bad ray arrays");
    int s;
    if ((s = staging_begin(ctx)) != BRE_OK) return s;
    if ((s = stage_tau_deferred(ctx, n, (const float4 *)ro, (const float4 *)rd, false, 0, 0)) != BRE_OK) return s;
    BRE_CUDA(cudaEventRecord(ctx->ev_stage, ctx->copy_stream));
    return BRE_OK;
}

static int stage_camera_impl(bre_ctx *ctx, float offx, float offy, long long halton);
int bre_stage_camera(bre_ctx *ctx, float offx, float offy) {
    CTX_CHECK();
    return stage_camera_impl(ctx, offx, offy, -1);
}
int bre_stage_camera_halton(bre_ctx *ctx, int64_t sample_num) {
    CTX_CHECK();
    if (sample_num < 0) return set_err(ctx, BRE_ERR_INVALID, "negative sample number");
    return stage_camera_impl(ctx, 0.f, 0.f, (long long)sample_num);
}
static int stage_camera_impl(bre_ctx *ctx, float offx, float offy, long long halton) {
    if (!ctx->have_camera) return set_err(ctx, BRE_ERR_INVALID, "bre_set_camera has not been called");
    int s;
    const uint32_t n = (uint32_t)ctx->camera.width * (uint32_t)ctx->camera.height;
    if ((s = ensure(ctx, ctx->rays_o, (size_t)n * 16)) != BRE_OK) return s;     // (before the streams fork: ensure may free)
    if ((s = ensure(ctx, ctx->rays_d, (size_t)n * 16)) != BRE_OK) return s;
    if ((s = staging_begin(ctx)) != BRE_OK) return s;
    if ((s = camera_rays(ctx, offx, offy, ctx->copy_stream, halton)) != BRE_OK) return s;
    s = stage_tau_deferred(ctx, n, (const float4 *)ctx->rays_o.p, (const float4 *)ctx->rays_d.p, true,
                           (uint32_t)ctx->camera.width, (uint32_t)ctx->camera.height);
    if (s != BRE_OK) return s;
    BRE_CUDA(cudaEventRecord(ctx->ev_stage, ctx->copy_stream));
    ctx->cam_staged = true; ctx->cam_offx = offx; ctx->cam_offy = offy; ctx->cam_halton = halton;
    return BRE_OK;
}

int bre_query_staged_to_film(bre_ctx *ctx, const uint32_t *pix, float nPathsNorm) {
    CTX_CHECK();
    int s = query_ready(ctx);
    if (s != BRE_OK) return s;
    if (!ctx->film.p) return set_err(ctx, BRE_ERR_INVALID, "bre_film_create has not been called");
    const uint32_t n = ctx->staged_rays;
    if (n == 0) return set_err(ctx, BRE_ERR_INVALID, "bre_stage_rays has not been called");
    ctx->staged_rays = 0;
    // the upload finished long ago in practice; waiting on the host too makes the caller's ray
    // arrays reusable as soon as this call returns
    BRE_CUDA(cudaEventSynchronize(ctx->ev_stage));
    BRE_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_stage, 0));
    if ((s = ensure(ctx, ctx->out_rgb, (size_t)n * 16)) != BRE_OK) return s;
    const uint32_t *pd = nullptr;
    if (pix) {
        if ((s = ensure(ctx, ctx->pix_idx, (size_t)n * 4)) != BRE_OK) return s;
        BRE_CUDA(cudaMemcpyAsync(ctx->pix_idx.p, pix, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
        pd = (const uint32_t *)ctx->pix_idx.p;
    }
    if ((s = run_query(ctx, n, (const float4 *)ctx->rays_o.p, (const float4 *)ctx->rays_d.p, nPathsNorm,
                       (float4 *)ctx->out_rgb.p)) != BRE_OK) return s;
    return film_accumulate(ctx, n, pd, (const float4 *)ctx->out_rgb.p);
}

static int camera_rays_impl(bre_ctx *ctx, float offx, float offy, long long halton, float *o4, float *d3);
int bre_camera_rays(bre_ctx *ctx, float offx, float offy, float *o4, float *d3) {
    CTX_CHECK();
    return camera_rays_impl(ctx, offx, offy, -1, o4, d3);
}
int bre_camera_rays_halton(bre_ctx *ctx, int64_t sample_num, float *o4, float *d3) {
    CTX_CHECK();
    if (sample_num < 0) return set_err(ctx, BRE_ERR_INVALID, "negative sample number");
    return camera_rays_impl(ctx, 0.f, 0.f, (long long)sample_num, o4, d3);
}
static int camera_rays_impl(bre_ctx *ctx, float offx, float offy, long long halton, float *o4, float *d3) {
    ctx->cam_staged = false;
    int s = camera_rays(ctx, offx, offy, nullptr, halton);
    if (s != BRE_OK) return s;
    const uint32_t n = (uint32_t)ctx->camera.width * (uint32_t)ctx->camera.height;
    if (o4) BRE_CUDA(cudaMemcpyAsync(o4, ctx->rays_o.p, (size_t)n * 16, cudaMemcpyDeviceToHost, ctx->stream));
    if (d3) {
        if ((s = ensure(ctx, ctx->scratch, (size_t)n * 12 + 64)) != BRE_OK) return s;
        k_unpack3<<<(n + 255) / 256, 256, 0, ctx->stream>>>((const float4 *)ctx->rays_d.p, n, (float *)ctx->scratch.p, nullptr);
        BRE_CUDA(cudaGetLastError());
        BRE_CUDA(cudaMemcpyAsync(d3, ctx->scratch.p, (size_t)n * 12, cudaMemcpyDeviceToHost, ctx->stream));
    }
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    return BRE_OK;
}

static int query_camera_impl(bre_ctx *ctx, float offx, float offy, long long halton, float nPathsNorm);
int bre_query_camera_to_film(bre_ctx *ctx, float offx, float offy, float nPathsNorm) {
    CTX_CHECK();
    return query_camera_impl(ctx, offx, offy, -1, nPathsNorm);
}
int bre_query_camera_halton_to_film(bre_ctx *ctx, int64_t sample_num, float nPathsNorm) {
    CTX_CHECK();
    if (sample_num < 0) return set_err(ctx, BRE_ERR_INVALID, "negative sample number");
    return query_camera_impl(ctx, 0.f, 0.f, (long long)sample_num, nPathsNorm);
}
static int query_camera_impl(bre_ctx *ctx, float offx, float offy, long long halton, float nPathsNorm) {
    int s = query_ready(ctx);
    if (s != BRE_OK) return s;
    if (!ctx->have_camera) return set_err(ctx, BRE_ERR_INVALID, "bre_set_camera has not been called");
    const uint32_t n = (uint32_t)ctx->camera.width * (uint32_t)ctx->camera.height;
    if (!ctx->film.p || ctx->n_pixels != n)
        return set_err(ctx, BRE_ERR_INVALID, "bre_film_create(width*height) has not been called");
    if (ctx->cam_staged && ctx->cam_halton == halton && (halton >= 0 || (ctx->cam_offx == offx && ctx->cam_offy == offy))) {
        // bre_stage_camera generated exactly these rays (and their optical-depth table) under the photon pass
        BRE_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_stage, 0));
    } else {
        ctx->tau_staged = false;
        if ((s = camera_rays(ctx, offx, offy, nullptr, halton)) != BRE_OK) return s;
    }
    ctx->cam_staged = false;
    if ((s = ensure(ctx, ctx->out_rgb, (size_t)n * 16)) != BRE_OK) return s;
    // K0's rays are a width x height image by construction: the tile path applies whatever the caller's grid hint says
    const uint32_t gw = ctx->grid_w, gh = ctx->grid_h;
    ctx->grid_w = (uint32_t)ctx->camera.width; ctx->grid_h = (uint32_t)ctx->camera.height;
    s = run_query(ctx, n, (const float4 *)ctx->rays_o.p, (const float4 *)ctx->rays_d.p, nPathsNorm,
                  (float4 *)ctx->out_rgb.p);
    ctx->grid_w = gw; ctx->grid_h = gh;
    if (s != BRE_OK) return s;
    return film_accumulate(ctx, n, nullptr, (const float4 *)ctx->out_rgb.p);
}

int bre_count(bre_ctx *ctx, uint32_t n, const float *o4, const float *d3) {
    CTX_CHECK();
    int s = query_ready(ctx);
    if (s != BRE_OK) return s;
    if (n > 0) {
        if (!o4 || !d3) return set_err(ctx, BRE_ERR_INVALID, "null ray arrays");
        if ((s = upload_rays(ctx, n, o4, d3)) != BRE_OK) return s;
    }
    return run_count(ctx, n, (const float4 *)ctx->rays_o.p, (const float4 *)ctx->rays_d.p);
}

int bre_count_d(bre_ctx *ctx, uint32_t n, const void *ro, const void *rd) {
    CTX_CHECK();
    int s = query_ready(ctx);
    if (s != BRE_OK) return s;
    return run_count(ctx, n, (const float4 *)ro, (const float4 *)rd);
}

int bre_get_counters(bre_ctx *ctx, bre_counters *out) {
    if (!ctx || !out) return BRE_ERR_INVALID;
    *out = ctx->counters;
    return BRE_OK;
}

// ---------------------------------------------------------------- film
int bre_film_create(bre_ctx *ctx, uint32_t nPixels) {
    CTX_CHECK();
    if (nPixels == 0) return set_err(ctx, BRE_ERR_INVALID, "empty film");
    if (ctx->film_peers_open && ctx->film.p && ctx->film.cap < (size_t)nPixels * 16) {
        // peers still map this film (photon-sharded gather): keep it alive until they have unmapped it
        ctx->p2p_zombies.push_back(ctx->film.p);
        ctx->film.p = nullptr; ctx->film.cap = 0;
    }
    int s = ensure(ctx, ctx->film, (size_t)nPixels * 16);
    if (s != BRE_OK) return s;
    ctx->n_pixels = nPixels;
    return bre_film_clear(ctx);
}
int bre_film_clear(bre_ctx *ctx) {
    CTX_CHECK();
    if (!ctx->film.p) return set_err(ctx, BRE_ERR_INVALID, "bre_film_create has not been called");
    BRE_CUDA(cudaMemsetAsync(ctx->film.p, 0, (size_t)ctx->n_pixels * 16, ctx->stream));
    return BRE_OK;
}
int bre_film_accumulate(bre_ctx *ctx, uint32_t n, const uint32_t *pix, const float *rgb) {
    CTX_CHECK();
    if (n == 0) return BRE_OK;
    if (!rgb) return set_err(ctx, BRE_ERR_INVALID, "null rgb");
    int s;
    if ((s = ensure(ctx, ctx->out_rgb, (size_t)n * 16)) != BRE_OK) return s;
    if ((s = upload_pack(ctx, rgb, n, nullptr, 0.f, (float4 *)ctx->out_rgb.p)) != BRE_OK) return s;
    const uint32_t *pd = nullptr;
    if (pix) {
        if ((s = ensure(ctx, ctx->pix_idx, (size_t)n * 4)) != BRE_OK) return s;
        for (uint32_t q = 0; q < n; ++q)
            if (pix[q] >= ctx->n_pixels) return set_err(ctx, BRE_ERR_INVALID, "pixel index out of range");
        BRE_CUDA(cudaMemcpyAsync(ctx->pix_idx.p, pix, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
        pd = (const uint32_t *)ctx->pix_idx.p;
    }
    s = film_accumulate(ctx, n, pd, (const float4 *)ctx->out_rgb.p);
    if (s != BRE_OK) return s;
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    return BRE_OK;
}
int bre_film_accumulate_d(bre_ctx *ctx, uint32_t n, const uint32_t *pix_d, const void *rgb_d) {
    CTX_CHECK();
    return film_accumulate(ctx, n, pix_d, (const float4 *)rgb_d);
}
int bre_film_resolve(bre_ctx *ctx, int iter, float scale, float *rgb_out) {
    CTX_CHECK();
    if (!rgb_out || iter < 0) return set_err(ctx, BRE_ERR_INVALID, "bad resolve arguments");
    return film_resolve(ctx, iter, scale, rgb_out, 0);
}
int bre_film_mean(bre_ctx *ctx, int iter, float *rgb_out) {
    CTX_CHECK();
    if (!rgb_out || iter < 0) return set_err(ctx, BRE_ERR_INVALID, "bad resolve arguments");
    return film_resolve(ctx, iter, 1.f, rgb_out, 1);
}

int bre_film_gather(bre_ctx *ctx, int iter, float scale, float *rgb_out, int mean_only) {
    CTX_CHECK();
    if (iter < 0) return set_err(ctx, BRE_ERR_INVALID, "bad resolve arguments");
    return film_gather(ctx, iter, scale, rgb_out, mean_only);
}

// ---------------------------------------------------------------- timing
int bre_set_timing(bre_ctx *ctx, int enabled) {
    CTX_CHECK();
    if (enabled && !ctx->ev[0][0])
        for (int i = 0; i < 8; ++i)
            for (int j = 0; j < 2; ++j) BRE_CUDA(cudaEventCreate(&ctx->ev[i][j]));
    ctx->timing = enabled != 0;
    return BRE_OK;
}
int bre_get_timings(bre_ctx *ctx, float ms[8]) {
    CTX_CHECK();
    if (!ms) return set_err(ctx, BRE_ERR_INVALID, "null output");
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    for (int i = 0; i < 8; ++i) {
        ms[i] = 0.f;
        if (ctx->ev_valid[i]) BRE_CUDA(cudaEventElapsedTime(&ms[i], ctx->ev[i][0], ctx->ev[i][1]));
    }
    return BRE_OK;
}

// ---------------------------------------------------------------- sharding helpers
void bre_shard_range(uint64_t n, int nRanks, int rank, uint64_t *first, uint64_t *count) {
    if (nRanks < 1) nRanks = 1;
    uint64_t base = n / (uint64_t)nRanks, rem = n % (uint64_t)nRanks;
    uint64_t r = (uint64_t)rank;
    uint64_t f = r * base + (r < rem ? r : rem);
    uint64_t c = base + (r < rem ? 1 : 0);
    if (first) *first = f;
    if (count) *count = c;
}
int bre_tile_owner(uint32_t tx, uint32_t ty, uint32_t tilesX, int nRanks) {
    if (nRanks < 1) nRanks = 1;
    return (int)(((uint64_t)ty * tilesX + tx) % (uint64_t)nRanks);
}

}  // extern "C"
