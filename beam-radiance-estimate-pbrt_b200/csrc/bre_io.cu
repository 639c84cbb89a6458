// bre_io.cu -- on-disk interchange: photon dump / load and the film checkpoint.
//
// The reference only gestures at resuming ("startiteration" fast-forwards the radius,
// integrators/photonbeam.cpp:354-356,594-595, but the accumulated image is lost).  These two files
// make a render resumable and give the parity path (photons dumped by one implementation, loaded
// into another) a stable format.  Little endian, fixed 64-byte header, then the device SoA arrays
// as they sit in HBM (float4 records), streamed through a pinned staging buffer in 32 MiB pieces.
//
//   photons: "BREPHOT\0" | u32 version=1 | u32 0 | u64 n | u64 n_paths | 32 bytes 0
//            | pos_r[n] {x,y,z,radius} | pow[n] {r,g,b,0} | dir[n] {wx,wy,wz,0}
//   film:    "BREFILM\0" | u32 version=1 | u32 n_pixels | u64 iterations_done | u64 0 | f32 radius | 28 bytes 0
//            | film[n_pixels] {sum r, sum g, sum b, 0}
#include "bre_internal.cuh"
#include <cstdio>
#include <cstring>

namespace bre {

int grow_photon_store(bre_ctx *ctx, uint64_t newCount);

static const size_t IO_CHUNK = 32u << 20;

struct Staging {
    void *p = nullptr;
    ~Staging() { if (p) cudaFreeHost(p); }
    bool alloc() { return cudaMallocHost(&p, IO_CHUNK) == cudaSuccess; }
};

static int write_device(bre_ctx *ctx, FILE *fp, const void *dev, size_t bytes, Staging &st) {
    for (size_t off = 0; off < bytes; off += IO_CHUNK) {
        size_t len = bytes - off < IO_CHUNK ? bytes - off : IO_CHUNK;
        BRE_CUDA(cudaMemcpyAsync(st.p, (const char *)dev + off, len, cudaMemcpyDeviceToHost, ctx->stream));
        BRE_CUDA(cudaStreamSynchronize(ctx->stream));
        if (fwrite(st.p, 1, len, fp) != len) return set_err(ctx, BRE_ERR_INVALID, "short write");
    }
    return BRE_OK;
}
static int read_device(bre_ctx *ctx, FILE *fp, void *dev, size_t bytes, Staging &st) {
    for (size_t off = 0; off < bytes; off += IO_CHUNK) {
        size_t len = bytes - off < IO_CHUNK ? bytes - off : IO_CHUNK;
        if (fread(st.p, 1, len, fp) != len) return set_err(ctx, BRE_ERR_INVALID, "file is truncated");
        BRE_CUDA(cudaMemcpyAsync((char *)dev + off, st.p, len, cudaMemcpyHostToDevice, ctx->stream));
        BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return BRE_OK;
}

struct Header {
    char magic[8];
    uint32_t version, u32a;
    uint64_t u64a, u64b;
    float f32a;
    char pad[28];
};
static_assert(sizeof(Header) == 64, "header layout");

}  // namespace bre

using namespace bre;

extern "C" {

int bre_save_photons(bre_ctx *ctx, const char *path, uint64_t n_paths) {
    if (!ctx || !path) return BRE_ERR_INVALID;
    cudaSetDevice(ctx->device);
    FILE *fp = fopen(path, "wb");
    if (!fp) return set_err(ctx, BRE_ERR_INVALID, std::string("cannot open ") + path);
    Staging st;
    if (!st.alloc()) { fclose(fp); return set_err(ctx, BRE_ERR_NOMEM, "pinned staging buffer"); }
    Header h;
    memset(&h, 0, sizeof(h));
    memcpy(h.magic, "BREPHOT", 8);
    h.version = 1; h.u64a = ctx->n_photons; h.u64b = n_paths;
    int s = fwrite(&h, 1, 64, fp) == 64 ? BRE_OK : set_err(ctx, BRE_ERR_INVALID, "short write");
    const size_t bytes = (size_t)ctx->n_photons * sizeof(float4);
    const void *arrays[3] = {ctx->pos_r.p, ctx->pow.p, ctx->dir.p};
    for (int a = 0; a < 3 && s == BRE_OK && bytes; ++a) s = write_device(ctx, fp, arrays[a], bytes, st);
    if (fclose(fp) != 0 && s == BRE_OK) s = set_err(ctx, BRE_ERR_INVALID, "close failed");
    return s;
}

int bre_load_photons(bre_ctx *ctx, const char *path, int append, uint64_t *n_loaded, uint64_t *n_paths) {
    if (!ctx || !path) return BRE_ERR_INVALID;
    cudaSetDevice(ctx->device);
    FILE *fp = fopen(path, "rb");
    if (!fp) return set_err(ctx, BRE_ERR_INVALID, std::string("cannot open ") + path);
    Header h;
    int s = BRE_OK;
    if (fread(&h, 1, 64, fp) != 64 || memcmp(h.magic, "BREPHOT", 8) != 0)
        s = set_err(ctx, BRE_ERR_INVALID, "not a photon file (bad magic)");
    else if (h.version != 1)
        s = set_err(ctx, BRE_ERR_UNSUPPORTED, "unknown photon file version");
    Staging st;
    if (s == BRE_OK && !st.alloc()) s = set_err(ctx, BRE_ERR_NOMEM, "pinned staging buffer");
    if (s == BRE_OK) {
        // refuse before touching the store when the file is shorter than its header claims
        long here = ftell(fp);
        fseek(fp, 0, SEEK_END);
        uint64_t avail = (uint64_t)(ftell(fp) - here);
        fseek(fp, here, SEEK_SET);
        if (h.u64a > (1ull << 40) || avail < h.u64a * 48ull) s = set_err(ctx, BRE_ERR_INVALID, "file is truncated");
    }
    if (s == BRE_OK) {
        const uint64_t n = h.u64a, base = append ? ctx->n_photons : 0;
        if (!append) ctx->n_photons = 0;
        ctx->built = false; ctx->topo_valid = false;
        s = grow_photon_store(ctx, base + n);
        const size_t bytes = (size_t)n * sizeof(float4);
        void *arrays[3] = {ctx->pos_r.p, ctx->pow.p, ctx->dir.p};
        for (int a = 0; a < 3 && s == BRE_OK && bytes; ++a)
            s = read_device(ctx, fp, (float4 *)arrays[a] + base, bytes, st);
        if (s == BRE_OK) {
            ctx->n_photons = base + n;
            if (n_loaded) *n_loaded = n;
            if (n_paths) *n_paths = h.u64b;
        }
    }
    fclose(fp);
    return s;
}

int bre_film_save(bre_ctx *ctx, const char *path, uint64_t iterations_done, float radius) {
    if (!ctx || !path) return BRE_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (!ctx->film.p) return set_err(ctx, BRE_ERR_INVALID, "bre_film_create has not been called");
    FILE *fp = fopen(path, "wb");
    if (!fp) return set_err(ctx, BRE_ERR_INVALID, std::string("cannot open ") + path);
    Staging st;
    if (!st.alloc()) { fclose(fp); return set_err(ctx, BRE_ERR_NOMEM, "pinned staging buffer"); }
    Header h;
    memset(&h, 0, sizeof(h));
    memcpy(h.magic, "BREFILM", 8);
    h.version = 1; h.u32a = ctx->n_pixels; h.u64a = iterations_done; h.f32a = radius;
    int s = fwrite(&h, 1, 64, fp) == 64 ? BRE_OK : set_err(ctx, BRE_ERR_INVALID, "short write");
    if (s == BRE_OK) s = write_device(ctx, fp, ctx->film.p, (size_t)ctx->n_pixels * 16, st);
    if (fclose(fp) != 0 && s == BRE_OK) s = set_err(ctx, BRE_ERR_INVALID, "close failed");
    return s;
}

int bre_film_load(bre_ctx *ctx, const char *path, uint64_t *iterations_done, float *radius) {
    if (!ctx || !path) return BRE_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (!ctx->film.p) return set_err(ctx, BRE_ERR_INVALID, "bre_film_create has not been called");
    FILE *fp = fopen(path, "rb");
    if (!fp) return set_err(ctx, BRE_ERR_INVALID, std::string("cannot open ") + path);
    Header h;
    int s = BRE_OK;
    if (fread(&h, 1, 64, fp) != 64 || memcmp(h.magic, "BREFILM", 8) != 0)
        s = set_err(ctx, BRE_ERR_INVALID, "not a film checkpoint (bad magic)");
    else if (h.version != 1)
        s = set_err(ctx, BRE_ERR_UNSUPPORTED, "unknown film checkpoint version");
    else if (h.u32a != ctx->n_pixels)
        s = set_err(ctx, BRE_ERR_INVALID, "film checkpoint has a different resolution");
    Staging st;
    if (s == BRE_OK && !st.alloc()) s = set_err(ctx, BRE_ERR_NOMEM, "pinned staging buffer");
    if (s == BRE_OK) {
        // stage into scratch first so a truncated file leaves the film untouched
        s = ensure(ctx, ctx->scratch, (size_t)ctx->n_pixels * 16);
        if (s == BRE_OK) s = read_device(ctx, fp, ctx->scratch.p, (size_t)ctx->n_pixels * 16, st);
        if (s == BRE_OK) {
            s = check_cuda(ctx, cudaMemcpyAsync(ctx->film.p, ctx->scratch.p, (size_t)ctx->n_pixels * 16,
                                                cudaMemcpyDeviceToDevice, ctx->stream), "film restore");
            if (s == BRE_OK) s = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "film restore");
        }
        if (s == BRE_OK) {
            if (iterations_done) *iterations_done = h.u64a;
            if (radius) *radius = h.f32a;
        }
    }
    fclose(fp);
    return s;
}

}  // extern "C"
