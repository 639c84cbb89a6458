// bre_film.cu -- K7: per-pixel film accumulation and resolve.
//
// Replaces (reference src/): PhotonBeamPixel::Ld += ... (integrators/photonbeam.cpp:504),
// L = Ld / (iter+1) (:578), Film::SetImage (core/film.cpp:132-140: RGB->XYZ, weight 1) and the
// arithmetic of Film::WriteImage (core/film.cpp:168-204: XYZ->RGB, /weight, max(0,.), *scale)
// with RGBToXYZ / XYZToRGB of core/spectrum.h:56-66.
#include "bre_internal.cuh"

#include <mutex>
#include <string>

namespace bre {

__global__ void __launch_bounds__(256) k_film_add(float4 *__restrict__ film, uint32_t n, uint32_t nPixels,
                                                  const uint32_t *__restrict__ pix,
                                                  const float4 *__restrict__ rgb) {
    uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    float4 v = __ldg(rgb + q);
    if (pix == nullptr) {
        float4 f = film[q];
        f.x = fadd(f.x, v.x); f.y = fadd(f.y, v.y); f.z = fadd(f.z, v.z);
        film[q] = f;
    } else {
        const uint32_t p = __ldg(pix + q);
        if (p >= nPixels) return;             // a caller's index beyond the film is dropped, never written
        float *f = reinterpret_cast<float *>(film + p);
        atomicAdd(f + 0, v.x); atomicAdd(f + 1, v.y); atomicAdd(f + 2, v.z);
    }
}

__global__ void __launch_bounds__(256) k_film_resolve(const float4 *__restrict__ film, uint32_t n,
                                                      float div, float scale, int meanOnly,
                                                      float *__restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 f = __ldg(film + i);
    // Ld / (iter + 1): the reference divides every channel (spectrum.h:181-188), no reciprocal
    float r = fdiv(f.x, div), g = fdiv(f.y, div), b = fdiv(f.z, div);
    if (meanOnly) {   // L = Ld / (iter+1), photonbeam.cpp:578: what the host hands to Film::SetImage
        out[3 * (size_t)i + 0] = r; out[3 * (size_t)i + 1] = g; out[3 * (size_t)i + 2] = b;
        return;
    }
    // RGBToXYZ spectrum.h:62-66
    float X = fadd(fadd(fmul(0.412453f, r), fmul(0.357580f, g)), fmul(0.180423f, b));
    float Y = fadd(fadd(fmul(0.212671f, r), fmul(0.715160f, g)), fmul(0.072169f, b));
    float Z = fadd(fadd(fmul(0.019334f, r), fmul(0.119193f, g)), fmul(0.950227f, b));
    // XYZToRGB spectrum.h:56-60
    float R = fsub(fsub(fmul(3.240479f, X), fmul(1.537150f, Y)), fmul(0.498535f, Z));
    float G = fadd(fadd(fmul(-0.969256f, X), fmul(1.875991f, Y)), fmul(0.041556f, Z));
    float B = fadd(fsub(fmul(0.055648f, X), fmul(0.204043f, Y)), fmul(1.057311f, Z));
    // weight = 1, no splats: max(0, rgb * 1) * scale   film.cpp:178-201
    out[3 * (size_t)i + 0] = fmul(fmaxf(0.f, R), scale);
    out[3 * (size_t)i + 1] = fmul(fmaxf(0.f, G), scale);
    out[3 * (size_t)i + 2] = fmul(fmaxf(0.f, B), scale);
}

int film_accumulate(bre_ctx *ctx, uint32_t n, const uint32_t *pix_d, const float4 *rgb_d) {
    if (!ctx->film.p) return set_err(ctx, BRE_ERR_INVALID, "bre_film_create has not been called");
    if (pix_d == nullptr && n > ctx->n_pixels)
        return set_err(ctx, BRE_ERR_INVALID, "more rays than pixels and no pixel index");
    if (n == 0) return BRE_OK;
    {
        PhaseTimer pt(ctx, PH_FILM);
        k_film_add<<<(n + 255) / 256, 256, 0, ctx->stream>>>((float4 *)ctx->film.p, n, ctx->n_pixels, pix_d, rgb_d);
    }
    BRE_CUDA(cudaGetLastError());
    return BRE_OK;
}

int film_resolve(bre_ctx *ctx, int iter, float scale, float *rgb_out_host, int meanOnly) {
    if (!ctx->film.p) return set_err(ctx, BRE_ERR_INVALID, "bre_film_create has not been called");
    uint32_t n = ctx->n_pixels;
    int s;
    if ((s = ensure(ctx, ctx->out_rgb, (size_t)n * 12)) != BRE_OK) return s;
    const float div = (float)(iter + 1);     // the divisor (the kernel divides, like the reference)
    k_film_resolve<<<(n + 255) / 256, 256, 0, ctx->stream>>>((const float4 *)ctx->film.p, n, div,
                                                             scale, meanOnly, (float *)ctx->out_rgb.p);
    BRE_CUDA(cudaGetLastError());
    BRE_CUDA(cudaMemcpyAsync(rgb_out_host, ctx->out_rgb.p, (size_t)n * 12, cudaMemcpyDeviceToHost,
                             ctx->stream));
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    return BRE_OK;
}

}  // namespace bre

// ---------------------------------------------------------------- K7 without any allocation
// bre_film_resolve_static: the resolve kernel above with its device staging buffer (the RGB triples before the
// device -> host copy; bre_film_resolve takes it from a grow-only cudaMalloc) in a file-scope device array.  For
// callers that must not allocate: CUDA graph capture, serving loops, harnesses that forbid cudaMalloc.  No
// counterpart in the reference.
//   * BRE_STATIC_RESOLVE_MAX_PIXELS pixels x 12 bytes = 24 MiB, held by every GPU that loads libbre.so until exit.
//   * In host code `g_resolve_scratch` is a host-side stand-in, not a device address: the address the kernel gets
//     comes from cudaGetSymbolAddress, once per device, with that device current.
//   * Every call on a device shares the array; it keeps the last call's contents.  Nothing has to start at zero
//     (the kernel writes all 3 n floats before the copy reads them), but two calls that overlap on one GPU (two
//     contexts / streams / threads) race on it: serialise them.
static __device__ __align__(256) float g_resolve_scratch[3u * BRE_STATIC_RESOLVE_MAX_PIXELS];

namespace bre {
constexpr int STATIC_MAX_DEVICES = 64;
static float *g_resolve_scratch_addr[STATIC_MAX_DEVICES] = {};
static std::mutex g_resolve_scratch_mu;

// device address of g_resolve_scratch on ctx->device (which is current: CTX_CHECK-style cudaSetDevice by the caller)
static int resolve_scratch_ptr(bre_ctx *ctx, float **out) {
    if (ctx->device < 0 || ctx->device >= STATIC_MAX_DEVICES)
        return set_err(ctx, BRE_ERR_UNSUPPORTED, "bre_film_resolve_static: device ordinal beyond 63");
    std::lock_guard<std::mutex> lock(g_resolve_scratch_mu);
    float *&slot = g_resolve_scratch_addr[ctx->device];
    if (!slot) {
        float *ptr = nullptr;
        BRE_CUDA(cudaGetSymbolAddress((void **)&ptr, g_resolve_scratch));
        if (!ptr) return set_err(ctx, BRE_ERR_CUDA, "cudaGetSymbolAddress(g_resolve_scratch) returned a null address");
        slot = ptr;
    }
    *out = slot;
    return BRE_OK;
}
}  // namespace bre

extern "C" int bre_film_resolve_static(bre_ctx *ctx, int iter, float scale, float *rgb_out, int mean_only) {
    using namespace bre;
    if (!ctx) return BRE_ERR_INVALID;
    BRE_CUDA(cudaSetDevice(ctx->device));
    if (!rgb_out || iter < 0) return set_err(ctx, BRE_ERR_INVALID, "bad resolve arguments");
    if (!ctx->film.p) return set_err(ctx, BRE_ERR_INVALID, "bre_film_create has not been called");
    const uint32_t n = ctx->n_pixels;
    if (n > BRE_STATIC_RESOLVE_MAX_PIXELS)
        return set_err(ctx, BRE_ERR_UNSUPPORTED,
                       "bre_film_resolve_static: film of " + std::to_string(n) + " pixels exceeds the static scratch of " +
                           std::to_string(BRE_STATIC_RESOLVE_MAX_PIXELS) + " pixels (BRE_STATIC_RESOLVE_MAX_PIXELS); use bre_film_resolve");
    float *scratch = nullptr;
    int s = resolve_scratch_ptr(ctx, &scratch);
    if (s != BRE_OK) return s;
    const float div = (float)(iter + 1);     // the divisor (the kernel divides, like the reference)
    k_film_resolve<<<(n + 255) / 256, 256, 0, ctx->stream>>>((const float4 *)ctx->film.p, n, div, scale, mean_only ? 1 : 0,
                                                             scratch);
    BRE_CUDA(cudaGetLastError());
    BRE_CUDA(cudaMemcpyAsync(rgb_out, scratch, (size_t)n * 12, cudaMemcpyDeviceToHost, ctx->stream));
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    return BRE_OK;
}
