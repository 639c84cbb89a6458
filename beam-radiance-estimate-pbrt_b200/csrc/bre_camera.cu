// bre_camera.cu -- K0: camera rays on the device, and the surface (triangle) inputs of the C ABI.
//
// Replaces (reference src/) the ray feed of the camera pass, integrators/photonbeam.cpp:444-481:
// PerspectiveCamera::GenerateRayDifferential for a pinhole (cameras/perspective.cpp:95-103),
// Transform::operator()(Ray) with its error-bound origin shift (core/transform.h:251-264, :278-300)
// and Scene::Intersect shortening ray.tMax to the first surface (core/primitive.cpp:97-101).
// One thread per pixel; rays are written as the two float4 arrays K5 reads, so a camera pass moves
// no ray data over PCIe.
#include "bre_internal.cuh"
#include "bre_surface.cuh"
#include <algorithm>
#include <cstring>
#include <vector>

namespace bre {

struct CameraDev {
    float r2c[16], c2w[12];
    float oErr[3];
    int x0, y0, w, h, inMedium;
    int haveBox;             // no triangles: the absorbing scene box of bre_set_bounds ends the camera rays too
    float blo[3], bhi[3];
};

// Camera samples of the reference's HaltonSampler (samplers/halton.cpp): the film position of pixel p in iteration
// `sample` is p + (dimension 0, dimension 1) of HaltonSampler(nIterations, pixelBounds) after StartPixel(p),
// SetSampleNumber(sample) (integrators/photonbeam.cpp:341,459-466; GetCameraSample core/sampler.cpp:46-52).
// Host part = the constructor (halton.cpp:64-91); device part = GetIndexForSample (:94-114) and SampleDimension for the
// unscrambled dimensions 0 / 1 (:116-125; InverseRadicalInverse lowdiscrepancy.h:82-91, RadicalInverse
// lowdiscrepancy.cpp:437-447, RadicalInverseSpecialized<3> :395-408).  Bit-exact with the oracle's restatement,
// which is pinned to the reference class (tests/golden/ref_halton.npz).
struct HaltonDev {
    int on;                       // 0: the uniform (offx, offy) of the launch
    int exp0, exp1;               // baseExponents
    uint32_t scale0, scale1;      // baseScales
    uint64_t stride, mi0, mi1;    // sampleStride, multInverse
    uint64_t sample;              // sample number (= iteration)
};
__device__ __forceinline__ uint64_t inv_radical_inverse(uint64_t inverse, uint32_t base, int nDigits) {
    uint64_t index = 0;
    for (int i = 0; i < nDigits; ++i) {
        const uint64_t digit = inverse % base;
        inverse /= base;
        index = index * base + digit;
    }
    return index;
}
__device__ __forceinline__ void halton_camera_sample(const HaltonDev &hs, int px, int py, float &u, float &v) {
    uint64_t offset = 0;
    if (hs.stride > 1) {
        const int mx = ((px % 128) + 128) % 128, my = ((py % 128) + 128) % 128;      // Mod(p, kMaxResolution)
        const uint64_t d0 = inv_radical_inverse((uint64_t)mx, 2u, hs.exp0), d1 = inv_radical_inverse((uint64_t)my, 3u, hs.exp1);
        offset = (d0 * (hs.stride / hs.scale0) * hs.mi0 + d1 * (hs.stride / hs.scale1) * hs.mi1) % hs.stride;
    }
    const uint64_t index = offset + hs.sample * hs.stride;
    // base 2: ReverseBits64(a) * 0x1p-64 is a uint64 x DOUBLE product narrowed to float
    u = __double2float_rn(__ull2double_rn(__brevll(index >> hs.exp0)) * 0x1p-64);
    // base 3: digits reversed into an integer, scaled by (1/3)^n accumulated in float
    uint64_t a = index / hs.scale1, rev = 0;
    const float invBase = 1.f / 3.f;
    float invBaseN = 1.f;
    while (a) {
        const uint64_t next = a / 3u;
        rev = rev * 3u + (a - next * 3u);
        invBaseN = __fmul_rn(invBaseN, invBase);
        a = next;
    }
    v = fminf(__fmul_rn(__ull2float_rn(rev), invBaseN), 0x1.fffffep-1f);     // min(., OneMinusEpsilon)
}

__global__ void __launch_bounds__(128) k_camera_rays(CameraDev cam, MeshDev mesh, float offx, float offy, HaltonDev hs,
                                                     float4 *__restrict__ ro, float4 *__restrict__ rd) {
    uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= (uint32_t)(cam.w * cam.h)) return;
    int x = (int)(q % (uint32_t)cam.w), y = (int)(q / (uint32_t)cam.w);
    if (hs.on) halton_camera_sample(hs, cam.x0 + x, cam.y0 + y, offx, offy);
    float fx = (float)(cam.x0 + x) + offx, fy = (float)(cam.y0 + y) + offy, fz = 0.f;
    float pc[3];
#pragma unroll
    for (int r = 0; r < 3; ++r)
        pc[r] = cam.r2c[4 * r] * fx + cam.r2c[4 * r + 1] * fy + cam.r2c[4 * r + 2] * fz + cam.r2c[4 * r + 3];
    float wp = cam.r2c[12] * fx + cam.r2c[13] * fy + cam.r2c[14] * fz + cam.r2c[15];
    if (wp != 1.f) { float inv = 1 / wp; pc[0] *= inv; pc[1] *= inv; pc[2] *= inv; }
    normalize3(pc);
    float o[3], d[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        o[r] = cam.c2w[4 * r] * 0.f + cam.c2w[4 * r + 1] * 0.f + cam.c2w[4 * r + 2] * 0.f + cam.c2w[4 * r + 3];
        d[r] = cam.c2w[4 * r] * pc[0] + cam.c2w[4 * r + 1] * pc[1] + cam.c2w[4 * r + 2] * pc[2];
    }
    float l2 = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
    float tMax = __int_as_float(0x7f800000);
    if (l2 > 0) {
        float dt = (fabsf(d[0]) * cam.oErr[0] + fabsf(d[1]) * cam.oErr[1] + fabsf(d[2]) * cam.oErr[2]) / l2;
#pragma unroll
        for (int r = 0; r < 3; ++r) o[r] += d[r] * dt;
        tMax -= dt;
    }
    SurfHit h;
    if (mesh.n) { if (mesh_intersect(mesh, o, d, tMax, h, false)) tMax = h.t; }
    else if (cam.haveBox) {
        // Bounds3::IntersectP(ray, &t0, &t1), geometry.h:1386-1408: the ray ends where it leaves the box
        float t0 = 0.f, t1 = tMax;
        bool hit = true;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            const float invD = 1.f / d[i];
            float tNear = (cam.blo[i] - o[i]) * invD, tFar = (cam.bhi[i] - o[i]) * invD;
            if (tNear > tFar) { const float sw = tNear; tNear = tFar; tFar = sw; }
            tFar *= BRE_SLAB_WIDEN;
            t0 = tNear > t0 ? tNear : t0;
            t1 = tFar < t1 ? tFar : t1;
            if (t0 > t1) hit = false;
        }
        tMax = hit ? t1 : 0.f;
    }
    float len = sqrtf(l2), inv = 1 / len;
    float t = cam.inMedium ? fminf(tMax * len, 3.402823466e+38f) : 0.f;
    ro[q] = make_float4(o[0], o[1], o[2], t);
    rd[q] = make_float4(d[0] * inv, d[1] * inv, d[2] * inv, 0.f);
}

// halton.cpp:45-61
static void extended_gcd(uint64_t a, uint64_t b, int64_t *x, int64_t *y) {
    if (b == 0) { *x = 1; *y = 0; return; }
    int64_t d = (int64_t)(a / b), xp, yp;
    extended_gcd(b, a % b, &xp, &yp);
    *x = yp;
    *y = xp - (d * yp);
}
static uint64_t multiplicative_inverse(int64_t a, int64_t n) {
    int64_t x, y;
    extended_gcd((uint64_t)a, (uint64_t)n, &x, &y);
    int64_t r = x - (x / n) * n;
    return (uint64_t)(r < 0 ? r + n : r);
}

int camera_rays(bre_ctx *ctx, float offx, float offy, cudaStream_t st, long long haltonSample) {
    if (!st) st = ctx->stream;
    if (!ctx->have_camera) return set_err(ctx, BRE_ERR_INVALID, "bre_set_camera has not been called");
    const bre_camera &c = ctx->camera;
    const uint32_t n = (uint32_t)c.width * (uint32_t)c.height;
    int s;
    if ((s = ensure(ctx, ctx->rays_o, (size_t)n * 16)) != BRE_OK) return s;
    if ((s = ensure(ctx, ctx->rays_d, (size_t)n * 16)) != BRE_OK) return s;
    CameraDev cd;
    memcpy(cd.r2c, c.raster_to_camera, sizeof(cd.r2c));
    memcpy(cd.c2w, c.camera_to_world, sizeof(cd.c2w));
    const float g3 = (3 * 5.9604644775390625e-8f) / (1 - 3 * 5.9604644775390625e-8f);
    for (int r = 0; r < 3; ++r) {   // transform.h:288-294 for the camera-space origin (0,0,0)
        const float *m = c.camera_to_world + 4 * r;
        cd.oErr[r] = g3 * (std::fabs(m[0] * 0.f) + std::fabs(m[1] * 0.f) + std::fabs(m[2] * 0.f) + std::fabs(m[3]));
    }
    cd.x0 = c.x0; cd.y0 = c.y0; cd.w = c.width; cd.h = c.height; cd.inMedium = c.in_medium;
    cd.haveBox = (ctx->n_tri == 0 && ctx->have_bounds) ? 1 : 0;
    for (int r = 0; r < 3; ++r) { cd.blo[r] = ctx->blo[r]; cd.bhi[r] = ctx->bhi[r]; }
    MeshDev mesh{ctx->n_tri, (const float4 *)ctx->tri.p, (const float *)ctx->tri_kdb.p};
    HaltonDev hs{};
    if (haltonSample >= 0) {
        // HaltonSampler constructor for sampleBounds = the camera's pixel window (halton.cpp:73-91)
        const int res[2] = {c.width, c.height};
        uint32_t scale[2];
        int ex[2];
        for (int i = 0; i < 2; ++i) {
            const int base = i == 0 ? 2 : 3, lim = std::min(res[i], 128);
            int sc = 1, e = 0;
            while (sc < lim) { sc *= base; ++e; }
            scale[i] = (uint32_t)sc; ex[i] = e;
        }
        hs.on = 1; hs.exp0 = ex[0]; hs.exp1 = ex[1]; hs.scale0 = scale[0]; hs.scale1 = scale[1];
        hs.stride = (uint64_t)scale[0] * scale[1];
        hs.mi0 = multiplicative_inverse(scale[1], scale[0]);
        hs.mi1 = multiplicative_inverse(scale[0], scale[1]);
        hs.sample = (uint64_t)haltonSample;
    }
    k_camera_rays<<<(n + 127) / 128, 128, 0, st>>>(cd, mesh, offx, offy, hs, (float4 *)ctx->rays_o.p,
                                                           (float4 *)ctx->rays_d.p);
    BRE_CUDA(cudaGetLastError());
    return BRE_OK;
}

}  // namespace bre

using namespace bre;

extern "C" {

int bre_set_triangles(bre_ctx *ctx, uint32_t nTri, const float *P, const uint32_t *idx, const float *Kd) {
    if (!ctx) return BRE_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (nTri == 0) { ctx->n_tri = 0; return BRE_OK; }
    if (!P || !idx || !Kd) return set_err(ctx, BRE_ERR_INVALID, "null triangle arrays");
    std::vector<float4> tri(3 * (size_t)nTri);
    std::vector<float> kdb(nTri);
    for (uint32_t i = 0; i < nTri; ++i) {
        const float *p0 = P + 3 * (size_t)idx[3 * i], *p1 = P + 3 * (size_t)idx[3 * i + 1], *p2 = P + 3 * (size_t)idx[3 * i + 2];
        tri[3 * i] = make_float4(p0[0], p0[1], p0[2], 0.f);   // attr bits = 0
        tri[3 * i + 1] = make_float4(p1[0], p1[1], p1[2], Kd[3 * i]);
        tri[3 * i + 2] = make_float4(p2[0], p2[1], p2[2], Kd[3 * i + 1]);
        kdb[i] = Kd[3 * i + 2];
    }
    int s;
    if ((s = ensure(ctx, ctx->tri, tri.size() * 16)) != BRE_OK) return s;
    if ((s = ensure(ctx, ctx->tri_kdb, kdb.size() * 4)) != BRE_OK) return s;
    BRE_CUDA(cudaMemcpyAsync(ctx->tri.p, tri.data(), tri.size() * 16, cudaMemcpyHostToDevice, ctx->stream));
    BRE_CUDA(cudaMemcpyAsync(ctx->tri_kdb.p, kdb.data(), kdb.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->n_tri = nTri;
    return BRE_OK;
}

int bre_set_triangle_attrs(bre_ctx *ctx, uint32_t nTri, const uint8_t *attrs) {
    if (!ctx) return BRE_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (nTri != ctx->n_tri || !attrs) return set_err(ctx, BRE_ERR_INVALID, "attrs must follow bre_set_triangles with the same n_tri");
    std::vector<int32_t> a(nTri);
    for (uint32_t i = 0; i < nTri; ++i) a[i] = attrs[i];
    // the attribute word is the .w of the first float4 of each triangle (stride 48 bytes)
    BRE_CUDA(cudaMemcpy2DAsync((char *)ctx->tri.p + 12, 48, a.data(), 4, 4, nTri, cudaMemcpyHostToDevice, ctx->stream));
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    return BRE_OK;
}

int bre_set_lights_in_medium(bre_ctx *ctx, int inMedium) {
    if (!ctx) return BRE_ERR_INVALID;
    ctx->lights_in_medium = inMedium ? 1 : 0;
    return BRE_OK;
}

int bre_set_camera(bre_ctx *ctx, const bre_camera *cam) {
    if (!ctx) return BRE_ERR_INVALID;
    if (!cam || cam->width < 1 || cam->height < 1 || (uint64_t)cam->width * (uint64_t)cam->height > 0x7fffffffull)
        return set_err(ctx, BRE_ERR_INVALID, "bad camera window");
    if (cam->camera_to_world[12] != 0.f || cam->camera_to_world[13] != 0.f || cam->camera_to_world[14] != 0.f ||
        cam->camera_to_world[15] != 1.f)
        return set_err(ctx, BRE_ERR_UNSUPPORTED, "camera_to_world must be affine");
    ctx->camera = *cam;
    ctx->have_camera = true;
    ctx->cam_staged = false;
    return BRE_OK;
}

}  // extern "C"
