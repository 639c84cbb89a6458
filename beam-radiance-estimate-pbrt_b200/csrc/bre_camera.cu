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

__global__ void __launch_bounds__(128) k_camera_rays(CameraDev cam, MeshDev mesh, float offx, float offy,
                                                     float4 *__restrict__ ro, float4 *__restrict__ rd) {
    uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= (uint32_t)(cam.w * cam.h)) return;
    int x = (int)(q % (uint32_t)cam.w), y = (int)(q / (uint32_t)cam.w);
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

int camera_rays(bre_ctx *ctx, float offx, float offy, cudaStream_t st) {
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
    k_camera_rays<<<(n + 127) / 128, 128, 0, st>>>(cd, mesh, offx, offy, (float4 *)ctx->rays_o.p,
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
