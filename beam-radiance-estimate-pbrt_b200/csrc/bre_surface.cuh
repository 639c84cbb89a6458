// bre_surface.cuh -- device-side surfaces for K1 (photon tracing) and the camera-ray kernel.
//
// Replaces (reference src/): Scene::Intersect (core/scene.cpp:50-54) over triangle shapes with the
// watertight ray-triangle test of Triangle::Intersect (shapes/triangle.cpp:177-392), the hit's medium
// interface (core/primitive.cpp:97-109), Interaction::SpawnRay / OffsetRayOrigin
// (core/interaction.h:64-67, core/geometry.h:1438-1458) and the single-lobe Lambertian
// BSDF::Sample_f of MatteMaterial (materials/matte.cpp:55-60, core/reflection.cpp:703-760, :378-389,
// core/sampling.cpp:113-130).  The library is compiled with -fmad=false, so the plain float
// expressions below round operation by operation exactly like the CPU oracle (-ffp-contract=off).
//
// Triangles are tested by brute force in index order against the shrinking tMax (closest hit): the
// scenes of this path are boxes and a few blockers (tens of triangles).  All
// lanes of a warp read the same triangle, so every load is an L1 broadcast.
#pragma once
#include "bre_internal.cuh"

namespace bre {

// tri[3*i+0] = {p0, attr bits}, tri[3*i+1] = {p1, Kd.r}, tri[3*i+2] = {p2, Kd.g}; kdb[i] = Kd.b
struct MeshDev {
    uint32_t n;
    const float4 *tri;
    const float *kdb;
};

struct SurfHit {
    float t;
    uint32_t tri;
    float p[3], pError[3], n[3], ss[3];
    uint32_t attr;
    float Kd[3];
};

__device__ __forceinline__ float gamma_n(int n) {   // core/pbrt.h:263-265
    const float eps = 5.9604644775390625e-8f;
    return (n * eps) / (1 - n * eps);
}
__device__ __forceinline__ float dot3(const float a[3], const float b[3]) {
    return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}
__device__ __forceinline__ void cross3(const float a[3], const float b[3], float out[3]) {
    double ax = a[0], ay = a[1], az = a[2], bx = b[0], by = b[1], bz = b[2];
    out[0] = (float)(__dsub_rn(__dmul_rn(ay, bz), __dmul_rn(az, by)));
    out[1] = (float)(__dsub_rn(__dmul_rn(az, bx), __dmul_rn(ax, bz)));
    out[2] = (float)(__dsub_rn(__dmul_rn(ax, by), __dmul_rn(ay, bx)));
}
__device__ __forceinline__ void normalize3(float v[3]) {
    float inv = 1.f / sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    v[0] *= inv; v[1] *= inv; v[2] *= inv;
}
__device__ __forceinline__ float max3abs(float a, float b, float c) {
    return fmaxf(fabsf(a), fmaxf(fabsf(b), fabsf(c)));
}

// The ray-dependent part of the watertight test (permutation + shear), computed once per ray.
struct TriRay {
    int kx, ky, kz;
    float ox, oy, oz, Sx, Sy, Sz;
    __device__ __forceinline__ void init(const float o[3], const float d[3]) {
        float ax = fabsf(d[0]), ay = fabsf(d[1]), az = fabsf(d[2]);
        kz = ax > ay ? (ax > az ? 0 : 2) : (ay > az ? 1 : 2);
        kx = kz + 1 == 3 ? 0 : kz + 1;
        ky = kx + 1 == 3 ? 0 : kx + 1;
        float dz = d[kz];
        Sx = -d[kx] / dz; Sy = -d[ky] / dz; Sz = 1.f / dz;
        ox = o[kx]; oy = o[ky]; oz = o[kz];
    }
};

__device__ __forceinline__ float comp(const float4 &v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : v.z); }

__device__ __forceinline__ bool tri_test(const TriRay &r, const float4 &P0, const float4 &P1,
                                         const float4 &P2, float tMax, float &tHit, float b[3]) {
    float a0 = comp(P0, r.kx) - r.ox, a1 = comp(P0, r.ky) - r.oy, a2 = comp(P0, r.kz) - r.oz;
    float b0 = comp(P1, r.kx) - r.ox, b1 = comp(P1, r.ky) - r.oy, b2 = comp(P1, r.kz) - r.oz;
    float c0 = comp(P2, r.kx) - r.ox, c1 = comp(P2, r.ky) - r.oy, c2 = comp(P2, r.kz) - r.oz;
    a0 += r.Sx * a2; a1 += r.Sy * a2;
    b0 += r.Sx * b2; b1 += r.Sy * b2;
    c0 += r.Sx * c2; c1 += r.Sy * c2;
    float e0 = b0 * c1 - b1 * c0;
    float e1 = c0 * a1 - c1 * a0;
    float e2 = a0 * b1 - a1 * b0;
    if (e0 == 0.f || e1 == 0.f || e2 == 0.f) {   // exactly on an edge: decide in double (:222-234)
        e0 = (float)__dsub_rn(__dmul_rn((double)c1, (double)b0), __dmul_rn((double)c0, (double)b1));
        e1 = (float)__dsub_rn(__dmul_rn((double)a1, (double)c0), __dmul_rn((double)a0, (double)c1));
        e2 = (float)__dsub_rn(__dmul_rn((double)b1, (double)a0), __dmul_rn((double)b0, (double)a1));
    }
    if ((e0 < 0 || e1 < 0 || e2 < 0) && (e0 > 0 || e1 > 0 || e2 > 0)) return false;
    float det = e0 + e1 + e2;
    if (det == 0) return false;
    a2 *= r.Sz; b2 *= r.Sz; c2 *= r.Sz;
    float tScaled = e0 * a2 + e1 * b2 + e2 * c2;
    if (det < 0 && (tScaled >= 0 || tScaled < tMax * det)) return false;
    else if (det > 0 && (tScaled <= 0 || tScaled > tMax * det)) return false;
    float invDet = 1 / det;
    float t = tScaled * invDet;
    float maxZt = max3abs(a2, b2, c2);
    float deltaZ = gamma_n(3) * maxZt;
    float maxXt = max3abs(a0, b0, c0);
    float maxYt = max3abs(a1, b1, c1);
    float deltaX = gamma_n(5) * (maxXt + maxZt);
    float deltaY = gamma_n(5) * (maxYt + maxZt);
    float deltaE = 2 * (gamma_n(2) * maxXt * maxYt + deltaY * maxXt + deltaX * maxYt);
    float maxE = max3abs(e0, e1, e2);
    float deltaT = 3 * (gamma_n(3) * maxE * maxZt + deltaE * maxZt + deltaZ * maxE) * fabsf(invDet);
    if (t <= deltaT) return false;
    b[0] = e0 * invDet; b[1] = e1 * invDet; b[2] = e2 * invDet;
    tHit = t;
    return true;
}

// closest hit; `full` = also fill hit point, error bound, normal and tangent
__device__ inline bool mesh_intersect(const MeshDev &mesh, const float o[3], const float d[3],
                                      float tMax, SurfHit &h, bool full) {
    TriRay r;
    r.init(o, d);
    int found = -1;
    float bb[3] = {0.f, 0.f, 0.f};
    for (uint32_t i = 0; i < mesh.n; ++i) {
        float4 P0 = __ldg(mesh.tri + 3 * i), P1 = __ldg(mesh.tri + 3 * i + 1), P2 = __ldg(mesh.tri + 3 * i + 2);
        float t, b[3];
        if (tri_test(r, P0, P1, P2, tMax, t, b)) {
            tMax = t; found = (int)i;
            bb[0] = b[0]; bb[1] = b[1]; bb[2] = b[2];
        }
    }
    if (found < 0) return false;
    h.t = tMax;
    h.tri = (uint32_t)found;
    if (!full) return true;
    float4 P0 = __ldg(mesh.tri + 3 * found), P1 = __ldg(mesh.tri + 3 * found + 1), P2 = __ldg(mesh.tri + 3 * found + 2);
    const float p0[3] = {P0.x, P0.y, P0.z}, p1[3] = {P1.x, P1.y, P1.z}, p2[3] = {P2.x, P2.y, P2.z};
    h.attr = (uint32_t)__float_as_int(P0.w);
    h.Kd[0] = P1.w; h.Kd[1] = P2.w; h.Kd[2] = __ldg(mesh.kdb + found);
    float dp02[3], dp12[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        h.p[k] = bb[0] * p0[k] + bb[1] * p1[k] + bb[2] * p2[k];                                   // :311
        h.pError[k] = gamma_n(7) * (fabsf(bb[0] * p0[k]) + fabsf(bb[1] * p1[k]) + fabsf(bb[2] * p2[k]));  // :302-308
        dp02[k] = p0[k] - p2[k];
        dp12[k] = p1[k] - p2[k];
        h.ss[k] = (-1.f * dp02[k] - -1.f * dp12[k]) * 1.f;   // dpdu for the default uv set (:288-295)
    }
    cross3(dp02, dp12, h.n);
    normalize3(h.n);
    if (h.attr & BRE_TRI_FLIP_NORMAL) { h.n[0] = -h.n[0]; h.n[1] = -h.n[1]; h.n[2] = -h.n[2]; }
    normalize3(h.ss);
    return true;
}

__device__ __forceinline__ void offset_ray_origin(const float p[3], const float pError[3], const float n[3],
                                                  const float w[3], float out[3]) {
    float dd = fabsf(n[0]) * pError[0] + fabsf(n[1]) * pError[1] + fabsf(n[2]) * pError[2];
    bool neg = dot3(w, n) < 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float off = dd * n[k];
        if (neg) off = -off;
        float po = p[k] + off;
        if (off > 0) po = next_up(po);
        else if (off < 0) po = next_down(po);
        out[k] = po;
    }
}

// Interaction::GetMedium over the interface GeometricPrimitive::Intersect picked
__device__ __forceinline__ bool medium_after(uint32_t attr, const float w[3], const float n[3], bool current) {
    uint32_t side = attr & 3u;
    if (side == BRE_TRI_MEDIUM_INHERIT) return current;
    bool outside = dot3(w, n) > 0;
    return side == BRE_TRI_MEDIUM_OUTSIDE ? outside : !outside;
}

// single LambertianReflection lobe; false = the sample carries nothing
__device__ inline bool lambert_sample(const SurfHit &h, const float R[3], const float woW[3], float u0,
                                      float u1, float wiW[3], float f[3], float &pdf) {
    if (R[0] == 0.f && R[1] == 0.f && R[2] == 0.f) return false;
    float ts[3];
    cross3(h.n, h.ss, ts);
    float woz = dot3(woW, h.n);
    if (woz == 0.f) return false;
    u0 = fminf(u0 * 1 - 0, BRE_ONE_MINUS_EPS);
    float ux = 2.f * u0 - 1, uy = 2.f * u1 - 1, dx = 0.f, dy = 0.f;
    if (!(ux == 0.f && uy == 0.f)) {   // sampling.cpp:113-130
        float theta, r;
        if (fabsf(ux) > fabsf(uy)) { r = ux; theta = (BRE_PI / 4) * (uy / ux); }
        else { r = uy; theta = (BRE_PI / 2) - (BRE_PI / 4) * (ux / uy); }
        dx = r * cosf(theta); dy = r * sinf(theta);
    }
    float z = sqrtf(fmaxf(0.f, 1 - dx * dx - dy * dy));
    if (woz < 0) z *= -1;
    pdf = (woz * z > 0) ? fabsf(z) * BRE_INV_PI : 0.f;
    if (pdf == 0.f) return false;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        wiW[k] = h.ss[k] * dx + ts[k] * dy + h.n[k] * z;
        f[k] = R[k] * BRE_INV_PI;
    }
    return true;
}

}  // namespace bre
