// bre_trace.cu -- K1: volume photon tracing into SoA photon records.
//
// Replaces (reference src/): the photon pass of PhotonBeamIntegrator::Render
// (integrators/photonbeam.cpp:365-437; sampler convention :386-391: every draw is PCG32
// seeded with pathIndex+1, core/rng.h:60-145) using the volume-photon deposit rule of
// PhotonIntegrator::ShootPhotons (integrators/photonmap.cpp:696-724: Photon(p, alpha, wo) at every
// medium scattering event), HomogeneousMedium::Sample (media/homogeneous.cpp:50-77),
// GridDensityMedium::Sample/Density (media/grid.cpp:45-87), HenyeyGreenstein::Sample_p
// (core/medium.cpp:186-205), PointLight/SpotLight::Sample_Le (lights/point.cpp:61-71,
// lights/spot.cpp:83-93) and the light power distribution (core/integrator.cpp:217-225).
//
// One thread per photon path.  Deposits go to per-path slots (path*maxDepth + k); a scan of the
// per-path counts and a compaction kernel then append them to the photon store in path-index
// order, so the store is deterministic and identical on every rank / run.
#include "bre_internal.cuh"
#include "bre_surface.cuh"

namespace bre {

struct Pcg {
    uint64_t state, inc;
    __device__ __forceinline__ uint32_t u32() {  // rng.h:138-144
        uint64_t old = state;
        state = old * 0x5851f42d4c957f2dULL + inc;
        uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
        uint32_t rot = (uint32_t)(old >> 59u);
        return (xs >> rot) | (xs << ((~rot + 1u) & 31));
    }
    __device__ __forceinline__ void seed(uint64_t seq) {  // rng.h:130-136
        state = 0u;
        inc = (seq << 1u) | 1u;
        u32();
        state += 0x853c49e6748fea9bULL;
        u32();
    }
    __device__ __forceinline__ float f() {  // rng.h:78-85
        return fminf(BRE_ONE_MINUS_EPS, fmul((float)u32(), 0x1p-32f));
    }
};

__device__ __forceinline__ void coordinate_system(const float v1[3], float v2[3], float v3[3]) {
    // geometry.h:1020-1027 (Vector3 / scalar multiplies by the reciprocal; Cross in double)
    if (fabsf(v1[0]) > fabsf(v1[1])) {
        float inv = fdiv(1.f, sqrtf(fadd(fmul(v1[0], v1[0]), fmul(v1[2], v1[2]))));
        v2[0] = fmul(-v1[2], inv); v2[1] = 0.f; v2[2] = fmul(v1[0], inv);
    } else {
        float inv = fdiv(1.f, sqrtf(fadd(fmul(v1[1], v1[1]), fmul(v1[2], v1[2]))));
        v2[0] = 0.f; v2[1] = fmul(v1[2], inv); v2[2] = fmul(-v1[1], inv);
    }
    double ax = v1[0], ay = v1[1], az = v1[2], bx = v2[0], by = v2[1], bz = v2[2];
    v3[0] = (float)(__dsub_rn(__dmul_rn(ay, bz), __dmul_rn(az, by)));
    v3[1] = (float)(__dsub_rn(__dmul_rn(az, bx), __dmul_rn(ax, bz)));
    v3[2] = (float)(__dsub_rn(__dmul_rn(ax, by), __dmul_rn(ay, bx)));
}

__device__ __forceinline__ float phase_hg_t(float cosTheta, float g) {
    float denom = fadd(fadd(1.f, fmul(g, g)), fmul(fmul(2.f, g), cosTheta));
    return fdiv(fmul(BRE_INV4PI, fsub(1.f, fmul(g, g))), fmul(denom, sqrtf(denom)));
}

__device__ void hg_sample(const float wo[3], float g, float u0, float u1, float wi[3]) {
    // medium.cpp:186-205
    float cosTheta;
    if (fabsf(g) < 1e-3f) cosTheta = fsub(1.f, fmul(2.f, u0));
    else {
        float sqr = fdiv(fsub(1.f, fmul(g, g)), fadd(fsub(1.f, g), fmul(fmul(2.f, g), u0)));
        cosTheta = fdiv(fsub(fadd(1.f, fmul(g, g)), fmul(sqr, sqr)), fmul(2.f, g));
    }
    float sinTheta = sqrtf(fmaxf(0.f, fsub(1.f, fmul(cosTheta, cosTheta))));
    float phi = fmul(fmul(2.f, BRE_PI), u1);
    float v1[3], v2[3];
    coordinate_system(wo, v1, v2);
    float a = fmul(sinTheta, cosf(phi)), b = fmul(sinTheta, sinf(phi));
#pragma unroll
    for (int k = 0; k < 3; ++k)
        wi[k] = fadd(fadd(fmul(a, v1[k]), fmul(b, v2[k])), fmul(cosTheta, -wo[k]));
}

// Bounds3::IntersectP(ray,&t0,&t1), geometry.h:1386-1408
__device__ __forceinline__ bool ray_box(const float o[3], const float d[3], float tMaxRay,
                                        const float lo[3], const float hi[3], float &t0, float &t1) {
    t0 = 0.f; t1 = tMaxRay;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float inv = fdiv(1.f, d[i]);
        float tNear = fmul(fsub(lo[i], o[i]), inv);
        float tFar = fmul(fsub(hi[i], o[i]), inv);
        if (tNear > tFar) { float s = tNear; tNear = tFar; tFar = s; }
        tFar = fmul(tFar, BRE_SLAB_WIDEN);
        t0 = tNear > t0 ? tNear : t0;
        t1 = tFar < t1 ? tFar : t1;
        if (t0 > t1) return false;
    }
    return true;
}

__device__ __forceinline__ float gridD(const MediumDev &m, int x, int y, int z) {
    if (x < 0 || y < 0 || z < 0 || x >= m.nx || y >= m.ny || z >= m.nz) return 0.f;
    return __ldg(m.density + ((size_t)z * m.ny + y) * m.nx + x);
}
__device__ __forceinline__ float lerpf(float t, float a, float b) {
    return fadd(fmul(fsub(1.f, t), a), fmul(t, b));
}
__device__ float grid_density_t(const MediumDev &m, float px, float py, float pz) {
    float sx = fsub(fmul(px, (float)m.nx), .5f), sy = fsub(fmul(py, (float)m.ny), .5f),
          sz = fsub(fmul(pz, (float)m.nz), .5f);
    float fx = floorf(sx), fy = floorf(sy), fz = floorf(sz);
    int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    float dx = fsub(sx, fx), dy = fsub(sy, fy), dz = fsub(sz, fz);
    float v000, v100, v010, v110, v001, v101, v011, v111;
    if ((unsigned)ix < (unsigned)(m.nx - 1) && (unsigned)iy < (unsigned)(m.ny - 1) &&
        (unsigned)iz < (unsigned)(m.nz - 1)) {   // interior: no bounds compares
        grid_cell_corners(m, ix, iy, iz, v000, v100, v010, v110, v001, v101, v011, v111);
    } else {
        v000 = gridD(m, ix, iy, iz); v100 = gridD(m, ix + 1, iy, iz);
        v010 = gridD(m, ix, iy + 1, iz); v110 = gridD(m, ix + 1, iy + 1, iz);
        v001 = gridD(m, ix, iy, iz + 1); v101 = gridD(m, ix + 1, iy, iz + 1);
        v011 = gridD(m, ix, iy + 1, iz + 1); v111 = gridD(m, ix + 1, iy + 1, iz + 1);
    }
    float d00 = lerpf(dx, v000, v100);
    float d10 = lerpf(dx, v010, v110);
    float d01 = lerpf(dx, v001, v101);
    float d11 = lerpf(dx, v011, v111);
    return lerpf(dz, lerpf(dy, d00, d10), lerpf(dy, d01, d11));
}

// Tracking step of k_trace_flat (TS_FAST, default): the same trilinear interpolant with fused
// lerps a + t (b - a) and fused coordinate arithmetic.  K1 is compared with the reference
// statistically and path by path within a tolerance (its logf / sincosf already differ from the
// host's), so only the rounding of the density changes (~1e-7 relative); the hit DECISIONS of the
// query are not involved.  k_trace<MESH> keeps the reference-order form.
#ifndef TS_FAST
#define TS_FAST 1
#endif
__device__ __forceinline__ float lerp_fused(float t, float a, float b) { return __fmaf_rn(t, b - a, a); }
__device__ __forceinline__ float grid_density_fused(const MediumDev &m, float px, float py, float pz) {
    const float sx = __fmaf_rn(px, (float)m.nx, -.5f), sy = __fmaf_rn(py, (float)m.ny, -.5f),
                sz = __fmaf_rn(pz, (float)m.nz, -.5f);
    const float fx = floorf(sx), fy = floorf(sy), fz = floorf(sz);
    const int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    const float dx = sx - fx, dy = sy - fy, dz = sz - fz;
    float v000, v100, v010, v110, v001, v101, v011, v111;
    if ((unsigned)ix < (unsigned)(m.nx - 1) && (unsigned)iy < (unsigned)(m.ny - 1) &&
        (unsigned)iz < (unsigned)(m.nz - 1)) {
        grid_cell_corners(m, ix, iy, iz, v000, v100, v010, v110, v001, v101, v011, v111);
    } else {
        v000 = gridD(m, ix, iy, iz); v100 = gridD(m, ix + 1, iy, iz);
        v010 = gridD(m, ix, iy + 1, iz); v110 = gridD(m, ix + 1, iy + 1, iz);
        v001 = gridD(m, ix, iy, iz + 1); v101 = gridD(m, ix + 1, iy, iz + 1);
        v011 = gridD(m, ix, iy + 1, iz + 1); v111 = gridD(m, ix + 1, iy + 1, iz + 1);
    }
    const float d00 = lerp_fused(dx, v000, v100), d10 = lerp_fused(dx, v010, v110);
    const float d01 = lerp_fused(dx, v001, v101), d11 = lerp_fused(dx, v011, v111);
    return lerp_fused(dz, lerp_fused(dy, d00, d10), lerp_fused(dy, d01, d11));
}

// Medium::Sample for a unit-direction ray segment [0,tMax).  Returns true when a scattering
// event was sampled at t; w multiplies the path weight.
__device__ bool medium_sample(const MediumDev &m, Pcg &rng, const float o[3], const float d[3],
                              float tMax, float &t, float w[3]) {
    if (m.kind == 0) {  // homogeneous.cpp:50-77
        int channel = min((int)fmul(rng.f(), 3.f), 2);
        float dist = fdiv(-logf(fsub(1.f, rng.f())), m.sigma_t[channel]);
        float len = sqrtf(fadd(fadd(fmul(d[0], d[0]), fmul(d[1], d[1])), fmul(d[2], d[2])));
        t = fminf(fmul(dist, len), tMax);
        bool sampled = t < tMax;
        float Tr[3], pdf = 0.f;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            Tr[c] = expf(fmul(fmul(-m.sigma_t[c], t), len));
            pdf = fadd(pdf, sampled ? fmul(m.sigma_t[c], Tr[c]) : Tr[c]);
        }
        pdf = fmul(pdf, 1.f / 3.f);
        if (pdf == 0.f) pdf = 1.f;
#pragma unroll
        for (int c = 0; c < 3; ++c)
            w[c] = sampled ? fdiv(fmul(Tr[c], m.sigma_s[c]), pdf) : fdiv(Tr[c], pdf);
        return sampled;
    }
    // grid.cpp:62-87, delta tracking
    float om[3], dm[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const float *a = m.w2m + 4 * r;
        om[r] = fadd(fadd(fadd(fmul(a[0], o[0]), fmul(a[1], o[1])), fmul(a[2], o[2])), a[3]);
        dm[r] = fadd(fadd(fmul(a[0], d[0]), fmul(a[1], d[1])), fmul(a[2], d[2]));
    }
    const float lo[3] = {0.f, 0.f, 0.f}, hi[3] = {1.f, 1.f, 1.f};
    float tMin, tEnd;
    w[0] = w[1] = w[2] = 1.f;
    if (!ray_box(om, dm, tMax, lo, hi, tMin, tEnd)) return false;
    float tt = tMin;
    const float st = m.sigma_t[0];
    for (;;) {
        tt = fsub(tt, fdiv(fmul(logf(fsub(1.f, rng.f())), m.invMaxDensity), st));
        if (tt >= tEnd) break;
        float dens = grid_density_t(m, fadd(om[0], fmul(dm[0], tt)), fadd(om[1], fmul(dm[1], tt)),
                                    fadd(om[2], fmul(dm[2], tt)));
        if (fmul(dens, m.invMaxDensity) > rng.f()) {
#pragma unroll
            for (int c = 0; c < 3; ++c) w[c] = fdiv(m.sigma_s[c], st);
            t = tt;
            return true;
        }
    }
    return false;
}

__device__ __forceinline__ float spot_falloff(const LightDev &L, const float w[3]) {  // spot.cpp:64-73
    float len = sqrtf(fadd(fadd(fmul(w[0], w[0]), fmul(w[1], w[1])), fmul(w[2], w[2])));
    float c = fdiv(fadd(fadd(fmul(w[0], L.axis[0]), fmul(w[1], L.axis[1])), fmul(w[2], L.axis[2])), len);
    if (c < L.cosTotalWidth) return 0.f;
    if (c > L.cosFalloffStart) return 1.f;
    float delta = fdiv(fsub(c, L.cosTotalWidth), fsub(L.cosFalloffStart, L.cosTotalWidth));
    return fmul(fmul(delta, delta), fmul(delta, delta));
}

// Light choice + Sample_Le + initial weight (photonbeam.cpp:394-414).  Returns false when the
// path carries nothing.  Draw order: light pick, uLight0 (2), uLight1 (2, unused), time (1).
__device__ bool emit_photon(const LightsDev *__restrict__ lights, Pcg &rng, float o[3], float d[3],
                            float alpha[3]) {
    const int nL = lights->n;
    float uPick = rng.f();
    int li = 0;
    for (int i = 0; i < nL; ++i) if (lights->cdf[i] <= uPick) li = i;  // sampling.h:90-100
    float funcInt = lights->funcInt;
    float lightPdf = funcInt > 0.f ? fdiv(lights->func[li], fmul(funcInt, (float)nL)) : fdiv(1.f, (float)nL);
    float u0 = rng.f(), u1 = rng.f();
    rng.f(); rng.f();  // uLight1
    rng.f();           // uLightTime
    const LightDev &L = lights->l[li];
    float Le[3], pdfDir;
    o[0] = L.pos[0]; o[1] = L.pos[1]; o[2] = L.pos[2];
    if (L.kind == 0) {  // point.cpp:61-71, sampling.cpp:98-103
        float z = fsub(1.f, fmul(2.f, u0));
        float r = sqrtf(fmaxf(0.f, fsub(1.f, fmul(z, z))));
        float phi = fmul(fmul(2.f, BRE_PI), u1);
        d[0] = fmul(r, cosf(phi)); d[1] = fmul(r, sinf(phi)); d[2] = z;
        pdfDir = BRE_INV4PI;
        Le[0] = L.I[0]; Le[1] = L.I[1]; Le[2] = L.I[2];
    } else {  // spot.cpp:83-93, sampling.cpp:136-142
        float cosTheta = fadd(fsub(1.f, u0), fmul(u0, L.cosTotalWidth));
        float sinTheta = sqrtf(fsub(1.f, fmul(cosTheta, cosTheta)));
        float phi = fmul(fmul(u1, 2.f), BRE_PI);
        float w0 = fmul(cosf(phi), sinTheta), w1 = fmul(sinf(phi), sinTheta);
#pragma unroll
        for (int k = 0; k < 3; ++k)
            d[k] = fadd(fadd(fmul(w0, L.ax[k]), fmul(w1, L.ay[k])), fmul(cosTheta, L.axis[k]));
        pdfDir = fdiv(1.f, fmul(fmul(2.f, BRE_PI), fsub(1.f, L.cosTotalWidth)));
        float f = spot_falloff(L, d);
        Le[0] = fmul(L.I[0], f); Le[1] = fmul(L.I[1], f); Le[2] = fmul(L.I[2], f);
    }
    bool ok = !(pdfDir == 0.f || (Le[0] == 0.f && Le[1] == 0.f && Le[2] == 0.f));
    alpha[0] = alpha[1] = alpha[2] = 0.f;
    if (ok) {
        float nd = fabsf(fadd(fadd(fmul(d[0], d[0]), fmul(d[1], d[1])), fmul(d[2], d[2])));
        float denom = fmul(fmul(lightPdf, 1.f), pdfDir);  // photonbeam.cpp:413-414
#pragma unroll
        for (int c = 0; c < 3; ++c) alpha[c] = fdiv(fmul(nd, Le[c]), denom);
        ok = !(alpha[0] == 0.f && alpha[1] == 0.f && alpha[2] == 0.f);
    }
    return ok;
}

template <bool MESH>
__global__ void __launch_bounds__(128) k_trace(MediumDev m, const LightsDev *__restrict__ lights,
                                               MeshDev mesh, int lightsInMedium,
                                               float blo0, float blo1, float blo2, float bhi0,
                                               float bhi1, float bhi2, uint64_t first,
                                               uint32_t nPaths, int maxDepth, float radius,
                                               float4 *__restrict__ slot_pos,
                                               float4 *__restrict__ slot_pow,
                                               float4 *__restrict__ slot_dir,
                                               uint32_t *__restrict__ counts) {
    uint32_t pid = blockIdx.x * blockDim.x + threadIdx.x;
    if (pid >= nPaths) return;
    const float blo[3] = {blo0, blo1, blo2}, bhi[3] = {bhi0, bhi1, bhi2};
    uint32_t stored = 0;
    Pcg rng;
    rng.seed(first + pid + 1);  // photonbeam.cpp:386-391
    float o[3], d[3], alpha[3];
    bool ok = emit_photon(lights, rng, o, d, alpha);
    const size_t base = (size_t)pid * maxDepth;
    if (MESH) {
        // photonbeam.cpp:258-325 as a loop: Scene::Intersect, Medium::Sample, deposit + phase
        // scatter, or surface bounce (:296-323).  `guard` bounds the interface pass-throughs.
        bool inMedium = lightsInMedium != 0;
        int guard = 64 + 8 * maxDepth;
        for (int depth = 0; ok && depth < maxDepth && guard > 0; ++depth, --guard) {
            SurfHit h;
            if (!mesh_intersect(mesh, o, d, __int_as_float(0x7f800000), h, true)) break;
            bool scattered = false;
            float ts = 0.f, w[3];
            if (inMedium) {
                scattered = medium_sample(m, rng, o, d, h.t, ts, w);
#pragma unroll
                for (int c = 0; c < 3; ++c) alpha[c] = fmul(alpha[c], w[c]);
            }
            if (alpha[0] == 0.f && alpha[1] == 0.f && alpha[2] == 0.f) break;
            if (scattered) {
                float p[3] = {fadd(o[0], fmul(d[0], ts)), fadd(o[1], fmul(d[1], ts)), fadd(o[2], fmul(d[2], ts))};
                float wo[3] = {-d[0], -d[1], -d[2]}, wi[3];
                slot_pos[base + stored] = make_float4(p[0], p[1], p[2], radius);
                slot_pow[base + stored] = make_float4(alpha[0], alpha[1], alpha[2], 0.f);
                slot_dir[base + stored] = make_float4(wo[0], wo[1], wo[2], 0.f);
                ++stored;
                float ua = rng.f(), ub = rng.f();
                hg_sample(wo, m.g, ua, ub, wi);
#pragma unroll
                for (int k = 0; k < 3; ++k) { o[k] = p[k]; d[k] = wi[k]; }
                if (depth + 1 >= maxDepth) break;
                rng.f();   // roulette draw with continue probability 1
                continue;
            }
            if ((h.attr & BRE_TRI_NO_MATERIAL) || h.Kd[0] < 0.f) {   // photonbeam.cpp:300-304
                --depth;
                float no[3];
                offset_ray_origin(h.p, h.pError, h.n, d, no);
                inMedium = medium_after(h.attr, d, h.n, inMedium);
                o[0] = no[0]; o[1] = no[1]; o[2] = no[2];
                continue;
            }
            float R[3], woW[3] = {-d[0], -d[1], -d[2]}, wiW[3], f[3], pdf;
#pragma unroll
            for (int c = 0; c < 3; ++c) R[c] = h.Kd[c] < 0.f ? 0.f : h.Kd[c];
            float ub0 = rng.f(), ub1 = rng.f();
            if (!lambert_sample(h, R, woW, ub0, ub1, wiW, f, pdf)) break;
            float betaNew[3], absdot = fabsf(dot3(wiW, h.n));
#pragma unroll
            for (int c = 0; c < 3; ++c) betaNew[c] = alpha[c] * f[c] * absdot / pdf;
            float no[3];
            offset_ray_origin(h.p, h.pError, h.n, wiW, no);
            inMedium = medium_after(h.attr, wiW, h.n, inMedium);
#pragma unroll
            for (int k = 0; k < 3; ++k) { o[k] = no[k]; d[k] = wiW[k]; }
            float yNew = 0.212671f * betaNew[0] + 0.715160f * betaNew[1] + 0.072169f * betaNew[2];
            float yOld = 0.212671f * alpha[0] + 0.715160f * alpha[1] + 0.072169f * alpha[2];
            float q = fmaxf(0.f, 1 - yNew / yOld);   // photonbeam.cpp:321-323
            if (rng.f() < q) break;
#pragma unroll
            for (int c = 0; c < 3; ++c) alpha[c] = betaNew[c] / (1 - q);
        }
        counts[pid] = stored;
        return;
    }
    for (int nInt = 1; ok; ++nInt) {
        float t0, t1;
        if (!ray_box(o, d, __int_as_float(0x7f800000), blo, bhi, t0, t1) || t0 > 0.f) break;
        float ts = 0.f, w[3];
        bool scattered = medium_sample(m, rng, o, d, t1, ts, w);
#pragma unroll
        for (int c = 0; c < 3; ++c) alpha[c] = fmul(alpha[c], w[c]);
        if (alpha[0] == 0.f && alpha[1] == 0.f && alpha[2] == 0.f) break;
        if (!scattered) break;  // reached the wall: absorbed
        float p[3] = {fadd(o[0], fmul(d[0], ts)), fadd(o[1], fmul(d[1], ts)), fadd(o[2], fmul(d[2], ts))};
        float wo[3] = {-d[0], -d[1], -d[2]}, wi[3];
        // Photon(misect.p, alpha, wo)  photonmap.cpp:720-723
        slot_pos[base + stored] = make_float4(p[0], p[1], p[2], radius);
        slot_pow[base + stored] = make_float4(alpha[0], alpha[1], alpha[2], 0.f);
        slot_dir[base + stored] = make_float4(wo[0], wo[1], wo[2], 0.f);
        ++stored;
        float ua = rng.f(), ub = rng.f();
        hg_sample(wo, m.g, ua, ub, wi);
#pragma unroll
        for (int k = 0; k < 3; ++k) { o[k] = p[k]; d[k] = wi[k]; }
        if (nInt >= maxDepth) break;      // photonmap.cpp:794
        if (rng.f() > 1.f) break;         // Russian roulette with continueProb == 1, :799-804
    }
    counts[pid] = stored;
}

// ---------------------------------------------------------------- K1, flattened (box-bounded media)
// Same paths, same draws, same records as k_trace<false>, restructured for SIMT efficiency
// (ncu on the nested-loop version: 8.7 of 32 lanes active on average).  Every lane is a small
// state machine -- NEW path -> start of a SEGment -> delta-TRACKing steps -- and the warp runs ONE
// loop in which each iteration advances every lane by one tracking step; a lane whose path ends
// claims the next path index from a global counter (persistent threads), so lanes neither wait for
// the longest rejection loop of a segment nor for the longest path of the warp.  Deposits still go
// to the slots of their path index, so the photon store is unchanged.
// The rare, long bodies (NEW: seed + emission, SEG: two ray-box clips, SCAT: deposit + phase sample)
// are BATCHED: a lane that needs one waits until TS_BATCH lanes of the warp need the same body (or
// fewer than TS_STARVE lanes are still tracking), so that each of them runs with many lanes instead
// of 1-5 per iteration; the tracking step runs every iteration with whoever is tracking.  Paths are
// independent (own PCG stream, own slots), so the schedule does not change any record.
enum { TS_NEW = 0, TS_SEG = 1, TS_TRACK = 2, TS_DONE = 3, TS_SCAT = 4 };
// measured (trace phase, ms): no batching 1.71-1.83 | BATCH 4 / 8 / 12 / 16 / 24 -> 1.71 / 1.58 / 1.51 / 1.61 / 1.81
// | STARVE 4 / 8 / 16 at BATCH 12 -> 1.50 / 1.50 / 1.58
#ifndef TS_BATCH
#define TS_BATCH 12
#endif
#ifndef TS_STARVE
#define TS_STARVE 8
#endif

__global__ void __launch_bounds__(128) k_trace_flat(MediumDev m, const LightsDev *__restrict__ lights,
                                                    float blo0, float blo1, float blo2, float bhi0,
                                                    float bhi1, float bhi2, uint64_t first,
                                                    uint32_t nPaths, int maxDepth, float radius,
                                                    float4 *__restrict__ slot_pos,
                                                    float4 *__restrict__ slot_pow,
                                                    float4 *__restrict__ slot_dir,
                                                    uint32_t *__restrict__ counts,
                                                    unsigned int *__restrict__ nextPath) {
    const float blo[3] = {blo0, blo1, blo2}, bhi[3] = {bhi0, bhi1, bhi2};
    const float INF = __int_as_float(0x7f800000);
    const unsigned FULL = 0xffffffffu;
    Pcg rng;
    rng.state = 0; rng.inc = 1;
    float o[3] = {0.f, 0.f, 0.f}, d[3] = {0.f, 0.f, 1.f}, alpha[3] = {0.f, 0.f, 0.f};
    float om[3] = {0.f, 0.f, 0.f}, dm[3] = {0.f, 0.f, 0.f};
    float tt = 0.f, tEnd = 0.f, ts = 0.f;
    uint32_t pid = 0, stored = 0;
    int nInt = 1, state = TS_NEW;
    const float stepScale = fdiv(m.invMaxDensity, m.sigma_t[0]);   // free-flight scale of a tracking step (TS_FAST)
    for (;;) {
        // one convergence point per iteration: who is where, which bodies run this time
        const unsigned mNew = __ballot_sync(FULL, state == TS_NEW), mSeg = __ballot_sync(FULL, state == TS_SEG);
        const unsigned mScat = __ballot_sync(FULL, state == TS_SCAT), mTrack = __ballot_sync(FULL, state == TS_TRACK);
        if ((mNew | mSeg | mScat | mTrack) == 0u) break;
        const bool starved = __popc(mTrack) < TS_STARVE;   // too few trackers: run whatever is waiting
        const bool runScat = mScat && (starved || __popc(mScat) >= TS_BATCH);
        const bool runNew = mNew && (starved || __popc(mNew) >= TS_BATCH);
        const bool runSeg = starved || __popc(mSeg) + (runScat ? __popc(mScat) : 0) + (runNew ? __popc(mNew) : 0) >= TS_BATCH;
        if (runScat && state == TS_SCAT) {
            const size_t slot = (size_t)pid * maxDepth + stored;
            float p[3] = {fadd(o[0], fmul(d[0], ts)), fadd(o[1], fmul(d[1], ts)), fadd(o[2], fmul(d[2], ts))};
            float wo[3] = {-d[0], -d[1], -d[2]}, wi[3];
            slot_pos[slot] = make_float4(p[0], p[1], p[2], radius);   // Photon(misect.p, alpha, wo)
            slot_pow[slot] = make_float4(alpha[0], alpha[1], alpha[2], 0.f);
            slot_dir[slot] = make_float4(wo[0], wo[1], wo[2], 0.f);
            ++stored;
            float ua = rng.f(), ub = rng.f();
            hg_sample(wo, m.g, ua, ub, wi);
#pragma unroll
            for (int k = 0; k < 3; ++k) { o[k] = p[k]; d[k] = wi[k]; }
            if (nInt >= maxDepth || rng.f() > 1.f) {           // photonmap.cpp:794; roulette with continueProb == 1, :799-804
                counts[pid] = stored; state = TS_NEW;          // (claims its next path in a later iteration)
            } else { ++nInt; state = TS_SEG; }
        }
        if (runNew && (mNew >> (threadIdx.x & 31u) & 1u)) {
            pid = atomicAdd(nextPath, 1u);   // (claiming one path ahead to hide the round trip: 1.454 -> 1.475 ms, not kept)
            if (pid >= nPaths) state = TS_DONE;
            else {
                rng.seed(first + pid + 1);  // photonbeam.cpp:386-391
                stored = 0; nInt = 1;
                if (emit_photon(lights, rng, o, d, alpha)) state = TS_SEG;
                else counts[pid] = 0;       // stays TS_NEW: next batch
            }
        }
        if (runSeg && state == TS_SEG) {
            bool endPath = false;
            // scene.Intersect stand-in: distance to the absorbing box wall
            float t0, t1;
            if (!ray_box(o, d, INF, blo, bhi, t0, t1) || t0 > 0.f) endPath = true;
            else if (m.kind == 0) {
                float w[3];
                const bool sc = medium_sample(m, rng, o, d, t1, ts, w);
#pragma unroll
                for (int c = 0; c < 3; ++c) alpha[c] = fmul(alpha[c], w[c]);
                if (alpha[0] == 0.f && alpha[1] == 0.f && alpha[2] == 0.f) endPath = true;
                else if (!sc) endPath = true;   // reached the wall: absorbed
                else state = TS_SCAT;
            } else {   // grid.cpp:62-75: ray into medium space, clip to the unit cube
#pragma unroll
                for (int r = 0; r < 3; ++r) {
                    const float *a = m.w2m + 4 * r;
                    om[r] = fadd(fadd(fadd(fmul(a[0], o[0]), fmul(a[1], o[1])), fmul(a[2], o[2])), a[3]);
                    dm[r] = fadd(fadd(fmul(a[0], d[0]), fmul(a[1], d[1])), fmul(a[2], d[2]));
                }
                const float lo[3] = {0.f, 0.f, 0.f}, hi[3] = {1.f, 1.f, 1.f};
                float tMin;
                if (!ray_box(om, dm, t1, lo, hi, tMin, tEnd)) endPath = true;
                else { tt = tMin; state = TS_TRACK; }
            }
            if (endPath) { counts[pid] = stored; state = TS_NEW; }
        }
        if (state == TS_TRACK) {   // one delta-tracking step, grid.cpp:76-86
            const float st = m.sigma_t[0];
#if TS_FAST
            tt = __fmaf_rn(-__logf(fsub(1.f, rng.f())), stepScale, tt);
#else
            tt = fsub(tt, fdiv(fmul(logf(fsub(1.f, rng.f())), m.invMaxDensity), st));
#endif
            if (tt >= tEnd) { counts[pid] = stored; state = TS_NEW; }
            else {
#if TS_FAST
                float dens = grid_density_fused(m, __fmaf_rn(dm[0], tt, om[0]), __fmaf_rn(dm[1], tt, om[1]),
                                                __fmaf_rn(dm[2], tt, om[2]));
#else
                float dens = grid_density_t(m, fadd(om[0], fmul(dm[0], tt)), fadd(om[1], fmul(dm[1], tt)),
                                            fadd(om[2], fmul(dm[2], tt)));
#endif
                if (fmul(dens, m.invMaxDensity) > rng.f()) {
#pragma unroll
                    for (int c = 0; c < 3; ++c) alpha[c] = fmul(alpha[c], fdiv(m.sigma_s[c], st));
                    if (alpha[0] == 0.f && alpha[1] == 0.f && alpha[2] == 0.f) { counts[pid] = stored; state = TS_NEW; }
                    else { ts = tt; state = TS_SCAT; }
                }
            }
        }
    }
}

// ---- exclusive scan of u32 (tiles of 2048: reduce, scan tile sums, apply) ----
constexpr int SC_T = 256, SC_I = 8, SC_TILE = SC_T * SC_I;

__global__ void __launch_bounds__(SC_T) k_scan_reduce(const uint32_t *__restrict__ in, uint32_t n,
                                                      uint32_t *__restrict__ tileSums) {
    __shared__ uint32_t ws[SC_T / 32];
    uint32_t base = blockIdx.x * SC_TILE, sum = 0;
#pragma unroll
    for (int k = 0; k < SC_I; ++k) {
        uint32_t i = base + k * SC_T + threadIdx.x;
        if (i < n) sum += __ldg(in + i);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t s = 0;
        for (int i = 0; i < SC_T / 32; ++i) s += ws[i];
        tileSums[blockIdx.x] = s;
    }
}
// single block: exclusive scan of tile sums (64-bit running totals into out64)
__global__ void __launch_bounds__(1024) k_scan_tiles(const uint32_t *__restrict__ tileSums,
                                                     uint32_t nTiles, uint64_t *__restrict__ tileOff,
                                                     uint64_t *__restrict__ total) {
    __shared__ uint64_t sh[1024];
    uint32_t per = (nTiles + 1023) / 1024;
    uint32_t b = threadIdx.x * per, e = min(b + per, nTiles);
    uint64_t sum = 0;
    for (uint32_t i = b; i < e; ++i) sum += tileSums[i];
    sh[threadIdx.x] = sum;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        uint64_t v = threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
        __syncthreads();
        sh[threadIdx.x] += v;
        __syncthreads();
    }
    uint64_t run = sh[threadIdx.x] - sum;
    for (uint32_t i = b; i < e; ++i) { tileOff[i] = run; run += tileSums[i]; }
    if (threadIdx.x == 1023) *total = sh[1023];
}
__global__ void __launch_bounds__(SC_T) k_scan_apply(const uint32_t *__restrict__ in, uint32_t n,
                                                     const uint64_t *__restrict__ tileOff,
                                                     uint64_t *__restrict__ out) {
    // thread handles SC_I consecutive elements
    __shared__ uint32_t ws[SC_T / 32];
    uint32_t base = blockIdx.x * SC_TILE + threadIdx.x * SC_I;
    uint32_t v[SC_I], sum = 0;
#pragma unroll
    for (int k = 0; k < SC_I; ++k) { v[k] = (base + k < n) ? __ldg(in + base + k) : 0; sum += v[k]; }
    uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5, incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) ws[w] = incl;
    __syncthreads();
    uint32_t woff = 0;
    for (uint32_t i = 0; i < w; ++i) woff += ws[i];
    uint64_t run = tileOff[blockIdx.x] + woff + incl - sum;
#pragma unroll
    for (int k = 0; k < SC_I; ++k) {
        if (base + k < n) out[base + k] = run;
        run += v[k];
    }
}

__global__ void __launch_bounds__(256) k_compact(uint32_t nPaths, int maxDepth,
                                                 const uint32_t *__restrict__ counts,
                                                 const uint64_t *__restrict__ offs,
                                                 const float4 *__restrict__ slot_pos,
                                                 const float4 *__restrict__ slot_pow,
                                                 const float4 *__restrict__ slot_dir,
                                                 float4 *__restrict__ pos_r, float4 *__restrict__ pw,
                                                 float4 *__restrict__ dr, uint64_t storeBase) {
    uint32_t pid = blockIdx.x * blockDim.x + threadIdx.x;
    if (pid >= nPaths) return;
    uint32_t c = __ldg(counts + pid);
    uint64_t dst = storeBase + __ldg(offs + pid);
    size_t src = (size_t)pid * maxDepth;
    for (uint32_t k = 0; k < c; ++k) {
        pos_r[dst + k] = __ldg(slot_pos + src + k);
        pw[dst + k] = __ldg(slot_pow + src + k);
        dr[dst + k] = __ldg(slot_dir + src + k);
    }
}

static int grow_store(bre_ctx *ctx, uint64_t newCount) {
    // grow the three store arrays, preserving contents
    size_t need = (size_t)newCount * sizeof(float4);
    Buf *bufs[3] = {&ctx->pos_r, &ctx->pow, &ctx->dir};
    for (Buf *b : bufs) {
        if (b->cap >= need) continue;
        size_t cap = need + need / 4;
        void *np = nullptr;
        cudaError_t e = cudaMalloc(&np, cap);
        if (e != cudaSuccess) return set_err(ctx, BRE_ERR_NOMEM, std::string("cudaMalloc photon store: ") + cudaGetErrorString(e));
        if (b->p && ctx->n_photons)
            cudaMemcpyAsync(np, b->p, (size_t)ctx->n_photons * sizeof(float4), cudaMemcpyDeviceToDevice, ctx->stream);
        cudaStreamSynchronize(ctx->stream);
        // a store exported over CUDA IPC may still be mapped by the peers: free it only after the
        // next collective handle exchange has made everyone close it (bre_nccl.cpp)
        bool exported = false;
        for (int a = 0; a < 3; ++a) exported |= ctx->p2p_open && b->p && b->p == ctx->p2p_exported[a];
        if (exported) ctx->p2p_zombies.push_back(b->p);
        else if (b->p) cudaFree(b->p);
        b->p = np; b->cap = cap;
    }
    return BRE_OK;
}
int grow_photon_store(bre_ctx *ctx, uint64_t newCount) { return grow_store(ctx, newCount); }

// Trace paths [first, first + nPaths) into per-path slots and scan the per-path counts.  Returns a
// view of the scratch arrays (valid until the next call) and the number of photons deposited.
// nPaths <= TRACE_CHUNK bounds the slot scratch (nPaths * maxDepth * 48 bytes).
int trace_to_slots(bre_ctx *ctx, uint64_t first, uint32_t nPaths, int maxDepth, float radius, SlotView *out) {
    if (!ctx->have_medium) return set_err(ctx, BRE_ERR_INVALID, "no medium set");
    if (!ctx->lights_d || ctx->n_lights == 0) return set_err(ctx, BRE_ERR_INVALID, "no lights set");
    const bool useMesh = ctx->n_tri > 0;
    if (!useMesh && !ctx->have_bounds)
        return set_err(ctx, BRE_ERR_INVALID, "no surfaces set (bre_set_triangles or bre_set_bounds)");
    if (maxDepth < 1) return set_err(ctx, BRE_ERR_INVALID, "max_depth must be >= 1");
    MeshDev mesh{ctx->n_tri, (const float4 *)ctx->tri.p, (const float *)ctx->tri_kdb.p};
    size_t slots = (size_t)nPaths * maxDepth;
    uint32_t nTiles = (nPaths + SC_TILE - 1) / SC_TILE;
    int s;
    // scratch layout: 3 slot arrays | counts | offsets(u64) | tileSums | tileOff(u64) | total(u64) | nextPath
    size_t oSlots = 0, oCounts = oSlots + 3 * slots * sizeof(float4);
    size_t oOffs = (oCounts + (size_t)nPaths * 4 + 15) & ~(size_t)15;
    size_t oTs = oOffs + (size_t)nPaths * 8;
    size_t oTo = (oTs + (size_t)nTiles * 4 + 15) & ~(size_t)15;
    size_t oTot = oTo + (size_t)nTiles * 8;
    if ((s = ensure(ctx, ctx->path_counts, oTot + 32)) != BRE_OK) return s;
    char *base = (char *)ctx->path_counts.p;
    float4 *sp = (float4 *)(base + oSlots), *sw = sp + slots, *sd = sw + slots;
    uint32_t *counts = (uint32_t *)(base + oCounts);
    uint64_t *offs = (uint64_t *)(base + oOffs);
    uint32_t *tileSums = (uint32_t *)(base + oTs);
    uint64_t *tileOff = (uint64_t *)(base + oTo);
    uint64_t *total = (uint64_t *)(base + oTot);
    unsigned int *nextPath = (unsigned int *)(base + oTot + 16);
    if (useMesh)
        k_trace<true><<<(nPaths + 127) / 128, 128, 0, ctx->stream>>>(
            ctx->medium, ctx->lights_d, mesh, ctx->lights_in_medium, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f,
            first, nPaths, maxDepth, radius, sp, sw, sd, counts);
    else if (ctx->trace_nested)
        k_trace<false><<<(nPaths + 127) / 128, 128, 0, ctx->stream>>>(
            ctx->medium, ctx->lights_d, mesh, 1, ctx->blo[0], ctx->blo[1], ctx->blo[2], ctx->bhi[0],
            ctx->bhi[1], ctx->bhi[2], first, nPaths, maxDepth, radius, sp, sw, sd, counts);
    else {
        // persistent threads: as many blocks as fit, each lane claims paths from nextPath
        int perSM = 0;
        BRE_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, k_trace_flat, 128, 0));
        if (perSM < 1) perSM = 1;
        unsigned blocks = (unsigned)(ctx->sm_count * perSM);
        if (blocks > (nPaths + 127) / 128) blocks = (nPaths + 127) / 128;
        BRE_CUDA(cudaMemsetAsync(nextPath, 0, 4, ctx->stream));
        k_trace_flat<<<blocks, 128, 0, ctx->stream>>>(
            ctx->medium, ctx->lights_d, ctx->blo[0], ctx->blo[1], ctx->blo[2], ctx->bhi[0], ctx->bhi[1],
            ctx->bhi[2], first, nPaths, maxDepth, radius, sp, sw, sd, counts, nextPath);
    }
    k_scan_reduce<<<nTiles, SC_T, 0, ctx->stream>>>(counts, nPaths, tileSums);
    k_scan_tiles<<<1, 1024, 0, ctx->stream>>>(tileSums, nTiles, tileOff, total);
    k_scan_apply<<<nTiles, SC_T, 0, ctx->stream>>>(counts, nPaths, tileOff, offs);
    BRE_CUDA(cudaGetLastError());
    uint64_t hTotal = 0;
    BRE_CUDA(cudaMemcpyAsync(&hTotal, total, 8, cudaMemcpyDeviceToHost, ctx->stream));
    BRE_CUDA(cudaStreamSynchronize(ctx->stream));
    out->sp = sp; out->sw = sw; out->sd = sd; out->counts = counts; out->offs = offs;
    out->total = hTotal; out->nPaths = nPaths; out->maxDepth = maxDepth;
    return BRE_OK;
}

int trace_photons(bre_ctx *ctx, uint64_t first, uint64_t nPathsTotal, int maxDepth, float radius) {
    ctx->built = false; ctx->topo_valid = false;
    PhaseTimer pt(ctx, PH_TRACE);
    for (uint64_t done = 0; done < nPathsTotal; done += TRACE_CHUNK) {
        uint32_t nPaths = (uint32_t)((nPathsTotal - done) < TRACE_CHUNK ? (nPathsTotal - done) : TRACE_CHUNK);
        SlotView v;
        int s = trace_to_slots(ctx, first + done, nPaths, maxDepth, radius, &v);
        if (s != BRE_OK) return s;
        if ((s = grow_store(ctx, ctx->n_photons + v.total)) != BRE_OK) return s;
        // tile-in-shared-memory compaction (the kernel of the multi-GPU push, writing to this rank only): coalesced runs
        if (ctx->compact_tiled && compact_tiled_local(ctx, v, ctx->n_photons) == BRE_OK) {
        } else {
            k_compact<<<(nPaths + 255) / 256, 256, 0, ctx->stream>>>(
                nPaths, maxDepth, v.counts, v.offs, v.sp, v.sw, v.sd, (float4 *)ctx->pos_r.p, (float4 *)ctx->pow.p,
                (float4 *)ctx->dir.p, ctx->n_photons);
            BRE_CUDA(cudaGetLastError());
        }
        ctx->n_photons += v.total;
    }
    return BRE_OK;
}

}  // namespace bre
