// oracle/ref_harness.cpp -- TEST INFRASTRUCTURE ONLY.
//
// Links against the UNMODIFIED reference sources compiled by oracle/Makefile (`make ref`) and
// produces the golden data that pins the CPU oracle (oracle/bre_oracle.c) to the reference:
//
//   ref_harness primitives <out.json>
//        known-answer vectors computed BY REFERENCE CODE for every primitive the hot path is
//        made of: PCG32 (core/rng.h), sphere/cone sampling (core/sampling.cpp), PhaseHG and
//        HenyeyGreenstein::p/Sample_p (core/medium.{h,cpp}), HomogeneousMedium::Tr,
//        GridDensityMedium::Density and ::Tr (media/*.cpp), Bounds3::IntersectP (core/geometry.h),
//        EncodeMorton3 / RadixSort (core/photonbeambvh.cpp), kernel() (integrators/photonmap.cpp),
//        Film::SetImage + Film::WriteImage (core/film.cpp).
//   ref_harness fogbox <out.bin> <nVolumePhotons> <res> <radius> [keepFirst]
//        builds the fog-box scene of SURVEY.md Appendix A with reference classes, runs the
//        reference's own PhotonIntegrator::ShootPhotons, generates camera rays with the reference's
//        PerspectiveCamera + Scene::Intersect, and evaluates the beam radiance estimate over the
//        dumped photons by BRUTE FORCE with reference vector/phase/medium/kernel code.
//        out.bin = photons, rays, per-ray hit lists and radiance (format: write_fogbox below).
//        A trailing "digest" stores per-ray hit counts + set digests instead of the lists (configs[0]-size fixture).
//   ref_harness gridspot <out.bin> <nPaths> <res> <radius>
//        query fixture for a GRID medium lit by a SPOT light: photon paths walked by reference classes through a
//        GridDensityMedium (as mode trace), camera rays of the reference's PerspectiveCamera, brute-force estimate
//        with transmittance from a fine integral of the reference's own Density() (its Tr is a stochastic
//        ratio-tracking estimate, media/grid.cpp:89-121; the integral is what it estimates).
//   ref_harness trace <out.bin> <homogeneous|grid> <firstPath> <nPaths> <maxDepth>
//        the volume-photon deposit loop (integrators/photonmap.cpp:696-724) driven with the
//        sampler stream of integrators/photonbeam.cpp:386-391 inside an absorbing triangle box,
//        every step executed by reference classes (Light::Sample_Le, Scene::Intersect,
//        Medium::Sample, HenyeyGreenstein::Sample_p, MediumInteraction::SpawnRay).
//
//   ref_harness surfaces <out.bin> <firstPath> <nPaths> <maxDepth> <resX> <resY>
//        photon paths through a fog box with matte walls (Lambertian bounce + Russian roulette,
//        integrators/photonbeam.cpp:296-323), a reverse-orientation wall and an interface-only
//        medium boundary, plus camera rays of a cropped PerspectiveCamera with first-hit tMax.
//
// All floats are written as raw IEEE-754 bit patterns so the comparison can be bit-exact.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <string>
#include <vector>

#include "pbrt.h"
#include "accelerators/bvh.h"
#include "samplers/halton.h"
#include "cameras/perspective.h"
#include "film.h"
#include "filters/box.h"
#include "geometry.h"
#include "integrator.h"
#include "lights/point.h"
#include "lights/spot.h"
#include "materials/matte.h"
#include "media/grid.h"
#include "media/homogeneous.h"
#include "medium.h"
#include "paramset.h"
#include "memory.h"
#include "parallel.h"
#include "rng.h"
#include "sampler.h"
#include "sampling.h"
#include "scene.h"
#include "shapes/triangle.h"
#include "spectrum.h"
#include "textures/constant.h"
#include "transform.h"

#include "ref_access.h"
#include "photonbeambvh.h"
#include "parallel.h"
#include <atomic>
#include <chrono>
#include <thread>

using namespace pbrt;

namespace {

uint32_t fbits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

// ---------------------------------------------------------------- tiny JSON writer
struct Json {
    FILE *fp;
    bool firstSection = true;
    explicit Json(const char *path) : fp(fopen(path, "w")) {
        if (!fp) { perror(path); exit(1); }
        fprintf(fp, "{\n");
    }
    ~Json() { fprintf(fp, "\n}\n"); fclose(fp); }
    bool firstKey = true;
    void begin(const char *name) {
        fprintf(fp, "%s \"%s\": {", firstSection ? "" : ",\n", name);
        firstSection = false; firstKey = true;
    }
    void end() { fprintf(fp, "}"); }
    void key(const std::string &k) { fprintf(fp, "%s\n  \"%s\": ", firstKey ? "" : ",", k.c_str()); firstKey = false; }
    // float arrays are stored as bit patterns under "f:<name>"
    void floats(const char *name, const std::vector<float> &v) {
        key(std::string("f:") + name);
        fputc('[', fp);
        for (size_t i = 0; i < v.size(); ++i) fprintf(fp, "%s%u", i ? "," : "", fbits(v[i]));
        fputc(']', fp);
    }
    void ints(const char *name, const std::vector<int64_t> &v) {
        key(std::string("i:") + name);
        fputc('[', fp);
        for (size_t i = 0; i < v.size(); ++i) fprintf(fp, "%s%lld", i ? "," : "", (long long)v[i]);
        fputc(']', fp);
    }
};

// A Sampler whose every value is RNG(seq).UniformFloat(): what AwesomeHaltonSampler
// (integrators/photonbeam.cpp:188-256) degenerates to after its 1000 Halton dimensions have been
// burnt (photonbeam.cpp:390-391).
class StreamSampler : public Sampler {
  public:
    explicit StreamSampler(uint64_t seq) : Sampler(1), rng(seq) {}
    Float Get1D() override { return rng.UniformFloat(); }
    Point2f Get2D() override {
        Float a = rng.UniformFloat();
        Float b = rng.UniformFloat();
        return Point2f(a, b);
    }
    std::unique_ptr<Sampler> Clone(int) override { return nullptr; }
    RNG rng;
};

void push3(std::vector<float> &v, const Vector3f &a) { v.push_back(a.x); v.push_back(a.y); v.push_back(a.z); }
void push3(std::vector<float> &v, const Point3f &a) { v.push_back(a.x); v.push_back(a.y); v.push_back(a.z); }
void pushS(std::vector<float> &v, const Spectrum &s) {
    Float rgb[3];
    s.ToRGB(rgb);
    v.push_back(rgb[0]); v.push_back(rgb[1]); v.push_back(rgb[2]);
}
Spectrum rgbS(float r, float g, float b) {
    Float c[3] = {r, g, b};
    return Spectrum::FromRGB(c);
}
Vector3f unit_from(RNG &r) { return UniformSampleSphere(Point2f(r.UniformFloat(), r.UniformFloat())); }

// ---------------------------------------------------------------- scenes
// 5x4x3 test grid used by the density / Tr vectors and by `trace grid`
std::vector<float> test_grid(int nx, int ny, int nz, uint64_t seq) {
    RNG r(seq);
    std::vector<float> d((size_t)nx * ny * nz);
    for (float &x : d) x = r.UniformFloat();
    return d;
}

const float kBoxLo[3] = {-4.f, -1.f, -4.f}, kBoxHi[3] = {4.f, 2.f, 4.f};

// Six quads (12 triangles) of the box [kBoxLo, kBoxHi]  (SURVEY.md Appendix A).
std::vector<std::shared_ptr<Primitive>> box_prims(const std::shared_ptr<Material> &mat,
                                                  const MediumInterface &mi) {
    static Transform id;
    const float *L = kBoxLo, *H = kBoxHi;
    const Point3f quads[6][4] = {
        {{L[0], L[1], L[2]}, {H[0], L[1], L[2]}, {H[0], L[1], H[2]}, {L[0], L[1], H[2]}},   // floor
        {{L[0], H[1], L[2]}, {L[0], H[1], H[2]}, {H[0], H[1], H[2]}, {H[0], H[1], L[2]}},   // ceiling
        {{L[0], L[1], H[2]}, {H[0], L[1], H[2]}, {H[0], H[1], H[2]}, {L[0], H[1], H[2]}},   // back
        {{L[0], L[1], L[2]}, {L[0], H[1], L[2]}, {H[0], H[1], L[2]}, {H[0], L[1], L[2]}},   // front
        {{L[0], L[1], L[2]}, {L[0], L[1], H[2]}, {L[0], H[1], H[2]}, {L[0], H[1], L[2]}},   // left
        {{H[0], L[1], L[2]}, {H[0], H[1], L[2]}, {H[0], H[1], H[2]}, {H[0], L[1], H[2]}}};  // right
    const int idx[6] = {0, 1, 2, 0, 2, 3};
    std::vector<std::shared_ptr<Primitive>> prims;
    for (int q = 0; q < 6; ++q) {
        auto tris = CreateTriangleMesh(&id, &id, false, 2, idx, 4, quads[q], nullptr, nullptr,
                                       nullptr, nullptr, nullptr);
        for (auto &t : tris) prims.push_back(std::make_shared<GeometricPrimitive>(t, mat, nullptr, mi));
    }
    return prims;
}

std::shared_ptr<Material> matte(float kd) {
    std::shared_ptr<Texture<Spectrum>> Kd = std::make_shared<ConstantTexture<Spectrum>>(Spectrum(kd));
    std::shared_ptr<Texture<Float>> sigma = std::make_shared<ConstantTexture<Float>>(0.f);
    return std::make_shared<MatteMaterial>(Kd, sigma, nullptr);
}

// ---------------------------------------------------------------- mode: primitives
void primitives(const char *path) {
    Json js(path);
    {   // core/rng.h:60-145
        js.begin("rng");
        const uint64_t seqs[] = {0, 1, 2, 12345, (1ull << 40) + 7};
        std::vector<int64_t> s, u;
        std::vector<float> f;
        for (uint64_t q : seqs) {
            RNG r(q);
            s.push_back((int64_t)q);
            for (int i = 0; i < 6; ++i) u.push_back(r.UniformUInt32());
            for (int i = 0; i < 6; ++i) f.push_back(r.UniformFloat());
        }
        js.ints("seq", s); js.ints("u32", u); js.floats("float", f);
        js.end();
    }
    {   // core/sampling.cpp:98-103,136-142
        js.begin("sampling");
        RNG r(7);
        std::vector<float> u, sph, cone;
        for (int i = 0; i < 64; ++i) {
            Point2f p(r.UniformFloat(), r.UniformFloat());
            u.push_back(p.x); u.push_back(p.y);
            push3(sph, UniformSampleSphere(p));
            push3(cone, UniformSampleCone(p, 0.8f));
        }
        js.floats("u", u); js.floats("sphere", sph); js.floats("cone", cone);
        js.floats("cosThetaMax", {0.8f});
        js.end();
    }
    {   // core/medium.h:66-69, core/medium.cpp:186-210
        js.begin("phase");
        const float gs[] = {-0.7f, 0.f, 0.0005f, 0.3f, 0.9f};
        std::vector<float> g, c, v, wo, wi, pv, su, swo, swi, sret;
        RNG r(11);
        for (float gg : gs)
            for (int i = 0; i <= 20; ++i) {
                float ct = -1.f + 0.1f * i;
                g.push_back(gg); c.push_back(ct); v.push_back(PhaseHG(ct, gg));
                HenyeyGreenstein hg(gg);
                Vector3f a = unit_from(r), b = unit_from(r);
                push3(wo, a); push3(wi, b); pv.push_back(hg.p(a, b));
                Point2f uu(r.UniformFloat(), r.UniformFloat());
                Vector3f o2 = unit_from(r), w2;
                float ret = hg.Sample_p(o2, &w2, uu);
                su.push_back(uu.x); su.push_back(uu.y);
                push3(swo, o2); push3(swi, w2); sret.push_back(ret);
            }
        js.floats("g", g); js.floats("cosTheta", c); js.floats("PhaseHG", v);
        js.floats("p_wo", wo); js.floats("p_wi", wi); js.floats("p", pv);
        js.floats("s_u", su); js.floats("s_wo", swo); js.floats("s_wi", swi); js.floats("s_ret", sret);
        js.end();
    }
    {   // media/homogeneous.cpp:44-48
        js.begin("homogeneous_tr");
        const float sa[3] = {0.05f, 0.1f, 0.2f}, ss[3] = {0.4f, 0.4f, 0.5f};
        HomogeneousMedium med(rgbS(sa[0], sa[1], sa[2]), rgbS(ss[0], ss[1], ss[2]), 0.3f);
        StreamSampler smp(1);
        std::vector<float> t, tr;
        RNG r(13);
        for (int i = 0; i < 64; ++i) {
            float tt = 12.f * r.UniformFloat() * r.UniformFloat();
            Ray ray(Point3f(0.1f, 0.2f, 0.3f), unit_from(r), tt);
            t.push_back(tt * ray.d.Length());
            pushS(tr, med.Tr(ray, smp));
        }
        js.floats("sigma_a", {sa[0], sa[1], sa[2]}); js.floats("sigma_s", {ss[0], ss[1], ss[2]});
        js.floats("t", t); js.floats("Tr", tr);
        js.end();
    }
    {   // media/grid.cpp:45-60 (Density), :89-121 (Tr, stochastic -> mean and standard error)
        js.begin("grid");
        const int nx = 5, ny = 4, nz = 3;
        std::vector<float> d = test_grid(nx, ny, nz, 17);
        Transform m2w = Translate(Vector3f(-1.f, 0.5f, 2.f)) * Scale(2.f, 1.5f, 3.f);
        Transform w2m = Inverse(m2w);
        GridDensityMedium med(Spectrum(0.5f), Spectrum(1.5f), 0.3f, nx, ny, nz, m2w, d.data());
        std::vector<float> w2mf;
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) w2mf.push_back(w2m.GetMatrix().m[i][j]);
        std::vector<float> p, dens;
        RNG r(19);
        for (int i = 0; i < 256; ++i) {
            Point3f q(-0.2f + 1.4f * r.UniformFloat(), -0.2f + 1.4f * r.UniformFloat(),
                      -0.2f + 1.4f * r.UniformFloat());
            push3(p, q); dens.push_back(med.Density(q));
        }
        std::vector<float> ro, rd, rt, trMean, trErr;
        for (int i = 0; i < 8; ++i) {
            Point3f o(-1.5f + 0.1f * i, 0.4f + 0.2f * i, 1.5f);
            Vector3f dir = Normalize(Vector3f(0.9f - 0.1f * i, 0.15f, 1.f));
            float tMax = 1.f + 0.6f * i;
            double sum = 0, sum2 = 0;
            const int N = 40000;
            for (int k = 0; k < N; ++k) {
                StreamSampler smp(100000ull * (i + 1) + k);
                Float v = med.Tr(Ray(o, dir, tMax), smp)[0];
                sum += v; sum2 += (double)v * v;
            }
            double mean = sum / N, var = sum2 / N - mean * mean;
            push3(ro, o); push3(rd, dir); rt.push_back(tMax);
            trMean.push_back((float)mean); trErr.push_back((float)std::sqrt(var / N));
        }
        js.ints("dims", {nx, ny, nz}); js.floats("density", d); js.floats("w2m", w2mf);
        js.floats("sigma_a", {0.5f}); js.floats("sigma_s", {1.5f});
        js.floats("p", p); js.floats("Density", dens);
        js.floats("ray_o", ro); js.floats("ray_d", rd); js.floats("ray_tMax", rt);
        js.floats("Tr_mean", trMean); js.floats("Tr_stderr", trErr);
        js.end();
    }
    {   // core/geometry.h:1386-1408 and :1410-1438
        js.begin("bounds");
        RNG r(23);
        std::vector<float> lo, hi, o, d, tm, t0s, t1s;
        std::vector<int64_t> hitP, hitSlab;
        for (int i = 0; i < 2000; ++i) {
            Point3f a(r.UniformFloat(), r.UniformFloat(), r.UniformFloat());
            float sz = (i % 4 == 0) ? 0.5f : 0.02f;
            Point3f b = a + sz * Vector3f(r.UniformFloat(), r.UniformFloat(), r.UniformFloat());
            Bounds3f box(a, b);
            Point3f org(2.f * r.UniformFloat() - 0.5f, 2.f * r.UniformFloat() - 0.5f, -1.5f * r.UniformFloat());
            // aim near the box so that a useful fraction of the tests are close calls
            Point3f target = a + Vector3f(r.UniformFloat() * 1.4f - 0.2f, r.UniformFloat() * 1.4f - 0.2f,
                                          r.UniformFloat() * 1.4f - 0.2f) * sz;
            Vector3f dir = Normalize(target - org);
            if (i % 7 == 0) dir.y = 0.f, dir = Normalize(dir);   // axis-parallel component (1/0 = inf)
            float tMax = (i % 3 == 0) ? (target - org).Length() * (0.5f + r.UniformFloat()) : Infinity;
            Ray ray(org, dir, tMax);
            Float t0 = 0, t1 = 0;
            bool h = box.IntersectP(ray, &t0, &t1);
            Vector3f invDir(1 / dir.x, 1 / dir.y, 1 / dir.z);
            int neg[3] = {invDir.x < 0, invDir.y < 0, invDir.z < 0};
            bool hs = box.IntersectP(ray, invDir, neg);
            push3(lo, box.pMin); push3(hi, box.pMax); push3(o, org); push3(d, dir); tm.push_back(tMax);
            hitP.push_back(h); t0s.push_back(h ? t0 : 0.f); t1s.push_back(h ? t1 : 0.f);
            hitSlab.push_back(hs);
        }
        js.floats("lo", lo); js.floats("hi", hi); js.floats("o", o); js.floats("d", d); js.floats("tMax", tm);
        js.ints("IntersectP", hitP); js.floats("t0", t0s); js.floats("t1", t1s);
        js.ints("IntersectP_invDir", hitSlab);
        js.end();
    }
    {   // core/photonbeambvh.cpp:109-139, :141-182, :438-445
        js.begin("morton");
        RNG r(29);
        const float lo[3] = {-0.3f, 0.1f, 2.f}, hi[3] = {1.7f, 0.9f, 2.5f};
        std::vector<float> p;
        std::vector<int64_t> code;
        for (int i = 0; i < 512; ++i) {
            float q[3];
            for (int k = 0; k < 3; ++k) q[k] = lo[k] + (hi[k] - lo[k]) * r.UniformFloat();
            if (i < 3) q[i] = hi[i];           // offset exactly 1 -> 1024 -> clamped to 1023
            if (i == 3) { q[0] = lo[0]; q[1] = lo[1]; q[2] = lo[2]; }
            p.push_back(q[0]); p.push_back(q[1]); p.push_back(q[2]);
            code.push_back(refx::morton_of_point(q, lo, hi));
        }
        // radix sort of codes with many duplicates: the reference sort is stable
        std::vector<int> prim(4096);
        std::vector<uint32_t> codes(4096);
        std::vector<int64_t> inCodes, outPrim;
        for (int i = 0; i < 4096; ++i) { prim[i] = i; codes[i] = r.UniformUInt32(1u << 30) & 0x3ff003ffu; inCodes.push_back(codes[i]); }
        refx::radix_sort(4096, prim.data(), codes.data());
        for (int i = 0; i < 4096; ++i) outPrim.push_back(prim[i]);
        js.floats("lo", {lo[0], lo[1], lo[2]}); js.floats("hi", {hi[0], hi[1], hi[2]});
        js.floats("p", p); js.ints("code", code);
        js.ints("sort_in_codes", inCodes); js.ints("sort_out_prim", outPrim);
        js.end();
    }
    {   // integrators/photonmap.cpp:170-174
        js.begin("kernel");
        RNG r(31);
        std::vector<float> a, b, md2, k;
        for (int i = 0; i < 64; ++i) {
            float pa[3] = {r.UniformFloat(), r.UniformFloat(), r.UniformFloat()};
            float rad = 0.05f + 0.1f * r.UniformFloat();
            float pb[3] = {pa[0] + rad * 0.5f * r.UniformFloat(), pa[1] - rad * 0.5f * r.UniformFloat(), pa[2]};
            for (int c = 0; c < 3; ++c) { a.push_back(pa[c]); b.push_back(pb[c]); }
            md2.push_back(rad * rad);
            k.push_back(refx::kernel(pa, pb, rad * rad));
        }
        js.floats("photon_p", a); js.floats("p", b); js.floats("maxDist2", md2); js.floats("kernel", k);
        js.end();
    }
    {   // core/film.cpp:132-140 (SetImage), :168-204 (WriteImage), via the PFM writer of ref_glue.cpp
        js.begin("film");
        const int W = 8, H = 4;
        const char *tmp = "/tmp/bre_ref_film.pfm";
        Film film(Point2i(W, H), Bounds2f(Point2f(0, 0), Point2f(1, 1)),
                  std::unique_ptr<Filter>(new BoxFilter(Vector2f(0.5f, 0.5f))), 35.f, tmp, 1.5f);
        RNG r(37);
        std::vector<Spectrum> img(W * H);
        std::vector<float> in;
        for (int i = 0; i < W * H; ++i) {
            float c[3] = {4.f * r.UniformFloat(), 4.f * r.UniformFloat(), 4.f * r.UniformFloat()};
            if (i == 5) c[0] = -0.25f;     // negative radiance is clamped by WriteImage
            if (i == 6) c[1] = 0.f;
            img[i] = rgbS(c[0], c[1], c[2]);
            in.push_back(c[0]); in.push_back(c[1]); in.push_back(c[2]);
        }
        film.SetImage(img.data());
        film.WriteImage();
        // read the PFM back (bottom-to-top scanlines)
        std::ifstream f(tmp, std::ios::binary);
        std::string magic; int w, h; float sc;
        f >> magic >> w >> h >> sc; f.get();
        std::vector<float> raw((size_t)w * h * 3), out((size_t)w * h * 3);
        f.read((char *)raw.data(), raw.size() * 4);
        for (int y = 0; y < h; ++y)
            memcpy(&out[(size_t)y * w * 3], &raw[(size_t)(h - 1 - y) * w * 3], (size_t)w * 12);
        js.ints("dims", {W, H}); js.floats("scale", {1.5f});
        js.floats("rgb_in", in); js.floats("rgb_out", out);
        // the same accumulators after three iterations: L = Ld / (iter + 1) with iter = 2 (integrators/photonbeam.cpp:578,
        // CoefficientSpectrum::operator/(Float) core/spectrum.h:181-188 -- a division per channel, not a reciprocal)
        std::vector<Spectrum> img3(W * H);
        for (int i = 0; i < W * H; ++i) img3[i] = img[i] / (2 + 1);
        film.SetImage(img3.data());
        film.WriteImage();
        std::ifstream f3(tmp, std::ios::binary);
        f3 >> magic >> w >> h >> sc; f3.get();
        f3.read((char *)raw.data(), raw.size() * 4);
        for (int y = 0; y < h; ++y)
            memcpy(&out[(size_t)y * w * 3], &raw[(size_t)(h - 1 - y) * w * 3], (size_t)w * 12);
        js.floats("rgb_out_iter2", out);
        js.end();
        remove(tmp);
    }
}

// ---------------------------------------------------------------- binary container
// "BREF1\n" then records: name[24] (zero padded), dtype[8] ("f4","u4","u8","i4"), count u64, data.
struct BinWriter {
    FILE *fp;
    explicit BinWriter(const char *path) : fp(fopen(path, "wb")) {
        if (!fp) { perror(path); exit(1); }
        fwrite("BREF1\n", 1, 6, fp);
    }
    ~BinWriter() { fclose(fp); }
    void add(const char *name, const char *dtype, const void *data, uint64_t count, size_t elem) {
        char n[24] = {0}, t[8] = {0};
        strncpy(n, name, 23); strncpy(t, dtype, 7);
        fwrite(n, 1, 24, fp); fwrite(t, 1, 8, fp); fwrite(&count, 8, 1, fp);
        if (count) fwrite(data, elem, count, fp);
    }
    void f4(const char *name, const std::vector<float> &v) { add(name, "f4", v.data(), v.size(), 4); }
    void u4(const char *name, const std::vector<uint32_t> &v) { add(name, "u4", v.data(), v.size(), 4); }
    void u8(const char *name, const std::vector<uint64_t> &v) { add(name, "u8", v.data(), v.size(), 8); }
};

// ---------------------------------------------------------------- mode: fogbox
// order-independent hit-set digest of the large fixtures: sum over the set of splitmix64(index)
static inline uint64_t hit_hash(uint32_t i) {
    uint64_t z = (uint64_t)i + 0x9e3779b97f4a7c15ull;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}
void fogbox(const char *path, int nVolumeWanted, int res, float radius, uint32_t keep, bool digestOnly) {
    const float sa = 0.05f, ss = 0.4f, g = 0.3f;
    HomogeneousMedium fog{Spectrum(sa), Spectrum(ss), g};
    MediumInterface inFog(&fog, &fog);
    std::vector<std::shared_ptr<Primitive>> prims = box_prims(matte(0.6f), inFog);
    std::shared_ptr<BVHAccel> bvh = std::make_shared<BVHAccel>(prims);
    std::vector<std::shared_ptr<Light>> lights;
    lights.push_back(std::make_shared<PointLight>(Translate(Vector3f(0.f, 0.9f, 0.f)), inFog, Spectrum(5.f)));
    Scene scene(bvh, lights);

    Film *film = new Film(Point2i(res, res), Bounds2f(Point2f(0, 0), Point2f(1, 1)),
                          std::unique_ptr<Filter>(new BoxFilter(Vector2f(0.5f, 0.5f))), 35.f,
                          "/tmp/bre_ref_fogbox.pfm", 1.f);
    Transform c2w = Inverse(LookAt(Point3f(0, 0, -3.5f), Point3f(0, 0, 0), Vector3f(0, 1, 0)));
    AnimatedTransform ac2w(&c2w, 0.f, &c2w, 1.f);
    std::shared_ptr<const Camera> camera = std::make_shared<PerspectiveCamera>(
        ac2w, Bounds2f(Point2f(-1, -1), Point2f(1, 1)), 0.f, 1.f, 0.f, 1e6f, 40.f, film, &fog);

    // ---- photon pass: the reference's own ShootPhotons
    refx::PhotonDump dump;
    refx::shoot_volume_photons(scene, camera, nVolumeWanted, 5, &dump);
    // ShootPhotons runs NumSystemCores() tasks of >= 4096 paths each, so the dump is large; the
    // fixture keeps a PREFIX of it (reference storage order).  n_paths stays the reference's nShot.
    if (keep && dump.p.size() / 3 > keep) { dump.p.resize(3 * (size_t)keep); dump.alpha.resize(3 * (size_t)keep); dump.wi.resize(3 * (size_t)keep); }
    const uint32_t n = (uint32_t)(dump.p.size() / 3);

    // ---- camera rays: pixel centres through the reference camera, tMax from Scene::Intersect
    // (integrators/photonbeam.cpp:465-481)
    std::vector<float> ro, rd;
    std::vector<Ray> rays;
    for (int y = 0; y < res; ++y)
        for (int x = 0; x < res; ++x) {
            CameraSample cs;
            cs.pFilm = Point2f(x + 0.5f, y + 0.5f);
            cs.pLens = Point2f(0.5f, 0.5f);
            cs.time = 0.f;
            RayDifferential ray;
            camera->GenerateRayDifferential(cs, &ray);
            SurfaceInteraction isect;
            scene.Intersect(ray, &isect);   // shortens ray.tMax (core/primitive.cpp:97-101)
            rays.push_back(ray);
            ro.push_back(ray.o.x); ro.push_back(ray.o.y); ro.push_back(ray.o.z); ro.push_back(ray.tMax);
            push3(rd, ray.d);
        }

    // ---- beam radiance estimate, brute force, reference arithmetic classes (rays over all host threads)
    const float nPaths = (float)dump.nShot;
    const float r2 = radius * radius;
    HenyeyGreenstein hg(g);
    const size_t nRays = rays.size();
    std::vector<std::vector<uint32_t>> hitsOf(nRays);
    std::vector<float> rgb(3 * nRays);
    {
        std::atomic<size_t> next(0);
        std::vector<std::thread> pool;
        const unsigned nThreads = std::max(1u, std::thread::hardware_concurrency());
        for (unsigned tIdx = 0; tIdx < nThreads; ++tIdx)
            pool.emplace_back([&]() {
                StreamSampler smp(1);
                for (size_t q = next++; q < nRays; q = next++) {
                    const Ray &ray = rays[q];
                    Spectrum L(0.f);
                    for (uint32_t i = 0; i < n; ++i) {
                        Point3f p(dump.p[3 * i], dump.p[3 * i + 1], dump.p[3 * i + 2]);
                        Vector3f v = p - ray.o;
                        Float t = Dot(v, ray.d);
                        if (!(t > 0 && t < ray.tMax)) continue;
                        Vector3f w = v - t * ray.d;
                        Float d2 = w.LengthSquared();
                        if (!(d2 < r2)) continue;            // strict <, core/kdtree.h:182-184
                        hitsOf[q].push_back(i);
                        float pp[3] = {p.x, p.y, p.z};
                        Point3f c = ray(t);
                        float cc[3] = {c.x, c.y, c.z};
                        Float k = refx::kernel(pp, cc, r2);   // integrators/photonmap.cpp:170-174
                        Vector3f wi(dump.wi[3 * i], dump.wi[3 * i + 1], dump.wi[3 * i + 2]);
                        Spectrum alpha = rgbS(dump.alpha[3 * i], dump.alpha[3 * i + 1], dump.alpha[3 * i + 2]);
                        Spectrum Tr = fog.Tr(Ray(ray.o, ray.d, t), smp);
                        // photonmap.cpp:198 normalisation, vsppm.cpp:488-489 phase orientation
                        L += (k / (nPaths * r2)) * hg.p(-ray.d, wi) * Tr * alpha;
                    }
                    Float c3[3];
                    L.ToRGB(c3);
                    rgb[3 * q] = c3[0]; rgb[3 * q + 1] = c3[1]; rgb[3 * q + 2] = c3[2];
                }
            });
        for (auto &th : pool) th.join();
    }
    std::vector<uint64_t> offs(1, 0);
    std::vector<uint32_t> idx, counts;
    std::vector<uint64_t> digest;
    for (size_t q = 0; q < nRays; ++q) {
        uint64_t dg = 0;
        for (uint32_t i : hitsOf[q]) { idx.push_back(i); dg += hit_hash(i); }
        offs.push_back(idx.size());
        counts.push_back((uint32_t)hitsOf[q].size());
        digest.push_back(dg);
    }

    BinWriter bw(path);
    bw.f4("medium", {sa, sa, sa, ss, ss, ss, g});
    bw.f4("radius", {radius});
    bw.f4("n_paths", {nPaths});
    bw.u4("res", {(uint32_t)res, (uint32_t)res});
    bw.f4("photon_p", dump.p); bw.f4("photon_alpha", dump.alpha); bw.f4("photon_wi", dump.wi);
    bw.f4("ray_o_tmax", ro); bw.f4("ray_d", rd);
    if (digestOnly) {
        // large fixtures: per-ray hit count + order-independent digest (sum of splitmix64(photon index)) instead of lists
        bw.u4("hit_counts", counts); bw.u8("hit_digest", digest);
    } else {
        bw.u8("hit_offsets", offs); bw.u4("hit_indices", idx);
    }
    bw.f4("radiance", rgb);
    fprintf(stderr, "fogbox: %u volume photons from %u paths, %zu rays, %zu hits\n", n, dump.nShot,
            rays.size(), idx.size());
}

// The volume-photon deposit loop of modes trace / gridspot: sampler stream of photonbeam.cpp:386-391, deposit rule of
// photonmap.cpp:696-724, every step executed by reference classes.
static void walk_volume_paths(const Scene &scene, const Distribution1D *lightDistr, uint64_t first, uint64_t nPaths,
                              int maxDepth, std::vector<float> &pp, std::vector<float> &pa, std::vector<float> &pw,
                              std::vector<uint32_t> &pid) {
    MemoryArena arena;
    for (uint64_t pi = first; pi < first + nPaths; ++pi) {
        StreamSampler sampler(pi + 1);                       // photonbeam.cpp:386-391
        Float lightPdf;
        int lightNum = lightDistr->SampleDiscrete(sampler.Get1D(), &lightPdf);   // :394-398
        const std::shared_ptr<Light> &light = scene.lights[lightNum];
        Point2f uLight0 = sampler.Get2D();
        Point2f uLight1 = sampler.Get2D();
        Float uLightTime = sampler.Get1D();
        RayDifferential photonRay;
        Normal3f nLight;
        Float pdfPos, pdfDir;
        Spectrum Le = light->Sample_Le(uLight0, uLight1, uLightTime, &photonRay, &nLight, &pdfPos, &pdfDir);
        if (pdfPos == 0 || pdfDir == 0 || Le.IsBlack()) continue;
        Spectrum alpha = (AbsDot(nLight, photonRay.d) * Le) / (lightPdf * pdfPos * pdfDir);   // :413-414
        if (alpha.IsBlack()) continue;
        SurfaceInteraction sisect;
        int nIntersections = 0;
        while (scene.Intersect(photonRay, &sisect)) {        // photonmap.cpp:696
            ++nIntersections;
            MediumInteraction misect;
            if (photonRay.medium) alpha *= photonRay.medium->Sample(photonRay, sampler, arena, &misect);
            if (alpha.IsBlack()) break;
            if (!misect.IsValid()) break;                    // reached the absorbing wall
            const Vector3f wo = -photonRay.d;
            Vector3f wi;
            misect.phase->Sample_p(wo, &wi, sampler.Get2D());
            push3(pp, misect.p); pushS(pa, alpha); push3(pw, wo);   // Photon(misect.p, alpha, wo) :720-723
            pid.push_back((uint32_t)pi);
            photonRay = misect.SpawnRay(wi);
            if (nIntersections >= maxDepth) break;           // :794
            Float sample = sampler.Get1D();                  // RR draw, continueProb == 1 (:799-804)
            if (sample > 1.f) break;
        }
        arena.Reset();
    }
}

// ---------------------------------------------------------------- mode: trace
void trace(const char *path, const char *kind, uint64_t first, uint64_t nPaths, int maxDepth) {
    const bool grid = strcmp(kind, "grid") == 0;
    const float sa[3] = {0.05f, 0.05f, 0.05f}, ss[3] = {0.4f, 0.4f, 0.4f};
    const float g = 0.3f;
    std::unique_ptr<Medium> med;
    const int nx = 6, ny = 5, nz = 4;
    std::vector<float> dens = test_grid(nx, ny, nz, 41);
    Transform m2w = Translate(Vector3f(kBoxLo[0], kBoxLo[1], kBoxLo[2])) *
                    Scale(kBoxHi[0] - kBoxLo[0], kBoxHi[1] - kBoxLo[1], kBoxHi[2] - kBoxLo[2]);
    Transform w2m = Inverse(m2w);
    if (grid) med.reset(new GridDensityMedium(Spectrum(sa[0]), Spectrum(ss[0] * 4), g, nx, ny, nz, m2w, dens.data()));
    else med.reset(new HomogeneousMedium(rgbS(sa[0], sa[1], sa[2]), rgbS(ss[0], ss[1], ss[2]), g));
    MediumInterface in(med.get(), med.get());
    std::vector<std::shared_ptr<Primitive>> prims = box_prims(nullptr, in);
    std::shared_ptr<BVHAccel> bvh = std::make_shared<BVHAccel>(prims);
    std::vector<std::shared_ptr<Light>> lights;
    lights.push_back(std::make_shared<PointLight>(Translate(Vector3f(0.f, 0.9f, 0.f)), in, rgbS(5.f, 4.f, 3.f)));
    // spot light aimed along (0.3,-1,0.2), built exactly as a scene file would (lights/spot.cpp:104-125)
    const Point3f spotFrom(1.f, 1.9f, -1.f), spotTo(1.3f, 0.9f, -0.8f);
    {
        ParamSet ps;
        std::unique_ptr<Point3f[]> from(new Point3f[1]), to(new Point3f[1]);
        from[0] = spotFrom; to[0] = spotTo;
        std::unique_ptr<Float[]> I(new Float[3]), ca(new Float[1]), cd(new Float[1]);
        I[0] = 30.f; I[1] = 30.f; I[2] = 20.f; ca[0] = 35.f; cd[0] = 10.f;
        ps.AddPoint3f("from", std::move(from), 1); ps.AddPoint3f("to", std::move(to), 1);
        ps.AddRGBSpectrum("I", std::move(I), 3);
        ps.AddFloat("coneangle", std::move(ca), 1); ps.AddFloat("conedeltaangle", std::move(cd), 1);
        lights.push_back(CreateSpotLight(Transform(), med.get(), ps));
    }
    Scene scene(bvh, lights);
    std::unique_ptr<Distribution1D> lightDistr = ComputeLightPowerDistribution(scene);

    std::vector<float> pp, pa, pw;
    std::vector<uint32_t> pid;
    walk_volume_paths(scene, lightDistr.get(), first, nPaths, maxDepth, pp, pa, pw, pid);
    std::vector<float> w2mf;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) w2mf.push_back(w2m.GetMatrix().m[i][j]);
    std::vector<float> lightAxis;
    push3(lightAxis, Normalize(spotTo - spotFrom));
    BinWriter bw(path);
    bw.u4("kind", {(uint32_t)grid});
    bw.f4("sigma_a", {sa[0], sa[1], sa[2]});
    if (grid) bw.f4("sigma_s", {ss[0] * 4, ss[0] * 4, ss[0] * 4}); else bw.f4("sigma_s", {ss[0], ss[1], ss[2]});
    bw.f4("g", {g});
    bw.u4("grid_dims", {(uint32_t)nx, (uint32_t)ny, (uint32_t)nz});
    bw.f4("grid_density", dens); bw.f4("w2m", w2mf);
    bw.f4("box_lo", {kBoxLo[0], kBoxLo[1], kBoxLo[2]}); bw.f4("box_hi", {kBoxHi[0], kBoxHi[1], kBoxHi[2]});
    bw.f4("spot_axis", lightAxis);
    bw.u8("range", {first, nPaths, (uint64_t)maxDepth});
    bw.f4("photon_p", pp); bw.f4("photon_alpha", pa); bw.f4("photon_wi", pw); bw.u4("path_id", pid);
    fprintf(stderr, "trace %s: %zu photons from %llu paths\n", kind, pid.size(), (unsigned long long)nPaths);
}

// ---------------------------------------------------------------- mode: gridspot
// Query fixture for a heterogeneous medium lit by a spot light.  Photons: reference classes walk paths from a
// SpotLight (lights/spot.cpp) through a GridDensityMedium filling the box (walk_volume_paths).  Rays: the reference's
// PerspectiveCamera inside the box + Scene::Intersect for tMax.  Estimate: brute force over all photons with the
// reference's vector / kernel / phase code; the transmittance to a hit at depth t is exp(-sigma_t * integral of the
// reference's Density() over [0, t]) with the integral taken by a 4096-step midpoint rule in double -- the quantity
// GridDensityMedium::Tr (ratio tracking, media/grid.cpp:89-121) estimates without bias (see mode gridtr).
void gridspot(const char *path, uint64_t nPaths, int res, float radius) {
    const float sa = 0.05f, ss = 1.6f, g = 0.3f;
    const int nx = 12, ny = 10, nz = 8;
    std::vector<float> dens = test_grid(nx, ny, nz, 43);
    Transform m2w = Translate(Vector3f(kBoxLo[0], kBoxLo[1], kBoxLo[2])) *
                    Scale(kBoxHi[0] - kBoxLo[0], kBoxHi[1] - kBoxLo[1], kBoxHi[2] - kBoxLo[2]);
    Transform w2m = Inverse(m2w);
    GridDensityMedium med(Spectrum(sa), Spectrum(ss), g, nx, ny, nz, m2w, dens.data());
    MediumInterface in(&med, &med);
    std::vector<std::shared_ptr<Primitive>> prims = box_prims(nullptr, in);
    std::shared_ptr<BVHAccel> bvh = std::make_shared<BVHAccel>(prims);
    std::vector<std::shared_ptr<Light>> lights;
    const Point3f spotFrom(-1.5f, 1.8f, -1.f), spotTo(0.4f, 0.2f, 0.6f);
    {
        ParamSet ps;
        std::unique_ptr<Point3f[]> from(new Point3f[1]), to(new Point3f[1]);
        from[0] = spotFrom; to[0] = spotTo;
        std::unique_ptr<Float[]> I(new Float[3]), ca(new Float[1]), cd(new Float[1]);
        I[0] = 30.f; I[1] = 25.f; I[2] = 20.f; ca[0] = 30.f; cd[0] = 8.f;
        ps.AddPoint3f("from", std::move(from), 1); ps.AddPoint3f("to", std::move(to), 1);
        ps.AddRGBSpectrum("I", std::move(I), 3);
        ps.AddFloat("coneangle", std::move(ca), 1); ps.AddFloat("conedeltaangle", std::move(cd), 1);
        lights.push_back(CreateSpotLight(Transform(), &med, ps));
    }
    Scene scene(bvh, lights);
    std::unique_ptr<Distribution1D> lightDistr = ComputeLightPowerDistribution(scene);
    std::vector<float> pp, pa, pw;
    std::vector<uint32_t> pid;
    walk_volume_paths(scene, lightDistr.get(), 0, nPaths, 5, pp, pa, pw, pid);
    const uint32_t n = (uint32_t)pid.size();

    Film *film = new Film(Point2i(res, res), Bounds2f(Point2f(0, 0), Point2f(1, 1)),
                          std::unique_ptr<Filter>(new BoxFilter(Vector2f(0.5f, 0.5f))), 35.f,
                          "/tmp/bre_ref_gridspot.pfm", 1.f);
    Transform c2w = Inverse(LookAt(Point3f(0.3f, 0.6f, -3.6f), Point3f(0.1f, 0.4f, 0.f), Vector3f(0, 1, 0)));
    AnimatedTransform ac2w(&c2w, 0.f, &c2w, 1.f);
    std::shared_ptr<const Camera> camera = std::make_shared<PerspectiveCamera>(
        ac2w, Bounds2f(Point2f(-1, -1), Point2f(1, 1)), 0.f, 1.f, 0.f, 1e6f, 50.f, film, &med);
    std::vector<float> ro, rd;
    std::vector<Ray> rays;
    for (int y = 0; y < res; ++y)
        for (int x = 0; x < res; ++x) {
            CameraSample cs;
            cs.pFilm = Point2f(x + 0.5f, y + 0.5f);
            cs.pLens = Point2f(0.5f, 0.5f);
            cs.time = 0.f;
            RayDifferential ray;
            camera->GenerateRayDifferential(cs, &ray);
            SurfaceInteraction isect;
            scene.Intersect(ray, &isect);   // shortens ray.tMax (core/primitive.cpp:97-101)
            rays.push_back(ray);
            ro.push_back(ray.o.x); ro.push_back(ray.o.y); ro.push_back(ray.o.z); ro.push_back(ray.tMax);
            push3(rd, ray.d);
        }
    const float nPathsF = (float)nPaths;
    const float r2 = radius * radius;
    const double sigT = (double)sa + (double)ss;
    HenyeyGreenstein hg(g);
    const size_t nRays = rays.size();
    std::vector<std::vector<uint32_t>> hitsOf(nRays);
    std::vector<std::vector<float>> tauOf(nRays);
    std::vector<float> rgb(3 * nRays);
    {
        std::atomic<size_t> next(0);
        std::vector<std::thread> pool;
        const unsigned nThreads = std::max(1u, std::thread::hardware_concurrency());
        for (unsigned tIdx = 0; tIdx < nThreads; ++tIdx)
            pool.emplace_back([&]() {
                for (size_t q = next++; q < nRays; q = next++) {
                    const Ray &ray = rays[q];
                    Spectrum L(0.f);
                    for (uint32_t i = 0; i < n; ++i) {
                        Point3f p(pp[3 * i], pp[3 * i + 1], pp[3 * i + 2]);
                        Vector3f v = p - ray.o;
                        Float t = Dot(v, ray.d);
                        if (!(t > 0 && t < ray.tMax)) continue;
                        Vector3f w = v - t * ray.d;
                        Float d2 = w.LengthSquared();
                        if (!(d2 < r2)) continue;            // strict <, core/kdtree.h:182-184
                        hitsOf[q].push_back(i);
                        float ppf[3] = {p.x, p.y, p.z};
                        Point3f c = ray(t);
                        float cc[3] = {c.x, c.y, c.z};
                        Float k = refx::kernel(ppf, cc, r2);   // integrators/photonmap.cpp:170-174
                        Vector3f wi(pw[3 * i], pw[3 * i + 1], pw[3 * i + 2]);
                        Spectrum alpha = rgbS(pa[3 * i], pa[3 * i + 1], pa[3 * i + 2]);
                        // optical depth over [0, t]: fine midpoint integral of the reference's Density() in medium space
                        Ray rm = w2m(Ray(ray.o, ray.d, t));
                        Float t0, t1;
                        double tau = 0;
                        if (Bounds3f(Point3f(0, 0, 0), Point3f(1, 1, 1)).IntersectP(rm, &t0, &t1)) {
                            const int N = 4096;
                            const double dt = ((double)t1 - (double)t0) / N;
                            for (int s2 = 0; s2 < N; ++s2) {
                                const double tt = (double)t0 + (s2 + 0.5) * dt;
                                Point3f pm((Float)(rm.o.x + rm.d.x * tt), (Float)(rm.o.y + rm.d.y * tt),
                                           (Float)(rm.o.z + rm.d.z * tt));
                                tau += (double)med.Density(pm) * dt;
                            }
                        }
                        tauOf[q].push_back((float)tau);
                        const Float Tr = (Float)std::exp(-sigT * tau);
                        L += (k / (nPathsF * r2)) * hg.p(-ray.d, wi) * Tr * alpha;
                    }
                    Float c3[3];
                    L.ToRGB(c3);
                    rgb[3 * q] = c3[0]; rgb[3 * q + 1] = c3[1]; rgb[3 * q + 2] = c3[2];
                }
            });
        for (auto &th : pool) th.join();
    }
    std::vector<uint64_t> offs(1, 0);
    std::vector<uint32_t> idx;
    std::vector<float> taus;
    for (size_t q = 0; q < nRays; ++q) {
        idx.insert(idx.end(), hitsOf[q].begin(), hitsOf[q].end());
        taus.insert(taus.end(), tauOf[q].begin(), tauOf[q].end());
        offs.push_back(idx.size());
    }
    std::vector<float> w2mf;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) w2mf.push_back(w2m.GetMatrix().m[i][j]);
    BinWriter bw(path);
    bw.f4("medium", {sa, sa, sa, ss, ss, ss, g});
    bw.u4("grid_dims", {(uint32_t)nx, (uint32_t)ny, (uint32_t)nz});
    bw.f4("grid_density", dens); bw.f4("w2m", w2mf);
    bw.f4("radius", {radius});
    bw.f4("n_paths", {nPathsF});
    bw.u4("res", {(uint32_t)res, (uint32_t)res});
    bw.f4("photon_p", pp); bw.f4("photon_alpha", pa); bw.f4("photon_wi", pw);
    bw.f4("ray_o_tmax", ro); bw.f4("ray_d", rd);
    bw.u8("hit_offsets", offs); bw.u4("hit_indices", idx); bw.f4("hit_tau", taus); bw.f4("radiance", rgb);
    fprintf(stderr, "gridspot: %u volume photons from %llu paths, %zu rays, %zu hits\n", n, (unsigned long long)nPaths,
            rays.size(), idx.size());
}

// ---------------------------------------------------------------- mode: halton
// Camera samples as the reference's camera pass draws them (integrators/photonbeam.cpp:341,459-466):
// HaltonSampler(nIterations, pixelBounds); per pixel StartPixel(p), SetSampleNumber(iter), then the first Get2D of
// GetCameraSample (core/sampler.cpp:46-52).  Several pixel windows (smaller and larger than kMaxResolution = 128,
// non-zero origin) x several iterations.
void halton(const char *path) {
    const int windows[4][4] = {{0, 0, 40, 30}, {3, 5, 150, 105}, {17, 9, 17 + 140, 9 + 131}, {100, 200, 101, 201}};
    // iterations stored per window (bit mask over iters[]): the large windows keep two / one of them
    const unsigned keep[4] = {15u, 5u, 8u, 15u};
    const int iters[4] = {0, 1, 7, 1000};
    BinWriter bw(path);
    std::vector<uint32_t> win;
    for (int wi = 0; wi < 4; ++wi) {
        const int *W = windows[wi];
        Bounds2i pb(Point2i(W[0], W[1]), Point2i(W[2], W[3]));
        for (int k = 0; k < 4; ++k) win.push_back((uint32_t)W[k]);
        HaltonSampler sampler(64, pb);
        for (int it = 0; it < 4; ++it) {
            if (!(keep[wi] >> it & 1u)) continue;
            std::vector<float> xy;
            for (Point2i p : pb) {
                sampler.StartPixel(p);
                sampler.SetSampleNumber(iters[it]);
                CameraSample cs = sampler.GetCameraSample(p);
                xy.push_back(cs.pFilm.x);
                xy.push_back(cs.pFilm.y);
            }
            char name[24];
            snprintf(name, sizeof(name), "w%d_i%d", wi, iters[it]);
            bw.f4(name, xy);
        }
    }
    bw.u4("windows", win);
    bw.u4("iters", {0, 1, 7, 1000});
    fprintf(stderr, "halton: 4 windows x 4 iterations\n");
}

// ---------------------------------------------------------------- mode: surfaces
// Fog box with matte walls of different reflectance, one reverse-orientation wall, and an
// interface-only quad (no material, MediumInterface inside = fog, outside = vacuum) at x = 1.5 that
// leaves the region x > 1.5 empty.  Photon paths are walked by REFERENCE classes with the loop of
// integrators/photonbeam.cpp:258-325 (surface bounce, interface skip, Russian roulette) and the
// volume-photon deposit rule of integrators/photonmap.cpp:696-724; camera rays come from the
// reference's PerspectiveCamera + Scene::Intersect over a cropped pixel window.
void surfaces(const char *path, uint64_t first, uint64_t nPaths, int maxDepth, int resX, int resY) {
    const float sa[3] = {0.05f, 0.06f, 0.07f}, ss[3] = {0.4f, 0.35f, 0.3f}, g = 0.3f;
    HomogeneousMedium fog{rgbS(sa[0], sa[1], sa[2]), rgbS(ss[0], ss[1], ss[2]), g};
    MediumInterface inFog(&fog, &fog);
    static Transform id;
    const float *L = kBoxLo, *H = kBoxHi;
    const Point3f quads[7][4] = {
        {{L[0], L[1], L[2]}, {H[0], L[1], L[2]}, {H[0], L[1], H[2]}, {L[0], L[1], H[2]}},
        {{L[0], H[1], L[2]}, {L[0], H[1], H[2]}, {H[0], H[1], H[2]}, {H[0], H[1], L[2]}},
        {{L[0], L[1], H[2]}, {H[0], L[1], H[2]}, {H[0], H[1], H[2]}, {L[0], H[1], H[2]}},
        {{L[0], L[1], L[2]}, {L[0], H[1], L[2]}, {H[0], H[1], L[2]}, {H[0], L[1], L[2]}},
        {{L[0], L[1], L[2]}, {L[0], L[1], H[2]}, {L[0], H[1], H[2]}, {L[0], H[1], L[2]}},
        {{H[0], L[1], L[2]}, {H[0], H[1], L[2]}, {H[0], H[1], H[2]}, {H[0], L[1], H[2]}},
        {{1.5f, L[1], L[2]}, {1.5f, H[1], L[2]}, {1.5f, H[1], H[2]}, {1.5f, L[1], H[2]}}};   // interface
    const float kd[7][3] = {{.6f, .6f, .6f}, {.7f, .7f, .7f}, {.2f, .5f, .8f}, {.5f, .5f, .5f},
                            {.8f, .3f, .2f}, {.3f, .8f, .3f}, {-1.f, -1.f, -1.f}};
    const int idx[6] = {0, 1, 2, 0, 2, 3};
    std::vector<std::shared_ptr<Primitive>> prims;
    std::vector<float> P, Kd;
    std::vector<uint32_t> I, attr;
    for (int q = 0; q < 7; ++q) {
        const bool reverse = (q == 4);
        const bool iface = (q == 6);
        auto tris = CreateTriangleMesh(&id, &id, reverse, 2, idx, 4, quads[q], nullptr, nullptr,
                                       nullptr, nullptr, nullptr);
        std::shared_ptr<Material> mat;
        if (!iface) {
            std::shared_ptr<Texture<Spectrum>> t = std::make_shared<ConstantTexture<Spectrum>>(rgbS(kd[q][0], kd[q][1], kd[q][2]));
            std::shared_ptr<Texture<Float>> sg = std::make_shared<ConstantTexture<Float>>(0.f);
            mat = std::make_shared<MatteMaterial>(t, sg, nullptr);
        }
        MediumInterface mi = iface ? MediumInterface(&fog, nullptr) : inFog;
        for (auto &t : tris) prims.push_back(std::make_shared<GeometricPrimitive>(t, mat, nullptr, mi));
        for (int v = 0; v < 4; ++v) push3(P, quads[q][v]);
        for (int t = 0; t < 2; ++t) {
            for (int k = 0; k < 3; ++k) I.push_back((uint32_t)(4 * q + idx[3 * t + k]));
            for (int c = 0; c < 3; ++c) Kd.push_back(kd[q][c]);
            // bits as include/bre.h BRE_TRI_*: 1 = medium inside, 4 = flip normal, 8 = no material
            attr.push_back((iface ? 1u : 0u) | (reverse ? 4u : 0u) | (iface ? 8u : 0u));
        }
    }
    std::shared_ptr<BVHAccel> bvh = std::make_shared<BVHAccel>(prims);
    std::vector<std::shared_ptr<Light>> lights;
    lights.push_back(std::make_shared<PointLight>(Translate(Vector3f(0.f, 0.9f, 0.f)), inFog, rgbS(5.f, 4.f, 3.f)));
    Scene scene(bvh, lights);
    std::unique_ptr<Distribution1D> lightDistr = ComputeLightPowerDistribution(scene);

    std::vector<float> pp, pa, pw;
    std::vector<uint32_t> pid;
    MemoryArena arena;
    for (uint64_t pi = first; pi < first + nPaths; ++pi) {
        StreamSampler sampler(pi + 1);
        Float lightPdf;
        int lightNum = lightDistr->SampleDiscrete(sampler.Get1D(), &lightPdf);
        const std::shared_ptr<Light> &light = scene.lights[lightNum];
        Point2f uLight0 = sampler.Get2D();
        Point2f uLight1 = sampler.Get2D();
        Float uLightTime = sampler.Get1D();
        RayDifferential ray;
        Normal3f nLight;
        Float pdfPos, pdfDir;
        Spectrum Le = light->Sample_Le(uLight0, uLight1, uLightTime, &ray, &nLight, &pdfPos, &pdfDir);
        if (pdfPos == 0 || pdfDir == 0 || Le.IsBlack()) continue;
        Spectrum beta = (AbsDot(nLight, ray.d) * Le) / (lightPdf * pdfPos * pdfDir);
        if (beta.IsBlack()) continue;
        SurfaceInteraction isect;
        for (int depth = 0; depth < maxDepth; ++depth) {
            if (!scene.Intersect(ray, &isect)) break;
            MediumInteraction mi;
            if (ray.medium) beta *= ray.medium->Sample(ray, sampler, arena, &mi);
            if (beta.IsBlack()) break;
            if (mi.IsValid()) {
                const Vector3f wo = -ray.d;
                Vector3f wi;
                mi.phase->Sample_p(wo, &wi, sampler.Get2D());
                push3(pp, mi.p); pushS(pa, beta); push3(pw, wo);
                pid.push_back((uint32_t)pi);
                ray = mi.SpawnRay(wi);
                if (depth + 1 >= maxDepth) break;
                (void)sampler.Get1D();
                continue;
            }
            isect.ComputeScatteringFunctions(ray, arena, true, TransportMode::Importance);
            if (!isect.bsdf) {                       // photonbeam.cpp:300-304
                --depth;
                ray = isect.SpawnRay(ray.d);
                continue;
            }
            Vector3f wo = -ray.d, wi;
            Float pdf;
            BxDFType flags;
            Spectrum fr = isect.bsdf->Sample_f(wo, &wi, sampler.Get2D(), &pdf, BSDF_ALL, &flags);   // :313
            if (fr.IsBlack() || pdf == 0.f) break;
            Spectrum betaNew = beta * fr * AbsDot(wi, isect.shading.n) / pdf;
            ray = (RayDifferential)isect.SpawnRay(wi);
            Float q = std::max((Float)0, 1 - betaNew.y() / beta.y());                               // :321
            if (sampler.Get1D() < q) break;
            beta = betaNew / (1 - q);
        }
        arena.Reset();
    }

    // ---- camera rays over a cropped window
    Film *film = new Film(Point2i(resX, resY), Bounds2f(Point2f(0.125f, 0.25f), Point2f(1, 0.875f)),
                          std::unique_ptr<Filter>(new BoxFilter(Vector2f(0.5f, 0.5f))), 35.f,
                          "/tmp/bre_ref_surfaces.pfm", 1.f);
    Transform c2w = Inverse(LookAt(Point3f(0.3f, 0.2f, -3.5f), Point3f(0, 0.1f, 0), Vector3f(0, 1, 0)));
    AnimatedTransform ac2w(&c2w, 0.f, &c2w, 1.f);
    const Float frame = (Float)resX / (Float)resY;
    std::shared_ptr<const PerspectiveCamera> camera = std::make_shared<PerspectiveCamera>(
        ac2w, Bounds2f(Point2f(-frame, -1), Point2f(frame, 1)), 0.f, 1.f, 0.f, 1e6f, 40.f, film, &fog);
    const Bounds2i pb = film->croppedPixelBounds;
    const Point2f off(0.25f, 0.625f);
    std::vector<float> ro, rd;
    for (int y = pb.pMin.y; y < pb.pMax.y; ++y)
        for (int x = pb.pMin.x; x < pb.pMax.x; ++x) {
            CameraSample cs;
            cs.pFilm = Point2f(x + off.x, y + off.y);
            cs.pLens = Point2f(0.5f, 0.5f);
            cs.time = 0.f;
            RayDifferential ray;
            Float w = camera->GenerateRayDifferential(cs, &ray);
            SurfaceInteraction si;
            scene.Intersect(ray, &si);
            const Float len = ray.d.Length();
            const Vector3f d = ray.d / len;
            ro.push_back(ray.o.x); ro.push_back(ray.o.y); ro.push_back(ray.o.z);
            ro.push_back((w > 0 && ray.medium) ? std::min(ray.tMax * len, MaxFloat) : 0.f);
            push3(rd, d);
        }
    std::vector<float> r2cf, c2wf;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) {
        r2cf.push_back(camera->RasterToCamera.GetMatrix().m[i][j]);
        c2wf.push_back(c2w.GetMatrix().m[i][j]);
    }
    BinWriter bw(path);
    bw.f4("sigma_a", {sa[0], sa[1], sa[2]}); bw.f4("sigma_s", {ss[0], ss[1], ss[2]}); bw.f4("g", {g});
    bw.f4("light_pos", {0.f, 0.9f, 0.f}); bw.f4("light_I", {5.f, 4.f, 3.f});
    bw.f4("mesh_P", P); bw.u4("mesh_idx", I); bw.f4("mesh_Kd", Kd); bw.u4("mesh_attr", attr);
    bw.u8("range", {first, nPaths, (uint64_t)maxDepth});
    bw.f4("photon_p", pp); bw.f4("photon_alpha", pa); bw.f4("photon_wi", pw); bw.u4("path_id", pid);
    bw.f4("raster_to_camera", r2cf); bw.f4("camera_to_world", c2wf);
    bw.u4("window", {(uint32_t)pb.pMin.x, (uint32_t)pb.pMin.y, (uint32_t)(pb.pMax.x - pb.pMin.x), (uint32_t)(pb.pMax.y - pb.pMin.y)});
    bw.f4("pixel_offset", {off.x, off.y});
    bw.f4("ray_o_tmax", ro); bw.f4("ray_d", rd);
    fprintf(stderr, "surfaces: %zu photons from %llu paths, %zu camera rays\n", pid.size(),
            (unsigned long long)nPaths, ro.size() / 4);
}

// ---------------------------------------------------------------- mode: gridtr
// Pins the deterministic optical-depth table of the product (K6) at the grid sizes the benchmarks run:
// for rays {o, d, t} read from <rays.bin> (8 floats per ray: o, d, t, 0) through the density grid read
// from <density.bin> (nx*ny*nz floats, x fastest) it stores
//   Tr_mean / Tr_stderr   nSamples evaluations of the reference's GridDensityMedium::Tr (media/grid.cpp:89-121,
//                         ratio tracking + Russian roulette) of Ray(o, d, tMax = t)
//   tau_fine              the optical depth sum Density(p) dt over 65536 midpoint steps of the ray's overlap with
//                         the medium box, Density = the reference's own trilinear lookup (media/grid.cpp:45-60),
//                         accumulated in double: what ratio tracking estimates without bias
void gridtr(const char *path, const char *densPath, int nx, int ny, int nz, const char *raysPath, int nSamples,
            float sigA, float sigS, float aspect, float boxLoX) {
    std::vector<float> dens((size_t)nx * ny * nz);
    {
        FILE *f = fopen(densPath, "rb");
        if (!f || fread(dens.data(), 4, dens.size(), f) != dens.size()) { perror(densPath); exit(1); }
        fclose(f);
    }
    std::vector<float> rays;
    {
        FILE *f = fopen(raysPath, "rb");
        if (!f) { perror(raysPath); exit(1); }
        float v[8];
        while (fread(v, 4, 8, f) == 8) rays.insert(rays.end(), v, v + 8);
        fclose(f);
    }
    const size_t nRays = rays.size() / 8;
    // medium box = [boxLoX, boxLoX + aspect] x [0,1] x [0,1]   (bench.py world_to_medium)
    Transform m2w = Translate(Vector3f(boxLoX, 0.f, 0.f)) * Scale(aspect, 1.f, 1.f);
    Transform w2m = Inverse(m2w);
    GridDensityMedium med(Spectrum(sigA), Spectrum(sigS), 0.3f, nx, ny, nz, m2w, dens.data());
    std::vector<float> trMean(nRays), trErr(nRays), tauFine(nRays), w2mf;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) w2mf.push_back(w2m.GetMatrix().m[i][j]);
    std::vector<std::thread> pool;
    std::atomic<size_t> next(0);
    const unsigned nThreads = std::max(1u, std::thread::hardware_concurrency());
    for (unsigned tIdx = 0; tIdx < nThreads; ++tIdx)
        pool.emplace_back([&]() {
            for (size_t i = next++; i < nRays; i = next++) {
                const float *v = &rays[8 * i];
                Point3f o(v[0], v[1], v[2]);
                Vector3f dir(v[3], v[4], v[5]);
                const float t = v[6];
                double sum = 0, sum2 = 0;
                for (int k = 0; k < nSamples; ++k) {
                    StreamSampler smp(1000000007ull * (i + 1) + (uint64_t)k);
                    Float tr = med.Tr(Ray(o, dir, t), smp)[0];
                    sum += tr; sum2 += (double)tr * tr;
                }
                const double mean = sum / nSamples, var = std::max(0.0, sum2 / nSamples - mean * mean);
                trMean[i] = (float)mean; trErr[i] = (float)std::sqrt(var / nSamples);
                // fine integral of the reference's Density() over the overlap with the unit cube in medium space
                Ray rm = w2m(Ray(o, dir, t));
                Float t0, t1;
                double tau = 0;
                if (Bounds3f(Point3f(0, 0, 0), Point3f(1, 1, 1)).IntersectP(rm, &t0, &t1)) {
                    const int N = 65536;
                    const double dt = ((double)t1 - (double)t0) / N;
                    for (int k = 0; k < N; ++k) {
                        const double tt = (double)t0 + (k + 0.5) * dt;
                        Point3f pm((Float)(rm.o.x + rm.d.x * tt), (Float)(rm.o.y + rm.d.y * tt), (Float)(rm.o.z + rm.d.z * tt));
                        tau += (double)med.Density(pm) * dt;
                    }
                }
                tauFine[i] = (float)tau;
            }
        });
    for (auto &th : pool) th.join();
    BinWriter bw(path);
    bw.u4("dims", {(uint32_t)nx, (uint32_t)ny, (uint32_t)nz});
    bw.f4("sigma", {sigA, sigS});
    bw.f4("w2m", w2mf);
    bw.f4("rays", rays);
    bw.f4("Tr_mean", trMean); bw.f4("Tr_stderr", trErr); bw.f4("tau_fine", tauFine);
    bw.u4("nSamples", {(uint32_t)nSamples});
    fprintf(stderr, "gridtr: %zu rays, %d samples each, grid %dx%dx%d\n", nRays, nSamples, nx, ny, nz);
}

// ---------------------------------------------------------------- mode: bvhquery
// the reference's OWN photon-BVH code timed on the benchmark's photons.  Every photon becomes a
// short PhotonBeam (length 1e-3 r along its direction, radius r) so that the unmodified PhotonBeamBVH (SAH build,
// core/photonbeambvh.cpp:204-425) can hold it; camera rays walk it with the unmodified PhotonBeamBVH::Intersect
// (collect-all traversal, :685-723, incl. its vector<shared_ptr> result) on all host threads (ParallelFor,
// core/parallel.h:101-102), and every candidate runs the photon-disk predicate of the oracle.  A TIMING of
// reference code, not a parity vector: PhotonBeam::WorldBound is not conservative (core/photonbeambvh.h:60-72,
// SURVEY.md 7), so a few hits are culled -- the harness prints how many of the brute-force hits it found.
// photons.bin: n x {p.xyz, r}; rays.bin: n x {o.xyz, d.xyz, tMax, 0}.
void bvhquery(const char *photonPath, const char *rayPath, int nThreads, double maxSeconds) {
    auto slurp = [](const char *path) {
        std::vector<float> v;
        FILE *f = fopen(path, "rb");
        if (!f) { perror(path); exit(1); }
        float buf[4096];
        size_t got;
        while ((got = fread(buf, 4, 4096, f)) > 0) v.insert(v.end(), buf, buf + got);
        fclose(f);
        return v;
    };
    const std::vector<float> ph = slurp(photonPath), ry = slurp(rayPath);
    const size_t n = ph.size() / 4, nq = ry.size() / 8;
    PbrtOptions.nThreads = nThreads;
    ParallelInit();
    std::vector<std::shared_ptr<PhotonBeam>> beams;
    beams.reserve(n);
    RNG rng(7);
    for (size_t i = 0; i < n; ++i) {
        const Point3f p(ph[4 * i], ph[4 * i + 1], ph[4 * i + 2]);
        const float r = ph[4 * i + 3];
        const Vector3f w = unit_from(rng) * (0.5e-3f * r);
        beams.push_back(std::make_shared<PhotonBeam>(p - w, p + w, r, Spectrum(0.f), Spectrum(1.f)));
    }
    auto t0 = std::chrono::steady_clock::now();
    PhotonBeamBVH bvh(std::move(beams));
    const double buildS = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::atomic<uint64_t> cand(0), hits(0), done(0);
    const int chunk = 256;
    t0 = std::chrono::steady_clock::now();
    double spent = 0;
    size_t q0 = 0;
    while (spent < maxSeconds) {
        if (q0 >= nq) q0 = 0;                    // a short ray file is walked again until the time is up
        const size_t q1 = std::min(nq, q0 + (size_t)chunk * 64);
        ParallelFor([&](int64_t c) {
            uint64_t lc = 0, lh = 0;
            for (size_t q = q0 + (size_t)c * chunk; q < std::min(q1, q0 + (size_t)(c + 1) * chunk); ++q) {
                const float *v = &ry[8 * q];
                Ray ray(Point3f(v[0], v[1], v[2]), Vector3f(v[3], v[4], v[5]), v[6]);
                auto got = bvh.Intersect(ray);
                lc += got.size();
                for (auto &b : got) {
                    const Point3f p = b->start + (b->end - b->start) * 0.5f;
                    const Vector3f pv = p - ray.o;
                    const float t = Dot(pv, ray.d);
                    if (!(t > 0 && t < ray.tMax)) continue;
                    const Vector3f wv = pv - t * ray.d;
                    if (Dot(wv, wv) < b->radius * b->radius) ++lh;
                }
            }
            cand += lc; hits += lh;
        }, (int64_t)((q1 - q0 + chunk - 1) / chunk), 1);
        done += q1 - q0;
        q0 = q1;
        spent = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    }
    ParallelCleanup();
    printf("{\"photons\": %zu, \"rays\": %llu, \"threads\": %d, \"build_s\": %.3f, \"query_s\": %.3f, "
           "\"queries_per_s\": %.1f, \"candidates\": %llu, \"hits\": %llu}\n",
           n, (unsigned long long)done.load(), nThreads, buildS, spent, (double)done.load() / spent,
           (unsigned long long)cand.load(), (unsigned long long)hits.load());
}

}  // namespace

int main(int argc, char **argv) {
    PbrtOptions.quiet = true;
    PbrtOptions.nThreads = 1;   // ShootPhotons claims blocks under a mutex: one thread = one order
    if (argc >= 3 && !strcmp(argv[1], "primitives")) { primitives(argv[2]); return 0; }
    if (argc >= 6 && !strcmp(argv[1], "fogbox")) {
        fogbox(argv[2], atoi(argv[3]), atoi(argv[4]), (float)atof(argv[5]), argc >= 7 ? (uint32_t)atoi(argv[6]) : 0u,
               argc >= 8 && !strcmp(argv[7], "digest"));
        return 0;
    }
    if (argc >= 3 && !strcmp(argv[1], "halton")) { halton(argv[2]); return 0; }
    if (argc >= 6 && !strcmp(argv[1], "gridspot")) {
        gridspot(argv[2], (uint64_t)atoll(argv[3]), atoi(argv[4]), (float)atof(argv[5]));
        return 0;
    }
    if (argc >= 7 && !strcmp(argv[1], "trace")) {
        trace(argv[2], argv[3], strtoull(argv[4], 0, 10), strtoull(argv[5], 0, 10), atoi(argv[6]));
        return 0;
    }
    if (argc >= 8 && !strcmp(argv[1], "surfaces")) {
        surfaces(argv[2], strtoull(argv[3], 0, 10), strtoull(argv[4], 0, 10), atoi(argv[5]), atoi(argv[6]), atoi(argv[7]));
        return 0;
    }
    if (argc >= 6 && !strcmp(argv[1], "bvhquery")) {
        bvhquery(argv[2], argv[3], atoi(argv[4]), atof(argv[5]));
        return 0;
    }
    if (argc >= 13 && !strcmp(argv[1], "gridtr")) {
        gridtr(argv[2], argv[3], atoi(argv[4]), atoi(argv[5]), atoi(argv[6]), argv[7], atoi(argv[8]), (float)atof(argv[9]),
               (float)atof(argv[10]), (float)atof(argv[11]), (float)atof(argv[12]));
        return 0;
    }
    fprintf(stderr, "usage: ref_harness primitives out.json | fogbox out.bin nPhotons res radius [keepFirst] | "
                    "trace out.bin homogeneous|grid first nPaths maxDepth | surfaces out.bin first nPaths maxDepth resX resY | "
                    "gridtr out.bin density.bin nx ny nz rays.bin nSamples sigma_a sigma_s aspect boxLoX\n");
    return 2;
}
