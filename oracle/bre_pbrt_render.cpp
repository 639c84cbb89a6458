// oracle/bre_pbrt_render.cpp -- TEST INFRASTRUCTURE ONLY.
//
// Drop-in check of the pbrt-side adapter (beam-radiance-estimate-pbrt_b200/host/bre_integrator.cpp):
// the fog-box scene of SURVEY.md Appendix A is built with the UNMODIFIED reference classes, the
// integrator is created the way core/api.cpp:1444-1468 would (CreateBreIntegrator(ParamSet, camera)),
// Render() runs -- photon paths by pbrt host code or by the K1 kernel, everything else on the GPU
// through libbre.so -- and the reference's Film writes the image (PFM through oracle/shim).
//   bre_pbrt_render <out.pfm> <dump.bin> <host|device> <res> <nPaths> <radius> [iterations [camerarays host|device [key=value ...]]]
// dump.bin (BREF1 container, see ref_harness.cpp) holds the photons of the last iteration and the
// camera rays, so that tests can evaluate the same estimate with the CPU oracle.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "pbrt.h"
#include "accelerators/bvh.h"
#include "cameras/perspective.h"
#include "film.h"
#include "filters/box.h"
#include "lights/point.h"
#include "materials/matte.h"
#include "media/homogeneous.h"
#include "paramset.h"
#include "samplers/halton.h"
#include "parallel.h"
#include "scene.h"
#include "shapes/triangle.h"
#include "textures/constant.h"
#include "transform.h"

#include "bre_integrator.h"

using namespace pbrt;

static void add(FILE *fp, const char *name, const char *dtype, const void *data, uint64_t count, size_t elem) {
    char n[24] = {0}, t[8] = {0};
    strncpy(n, name, 23); strncpy(t, dtype, 7);
    fwrite(n, 1, 24, fp); fwrite(t, 1, 8, fp); fwrite(&count, 8, 1, fp);
    if (count) fwrite(data, elem, count, fp);
}

int main(int argc, char **argv) {
    if (argc < 7) {
        fprintf(stderr, "usage: bre_pbrt_render out.pfm dump.bin host|device res nPaths radius [iterations]\n");
        return 2;
    }
    const char *outPfm = argv[1], *dumpPath = argv[2], *tracer = argv[3];
    const int res = atoi(argv[4]), nPaths = atoi(argv[5]);
    const float radius = (float)atof(argv[6]);
    const int iterations = argc > 7 ? atoi(argv[7]) : 1;
    const char *camrays = argc > 8 ? argv[8] : "device";
    PbrtOptions.quiet = true;
    PbrtOptions.nThreads = 0;
    ParallelInit();

    const float sa = 0.05f, ss = 0.4f, g = 0.3f;
    HomogeneousMedium fog{Spectrum(sa), Spectrum(ss), g};
    MediumInterface inFog(&fog, &fog);
    static Transform id;
    const Point3f L(-4.f, -1.f, -4.f), H(4.f, 2.f, 4.f);
    const Point3f quads[6][4] = {
        {{L.x, L.y, L.z}, {H.x, L.y, L.z}, {H.x, L.y, H.z}, {L.x, L.y, H.z}},
        {{L.x, H.y, L.z}, {L.x, H.y, H.z}, {H.x, H.y, H.z}, {H.x, H.y, L.z}},
        {{L.x, L.y, H.z}, {H.x, L.y, H.z}, {H.x, H.y, H.z}, {L.x, H.y, H.z}},
        {{L.x, L.y, L.z}, {L.x, H.y, L.z}, {H.x, H.y, L.z}, {H.x, L.y, L.z}},
        {{L.x, L.y, L.z}, {L.x, L.y, H.z}, {L.x, H.y, H.z}, {L.x, H.y, L.z}},
        {{H.x, L.y, L.z}, {H.x, H.y, L.z}, {H.x, H.y, H.z}, {H.x, L.y, H.z}}};
    const int idx[6] = {0, 1, 2, 0, 2, 3};
    std::shared_ptr<Texture<Spectrum>> Kd = std::make_shared<ConstantTexture<Spectrum>>(Spectrum(0.6f));
    std::shared_ptr<Texture<Float>> sig = std::make_shared<ConstantTexture<Float>>(0.f);
    std::shared_ptr<Material> mat = std::make_shared<MatteMaterial>(Kd, sig, nullptr);
    std::vector<std::shared_ptr<Primitive>> prims;
    for (int q = 0; q < 6; ++q)
        for (auto &t : CreateTriangleMesh(&id, &id, false, 2, idx, 4, quads[q], nullptr, nullptr, nullptr, nullptr, nullptr))
            prims.push_back(std::make_shared<GeometricPrimitive>(t, mat, nullptr, inFog));
    std::shared_ptr<BVHAccel> bvh = std::make_shared<BVHAccel>(prims);
    std::vector<std::shared_ptr<Light>> lights;
    lights.push_back(std::make_shared<PointLight>(Translate(Vector3f(0.f, 0.9f, 0.f)), inFog, Spectrum(5.f)));
    Scene scene(bvh, lights);

    Film *film = new Film(Point2i(res, res), Bounds2f(Point2f(0, 0), Point2f(1, 1)),
                          std::unique_ptr<Filter>(new BoxFilter(Vector2f(0.5f, 0.5f))), 35.f, outPfm, 1.f);
    Transform c2w = Inverse(LookAt(Point3f(0, 0, -3.5f), Point3f(0, 0, 0), Vector3f(0, 1, 0)));
    AnimatedTransform ac2w(&c2w, 0.f, &c2w, 1.f);
    std::shared_ptr<const Camera> camera = std::make_shared<PerspectiveCamera>(
        ac2w, Bounds2f(Point2f(-1, -1), Point2f(1, 1)), 0.f, 1.f, 0.f, 1e6f, 40.f, film, &fog);

    // Integrator "bre" "integer photonsperiteration" [nPaths] "float initialbeamradius" [radius] ...
    ParamSet ps;
    auto addInt = [&](const char *n, int v) { std::unique_ptr<int[]> a(new int[1]); a[0] = v; ps.AddInt(n, std::move(a), 1); };
    addInt("iterations", iterations); addInt("photonsperiteration", nPaths); addInt("maxdepth", 5);
    { std::unique_ptr<Float[]> a(new Float[1]); a[0] = radius; ps.AddFloat("initialbeamradius", std::move(a), 1); }
    { std::unique_ptr<std::string[]> a(new std::string[1]); a[0] = tracer; ps.AddString("photontracer", std::move(a), 1); }
    { std::unique_ptr<std::string[]> a(new std::string[1]); a[0] = camrays; ps.AddString("camerarays", std::move(a), 1); }
    // trailing key=value arguments: further integer / string parameters of the scene-file surface
    // (startiteration=2 enditeration=4 imagewritefrequency=1 checkpoint=/tmp/film.bin rendersurfaces=false)
    bool uniformSamples = false;
    for (int i = 9; i < argc; ++i) {
        const char *eq = strchr(argv[i], '=');
        if (!eq) continue;
        const std::string key(argv[i], eq - argv[i]), val(eq + 1);
        if (key == "camerasamples") uniformSamples = val == "uniform";
        if (key == "checkpoint" || key == "sharding" || key == "camerasamples") { std::unique_ptr<std::string[]> a(new std::string[1]); a[0] = val; ps.AddString(key, std::move(a), 1); }
        else if (key == "rendersurfaces") { std::unique_ptr<bool[]> a(new bool[1]); a[0] = val == "true"; ps.AddBool(key, std::move(a), 1); }
        else addInt(key.c_str(), atoi(val.c_str()));
    }
    std::unique_ptr<Integrator> integ(CreateBreIntegrator(ps, camera));
    if (!integ) return 1;
    integ->Render(scene);

    // dump what the oracle needs to evaluate the same estimate
    BreIntegrator *bi = static_cast<BreIntegrator *>(integ.get());
    // (the rays of ITERATION 0: pixel centres with "camerasamples uniform", else sample 0 of the reference's HaltonSampler)
    std::vector<float> ro, rd;
    HaltonSampler hsamp(1, camera->film->croppedPixelBounds);
    for (int y = 0; y < res; ++y)
        for (int x = 0; x < res; ++x) {
            CameraSample cs;
            cs.pFilm = Point2f(x + 0.5f, y + 0.5f); cs.pLens = Point2f(0.5f, 0.5f); cs.time = 0.f;
            if (!uniformSamples) {
                hsamp.StartPixel(Point2i(x, y));
                hsamp.SetSampleNumber(0);
                cs = hsamp.GetCameraSample(Point2i(x, y));
            }
            RayDifferential ray;
            camera->GenerateRayDifferential(cs, &ray);
            SurfaceInteraction isect;
            scene.Intersect(ray, &isect);
            ro.push_back(ray.o.x); ro.push_back(ray.o.y); ro.push_back(ray.o.z); ro.push_back(ray.tMax);
            rd.push_back(ray.d.x); rd.push_back(ray.d.y); rd.push_back(ray.d.z);
        }
    FILE *fp = fopen(dumpPath, "wb");
    if (!fp) { perror(dumpPath); return 1; }
    fwrite("BREF1\n", 1, 6, fp);
    const float med[7] = {sa, sa, sa, ss, ss, ss, g}, rad = radius, np = (float)bi->lastPathsEmitted;
    add(fp, "medium", "f4", med, 7, 4); add(fp, "radius", "f4", &rad, 1, 4); add(fp, "n_paths", "f4", &np, 1, 4);
    const uint32_t devCam = bi->usedDeviceCamera ? 1u : 0u;
    add(fp, "used_device_camera", "u4", &devCam, 1, 4);
    add(fp, "photons9", "f4", bi->lastPhotons.data(), bi->lastPhotons.size(), 4);
    add(fp, "ray_o_tmax", "f4", ro.data(), ro.size(), 4); add(fp, "ray_d", "f4", rd.data(), rd.size(), 4);
    fclose(fp);
    ParallelCleanup();
    return 0;
}
