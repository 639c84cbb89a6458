// bre_integrator.cpp -- pbrt-side adapter of libbre.so (see bre_integrator.h).
//
// Needs read access to a few private members of the fork's media and lights (sigma_a, sigma_s, g,
// density grid, WorldToMedium; pLight, I, cone cosines): INTEGRATION.md adds
// `friend class BreIntegrator;` to media/homogeneous.h, media/grid.h, lights/point.h, lights/spot.h.
#include "bre_integrator.h"
#include "samplers/halton.h"

#include "bre.h"   // the C ABI (include/bre.h of the B200 library)

#include "accelerators/bvh.h"
#include "cameras/perspective.h"
#include "film.h"
#include "interaction.h"
#include "lowdiscrepancy.h"
#include "parallel.h"
#include "lights/point.h"
#include "materials/matte.h"
#include "primitive.h"
#include "shapes/triangle.h"
#include "textures/constant.h"
#include "lights/spot.h"
#include "media/grid.h"
#include "media/homogeneous.h"
#include "memory.h"
#include "paramset.h"
#include "progressreporter.h"
#include "reflection.h"
#include "rng.h"
#include "sampler.h"
#include "sampling.h"
#include "scene.h"
#include "stats.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <thread>

namespace pbrt {

namespace {

// Every sample of a photon path is PCG32 of stream pathIndex+1: what the reference's
// AwesomeHaltonSampler yields once its 1000 Halton dimensions are burnt (photonbeam.cpp:386-391).
class PathSampler : public Sampler {
  public:
    explicit PathSampler(uint64_t seq) : Sampler(1), rng(seq) {}
    Float Get1D() override { return rng.UniformFloat(); }
    Point2f Get2D() override {
        Float a = rng.UniformFloat();
        Float b = rng.UniformFloat();
        return Point2f(a, b);
    }
    std::unique_ptr<Sampler> Clone(int) override { return nullptr; }
    RNG rng;
};

// ---- multi-GPU launch environment: one pbrt process per GPU ----------------------------------------
int EnvInt(const char *const names[], int n, int dflt) {
    for (int i = 0; i < n; ++i) {
        const char *v = getenv(names[i]);
        if (v && *v) return atoi(v);
    }
    return dflt;
}
struct RankInfo { int rank = 0, nRanks = 1, localRank = 0; };
RankInfo FindRank() {
    static const char *const r[] = {"BRE_RANK", "OMPI_COMM_WORLD_RANK", "RANK"};
    static const char *const n[] = {"BRE_NRANKS", "OMPI_COMM_WORLD_SIZE", "WORLD_SIZE"};
    static const char *const l[] = {"BRE_LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK", "LOCAL_RANK"};
    RankInfo ri;
    ri.nRanks = std::max(1, EnvInt(n, 3, 1));
    ri.rank = std::min(std::max(0, EnvInt(r, 3, 0)), ri.nRanks - 1);
    ri.localRank = EnvInt(l, 3, ri.rank);
    return ri;
}
// rank 0 creates the NCCL id and publishes it through a file; the others wait for the complete file
bool ShareNcclId(const RankInfo &ri, unsigned char id[128]) {
    const char *path = getenv("BRE_NCCL_ID_FILE");
    if (!path || !*path) { Error("bre: %d ranks but BRE_NCCL_ID_FILE is not set", ri.nRanks); return false; }
    if (ri.rank == 0) {
        if (bre_nccl_unique_id(id) != BRE_OK) { Error("bre: bre_nccl_unique_id failed (libnccl.so.2 missing?)"); return false; }
        const std::string tmp = std::string(path) + ".tmp";
        FILE *f = fopen(tmp.c_str(), "wb");
        if (!f || fwrite(id, 1, 128, f) != 128) { Error("bre: cannot write %s", tmp.c_str()); if (f) fclose(f); return false; }
        fclose(f);
        return rename(tmp.c_str(), path) == 0;
    }
    for (int tries = 0; tries < 6000; ++tries) {          // up to ~60 s
        FILE *f = fopen(path, "rb");
        if (f) {
            const size_t got = fread(id, 1, 128, f);
            fclose(f);
            if (got == 128) return true;
        }
        std::this_thread::sleep_for(std::chrono::milliseconds(10));
    }
    Error("bre: rank %d never saw the NCCL id file %s", ri.rank, path);
    return false;
}

bool Check(bre_ctx *ctx, int status, const char *what) {
    if (status == BRE_OK) return true;
    Error("bre: %s failed (%d): %s", what, status, bre_last_error(ctx));
    return false;
}

void ToRGB3(const Spectrum &s, float out[3]) {
    Float rgb[3];
    s.ToRGB(rgb);
    out[0] = rgb[0]; out[1] = rgb[1]; out[2] = rgb[2];
}

}  // namespace

// ---- scene -> C ABI ------------------------------------------------------------------------
bool BreIntegrator::SetupMedium(bre_ctx *ctx, const Medium *medium) const {
    if (const HomogeneousMedium *h = dynamic_cast<const HomogeneousMedium *>(medium)) {
        float sa[3], ss[3];
        ToRGB3(h->sigma_a, sa); ToRGB3(h->sigma_s, ss);
        return Check(ctx, bre_set_medium_homogeneous(ctx, sa, ss, h->g), "bre_set_medium_homogeneous");
    }
    if (const GridDensityMedium *g = dynamic_cast<const GridDensityMedium *>(medium)) {
        float sa[3], ss[3], w2m[16];
        ToRGB3(g->sigma_a, sa); ToRGB3(g->sigma_s, ss);
        const Matrix4x4 &m = g->WorldToMedium.GetMatrix();
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) w2m[4 * i + j] = m.m[i][j];
        return Check(ctx, bre_set_medium_grid(ctx, sa, ss, g->g, g->nx, g->ny, g->nz,
                                              g->density.get(), w2m), "bre_set_medium_grid");
    }
    Error("bre: the camera must be inside a \"homogeneous\" or \"heterogeneous\" medium");
    return false;
}

bool BreIntegrator::SetupLights(bre_ctx *ctx, const Scene &scene) const {
    std::vector<bre_light> lights;
    for (const std::shared_ptr<Light> &l : scene.lights) {
        bre_light b = {};
        if (const PointLight *p = dynamic_cast<const PointLight *>(l.get())) {
            b.kind = 0;
            b.pos[0] = p->pLight.x; b.pos[1] = p->pLight.y; b.pos[2] = p->pLight.z;
            ToRGB3(p->I, b.I);
        } else if (const SpotLight *s = dynamic_cast<const SpotLight *>(l.get())) {
            b.kind = 1;
            b.pos[0] = s->pLight.x; b.pos[1] = s->pLight.y; b.pos[2] = s->pLight.z;
            ToRGB3(s->I, b.I);
            Vector3f axis = Normalize(s->LightToWorld(Vector3f(0, 0, 1)));
            b.axis[0] = axis.x; b.axis[1] = axis.y; b.axis[2] = axis.z;
            b.cos_total_width = s->cosTotalWidth;
            b.cos_falloff_start = s->cosFalloffStart;
        } else {
            Error("bre: the device photon tracer supports point and spot lights only "
                  "(use \"string photontracer\" \"host\")");
            return false;
        }
        lights.push_back(b);
    }
    return Check(ctx, bre_set_lights(ctx, (int)lights.size(), lights.data()), "bre_set_lights");
}

// Scene::aggregate -> triangles for the device photon tracer and the device camera rays.  Supported:
// a BVHAccel of GeometricPrimitives over Triangle shapes with MatteMaterial (constant Kd, sigma 0)
// or no material, no area lights; medium interfaces that are either no transition or a transition
// between the camera's medium and vacuum.  Anything else returns false (the caller then keeps pbrt's
// own host code for that stage).
static const size_t kMaxDeviceTriangles = 4096;
bool BreIntegrator::SetupSurfaces(bre_ctx *ctx, const Scene &scene, const Medium *medium, bool *tooMany) const {
    *tooMany = false;
    const BVHAccel *bvh = dynamic_cast<const BVHAccel *>(scene.aggregate.get());
    if (!bvh) return false;
    std::vector<float> P, Kd;
    std::vector<uint32_t> idx;
    std::vector<uint8_t> attrs;
    for (const std::shared_ptr<Primitive> &prim : bvh->primitives) {
        const GeometricPrimitive *gp = dynamic_cast<const GeometricPrimitive *>(prim.get());
        if (!gp || gp->areaLight) return false;
        const Triangle *tri = dynamic_cast<const Triangle *>(gp->shape.get());
        if (!tri || tri->mesh->alphaMask || tri->mesh->n || tri->mesh->s || tri->mesh->uv) return false;
        uint8_t a = 0;
        float kd[3] = {-1.f, -1.f, -1.f};
        if (gp->material) {
            const MatteMaterial *mm = dynamic_cast<const MatteMaterial *>(gp->material.get());
            if (!mm || mm->bumpMap) return false;
            const ConstantTexture<Spectrum> *k = dynamic_cast<const ConstantTexture<Spectrum> *>(mm->Kd.get());
            const ConstantTexture<Float> *sg = dynamic_cast<const ConstantTexture<Float> *>(mm->sigma.get());
            if (!k || !sg || sg->value != 0) return false;
            ToRGB3(k->value.Clamp(), kd);
        } else
            a |= BRE_TRI_NO_MATERIAL;
        const MediumInterface &mi = gp->mediumInterface;
        if (mi.IsMediumTransition()) {
            if (mi.inside == medium && !mi.outside) a |= BRE_TRI_MEDIUM_INSIDE;
            else if (mi.outside == medium && !mi.inside) a |= BRE_TRI_MEDIUM_OUTSIDE;
            else return false;
        }
        if (tri->reverseOrientation ^ tri->transformSwapsHandedness) a |= BRE_TRI_FLIP_NORMAL;
        for (int v = 0; v < 3; ++v) {
            const Point3f &p = tri->mesh->p[tri->v[v]];
            idx.push_back((uint32_t)(P.size() / 3));
            P.push_back(p.x); P.push_back(p.y); P.push_back(p.z);
        }
        Kd.push_back(kd[0]); Kd.push_back(kd[1]); Kd.push_back(kd[2]);
        attrs.push_back(a);
    }
    if (attrs.empty()) return false;
    // K0 / K1 test the triangles by brute force (pixels x triangles, paths x triangles): fine for the boxes and blockers
    // of this path's scenes, a hang for a real mesh.  Beyond the cap the host camera / host tracer of pbrt take over.
    if (attrs.size() > kMaxDeviceTriangles) {
        Warning("bre: %d triangles exceed the device paths' brute-force cap (%d): camera rays and photon paths are "
                "generated by pbrt's host code", (int)attrs.size(), (int)kMaxDeviceTriangles);
        *tooMany = true;
        return false;
    }
    return Check(ctx, bre_set_triangles(ctx, (uint32_t)attrs.size(), P.data(), idx.data(), Kd.data()), "bre_set_triangles") &&
           Check(ctx, bre_set_triangle_attrs(ctx, (uint32_t)attrs.size(), attrs.data()), "bre_set_triangle_attrs");
}

// Pinhole PerspectiveCamera -> bre_set_camera (rays are then generated by the K0 kernel).
bool BreIntegrator::SetupCamera(bre_ctx *ctx, const Bounds2i &pixelBounds) const {
    const PerspectiveCamera *pc = dynamic_cast<const PerspectiveCamera *>(camera.get());
    if (!pc || pc->lensRadius > 0 || pc->CameraToWorld.actuallyAnimated) return false;
    bre_camera cam = {};
    const Matrix4x4 &r2c = pc->RasterToCamera.GetMatrix();
    const Matrix4x4 &c2w = pc->CameraToWorld.startTransform->GetMatrix();
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            cam.raster_to_camera[4 * i + j] = r2c.m[i][j];
            cam.camera_to_world[4 * i + j] = c2w.m[i][j];
        }
    cam.x0 = pixelBounds.pMin.x; cam.y0 = pixelBounds.pMin.y;
    cam.width = pixelBounds.pMax.x - pixelBounds.pMin.x;
    cam.height = pixelBounds.pMax.y - pixelBounds.pMin.y;
    cam.in_medium = camera->medium != nullptr;
    return bre_set_camera(ctx, &cam) == BRE_OK;
}

// ---- host photon tracer ----------------------------------------------------------------------
// Photon paths walked by pbrt's own Scene / Medium / BSDF code; the volume photons they deposit are
// uploaded through bre_upload_photons.  Deposit rule: Photon(mi.p, beta, -ray.d) at every medium
// scattering event (integrators/photonmap.cpp:696-724, without its surface-cosine line :718);
// surface bounce and Russian roulette as integrators/photonbeam.cpp:296-323.
uint64_t BreIntegrator::TracePhotonsHost(bre_ctx *ctx, const Scene &scene,
                                         const Distribution1D &lightDistr, uint64_t firstIndex,
                                         uint64_t nPaths, Float radius) {
    std::vector<float> pp, pa, pw;
    MemoryArena arena;
    for (uint64_t pi = firstIndex; pi < firstIndex + nPaths; ++pi) {
        PathSampler sampler(pi + 1);
        Float lightPdf;
        int lightNum = lightDistr.SampleDiscrete(sampler.Get1D(), &lightPdf);
        const std::shared_ptr<Light> &light = scene.lights[lightNum];
        Point2f uLight0 = sampler.Get2D();
        Point2f uLight1 = sampler.Get2D();
        Float uLightTime = Lerp(sampler.Get1D(), camera->shutterOpen, camera->shutterClose);
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
                float a[3];
                ToRGB3(beta, a);
                pp.push_back(mi.p.x); pp.push_back(mi.p.y); pp.push_back(mi.p.z);
                pa.push_back(a[0]); pa.push_back(a[1]); pa.push_back(a[2]);
                pw.push_back(wo.x); pw.push_back(wo.y); pw.push_back(wo.z);
                ray = mi.SpawnRay(wi);
                if (depth + 1 >= maxDepth) break;
                (void)sampler.Get1D();   // roulette draw with continue probability 1 (photonmap.cpp:799-804)
                continue;
            }
            isect.ComputeScatteringFunctions(ray, arena, true, TransportMode::Importance);
            if (!isect.bsdf) {           // medium boundary without a material: photonbeam.cpp:300-304
                --depth;
                ray = isect.SpawnRay(ray.d);
                continue;
            }
            Vector3f wo = -ray.d, wi;
            Float pdf;
            BxDFType flags;
            Spectrum fr = isect.bsdf->Sample_f(wo, &wi, sampler.Get2D(), &pdf, BSDF_ALL, &flags);
            if (fr.IsBlack() || pdf == 0.f) break;
            Spectrum betaNew = beta * fr * AbsDot(wi, isect.shading.n) / pdf;
            ray = (RayDifferential)isect.SpawnRay(wi);
            Float q = std::max((Float)0, 1 - betaNew.y() / beta.y());
            if (sampler.Get1D() < q) break;
            beta = betaNew / (1 - q);
        }
        arena.Reset();
    }
    const uint64_t n = pp.size() / 3;
    if (n && !Check(ctx, bre_upload_photons(ctx, n, pp.data(), pa.data(), pw.data(), radius, nullptr, 1),
                    "bre_upload_photons"))
        return 0;
    lastPhotons.clear();
    lastPhotons.reserve(9 * n);
    for (uint64_t i = 0; i < n; ++i) {
        for (int k = 0; k < 3; ++k) lastPhotons.push_back(pp[3 * i + k]);
        for (int k = 0; k < 3; ++k) lastPhotons.push_back(pa[3 * i + k]);
        for (int k = 0; k < 3; ++k) lastPhotons.push_back(pw[3 * i + k]);
    }
    return n;
}

// ---- Render: replaces PhotonBeamIntegrator::Render (integrators/photonbeam.cpp:328-587) ---------
void BreIntegrator::Render(const Scene &scene) {
    ProfilePhase p(Prof::IntegratorRender);
    const Medium *medium = camera->medium;
    if (!medium || !renderMedia) {
        Error("bre: nothing to render (camera outside any medium, or rendermedia false)");
        return;
    }
    const RankInfo ri = FindRank();
    unsigned char ncclId[128];
    if (ri.nRanks > 1 && !ShareNcclId(ri, ncclId)) return;
    bre_ctx *ctx = nullptr;
    bre_config cfg = {};
    cfg.device = device >= 0 ? device : ri.localRank;
    cfg.n_ranks = ri.nRanks; cfg.rank = ri.rank;
    cfg.nccl_unique_id = ri.nRanks > 1 ? ncclId : nullptr;
    const bool sharded = ri.nRanks > 1;
    int st = bre_create(&ctx, &cfg);
    if (st != BRE_OK) {
        Error("bre: bre_create failed (%d): a CUDA device is required, there is no CPU fallback", st);
        return;
    }
    if (!SetupMedium(ctx, medium)) { bre_destroy(ctx); return; }
    std::unique_ptr<Distribution1D> lightDistr = ComputeLightPowerDistribution(scene);
    if (!lightDistr) { Error("bre: the scene has no lights"); bre_destroy(ctx); return; }
    // surfaces on the device: needed by the device photon tracer (else: absorbing world box) and by
    // the device camera rays (else: rays generated by pbrt's camera on the host)
    bool tooManyTriangles = false;
    const bool haveSurfaces = (this->deviceTracer || deviceCamera) && SetupSurfaces(ctx, scene, medium, &tooManyTriangles);
    // (shadows the member from here on: a mesh beyond the brute-force cap is traced by pbrt's host code)
    const bool deviceTracer = this->deviceTracer && !tooManyTriangles;
    if (deviceTracer) {
        const Bounds3f wb = scene.WorldBound();
        float lo[3] = {wb.pMin.x, wb.pMin.y, wb.pMin.z}, hi[3] = {wb.pMax.x, wb.pMax.y, wb.pMax.z};
        if (!haveSurfaces)
            Warning("bre: scene surfaces are not all matte triangles; the device photon tracer treats "
                    "the world bound as an absorbing box");
        if (!SetupLights(ctx, scene) || !Check(ctx, bre_set_bounds(ctx, lo, hi), "bre_set_bounds")) {
            bre_destroy(ctx);
            return;
        }
        for (const std::shared_ptr<Light> &l : scene.lights)
            if (l->mediumInterface.inside != medium) bre_set_lights_in_medium(ctx, 0);
    }
    const Bounds2i pixelBounds = camera->film->croppedPixelBounds;
    const bool cameraOnDevice = deviceCamera && haveSurfaces && SetupCamera(ctx, pixelBounds);
    usedDeviceCamera = cameraOnDevice;
    const Vector2i ext = pixelBounds.Diagonal();
    const uint32_t nPixels = (uint32_t)(ext.x * ext.y);
    const uint64_t nPaths = photonsPerIteration > 0 ? (uint64_t)photonsPerIteration : nPixels;
    std::vector<float> mean(3 * (size_t)nPixels);
    // host-generated rays live in page-locked memory so that bre_stage_rays overlaps the photon pass
    float *rayO = nullptr, *rayD = nullptr;
    if (!cameraOnDevice) {
        rayO = (float *)bre_alloc_pinned(sizeof(float) * 4 * (size_t)nPixels);
        rayD = (float *)bre_alloc_pinned(sizeof(float) * 3 * (size_t)nPixels);
        if (!rayO || !rayD) { Error("bre: cannot allocate page-locked ray buffers"); bre_destroy(ctx); return; }
    }
    std::unique_ptr<Spectrum[]> image(new Spectrum[nPixels]);
    if (!Check(ctx, bre_film_create(ctx, nPixels), "bre_film_create")) { bre_destroy(ctx); return; }
    // the camera pass visits the image tile by tile (photonbeam.cpp:345-347,444-477): tell the library the layout
    if (!Check(ctx, bre_set_ray_grid(ctx, (uint32_t)ext.x, (uint32_t)ext.y), "bre_set_ray_grid")) { bre_destroy(ctx); return; }
    // several ranks: "photons" (default) -- every rank keeps the photon paths it traces, builds and queries its own
    // 1/N of the map, and the partial films are summed on rank 0; "tiles" -- photons all-gathered, image tiles split
    const bool byPhoton = sharded && sharding != "tiles";
    if (byPhoton) bre_set_photon_sharding(ctx, 1);
    else if (sharded) bre_set_tile_sharding(ctx, 1);

    // radius schedule fast-forwarded to startiteration (photonbeam.cpp:354-356)
    Float radius = initialRadius;
    for (int i = 0; i < startIteration; ++i) radius *= (i + alpha) / (i + 1);
    if (startIteration > 0 && !checkpoint.empty()) {
        // resume the SAME estimate: accumulators of the iterations already done (the reference restores the radius only)
        uint64_t done = 0;
        float rSaved = 0;
        if (bre_film_load(ctx, checkpoint.c_str(), &done, &rSaved) == BRE_OK) {
            if ((int)done != startIteration)
                Warning("bre: checkpoint %s holds %d iterations, startiteration is %d", checkpoint.c_str(), (int)done, startIteration);
        } else
            Warning("bre: cannot load checkpoint %s (%s): starting from an empty film", checkpoint.c_str(), bre_last_error(ctx));
    }
    const bool haltonSamples = cameraSamples != "uniform";
    HaltonSampler haltonSampler(nIterations, pixelBounds);          // photonbeam.cpp:341
    ProgressReporter progress(std::max(0, endIteration - startIteration), "Rendering (bre)");
    for (int iter = startIteration; iter < endIteration; ++iter) {
        // ---- camera samples.  "halton" (default): what the reference draws -- pixel p of iteration iter looks through
        // p + the first two dimensions of HaltonSampler(nIterations, pixelBounds) after StartPixel(p),
        // SetSampleNumber(iter) (photonbeam.cpp:341,459-466).  "uniform": one (2,3) Halton offset per iteration for
        // every pixel (pixel centres in iteration 0): a regular ray grid.
        const Point2f off = iter == 0 ? Point2f(0.5f, 0.5f)
                                      : Point2f(RadicalInverse(0, iter), RadicalInverse(1, iter));
        // ---- camera rays first (photonbeam.cpp:444-481): they do not depend on the photons, so the
        // host generates them now and the upload overlaps the photon pass and the LBVH build
        if (!cameraOnDevice) {
            ParallelFor([&](int64_t y) {
                std::unique_ptr<Sampler> rowSampler = haltonSamples ? haltonSampler.Clone((int)y) : nullptr;
                for (int x = 0; x < ext.x; ++x) {
                    const size_t q = (size_t)y * ext.x + x;
                    CameraSample cs;
                    if (haltonSamples) {
                        const Point2i pPixel(pixelBounds.pMin.x + x, pixelBounds.pMin.y + (int)y);
                        rowSampler->StartPixel(pPixel);
                        rowSampler->SetSampleNumber(iter);
                        cs = rowSampler->GetCameraSample(pPixel);    // pFilm, time, pLens as the reference's sampler gives them
                    } else {
                        cs.pFilm = Point2f(pixelBounds.pMin.x + x + off.x, pixelBounds.pMin.y + y + off.y);
                        cs.pLens = Point2f(0.5f, 0.5f);
                        cs.time = camera->shutterOpen;
                    }
                    RayDifferential ray;
                    Float w = camera->GenerateRayDifferential(cs, &ray);
                    SurfaceInteraction isect;
                    scene.Intersect(ray, &isect);          // shortens ray.tMax (photonbeam.cpp:481)
                    const Float len = ray.d.Length();
                    const Vector3f d = ray.d / len;
                    rayO[4 * q + 0] = ray.o.x; rayO[4 * q + 1] = ray.o.y; rayO[4 * q + 2] = ray.o.z;
                    rayO[4 * q + 3] = (w > 0 && ray.medium) ? std::min(ray.tMax * len, MaxFloat) : 0.f;
                    rayD[3 * q + 0] = d.x; rayD[3 * q + 1] = d.y; rayD[3 * q + 2] = d.z;
                }
            }, ext.y, 4);
            if (!Check(ctx, bre_stage_rays(ctx, nPixels, rayO, rayD), "bre_stage_rays")) break;
        } else if (!(haltonSamples ? Check(ctx, bre_stage_camera_halton(ctx, iter), "bre_stage_camera_halton")
                                   : Check(ctx, bre_stage_camera(ctx, off.x, off.y), "bre_stage_camera")))   // K0 + K6 under the photon pass
            break;
        // ---- photon pass (photonbeam.cpp:365-437): path indices iter*nPaths .. as :386-388
        // "reusephotons": the photon map of iteration 0 is kept and only its radius shrinks, so
        // bre_build degenerates to a refit of the existing LBVH
        const bool retrace = iter == 0 || !reusePhotons;
        bool ok = retrace ? Check(ctx, bre_clear_photons(ctx), "bre_clear_photons")
                          : Check(ctx, bre_set_radius(ctx, radius), "bre_set_radius");
        uint64_t stored = 0;
        if (!retrace) {
        } else if (ok && deviceTracer && sharded && !byPhoton && iter == 0) {
            // fused: every rank traces its slice of [0, nPaths) and pushes it into everyone's store over NVLink.
            // (The collective traces path indices from 0; later iterations of a sharded render use the split below.)
            ok = Check(ctx, bre_trace_exchange_photons(ctx, nPaths, maxDepth, radius, &stored), "bre_trace_exchange_photons");
        } else if (ok) {
            uint64_t first = 0, count = nPaths;
            bre_shard_range(nPaths, ri.nRanks, ri.rank, &first, &count);
            if (deviceTracer)
                ok = Check(ctx, bre_trace_photons(ctx, (uint64_t)iter * nPaths + first, count, maxDepth, radius, &stored),
                           "bre_trace_photons");
            else
                stored = TracePhotonsHost(ctx, scene, *lightDistr, (uint64_t)iter * nPaths + first, count, radius);
            if (ok && sharded && !byPhoton) ok = Check(ctx, bre_allgather_photons(ctx), "bre_allgather_photons");
        }
        lastPathsEmitted = nPaths;
        // ---- photon structure (photonbeam.cpp:438)
        ok = ok && Check(ctx, bre_build(ctx), "bre_build");
        // ---- camera pass (photonbeam.cpp:444-557): one query segment per pixel
        if (ok) {
            if (cameraOnDevice && haltonSamples)
                ok = Check(ctx, bre_query_camera_halton_to_film(ctx, iter, (float)nPaths), "bre_query_camera_halton_to_film");
            else if (cameraOnDevice)
                ok = Check(ctx, bre_query_camera_to_film(ctx, off.x, off.y, (float)nPaths), "bre_query_camera_to_film");
            else
                ok = Check(ctx, bre_query_staged_to_film(ctx, nullptr, (float)nPaths), "bre_query_staged_to_film");
        }
        if (!ok) break;
        // ---- progressive radius (photonbeam.cpp:562)
        radius *= (iter + alpha) / (iter + 1);
        // ---- film (photonbeam.cpp:565-584)
        // writeFrequency is the reference's: its default 1 << 31 (INT_MIN) never divides iter + 1, so only the last
        // iteration writes; 0 would divide by zero there and is treated as "never" here
        if (iter + 1 == endIteration || (writeFrequency != 0 && ((iter + 1) % writeFrequency) == 0)) {
            // L = Ld / (iter + 1), photonbeam.cpp:578; sharded: the ranks' tiles are gathered on rank 0 (collective)
            if (!Check(ctx, bre_film_gather(ctx, iter, 1.f, ri.rank == 0 ? mean.data() : nullptr, 1), "bre_film_gather")) break;
            if (ri.rank == 0) {
                for (uint32_t i = 0; i < nPixels; ++i) {
                    Float rgb[3] = {mean[3 * i], mean[3 * i + 1], mean[3 * i + 2]};
                    image[i] = Spectrum::FromRGB(rgb);
                }
                camera->film->SetImage(image.get());
                camera->film->WriteImage();
            }
        }
        if (iter + 1 == endIteration && !checkpoint.empty() && !sharded)
            Check(ctx, bre_film_save(ctx, checkpoint.c_str(), (uint64_t)endIteration, (float)radius), "bre_film_save");
        progress.Update();
        (void)stored;
    }
    progress.Done();
    bre_destroy(ctx);
    bre_free_pinned(rayO);
    bre_free_pinned(rayD);
}

Integrator *CreateBreIntegrator(const ParamSet &params, std::shared_ptr<const Camera> camera) {
    int nIterations = params.FindOneInt("iterations", params.FindOneInt("numiterations", 64));
    int maxDepth = params.FindOneInt("maxdepth", params.FindOneInt("maxphotondepth", 5));
    int photonsPerIter = params.FindOneInt("photonsperiteration", params.FindOneInt("volumephotons", -1));
    int writeFreq = params.FindOneInt("imagewritefrequency", 1 << 31);
    int startIteration = params.FindOneInt("startiteration", 0);
    int endIteration = params.FindOneInt("enditeration", nIterations);
    if (!params.FindOneBool("rendersurfaces", true))
        ;   // what this integrator does
    else
        Warning("bre: \"rendersurfaces\" true is not supported: only in-scattered media radiance is gathered, "
                "surfaces stay black (set \"bool rendersurfaces\" [\"false\"] to silence this)");
    Float radius = params.FindOneFloat("initialbeamradius",
                                       params.FindOneFloat("radius", params.FindOneFloat("maxdist", 1.f)));
    Float alpha = params.FindOneFloat("alpha", 0.5f);
    if (PbrtOptions.quickRender) {
        nIterations = std::max(1, nIterations / 16);
        endIteration = std::min(endIteration, nIterations);
        startIteration = std::min(startIteration, endIteration);
    }
    bool renderMedia = params.FindOneBool("rendermedia", true);
    std::string tracer = params.FindOneString("photontracer", "host");
    int device = params.FindOneInt("device", -1);    // -1: the process's local rank (0 for a single process)
    std::string camrays = params.FindOneString("camerarays", "device");
    if ((tracer != "host" && tracer != "device") || (camrays != "host" && camrays != "device")) {
        Error("bre: photontracer and camerarays must be \"host\" or \"device\"");
        return nullptr;
    }
    if (startIteration < 0 || endIteration > nIterations || startIteration > endIteration) {
        Error("bre: need 0 <= startiteration <= enditeration <= iterations");
        return nullptr;
    }
    return new BreIntegrator(camera, nIterations, photonsPerIter, maxDepth, radius, alpha,
                             writeFreq, renderMedia, tracer == "device", camrays == "device",
                             params.FindOneBool("reusephotons", false), device, startIteration, endIteration,
                             params.FindOneString("checkpoint", ""), params.FindOneString("sharding", "photons"),
                             params.FindOneString("camerasamples", "halton"));
}

}  // namespace pbrt
