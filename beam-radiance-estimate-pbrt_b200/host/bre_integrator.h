// bre_integrator.h -- pbrt-side adapter of libbre.so: the drop-in Integrator plugin.
//
// Lives inside the pbrt-v3 fork (copy this file and bre_integrator.cpp to src/integrators/, see
// INTEGRATION.md).  It implements the reference's plugin surface
//     class Integrator { virtual void Render(const Scene &scene) = 0; }     core/integrator.h:53-58
//     Integrator *CreateXxxIntegrator(const ParamSet &, std::shared_ptr<const Camera>)
//                                                        (pattern: integrators/photonbeam.h:63-64)
// and replaces the body of PhotonBeamIntegrator::Render (integrators/photonbeam.cpp:328-587) by
// calls into the C ABI of include/bre.h.  All compute runs in hand-written CUDA kernels behind that
// ABI; this file only walks pbrt's scene objects, feeds plain arrays across the boundary and hands
// the resolved image to Film::SetImage / Film::WriteImage.  There is no CPU fallback: if
// bre_create fails the integrator reports the error (pbrt convention: Error(), no exceptions) and
// renders nothing.
//
// Scene-file surface: Integrator "bre" with the UNION of the reference's parameter names
//   photon count     "integer photonsperiteration" (photonbeam.cpp:596) | "integer volumephotons"
//                    (photonmap.cpp:946); -1 / absent = number of pixels (photonbeam.h:37-39)
//   gather radius    "float initialbeamradius" (photonbeam.cpp:598) | "float radius"
//                    (vsppm.cpp:667) | "float maxdist" (photonmap.cpp:953)
//   max depth        "integer maxdepth" (photonbeam.cpp:595) | "integer maxphotondepth" (photonmap.cpp:951)
//   "integer iterations"|"numiterations", "float alpha", "integer imagewritefrequency" (default 1<<31 =
//   INT_MIN: (iter+1) % it is never 0, so only the last iteration writes, photonbeam.cpp:565,598),
//   "integer startiteration" / "integer enditeration" (photonbeam.cpp:354-358,594-595: the radius schedule
//   is fast-forwarded, L is still divided by iter+1), "bool rendermedia" as in photonbeam.cpp:590-603;
//   "bool rendersurfaces" is accepted: only media radiance is gathered, surfaces stay black
//   (the reference's default true also shades surfaces, photonbeam.cpp:519-546) -- a warning says so.
//   new: "string checkpoint" [""]  film accumulators + iterations done + radius are loaded from this file
//        when startiteration > 0 and saved to it after the last iteration (bre_film_load / bre_film_save):
//        a resumed render continues the SAME estimate; the reference's startiteration restores the radius only.
//   new: "string camerasamples" ["halton"|"uniform"]  halton (default) = the reference's camera samples: pixel p of
//        iteration i looks through p + HaltonSampler(nIterations, pixelBounds) sample i (photonbeam.cpp:341,459-466),
//        computed per pixel on the device (bre_stage_camera_halton) or by pbrt's own sampler for host rays;
//        uniform = one sub-pixel offset per iteration for every pixel (a regular ray grid).
//   new: "string sharding" ["photons"|"tiles"]  several ranks (BRE_NRANKS / BRE_RANK / BRE_NCCL_ID_FILE): shard the
//        photon paths (each rank builds and queries its own 1/N of the map, bre_film_gather sums the films; default)
//        or the 16x16 image tiles (photons all-gathered); rank 0 writes the image either way.
//   new: "string photontracer" ["host"|"device"]  host = photon paths traced by pbrt's own
//        Scene/Medium/BSDF code (any scene), device = K1 kernel (matte triangle scenes with point /
//        spot lights; other scenes: media inside the absorbing world box).
//   new: "string camerarays" ["device"|"host"]  device (default) = K0 kernel for a pinhole
//        PerspectiveCamera over matte triangle scenes (bit-identical rays, no PCIe ray traffic);
//        falls back to pbrt's camera + Scene::Intersect on the host for anything else.
//   multi-GPU: one pbrt process per GPU.  Rank / world size come from BRE_RANK / BRE_NRANKS, else
//        OMPI_COMM_WORLD_RANK / _SIZE, else RANK / WORLD_SIZE; the device defaults to the local rank; the 128-byte
//        NCCL id travels through the file named by BRE_NCCL_ID_FILE (rank 0 writes it, the others wait for it).
//        Photon paths are split by index and exchanged over NVLink, every rank answers the 16x16 image tiles
//        bre_tile_owner gives it (photonbeam.cpp:345-347), the film is gathered on rank 0 and only rank 0
//        calls Film::SetImage / WriteImage (photonbeam.cpp:582-583).
//   new: "bool reusephotons" [false]  true = trace photons once, later iterations only shrink the
//        radius (photonbeam.cpp:562) and refit the LBVH instead of re-tracing and rebuilding.
#ifndef PBRT_INTEGRATORS_BRE_H
#define PBRT_INTEGRATORS_BRE_H

#include "pbrt.h"
#include "integrator.h"
#include "camera.h"
#include <string>

struct bre_ctx;

namespace pbrt {

class BreIntegrator : public Integrator {
  public:
    BreIntegrator(std::shared_ptr<const Camera> camera, int nIterations, int photonsPerIteration,
                  int maxDepth, Float initialRadius, Float alpha, int writeFrequency,
                  bool renderMedia, bool deviceTracer, bool deviceCamera, bool reusePhotons, int device,
                  int startIteration = 0, int endIteration = -1, std::string checkpoint = "",
                  std::string sharding = "photons", std::string cameraSamples = "halton")
        : camera(camera), nIterations(nIterations), photonsPerIteration(photonsPerIteration),
          maxDepth(maxDepth), initialRadius(initialRadius), alpha(alpha),
          writeFrequency(writeFrequency), renderMedia(renderMedia), deviceTracer(deviceTracer),
          deviceCamera(deviceCamera), reusePhotons(reusePhotons), device(device),
          startIteration(startIteration), endIteration(endIteration < 0 ? nIterations : endIteration),
          checkpoint(std::move(checkpoint)), sharding(std::move(sharding)), cameraSamples(std::move(cameraSamples)) {}
    void Render(const Scene &scene);

    // test hook: after Render, photons of the last iteration in reference layout (p, alpha, wi)
    std::vector<float> lastPhotons;
    uint64_t lastPathsEmitted = 0;
    bool usedDeviceCamera = false;

  private:
    bool SetupMedium(bre_ctx *ctx, const Medium *medium) const;
    bool SetupLights(bre_ctx *ctx, const Scene &scene) const;
    bool SetupSurfaces(bre_ctx *ctx, const Scene &scene, const Medium *medium, bool *tooMany) const;
    bool SetupCamera(bre_ctx *ctx, const Bounds2i &pixelBounds) const;
    uint64_t TracePhotonsHost(bre_ctx *ctx, const Scene &scene, const Distribution1D &lightDistr,
                              uint64_t firstIndex, uint64_t nPaths, Float radius);

    std::shared_ptr<const Camera> camera;
    const int nIterations, photonsPerIteration, maxDepth;
    const Float initialRadius, alpha;
    const int writeFrequency;
    const bool renderMedia, deviceTracer, deviceCamera, reusePhotons;
    const int device;
    const int startIteration, endIteration;
    const std::string checkpoint;
    const std::string sharding;
    const std::string cameraSamples;
};

Integrator *CreateBreIntegrator(const ParamSet &params, std::shared_ptr<const Camera> camera);

}  // namespace pbrt

#endif  // PBRT_INTEGRATORS_BRE_H
