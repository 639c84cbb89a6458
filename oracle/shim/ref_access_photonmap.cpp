// oracle/shim/ref_access_photonmap.cpp -- TEST INFRASTRUCTURE ONLY.
//
// `struct Photon`, `kernel()` and PhotonIntegrator::ShootPhotons are private to
// integrators/photonmap.cpp (:61-70, :170-174, :616-906) and "photonmap" is not dispatched by
// MakeIntegrator (core/api.cpp:1444-1468), so the volume-photon pass of the reference cannot be
// reached from a scene file or through a public header.  This translation unit compiles that
// reference source file WHERE IT LIES (nothing is copied) by including it -- built with
// -fno-access-control, see oracle/Makefile -- and exports accessors.  It is linked in place of
// integrators/photonmap.o.
#include "integrators/photonmap.cpp"

#include "ref_access.h"

namespace refx {

// kernel() of integrators/photonmap.cpp:170-174 for a photon at `photonP`.
float kernel(const float photonP[3], const float p[3], float maxDist2) {
    pbrt::Photon ph(pbrt::Point3f(photonP[0], photonP[1], photonP[2]), pbrt::Spectrum(0.f),
                    pbrt::Vector3f(0, 0, 1));
    return pbrt::kernel(&ph, pbrt::Point3f(p[0], p[1], p[2]), maxDist2);
}

// Runs the UNMODIFIED PhotonIntegrator::ShootPhotons (volume photons only) and copies out the
// volume photon records `Photon{p, alpha, wi}` in the order the reference stored them.
void shoot_volume_photons(const pbrt::Scene &scene, std::shared_ptr<const pbrt::Camera> camera,
                          int nVolumeWanted, int maxPhotonDepth, PhotonDump *out) {
    using namespace pbrt;
    PhotonIntegrator integ(camera, nullptr, camera->film->croppedPixelBounds,
                           /*ncaus*/ 0, /*nindir*/ 0, nVolumeWanted, /*reqphotons*/ false,
                           /*nLookup*/ 50, /*maxspecdepth*/ 5, maxPhotonDepth, /*maxdist*/ 0.1f,
                           /*finalGather*/ false, /*gatherSamples*/ 32, /*gatherAngle*/ 10.f);
    integ.ComputeLightPhotonDistrib(scene);
    int nDirectPaths = 0;
    uint32_t nShot = 0;
    std::vector<Photon> caustic, direct, indirect, volume;
    std::vector<RadiancePhoton> radiance;
    std::vector<Spectrum> rpR, rpT;
    integ.ShootPhotons(scene, caustic, direct, indirect, volume, radiance, rpR, rpT, nDirectPaths,
                       nShot);
    out->nShot = nShot;
    out->p.clear(); out->alpha.clear(); out->wi.clear();
    for (const Photon &ph : volume) {
        Float rgb[3];
        ph.alpha.ToRGB(rgb);
        for (int c = 0; c < 3; ++c) {
            out->p.push_back(ph.p[c]);
            out->alpha.push_back(rgb[c]);
            out->wi.push_back(ph.wi[c]);
        }
    }
}

}  // namespace refx
