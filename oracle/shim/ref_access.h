// oracle/shim/ref_access.h -- TEST INFRASTRUCTURE ONLY.
// Accessors to TU-local pieces of the reference (see ref_access_bvh.cpp, ref_access_photonmap.cpp).
#pragma once
#include <cstdint>
#include <memory>
#include <vector>

namespace pbrt {
class Scene;
class Camera;
}

namespace refx {

uint32_t encode_morton3(float x, float y, float z);
uint32_t morton_of_point(const float p[3], const float lo[3], const float hi[3]);
void radix_sort(int n, int *primitiveIndex, uint32_t *mortonCode);

float kernel(const float photonP[3], const float p[3], float maxDist2);

struct PhotonDump {
    std::vector<float> p, alpha, wi;   // n*3 each, reference storage order
    uint32_t nShot = 0;                // photon paths emitted (the estimator's normaliser)
};
void shoot_volume_photons(const pbrt::Scene &scene, std::shared_ptr<const pbrt::Camera> camera,
                          int nVolumeWanted, int maxPhotonDepth, PhotonDump *out);

}  // namespace refx
