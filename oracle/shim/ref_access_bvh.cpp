// oracle/shim/ref_access_bvh.cpp -- TEST INFRASTRUCTURE ONLY.
//
// The Morton-code helpers of the reference's photon BVH are TU-local (`inline` / `static` in
// core/photonbeambvh.cpp:109-182), so they cannot be linked against.  This translation unit
// compiles that reference source file WHERE IT LIES (nothing is copied) by including it, and
// exports thin accessors to the TU-local pieces.  oracle/Makefile links this object in place of
// core/photonbeambvh.o (it defines every symbol that object would).
#include "core/photonbeambvh.cpp"

namespace refx {

// EncodeMorton3 (core/photonbeambvh.cpp:134-139) of an already scaled offset vector.
uint32_t encode_morton3(float x, float y, float z) {
    return pbrt::EncodeMorton3(pbrt::Vector3f(x, y, z));
}

// The code HLBVH2Build gives a primitive centroid (core/photonbeambvh.cpp:438-445):
// Bounds3::Offset (core/geometry.h:801-807) * 2^10, then EncodeMorton3.
uint32_t morton_of_point(const float p[3], const float lo[3], const float hi[3]) {
    PBRT_CONSTEXPR int mortonBits = 10;
    PBRT_CONSTEXPR int mortonScale = 1 << mortonBits;
    pbrt::Bounds3f bounds(pbrt::Point3f(lo[0], lo[1], lo[2]), pbrt::Point3f(hi[0], hi[1], hi[2]));
    pbrt::Vector3f centroidOffset = bounds.Offset(pbrt::Point3f(p[0], p[1], p[2]));
    return pbrt::EncodeMorton3(centroidOffset * mortonScale);
}

// RadixSort (core/photonbeambvh.cpp:141-182) on (primitiveIndex, mortonCode) records, in place.
void radix_sort(int n, int *primitiveIndex, uint32_t *mortonCode) {
    std::vector<pbrt::MortonPhotonBeam> v(n);
    for (int i = 0; i < n; ++i) {
        v[i].primitiveIndex = primitiveIndex[i];
        v[i].mortonCode = mortonCode[i];
    }
    pbrt::RadixSort(&v);
    for (int i = 0; i < n; ++i) {
        primitiveIndex[i] = v[i].primitiveIndex;
        mortonCode[i] = v[i].mortonCode;
    }
}

}  // namespace refx
