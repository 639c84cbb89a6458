// oracle/shim/ref_imageio.cpp -- TEST INFRASTRUCTURE ONLY.
// Stand-in for the reference's core/imageio.cpp, which needs OpenEXR (cmake-generated config headers),
// lodepng and targa: pbrt::WriteImage / ReadImage (core/imageio.h:49-56) for float32 PFM output only.
// Linked into oracle/_ref/libpbrtref.a and into oracle/_ref/pbrt (the reference's own command line).
#include "pbrt.h"
#include "geometry.h"
#include "imageio.h"
#include "spectrum.h"
#include <cstdio>

namespace pbrt {

void WriteImage(const std::string &name, const Float *rgb,
                const Bounds2i &outputBounds, const Point2i &) {
    Vector2i res = outputBounds.Diagonal();
    FILE *fp = fopen(name.c_str(), "wb");
    if (!fp) { Error("Unable to open output PFM file \"%s\"", name.c_str()); return; }
    // PFM: bottom-to-top scanlines, negative scale = little endian.
    fprintf(fp, "PF\n%d %d\n-1.0\n", res.x, res.y);
    for (int y = res.y - 1; y >= 0; --y)
        fwrite(rgb + 3 * (size_t)y * res.x, sizeof(float), 3 * (size_t)res.x, fp);
    fclose(fp);
}

std::unique_ptr<RGBSpectrum[]> ReadImage(const std::string &name, Point2i *) {
    Error("ReadImage(\"%s\") is not available in the oracle subset build", name.c_str());
    return nullptr;
}

}  // namespace pbrt
