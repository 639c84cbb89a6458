// oracle/shim/ref_glue.cpp -- TEST INFRASTRUCTURE ONLY.
// Link-time glue for the reference subset built by oracle/Makefile.  The
// reference's core/imageio.cpp needs OpenEXR/lodepng/targa and core/api.cpp
// drags in every plugin; neither is on the beam-radiance hot path, so this
// file supplies the symbols the subset still needs from them:
//   pbrt::PbrtOptions               (defined in core/api.cpp:130 upstream)
//   pbrt::line_num / current_file   (parser state read by core/error.cpp; defined in core/parser.cpp)
//   pbrt::WriteImage / ReadImage    (core/imageio.h:49-56) -- float32 PFM only
#include "pbrt.h"
#include "geometry.h"
#include "imageio.h"
#include "spectrum.h"
#include <cstdio>

namespace pbrt {

Options PbrtOptions;
int line_num = 0;
std::string current_file;

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
