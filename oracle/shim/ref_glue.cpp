// oracle/shim/ref_glue.cpp -- TEST INFRASTRUCTURE ONLY.
// Link-time glue for the reference subset built by oracle/Makefile.  The
// reference's core/imageio.cpp needs OpenEXR/lodepng/targa and core/api.cpp
// drags in every plugin; neither is on the beam-radiance hot path, so this
// file supplies the symbols the subset still needs from them:
//   pbrt::PbrtOptions               (defined in core/api.cpp:130 upstream)
//   pbrt::line_num / current_file   (parser state read by core/error.cpp; defined in core/parser.cpp)
// (pbrt::WriteImage / ReadImage live in ref_imageio.cpp, which the full pbrt binary links too)
#include "pbrt.h"
#include "geometry.h"
#include "imageio.h"
#include "spectrum.h"
#include <cstdio>

namespace pbrt {

Options PbrtOptions;
int line_num = 0;
std::string current_file;

}  // namespace pbrt
