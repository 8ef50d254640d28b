// Test-infrastructure stub (NOT product code).  Shadows the reference's <pbrt/util/gui.h> (GLFW/OpenGL window;
// no graphics stack on this box) for the trimmed CPU build in oracle/build_ref_pbrt.sh.  Only the two members that
// non-GUI code references exist: GUI::Initialize (pbrt.cpp:128) and GUI::GetResolution (film.cpp:83).
#ifndef PBRT_UTIL_GUI_H
#define PBRT_UTIL_GUI_H
#include <pbrt/pbrt.h>
#include <pbrt/util/vecmath.h>
namespace pbrt {
class GUI {
  public:
    static void Initialize() {}
    static Point2i GetResolution() { return Point2i(1280, 720); }
};
}  // namespace pbrt
#endif
