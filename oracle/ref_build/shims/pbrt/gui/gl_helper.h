// Headless stand-in for the reference's GLFW/OpenGL preview helper (test infrastructure only).
// Shadows <pbrt/gui/gl_helper.h> when oracle/ref_build/shims is first on the include path, so the
// reference's integrators compile and link without glad/GLFW/OpenGL. No graphics stack exists on
// either machine; --preview therefore shows nothing. Only the members the reference's integrators
// touch are provided.
#pragma once

#include <iomanip>
#include <pbrt/euclidean_space/point2.h>
#include <pbrt/gpu/macro.h>
#include <sstream>
#include <string>

class GLHelper {
  public:
    uint8_t *gpu_frame_buffer = nullptr;

    void init(const std::string &, const Point2i &) {}

    void draw_frame(const std::string &) {}

    static std::string assemble_title(const Real progress_percentage) {
        std::stringstream ss;
        ss << std::fixed << std::setprecision(1) << (progress_percentage * 100.0) << "%";
        return ss.str();
    }
};
