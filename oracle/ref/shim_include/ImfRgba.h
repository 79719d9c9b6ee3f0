// TEST INFRASTRUCTURE.  Stand-in for OpenEXR's ImfRgba.h / ImfRgbaFile.h so that the reference's src/core/imageio.cpp compiles
// UNMODIFIED without the vendored OpenEXR tree (which needs cmake-generated headers): its PNG / TGA / PFM paths are the
// reference's own code; opening an .exr throws, which imageio.cpp reports through Error() ("Unable to read / Error writing").
#pragma once
#include <stdexcept>
namespace Imath {
struct V2i { int x, y; V2i(int x = 0, int y = 0) : x(x), y(y) {} };
struct Box2i { V2i min, max; Box2i() {} Box2i(V2i a, V2i b) : min(a), max(b) {} };
}  // namespace Imath
namespace Imf {
struct Rgba { float r, g, b, a; Rgba(float r = 0, float g = 0, float b = 0, float a = 1) : r(r), g(g), b(b), a(a) {} };
enum RgbaChannels { WRITE_RGB = 7 };
class RgbaInputFile {
  public:
    explicit RgbaInputFile(const char *) { throw std::runtime_error("OpenEXR is not part of this test build (use .pfm / .png)"); }
    Imath::Box2i dataWindow() const { return Imath::Box2i(); }
    Imath::Box2i displayWindow() const { return Imath::Box2i(); }
    void setFrameBuffer(Rgba *, size_t, size_t) {}
    void readPixels(int, int) {}
};
class RgbaOutputFile {
  public:
    RgbaOutputFile(const char *, const Imath::Box2i &, const Imath::Box2i &, RgbaChannels) { throw std::runtime_error("OpenEXR is not part of this test build (use .pfm / .png)"); }
    void setFrameBuffer(const Rgba *, size_t, size_t) {}
    void writePixels(int) {}
};
}  // namespace Imf
