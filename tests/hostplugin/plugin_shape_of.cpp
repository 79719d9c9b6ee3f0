// GeometricPrimitive::shape has no accessor (primitive.h:135-140); the test driver needs it to re-wrap the scene's
// shapes with a material.  TEST INFRASTRUCTURE ONLY.
#include "std_first.h"
#define private public
#include "core/primitive.h"
#undef private
std::shared_ptr<pbrt::Shape> shape_of(const pbrt::Primitive *p) {
    return static_cast<const pbrt::GeometricPrimitive *>(p)->shape;
}
