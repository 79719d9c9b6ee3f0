// Host-side tree builders: shared helpers.  Float semantics must match the reference's Release
// build (baseline x86-64, no FMA contraction, no fast-math): compiled with -ffp-contract=off.
#pragma once
#include <algorithm>
#include <cstdint>
#include <limits>

namespace b200rt {

struct Box {
    float lo[3], hi[3];
    // Bounds3() of the reference (core/geometry.h:910-915): inverted extremes.
    Box() {
        for (int i = 0; i < 3; ++i) {
            lo[i] = std::numeric_limits<float>::max();
            hi[i] = std::numeric_limits<float>::lowest();
        }
    }
    void grow(const Box &b) {  // Union(b1, b2), core/geometry.h:1600-1605
        for (int i = 0; i < 3; ++i) {
            lo[i] = std::min(lo[i], b.lo[i]);
            hi[i] = std::max(hi[i], b.hi[i]);
        }
    }
    // Bounds3::SurfaceArea, core/geometry.h:946-949
    float area() const {
        float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        return 2 * (dx * dy + dx * dz + dy * dz);
    }
};

// Triangle::WorldBound, shapes/triangle.cpp:180-186: Union(Bounds3f(p0, p1), p2).
inline Box tri_box(const float *v) {
    Box b;
    for (int i = 0; i < 3; ++i) {
        b.lo[i] = std::min(std::min(v[i], v[3 + i]), v[6 + i]);
        b.hi[i] = std::max(std::max(v[i], v[3 + i]), v[6 + i]);
    }
    return b;
}

}  // namespace b200rt
