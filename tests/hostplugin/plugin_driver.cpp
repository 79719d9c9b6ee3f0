// Test driver for the C++ drop-in (thesis-pbrt-v3_b200/host/b200accel.cpp): creates `Accelerator "b200"` over
// the reference's own Primitives and exposes it behind the same ref_accel handle the reference accelerators
// use in oracle/ref/ref_scene.cpp, so ONE harness (ref_intersect / ref_intersect_p, which only see the
// Primitive interface) drives both sides.  TEST INFRASTRUCTURE ONLY.
#include "std_first.h"
#include "ref_internal.h"
#include "core/primitive.h"
#include "core/paramset.h"
#include "core/interaction.h"
#include "b200accel.h"

using namespace pbrt;

extern "C" ref_accel *plugin_accel_create(ref_scene *s, const char *tree, const char *builder,
                                          const char *const *int_names, const int *int_vals,
                                          const char *const *float_names, const float *float_vals) {
    ParamSet ps;
    for (int i = 0; int_names && int_names[i]; ++i) {
        std::unique_ptr<int[]> v(new int[1]);
        v[0] = int_vals[i];
        ps.AddInt(int_names[i], std::move(v), 1);
    }
    for (int i = 0; float_names && float_names[i]; ++i) {
        std::unique_ptr<Float[]> v(new Float[1]);
        v[0] = float_vals[i];
        ps.AddFloat(float_names[i], std::move(v), 1);
    }
    std::unique_ptr<std::string[]> t(new std::string[1]), b(new std::string[1]);
    t[0] = tree; b[0] = builder;
    ps.AddString("tree", std::move(t), 1);
    ps.AddString("builder", std::move(b), 1);
    // what MakeAccelerator's new branch does (api.cpp:790-1015)
    std::shared_ptr<Primitive> acc = CreateB200Accelerator(s->prims, ps);
    if (!acc) return nullptr;
    ref_accel *a = new ref_accel;
    a->agg = acc;
    a->scene = s;
    return a;
}

extern "C" int plugin_is_current(const ref_accel *a) { return B200Accel::current() == a->agg.get(); }

extern "C" void plugin_world_bound(const ref_accel *a, float *out6) {
    Bounds3f b = a->agg->WorldBound();
    out6[0] = b.pMin.x; out6[1] = b.pMin.y; out6[2] = b.pMin.z; out6[3] = b.pMax.x; out6[4] = b.pMax.y; out6[5] = b.pMax.z;
}

// Batched path: B200Accel::IntersectBatch + FillInteraction -> the same ref_hit records ref_intersect writes.
extern "C" int plugin_intersect_batch(const ref_accel *a, const float *rays, uint64_t n, ref_hit *hits) {
    const B200Accel *acc = static_cast<const B200Accel *>(a->agg.get());
    std::vector<Ray> rs;
    rs.reserve(n);
    for (uint64_t i = 0; i < n; ++i) {
        const float *r = rays + 8 * i;
        rs.emplace_back(Point3f(r[0], r[1], r[2]), Vector3f(r[4], r[5], r[6]), r[3], r[7]);
    }
    std::vector<b200rt_hit> dh(n);
    std::vector<Float> tIn(n);
    for (uint64_t i = 0; i < n; ++i) tIn[i] = rs[i].tMax;
    if (!acc->IntersectBatch(rs.data(), n, dh.data())) return 1;
    for (uint64_t i = 0; i < n; ++i) {
        ref_hit h;
        std::memset(&h, 0, sizeof(h));
        h.prim = -1;
        h.t = rs[i].tMax;
        if (dh[i].prim != B200RT_MISS) {
            SurfaceInteraction isect;
            rs[i].tMax = tIn[i];
            if (!acc->FillInteraction(rs[i], dh[i], &isect)) return 2;
            auto it = a->scene->index.find(isect.primitive);
            h.prim = it == a->scene->index.end() ? -2 : it->second;
            h.t = rs[i].tMax;
            h.uv[0] = isect.uv.x; h.uv[1] = isect.uv.y;
            h.p[0] = isect.p.x; h.p[1] = isect.p.y; h.p[2] = isect.p.z;
        }
        hits[i] = h;
    }
    return 0;
}

extern "C" int plugin_intersect_p_batch(const ref_accel *a, const float *rays, uint64_t n, uint8_t *occ) {
    const B200Accel *acc = static_cast<const B200Accel *>(a->agg.get());
    std::vector<Ray> rs;
    rs.reserve(n);
    for (uint64_t i = 0; i < n; ++i) {
        const float *r = rays + 8 * i;
        rs.emplace_back(Point3f(r[0], r[1], r[2]), Vector3f(r[4], r[5], r[6]), r[3], r[7]);
    }
    return acc->IntersectPBatch(rs.data(), n, occ) ? 0 : 1;
}
