// Internal glue shared by the shim TUs (TEST INFRASTRUCTURE ONLY; see ref_api.h).
#ifndef B200RT_REF_INTERNAL_H
#define B200RT_REF_INTERNAL_H
#include <memory>
#include <unordered_map>
#include <vector>
#include "ref_api.h"

namespace pbrt { class Primitive; class ParamSet; class Transform; }

struct ref_scene {
    std::vector<std::shared_ptr<pbrt::Primitive>> prims;
    std::unordered_map<const pbrt::Primitive *, int> index;
    std::vector<std::unique_ptr<pbrt::Transform>> transforms;
};

struct ref_accel {
    std::shared_ptr<pbrt::Primitive> agg;
    const ref_scene *scene = nullptr;
    int (*info)(const ref_accel *, ref_tree_info *) = nullptr;
    int (*exp)(const ref_accel *, void *, uint32_t *, float *) = nullptr;
};

// one per shim TU (the reference's accelerator .cpp files cannot share a TU)
ref_accel *ref_make_bvh(ref_scene *, const pbrt::ParamSet &);
ref_accel *ref_make_kd(ref_scene *, const pbrt::ParamSet &);
ref_accel *ref_make_rbsp(ref_scene *, const pbrt::ParamSet &);
ref_accel *ref_make_rbspkd(ref_scene *, const pbrt::ParamSet &);
ref_accel *ref_make_bsp(ref_scene *, const char *kind, const pbrt::ParamSet &);
#endif
