// Shim TU: the reference's BVH (accelerators/bvh.cpp) compiled verbatim, with member access opened
// so the flattened LinearBVHNode array (bvh.cpp:123-152) and primNumMapping can be exported.
// TEST INFRASTRUCTURE ONLY.
#include "std_first.h"
#include "ref_internal.h"
#define private public
#define protected public
#include "accelerators/bvh.cpp"
#undef private
#undef protected

using namespace pbrt;
static_assert(sizeof(LinearBVHNode) == 32, "LinearBVHNode must be 32 bytes");

static int bvh_info(const ref_accel *a, ref_tree_info *ti) {
    const BVHAccel *b = static_cast<const BVHAccel *>(a->agg.get());
    std::memset(ti, 0, sizeof(*ti));
    // node count = index of the right-most leaf + 1 (depth-first layout)
    uint32_t i = 0;
    if (!b->nodes) return 0;
    while (!b->nodes[i].IsLeaf()) i = (uint32_t)b->nodes[i].secondChildOffset;
    ti->node_count = i + 1;
    ti->node_stride = 32;
    ti->n_indices = (uint32_t)b->primNumMapping.size();
    ti->n_prims = (uint32_t)b->primitives.size();
    Bounds3f wb = b->WorldBound();
    ti->bounds[0] = wb.pMin.x; ti->bounds[1] = wb.pMin.y; ti->bounds[2] = wb.pMin.z;
    ti->bounds[3] = wb.pMax.x; ti->bounds[4] = wb.pMax.y; ti->bounds[5] = wb.pMax.z;
    return 0;
}
static int bvh_export(const ref_accel *a, void *nodes, uint32_t *order, float *) {
    const BVHAccel *b = static_cast<const BVHAccel *>(a->agg.get());
    ref_tree_info ti; bvh_info(a, &ti);
    if (nodes) std::memcpy(nodes, b->nodes, (size_t)ti.node_count * 32);
    if (order) std::memcpy(order, b->primNumMapping.data(), b->primNumMapping.size() * 4);
    return 0;
}
ref_accel *ref_make_bvh(ref_scene *s, const ParamSet &ps) {
    ref_accel *a = new ref_accel;
    a->agg = CreateBVHAccelerator(s->prims, ps);
    a->scene = s; a->info = bvh_info; a->exp = bvh_export;
    return a;
}
