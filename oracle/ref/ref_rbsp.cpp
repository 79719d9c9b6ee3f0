// Shim TU: the reference's accelerators/rbsp.cpp compiled verbatim, with member access opened so
// the 8-byte node array, primitiveIndices, directions and bounds (genericBSP.h:185-193) can be
// exported.  TEST INFRASTRUCTURE ONLY.
#include "std_first.h"
#include "ref_internal.h"
#define private public
#define protected public
#include "accelerators/rbsp.cpp"
#undef private
#undef protected

using namespace pbrt;

static int t_info(const ref_accel *a, ref_tree_info *ti) {
    const RBSP *b = static_cast<const RBSP *>(a->agg.get());
    std::memset(ti, 0, sizeof(*ti));
    static_assert(sizeof(b->nodes[0]) == 8, "node must be 8 bytes");
    ti->node_count = b->nextFreeNode;
    ti->node_stride = 8;
    ti->n_indices = (uint32_t)b->primitiveIndices.size();
    ti->n_directions = (uint32_t)b->directions.size();
    ti->n_prims = (uint32_t)b->primitives.size();
    ti->bounds[0] = b->bounds.pMin.x; ti->bounds[1] = b->bounds.pMin.y; ti->bounds[2] = b->bounds.pMin.z;
    ti->bounds[3] = b->bounds.pMax.x; ti->bounds[4] = b->bounds.pMax.y; ti->bounds[5] = b->bounds.pMax.z;
    return 0;
}
static int t_export(const ref_accel *a, void *nodes, uint32_t *indices, float *directions) {
    const RBSP *b = static_cast<const RBSP *>(a->agg.get());
    if (nodes) std::memcpy(nodes, b->nodes, (size_t)b->nextFreeNode * 8);
    if (indices && !b->primitiveIndices.empty())
        std::memcpy(indices, b->primitiveIndices.data(), b->primitiveIndices.size() * 4);
    if (directions)
        for (size_t i = 0; i < b->directions.size(); ++i) {
            directions[3 * i] = b->directions[i].x;
            directions[3 * i + 1] = b->directions[i].y;
            directions[3 * i + 2] = b->directions[i].z;
        }
    return 0;
}
ref_accel *ref_make_rbsp(ref_scene *s, const ParamSet &ps) {
    ref_accel *a = new ref_accel;
    a->agg = CreateRBSPTreeAccelerator(s->prims, ps);
    a->scene = s; a->info = t_info; a->exp = t_export;
    return a;
}
