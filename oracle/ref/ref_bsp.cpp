// Shim TU for the general-BSP families: the reference's accelerators/BSP.cpp, BSPKd.cpp and the
// builders bspPaper(.Kd), bspNodeBased(.FastKd), bspCluster(.FastKd) are compiled VERBATIM as separate objects
// (oracle/ref/Makefile REF_SRCS); this file only opens member access on their HEADERS to export the 20-byte node
// arrays (BSP.h:122-184, BSPKd.h:11-174), primitiveIndices and bounds.  TEST INFRASTRUCTURE ONLY.
#include "std_first.h"
#include "ref_internal.h"
#define private public
#define protected public
#include "accelerators/bspPaper.h"
#include "accelerators/bspPaperKd.h"
#include "accelerators/bspCluster.h"
#include "accelerators/bspClusterFastKd.h"
#undef private
#undef protected
#include "core/paramset.h"

using namespace pbrt;
static_assert(sizeof(BSPNode) == 20 && sizeof(BSPKdNode) == 20, "general BSP nodes are 20 bytes");

template <class Tree>
static int t_info(const ref_accel *a, ref_tree_info *ti) {
    const Tree *b = static_cast<const Tree *>(a->agg.get());
    std::memset(ti, 0, sizeof(*ti));
    ti->node_count = b->nextFreeNode;
    ti->node_stride = 20;
    ti->n_indices = (uint32_t)b->primitiveIndices.size();
    ti->n_directions = 0;
    ti->n_prims = (uint32_t)b->primitives.size();
    ti->bounds[0] = b->bounds.pMin.x; ti->bounds[1] = b->bounds.pMin.y; ti->bounds[2] = b->bounds.pMin.z;
    ti->bounds[3] = b->bounds.pMax.x; ti->bounds[4] = b->bounds.pMax.y; ti->bounds[5] = b->bounds.pMax.z;
    return 0;
}
template <class Tree>
static int t_export(const ref_accel *a, void *nodes, uint32_t *indices, float *) {
    const Tree *b = static_cast<const Tree *>(a->agg.get());
    if (nodes) std::memcpy(nodes, b->nodes, (size_t)b->nextFreeNode * 20);
    if (indices && !b->primitiveIndices.empty())
        std::memcpy(indices, b->primitiveIndices.data(), b->primitiveIndices.size() * 4);
    return 0;
}
template <class Tree>
static ref_accel *wrap(ref_scene *s, std::shared_ptr<Tree> t) {
    ref_accel *a = new ref_accel;
    a->agg = t; a->scene = s; a->info = t_info<Tree>; a->exp = t_export<Tree>;
    return a;
}
// kind: "bsppaper" | "bsppaperkd" | "bspcluster" | "bspclusterfastkd"  (api.cpp:847-1010 names)
ref_accel *ref_make_bsp(ref_scene *s, const char *kind, const ParamSet &ps) {
    const std::string k(kind);
    if (k == "bsppaper") return wrap<BSP>(s, std::shared_ptr<BSP>(CreateBSPPaperTreeAccelerator(s->prims, ps)));
    if (k == "bsppaperkd") return wrap<BSPKd>(s, std::shared_ptr<BSPKd>(CreateBSPPaperKdTreeAccelerator(s->prims, ps)));
    if (k == "bspcluster") return wrap<BSP>(s, std::shared_ptr<BSP>(CreateBSPClusterTreeAccelerator(s->prims, ps)));
    if (k == "bspclusterfastkd")
        return wrap<BSPKd>(s, std::shared_ptr<BSPKd>(CreateBSPClusterFastKdTreeAccelerator(s->prims, ps)));
    return nullptr;
}
