// Sweep-SAH BVH builder producing the reference's flattened LinearBVHNode array node for node.
//
// Restates BVHAccel::BVHAccel / iterativeBuild / flattenBVHTree
// (reference src/accelerators/bvh.cpp:155-352): explicit LIFO work stack (second child built
// first), full SAH sweep over centroids sorted by (centroid, primitive number) on all three
// axes, split kept when cheaper than a leaf or when the node holds more than maxnodeprims, leaves
// filled in std::partition order, depth-first flattening with `first child = i+1`.
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>
#include "../../include/b200rt.h"
#include "build_common.h"

namespace b200rt {
namespace {

struct PrimInfo {  // bvh.cpp:52-65
    size_t prim;
    Box box;
    float centroid[3];
};
struct Centroid {  // bvh.cpp:111-121
    float t;
    uint32_t offset, prim;
};
struct BuildNode {
    Box box;
    int child[2] = {-1, -1};
    uint32_t axis = 0, first = 0, count = 0;
    bool leaf = false;
};
struct Todo { int node, start, end; };

struct LinearNode {  // bvh.cpp:123-152, 32 bytes
    float lo[3], hi[3];
    int32_t offset;
    uint32_t word;
};
static_assert(sizeof(LinearNode) == 32, "LinearBVHNode layout");

int flatten(const std::vector<BuildNode> &bn, int node, LinearNode *out, int *offset) {  // bvh.cpp:335-350
    // explicit stack instead of recursion (sweep-SAH trees can be deep)
    struct Frame { int node, my, stage; };
    std::vector<Frame> st;
    st.push_back({node, -1, 0});
    int rootOffset = -1;
    while (!st.empty()) {
        Frame &f = st.back();
        const BuildNode &n = bn[f.node];
        if (f.stage == 0) {
            f.my = (*offset)++;
            if (rootOffset < 0) rootOffset = f.my;
            LinearNode &ln = out[f.my];
            std::memcpy(ln.lo, n.box.lo, 12);
            std::memcpy(ln.hi, n.box.hi, 12);
            if (n.leaf) {
                ln.word = 3u | (n.count << 2);
                ln.offset = (int32_t)n.first;
                st.pop_back();
                continue;
            }
            ln.word = n.axis | (n.count << 2);
            ln.offset = 0;
            f.stage = 1;
            st.push_back({n.child[0], -1, 0});
        } else if (f.stage == 1) {
            out[f.my].offset = *offset;  // secondChildOffset
            f.stage = 2;
            st.push_back({n.child[1], -1, 0});
        } else {
            st.pop_back();
        }
    }
    return rootOffset;
}

}  // namespace
}  // namespace b200rt

using namespace b200rt;

extern "C" int b200rt_build_bvh(const float *tri_verts, uint32_t n_prims, int max_node_prims, int isect_cost,
                                int traversal_cost, void **out_nodes, uint32_t *out_node_count,
                                uint32_t **out_prim_order) {
    if (!tri_verts || !out_nodes || !out_node_count || !out_prim_order) return B200RT_EINVAL;
    *out_nodes = nullptr; *out_node_count = 0; *out_prim_order = nullptr;
    if (n_prims == 0) return B200RT_OK;
    const int maxPrimsInNode = std::min(255, max_node_prims);  // bvh.cpp:157
    const int isectCost = isect_cost, traversalCost = traversal_cost;
    const float Infinity = std::numeric_limits<float>::infinity();

    std::vector<PrimInfo> info(n_prims);
    for (size_t i = 0; i < n_prims; ++i) {
        info[i].prim = i;
        info[i].box = tri_box(tri_verts + 9 * i);
        for (int a = 0; a < 3; ++a) info[i].centroid[a] = .5f * info[i].box.lo[a] + .5f * info[i].box.hi[a];
    }
    std::unique_ptr<Centroid[]> cen[3];
    for (auto &c : cen) c.reset(new Centroid[n_prims]);
    std::unique_ptr<Box[]> r2l(new Box[n_prims]), l2r(new Box[n_prims]);

    std::vector<BuildNode> bn;
    bn.reserve(2 * (size_t)n_prims);
    std::vector<uint32_t> order;  // slot -> original primitive (primNumMapping)
    order.reserve(n_prims);
    bn.emplace_back();
    std::vector<Todo> stack;
    stack.push_back({0, 0, (int)n_prims});
    while (!stack.empty()) {
        Todo cur = stack.back();
        stack.pop_back();
        Box bounds;
        for (int i = cur.start; i < cur.end; ++i) bounds.grow(info[i].box);
        uint32_t nPrimitives = cur.end - cur.start;
        if (nPrimitives == 1) {
            BuildNode &n = bn[cur.node];
            n.leaf = true; n.first = (uint32_t)order.size(); n.count = 1; n.box = bounds;
            order.push_back((uint32_t)info[cur.start].prim);
            continue;
        }
        uint32_t bestAxis = -1, bestOffset = -1, bestPrimNum = -1;
        Box bestBounds;
        float bestCost = Infinity;
        float oldCost = isectCost * float(nPrimitives);
        float totalSA = bounds.area();
        float invTotalSA = 1 / totalSA;
        for (uint32_t dim = 0; dim < 3; dim++) {
            Centroid *c = cen[dim].get();
            for (uint32_t i = 0; i < nPrimitives; ++i)
                c[i] = {info[cur.start + i].centroid[dim], (uint32_t)(cur.start + i), (uint32_t)info[cur.start + i].prim};
            std::sort(c, c + nPrimitives, [](const Centroid &a, const Centroid &b) -> bool {
                if (a.t == b.t) return (int)a.prim < (int)b.prim;
                return a.t < b.t;
            });
            Box acc;
            for (int i = (int)nPrimitives - 1; i >= 0; i--) { acc.grow(info[c[i].offset].box); r2l[i] = acc; }
            Box acc2;
            for (int i = 0; i < (int)nPrimitives - 1; ++i) { acc2.grow(info[c[i].offset].box); l2r[i] = acc2; }
            for (int i = 0; i < (int)nPrimitives - 1; ++i) {
                float cost = traversalCost + isectCost *
                                                 ((i + 1) * l2r[i].area() + (nPrimitives - i - 1) * r2l[i + 1].area()) *
                                                 invTotalSA;
                if (cost < bestCost) {
                    bestCost = cost; bestAxis = dim; bestOffset = c[i].offset; bestPrimNum = c[i].prim;
                    bestBounds = r2l[0];
                }
            }
        }
        if (bestAxis != (uint32_t)-1 && (bestCost < oldCost || nPrimitives > (uint32_t)maxPrimsInNode)) {
            const float bestCentroid = info[bestOffset].centroid[bestAxis];
            PrimInfo *pmid = std::partition(&info[cur.start], &info[cur.end - 1] + 1, [&](const PrimInfo &pi) {
                return pi.centroid[bestAxis] < bestCentroid ||
                       (pi.centroid[bestAxis] == bestCentroid && pi.prim <= bestPrimNum);
            });
            int mid = (int)(pmid - &info[0]);
            int c0 = (int)bn.size(); bn.emplace_back();
            int c1 = (int)bn.size(); bn.emplace_back();
            BuildNode &n = bn[cur.node];
            n.leaf = false; n.axis = bestAxis; n.child[0] = c0; n.child[1] = c1; n.count = nPrimitives; n.box = bestBounds;
            stack.push_back({c0, cur.start, mid});
            stack.push_back({c1, mid, cur.end});
        } else {
            BuildNode &n = bn[cur.node];
            n.leaf = true; n.first = (uint32_t)order.size(); n.count = nPrimitives; n.box = bounds;
            for (int i = cur.start; i < cur.end; ++i) order.push_back((uint32_t)info[i].prim);
        }
    }
    const size_t total = bn.size();
    LinearNode *nodes = (LinearNode *)std::malloc(total * sizeof(LinearNode));
    uint32_t *ord = (uint32_t *)std::malloc(sizeof(uint32_t) * n_prims);
    if (!nodes || !ord) { std::free(nodes); std::free(ord); return B200RT_ENOMEM; }
    int offset = 0;
    flatten(bn, 0, nodes, &offset);
    std::memcpy(ord, order.data(), sizeof(uint32_t) * n_prims);
    *out_nodes = nodes; *out_node_count = (uint32_t)total; *out_prim_order = ord;
    return B200RT_OK;
}

extern "C" void b200rt_free(void *p) { std::free(p); }
