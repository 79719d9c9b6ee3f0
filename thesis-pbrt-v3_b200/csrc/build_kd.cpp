// SAH kd-tree builder producing the reference's 8-byte KdAccelNode array node for node.
//
// Restates KdTreeAccel::buildTree (reference src/accelerators/kdtreeaccel.cpp:208-379) on the
// GenericBSP base (genericBSP.h:61-193): explicit LIFO work stack (below child directly follows its
// parent, above child patched in when popped), edge sweep on all three axes with edges sorted by
// (t, Start<End), leaf when nPrims <= maxprims / depth exhausted / 3 bad refines / cost rule,
// primitive lists kept in one (maxDepth+1)*N scratch array exactly as the reference does so child
// lists come out in the same order.
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>
#include "../../include/b200rt.h"
#include "build_common.h"

namespace b200rt {
namespace {
struct Edge {  // genericBSP.h:47-58 ; type 0 = Start, 1 = End
    float t;
    uint32_t prim;
    int type;
};
struct KdTodo {  // kdtreeaccel.h:77-90
    uint32_t depth, nPrims, badRefines;
    Box box;
    uint32_t *prims;
    uint32_t parent;
};
struct KdNode { uint32_t w0, w1; };  // kdtreeaccel.cpp:146-159
}  // namespace
}  // namespace b200rt
using namespace b200rt;

extern "C" int b200rt_build_kdtree(const float *tri_verts, uint32_t n_prims, int isect_cost, int traversal_cost,
                                   float empty_bonus, int max_prims, int max_depth, void **out_nodes,
                                   uint32_t *out_node_count, uint32_t **out_prim_indices, uint32_t *out_n_indices,
                                   float out_bounds[6]) {
    if (!tri_verts || !out_nodes || !out_node_count || !out_prim_indices || !out_n_indices || !out_bounds)
        return B200RT_EINVAL;
    *out_nodes = nullptr; *out_node_count = 0; *out_prim_indices = nullptr; *out_n_indices = 0;
    const uint32_t isectCost = (uint32_t)isect_cost, traversalCost = (uint32_t)traversal_cost;
    const uint32_t maxPrims = (uint32_t)max_prims;
    const float emptyBonus = empty_bonus;
    const float Infinity = std::numeric_limits<float>::infinity();
    // genericBSP.h:75-78 + geometry.h:1845-1847: round(2 + 1.6f * Log2Int(N))
    uint32_t maxDepth = (uint32_t)max_depth;
    if (max_depth == -1)
        maxDepth = (uint32_t)std::round(2 + 1.6f * (n_prims ? (63 - __builtin_clzll((unsigned long long)n_prims)) : 0));

    std::vector<Box> primBox(n_prims);
    Box bounds;
    for (size_t i = 0; i < n_prims; ++i) {
        primBox[i] = tri_box(tri_verts + 9 * i);
        bounds.grow(primBox[i]);
    }
    std::unique_ptr<Edge[]> edges[3];
    for (auto &e : edges) e.reset(new Edge[2 * (size_t)n_prims + 1]);
    std::unique_ptr<uint32_t[]> prims(new uint32_t[((size_t)maxDepth + 1) * n_prims + 1]);
    for (uint32_t i = 0; i < n_prims; ++i) prims[i] = i;

    std::vector<KdNode> nodes;
    std::vector<uint32_t> indices;
    auto init_leaf = [&](KdNode &n, const uint32_t *pn, uint32_t np) {  // kdtreeaccel.cpp:188-201
        n.w1 = 3u | (np << 2u);
        if (np == 0) n.w0 = 0;
        else if (np == 1) n.w0 = pn[0];
        else {
            n.w0 = (uint32_t)indices.size();
            for (uint32_t i = 0; i < np; ++i) indices.push_back(pn[i]);
        }
    };

    std::vector<KdTodo> stack;
    stack.push_back({maxDepth, n_prims, 0, bounds, &prims[0], (uint32_t)-1});
    while (!stack.empty()) {
        KdTodo cur = stack.back();
        stack.pop_back();
        const uint32_t nodeNum = (uint32_t)nodes.size();
        if (cur.parent != (uint32_t)-1) nodes[cur.parent].w1 |= (nodeNum << 2u);  // setAboveChild
        nodes.push_back({0, 0});
        if (cur.nPrims <= maxPrims || cur.depth == 0) {
            init_leaf(nodes[nodeNum], cur.prims, cur.nPrims);
            continue;
        }
        uint32_t bestAxis = -1, bestOffset = -1;
        float bestCost = Infinity;
        float oldCost = isectCost * float(cur.nPrims);
        float totalSA = cur.box.area();
        const float invTotalSA = 1 / totalSA;
        const float d[3] = {cur.box.hi[0] - cur.box.lo[0], cur.box.hi[1] - cur.box.lo[1], cur.box.hi[2] - cur.box.lo[2]};
        for (uint32_t axis = 0; axis < 3; ++axis) {
            Edge *e = edges[axis].get();
            for (uint32_t i = 0; i < cur.nPrims; ++i) {
                const uint32_t pn = cur.prims[i];
                e[2 * i] = {primBox[pn].lo[axis], pn, 0};
                e[2 * i + 1] = {primBox[pn].hi[axis], pn, 1};
            }
            std::sort(e, e + 2 * (size_t)cur.nPrims, [](const Edge &a, const Edge &b) -> bool {
                if (a.t == b.t) return a.type < b.type;
                return a.t < b.t;
            });
            uint32_t nBelow = 0, nAbove = cur.nPrims;
            for (uint32_t i = 0; i < 2 * cur.nPrims; ++i) {
                if (e[i].type == 1) --nAbove;
                const float edgeT = e[i].t;
                if (edgeT > cur.box.lo[axis] && edgeT < cur.box.hi[axis]) {
                    const uint32_t o0 = (axis + 1) % 3, o1 = (axis + 2) % 3;
                    const float belowSA = 2 * (d[o0] * d[o1] + (edgeT - cur.box.lo[axis]) * (d[o0] + d[o1]));
                    const float aboveSA = 2 * (d[o0] * d[o1] + (cur.box.hi[axis] - edgeT) * (d[o0] + d[o1]));
                    const float pBelow = belowSA * invTotalSA;
                    const float pAbove = aboveSA * invTotalSA;
                    const float eb = (nAbove == 0 || nBelow == 0) ? emptyBonus : 0;
                    const float cost = traversalCost + isectCost * (1 - eb) * (pBelow * nBelow + pAbove * nAbove);
                    if (cost < bestCost) { bestCost = cost; bestAxis = axis; bestOffset = i; }
                }
                if (e[i].type == 0) ++nBelow;
            }
        }
        if (bestCost > oldCost) ++cur.badRefines;
        if ((bestCost > 4 * oldCost && cur.nPrims < 16) || bestAxis == (uint32_t)-1 || cur.badRefines == 3) {
            init_leaf(nodes[nodeNum], cur.prims, cur.nPrims);
            continue;
        }
        // classify: above list first (it overwrites the parent's list in place), then below
        const Edge *be = edges[bestAxis].get();
        uint32_t n0 = 0, n1 = 0;
        uint32_t *prims1 = cur.prims;
        for (uint32_t i = bestOffset + 1; i < 2 * cur.nPrims; ++i)
            if (be[i].type == 1) prims1[n1++] = be[i].prim;
        uint32_t *prims0 = prims1 + n1;
        for (uint32_t i = 0; i < bestOffset; ++i)
            if (be[i].type == 0) prims0[n0++] = be[i].prim;
        const float tSplit = be[bestOffset].t;
        Box b0 = cur.box, b1 = cur.box;
        b0.hi[bestAxis] = b1.lo[bestAxis] = tSplit;
        std::memcpy(&nodes[nodeNum].w0, &tSplit, 4);  // InitInterior
        nodes[nodeNum].w1 = bestAxis;
        stack.push_back({cur.depth - 1, n1, cur.badRefines, b1, prims1, nodeNum});
        stack.push_back({cur.depth - 1, n0, cur.badRefines, b0, prims0, (uint32_t)-1});
    }
    KdNode *on = (KdNode *)std::malloc(std::max<size_t>(1, nodes.size()) * sizeof(KdNode));
    uint32_t *oi = (uint32_t *)std::malloc(std::max<size_t>(1, indices.size()) * sizeof(uint32_t));
    if (!on || !oi) { std::free(on); std::free(oi); return B200RT_ENOMEM; }
    std::memcpy(on, nodes.data(), nodes.size() * sizeof(KdNode));
    if (!indices.empty()) std::memcpy(oi, indices.data(), indices.size() * sizeof(uint32_t));
    *out_nodes = on; *out_node_count = (uint32_t)nodes.size();
    *out_prim_indices = oi; *out_n_indices = (uint32_t)indices.size();
    for (int i = 0; i < 3; ++i) { out_bounds[i] = bounds.lo[i]; out_bounds[3 + i] = bounds.hi[i]; }
    return B200RT_OK;
}
