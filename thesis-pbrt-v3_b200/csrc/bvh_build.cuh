// BVHAccel's sweep-SAH build (src/accelerators/bvh.cpp:155-352: BVHAccel::BVHAccel, iterativeBuild, flattenBVHTree)
// re-designed for the B200: level-synchronous, sorted once.
//
// The reference pops one node at a time and, for each of the three axes, sorts the node's primitives by (centroid, primitive
// number), sweeps the boxes from both ends and evaluates the SAH at every position (bvh.cpp:228-275).  That comparator is a
// TOTAL order that does not depend on the node, so here every axis is sorted ONCE for the whole scene; a tree level then is
//
//   1. per axis: the boxes gathered in sorted order, two segmented scans (prefix and suffix unions, the node as segment key) —
//      unions are exact min / max, so the scan order cannot change a bit of them;
//   2. one thread per (axis, position): the SAH cost (bvh.cpp:262-266) and a 64-bit atomic arg-min per node keyed
//      (cost, axis, position): the reference's strict `cost < bestCost` scan order;
//   3. per node: split or leaf (bvh.cpp:278: cheaper than a leaf, or more than maxPrimsInNode primitives);
//   4. per axis: each node's sorted run split into its children's runs by a stable partition (one scan), which keeps them sorted;
//   5. the `primitiveInfo` array itself re-ordered exactly as std::partition leaves it (bvh.cpp:280-285) — libstdc++'s
//      two-pointer partition swaps the k-th misplaced element from the left with the k-th misplaced element from the right, so
//      the permutation is two scans and a scatter — because that order is what ends up in the leaves (primNumMapping).
//
// Afterwards nodes are numbered as flattenBVHTree does (first child = parent + 1, second child after the first subtree,
// bvh.cpp:335-350) and the leaves' primitive ranges are laid out in the order the reference's LIFO work stack reaches them
// (second child first, bvh.cpp:298-299): both from subtree sizes, bottom-up then top-down.  The result — 32-byte LinearBVHNodes
// and primNumMapping — equals the reference's byte for byte; there are no ties to break (the sort key is unique).
//
// As in rbsp_build.cuh the functors are the kernels (build_rbsp.cu launches them, tests/hostbuild loops over them).
#pragma once
#include <stdint.h>
#include <stddef.h>

#include "rbsp_build.cuh"

namespace b200rt {
namespace bvhb {

using rbsp::atomic_min_u64;
using rbsp::float_order;
using rbsp::float_unorder;

struct Box6 { float lo[3], hi[3]; };
WF_HD Box6 box_empty() {   // Bounds3(), geometry.h:910-915
    Box6 b;
    for (int i = 0; i < 3; ++i) { b.lo[i] = 3.402823466e+38f; b.hi[i] = -3.402823466e+38f; }
    return b;
}
struct BoxUnion {   // Union(b1, b2), geometry.h:1600-1605
    WF_HD Box6 operator()(const Box6 &a, const Box6 &b) const {
        Box6 r;
        for (int i = 0; i < 3; ++i) { r.lo[i] = smin(a.lo[i], b.lo[i]); r.hi[i] = smax(a.hi[i], b.hi[i]); }
        return r;
    }
};
WF_HD float box_area(const Box6 &b) {   // Bounds3::SurfaceArea, geometry.h:946-949
    const float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
    return 2 * (dx * dy + dx * dz + dy * dz);
}

struct BvhParams { int maxPrimsInNode, isectCost, traversalCost; };

enum : uint32_t { kLeaf = 0, kInterior = 1, kActive = 2, kNoNode = 0xffffffffu };

struct BLevel {
    uint32_t K = 0;
    uint32_t *start = nullptr, *end = nullptr, *kind = nullptr, *axis = nullptr, *bestI = nullptr, *child = nullptr;
    Box6 *box = nullptr;
    uint64_t *best = nullptr;
    uint32_t *size = nullptr, *pre = nullptr, *off = nullptr;
};

struct PrimInfoF {   // BVHPrimitiveInfo, bvh.cpp:52-65: bounds and centroid = .5f * pMin + .5f * pMax
    const float *tris;
    Box6 *box;
    float *centroid;   // n x 3
    WF_HD void operator()(size_t p) const {
        const float *v = tris + 9 * p;
        Box6 b;
        for (int i = 0; i < 3; ++i) {
            b.lo[i] = smin(smin(v[i], v[3 + i]), v[6 + i]);
            b.hi[i] = smax(smax(v[i], v[3 + i]), v[6 + i]);
            centroid[3 * p + i] = .5f * b.lo[i] + .5f * b.hi[i];
        }
        box[p] = b;
    }
};

struct CentroidKeyF {   // the comparator of bvh.cpp:236-241 as one 64-bit key: (centroid, primitive number)
    const float *centroid;
    uint32_t dim;
    uint64_t *keys;
    uint32_t *vals;
    WF_HD void operator()(size_t p) const {
        keys[p] = ((uint64_t)float_order(centroid[3 * p + dim] + 0.0f) << 32) | (uint32_t)p;
        vals[p] = (uint32_t)p;
    }
};

struct InitNodeF {
    BLevel L;
    WF_HD void operator()(size_t k) const {
        L.kind[k] = (L.end[k] - L.start[k] == 1) ? kLeaf : kActive;   // bvh.cpp:214
        L.best[k] = ~0ull;
    }
};

struct GatherBoxF {
    const uint32_t *sorted;
    const Box6 *box;
    Box6 *out;
    WF_HD void operator()(size_t p) const { out[p] = box[sorted[p]]; }
};

struct NodeBoxF {   // the node's bounds = the suffix union at its first position (r2l[0], bvh.cpp:246-250,270)
    BLevel L;
    const Box6 *r2l;
    WF_HD void operator()(size_t k) const { L.box[k] = r2l[L.start[k]]; }
};

struct SweepCostF {   // bvh.cpp:258-273
    BLevel L;
    BvhParams P;
    uint32_t dim;
    const uint32_t *posNode;
    const Box6 *l2r, *r2l;
    WF_HD void operator()(size_t p) const {
        const uint32_t k = posNode[p];
        if (k == kNoNode || L.kind[k] != kActive) return;
        const uint32_t nPrimitives = L.end[k] - L.start[k];
        const int i = (int)((uint32_t)p - L.start[k]);
        if (i >= (int)nPrimitives - 1) return;
        const float totalSA = box_area(L.box[k]);
        const float invTotalSA = 1 / totalSA;
        const float cost = P.traversalCost + P.isectCost * ((i + 1) * box_area(l2r[p]) + (nPrimitives - i - 1) * box_area(r2l[p + 1])) * invTotalSA;
        if (!(cost < wf_inf())) return;
        atomic_min_u64(L.best + k, ((uint64_t)float_order(cost + 0.0f) << 32) | ((uint64_t)dim << 28) | (uint32_t)i);
    }
};

struct DecideF {   // bvh.cpp:278
    BLevel L;
    BvhParams P;
    uint32_t *interior;   // K + 1
    WF_HD void operator()(size_t k) const {
        interior[k] = 0;
        if (k >= L.K || L.kind[k] != kActive) return;
        const uint64_t b = L.best[k];
        const uint32_t n = L.end[k] - L.start[k];
        const float oldCost = P.isectCost * float(n);
        const bool none = b == ~0ull;
        const float bestCost = none ? wf_inf() : float_unorder((uint32_t)(b >> 32));
        if (!none && (bestCost < oldCost || n > (uint32_t)P.maxPrimsInNode)) {
            L.kind[k] = kInterior;
            L.axis[k] = (uint32_t)(b >> 28) & 15u;
            L.bestI[k] = (uint32_t)b & 0x0fffffffu;
            interior[k] = 1;
        } else {
            L.kind[k] = kLeaf;
        }
    }
};

struct ChildrenF {   // ranges of the children: [start, mid) and [mid, end), mid = start + bestI + 1
    BLevel L, Nx;
    const uint32_t *rank;
    WF_HD void operator()(size_t k) const {
        if (L.kind[k] != kInterior) return;
        const uint32_t c = 2 * rank[k], mid = L.start[k] + L.bestI[k] + 1;
        L.child[k] = c;
        Nx.start[c] = L.start[k]; Nx.end[c] = mid;
        Nx.start[c + 1] = mid; Nx.end[c + 1] = L.end[k];
    }
};

struct SideF {   // which child a primitive goes to: its position in the winning axis' sorted run (the predicate of bvh.cpp:281-284)
    BLevel L;
    const uint32_t *posNode;
    const uint32_t *sorted0, *sorted1, *sorted2;
    uint8_t *side;
    WF_HD void operator()(size_t p) const {
        const uint32_t k = posNode[p];
        if (k == kNoNode || L.kind[k] != kInterior) return;
        const uint32_t a = L.axis[k];
        const uint32_t prim = (a == 0 ? sorted0 : a == 1 ? sorted1 : sorted2)[p];
        side[prim] = ((uint32_t)p - L.start[k] <= L.bestI[k]) ? 0 : 1;
    }
};

struct LeftFlagF {   // n + 1 flags for the stable partition of one axis' runs
    BLevel L;
    const uint32_t *posNode, *sorted;
    const uint8_t *side;
    uint32_t n;
    uint32_t *flag;
    WF_HD void operator()(size_t p) const {
        uint32_t f = 0;
        if (p < n) {
            const uint32_t k = posNode[p];
            if (k != kNoNode && L.kind[k] == kInterior) f = side[sorted[p]] == 0;
        }
        flag[p] = f;
    }
};

struct StablePartitionF {
    BLevel L;
    const uint32_t *posNode, *sorted, *gL;
    const uint8_t *side;
    uint32_t *out;
    WF_HD void operator()(size_t p) const {
        const uint32_t k = posNode[p], prim = sorted[p];
        if (k == kNoNode || L.kind[k] != kInterior) { out[p] = prim; return; }
        const uint32_t s = L.start[k], lefts = gL[p] - gL[s];
        if (side[prim] == 0) out[s + lefts] = prim;
        else out[s + L.bestI[k] + 1 + ((uint32_t)p - s - lefts)] = prim;
    }
};

// std::partition on primitiveInfo[start, end) (bvh.cpp:280-285) as libstdc++ does it for bidirectional iterators: the k-th
// element of the left zone that belongs right is swapped with the k-th element (from the END) of the right zone that belongs left
struct MisplacedFlagF {
    BLevel L;
    const uint32_t *posNode, *info;
    const uint8_t *side;
    uint32_t n;
    uint32_t *fF, *fT;   // n + 1 each
    WF_HD void operator()(size_t p) const {
        uint32_t f = 0, t = 0;
        if (p < n) {
            const uint32_t k = posNode[p];
            if (k != kNoNode && L.kind[k] == kInterior) {
                const bool leftZone = (uint32_t)p - L.start[k] <= L.bestI[k];
                const uint32_t s = side[info[p]];
                f = leftZone && s == 1;
                t = !leftZone && s == 0;
            }
        }
        fF[p] = f;
        fT[p] = t;
    }
};
struct PairUpF {
    BLevel L;
    const uint32_t *posNode, *fF, *fT, *gF, *gT;
    uint32_t *pairF, *pairT;
    WF_HD void operator()(size_t p) const {
        if (!fF[p] && !fT[p]) return;
        const uint32_t k = posNode[p], s = L.start[k], mid = s + L.bestI[k] + 1, base = gF[s];
        if (fF[p]) pairF[base + (gF[p] - gF[s])] = (uint32_t)p;
        else {
            const uint32_t m = gT[L.end[k]] - gT[mid];
            pairT[base + (m - 1 - (gT[p] - gT[mid]))] = (uint32_t)p;
        }
    }
};
struct SwapPairsF {
    const uint32_t *pairF, *pairT, *info;
    uint32_t *out;   // a copy of info
    WF_HD void operator()(size_t j) const {
        const uint32_t a = pairF[j], b = pairT[j];
        out[a] = info[b];
        out[b] = info[a];
    }
};

struct NextPosNodeF {
    BLevel L;
    const uint32_t *posNode;
    uint32_t *out;
    WF_HD void operator()(size_t p) const {
        const uint32_t k = posNode[p];
        if (k == kNoNode || L.kind[k] != kInterior) { out[p] = kNoNode; return; }
        out[p] = L.child[k] + (((uint32_t)p - L.start[k] <= L.bestI[k]) ? 0u : 1u);
    }
};

struct CopyU32F {
    const uint32_t *in;
    uint32_t *out;
    WF_HD void operator()(size_t i) const { out[i] = in[i]; }
};

// ---- numbering and emission -----------------------------------------------------------------------------------------------
struct SubtreeSizeF {
    BLevel L, Nx;
    WF_HD void operator()(size_t k) const {
        L.size[k] = L.kind[k] == kInterior ? 1 + Nx.size[L.child[k]] + Nx.size[L.child[k] + 1] : 1;
    }
};
struct AssignF {
    BLevel L, Nx;
    WF_HD void operator()(size_t k) const {
        if (L.kind[k] != kInterior) return;
        const uint32_t c = L.child[k];
        Nx.pre[c] = L.pre[k] + 1;                       // flattenBVHTree: first child right after its parent
        Nx.pre[c + 1] = L.pre[k] + 1 + Nx.size[c];      // second child after the first subtree
        Nx.off[c + 1] = L.off[k];                       // the work stack builds the SECOND child's subtree first (bvh.cpp:298-299)
        Nx.off[c] = L.off[k] + (Nx.end[c + 1] - Nx.start[c + 1]);
    }
};
struct EmitF {   // LinearBVHNode, bvh.cpp:123-152,335-350
    BLevel L, Nx;
    const uint32_t *info;
    uint32_t *nodes;   // 8 words per node
    uint32_t *order;   // primNumMapping
    WF_HD void operator()(size_t k) const {
        uint32_t *w = nodes + 8 * (size_t)L.pre[k];
        const Box6 &b = L.box[k];
        for (int i = 0; i < 3; ++i) { w[i] = f2u(b.lo[i]); w[3 + i] = f2u(b.hi[i]); }
        const uint32_t n = L.end[k] - L.start[k];
        if (L.kind[k] == kInterior) {
            w[6] = Nx.pre[L.child[k] + 1];
            w[7] = L.axis[k] | (n << 2);
        } else {
            w[6] = L.off[k];
            w[7] = 3u | (n << 2);
            for (uint32_t i = 0; i < n; ++i) order[L.off[k] + i] = info[L.start[k] + i];
        }
    }
};

}  // namespace bvhb
}  // namespace b200rt
