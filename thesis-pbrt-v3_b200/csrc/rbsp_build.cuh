// RBSP::buildTree (src/accelerators/rbsp.cpp:181-403) re-designed for the B200: level-synchronous instead of depth-first.
//
// The reference pops one node at a time and, for each of its 2n bound edges in each of the M directions, cuts the node's k-DOP
// and measures both halves (38-319 s per tree on the thesis machine, killerroo-simple-times:4).  Every one of those candidate
// planes is independent, and so are the nodes of one tree level, so here a LEVEL is processed at once:
//
//   1. per direction: (node, projected bound, start/end) keys of every primitive reference of the level, one radix sort —
//      the reference's per-node std::sort (rbsp.cpp:282-294) for all nodes at once;
//   2. one thread per DISTINCT candidate plane (equal bounds of neighbouring triangles share one cut): k-DOP cut + both
//      surface areas (EvalAreas), then one thread per bound edge: the SAH cost from the edge's running counts (two scans) and a
//      64-bit atomic arg-min per node keyed (cost, direction, offset) — the reference's strict `cost < bestCost` scan order;
//   3. per node: leaf / interior decision (badRefines and the 4x rule, rbsp.cpp:336-344), the winning cut redone to produce
//      the children's meshes, areas and per-direction extents;
//   4. primitive references scattered to the children with two scans (rbsp.cpp:347-356: below = starts before the offset,
//      above = ends after it).
//
// When the last level is done the nodes are numbered in the reference's order (depth-first, below child = parent + 1,
// rbsp.cpp:235-255,390-397) from subtree sizes, bottom-up then top-down, and emitted as 8-byte RBSPNodes (rbsp.cpp:15-135).
//
// Ties: std::sort is unstable, so where several bound edges of a node have the same (t, type) the reference's order among them
// is an accident of libstdc++'s introsort; the radix sort here is stable (reference order of the parent).  The sequence of
// (t, type) values — all the SAH sees — is the same, so split planes, child sets and node arrays come out identical; only the
// order of the primitive indices INSIDE a leaf can differ (tests compare leaves as sets, nodes byte for byte).
//
// The functors below are the kernels: build_rbsp.cu launches each as one grid (for_each), tests/hostbuild runs the same
// functors in loops on the host (test infrastructure) so the whole builder is pinned on the CPU container too.
#pragma once
#include <stdint.h>
#include <stddef.h>

#include "rbsp_core.cuh"

namespace b200rt {
namespace rbsp {

struct BuildParams {
    uint32_t M, isectCost, traversalCost, maxPrims, maxDepth;
    float emptyBonus;
    float dirs[kMaxDirs * 3];
    // kd = 1: KdTreeAccel::buildTree (src/accelerators/kdtreeaccel.cpp:208-379) — the same sweep over the three axes with the
    // node's BOX instead of a k-DOP: areas in closed form (:300-305), children's boxes by replacing one bound (:362-364),
    // the same 8-byte node layout (M = 3: two flag bits, leaf = 3).  No meshes, no EvalAreas pass.
    uint32_t kd;
    // rbspkd = 1: RBSPKd::buildTree (src/accelerators/rbspKd.cpp:187-470): the axis-aligned directions (d < 3) are charged
    // kdTraversalCost instead of traversalCost; the others BSP_ALPHA * isectCost * (n - 1) + kdTraversalCost, and a second
    // arg-min over them with the plain traversalCost ("fixed") only feeds the bad-refine and leaf rules (:403-406).
    uint32_t rbspkd, kdTraversalCost;
};
enum : uint32_t { kErrFixedOnly = 2 };   // rbspKd.cpp:417-428 indexes edges[-1] when only the "fixed" search found a plane

enum : uint32_t { kLeaf = 0, kInterior = 1, kActive = 2 };
enum : uint32_t { kErrMeshOverflow = 1 };

WF_HD void atomic_min_u64(uint64_t *p, uint64_t v) {
#ifdef __CUDA_ARCH__
    atomicMin((unsigned long long *)p, (unsigned long long)v);
#else
    if (v < *p) *p = v;
#endif
}
WF_HD void atomic_min_u32(uint32_t *p, uint32_t v) {
#ifdef __CUDA_ARCH__
    atomicMin(p, v);
#else
    if (v < *p) *p = v;
#endif
}
WF_HD void atomic_max_u32(uint32_t *p, uint32_t v) {
#ifdef __CUDA_ARCH__
    atomicMax(p, v);
#else
    if (v > *p) *p = v;
#endif
}

// One tree level.  Everything except `mesh` and the sorted arrays lives until the tree is emitted.
struct Level {
    uint32_t K = 0, N = 0;   // nodes, primitive references
    uint32_t *nPrims = nullptr, *depth = nullptr, *bad = nullptr, *refBase = nullptr;
    float *area = nullptr, *bounds = nullptr;   // bounds: K x M x (min, max)
    KdopMesh *mesh = nullptr;
    uint32_t *kind = nullptr, *bestD = nullptr, *bestOff = nullptr, *childBelow = nullptr;
    float *split = nullptr;
    uint64_t *best = nullptr, *bestFixed = nullptr;
    uint32_t *size = nullptr, *idxCount = nullptr, *pre = nullptr, *off = nullptr;
    uint32_t *refPrim = nullptr, *refNode = nullptr;
};

struct SortedDirs {
    const uint64_t *key[kMaxDirs];
    const uint32_t *ref[kMaxDirs];
};

// ---- kernels ------------------------------------------------------------------------------------------------------------

struct PrimBoundsF {   // allPrimBounds, rbsp.cpp:189-201
    const float *tris;
    float *primB;
    BuildParams P;
    WF_HD void operator()(size_t i) const {
        const size_t p = i / P.M, d = i % P.M;
        float mn, mx;
        if (P.kd) {   // Triangle::WorldBound, triangle.cpp:180-186: Union(Bounds3f(p0, p1), p2), one axis of it
            const float *v = tris + 9 * p + d;
            mn = smin(smin(v[0], v[3]), v[6]);
            mx = smax(smax(v[0], v[3]), v[6]);
        } else {
            tri_extent(tris + 9 * p, P.dirs + 3 * d, mn, mx);
        }
        primB[2 * i] = mn;
        primB[2 * i + 1] = mx;
    }
};

struct RootBoundsF {   // rootNodeMBounds, rbsp.cpp:198
    const float *primB;
    uint32_t *ord;
    uint32_t M;
    WF_HD void operator()(size_t i) const {
        const size_t d = i % M;
        atomic_min_u32(ord + 2 * d, float_order(primB[2 * i]));
        atomic_max_u32(ord + 2 * d + 1, float_order(primB[2 * i + 1]));
    }
};

struct IotaF {
    uint32_t *a;
    WF_HD void operator()(size_t i) const { a[i] = (uint32_t)i; }
};

struct FillF {
    uint32_t *a;
    uint32_t v;
    WF_HD void operator()(size_t i) const { a[i] = v; }
};

struct InitNodeF {   // termination criteria, rbsp.cpp:258
    Level L;
    BuildParams P;
    WF_HD void operator()(size_t k) const {
        L.kind[k] = (L.nPrims[k] <= P.maxPrims || L.depth[k] == 0) ? kLeaf : kActive;
        L.best[k] = ~0ull;
        L.bestFixed[k] = ~0ull;
        if (P.kd) {   // Bounds3::SurfaceArea, geometry.h:946-949
            const float *b = L.bounds + (size_t)k * 6;
            const float dx = b[1] - b[0], dy = b[3] - b[2], dz = b[5] - b[4];
            L.area[k] = 2 * (dx * dy + dx * dz + dy * dz);
        }
    }
};

struct MakeKeysF {   // the bound edges of rbsp.cpp:282-287 for every node of the level, as radix-sortable keys
    Level L;
    const float *primB;
    uint32_t M, d;
    uint64_t *keys;
    uint32_t *vals;
    WF_HD void operator()(size_t i) const {
        const uint32_t r = (uint32_t)(i >> 1), e = (uint32_t)(i & 1);
        const float t = primB[((size_t)L.refPrim[r] * M + d) * 2 + e];
        keys[i] = ((uint64_t)L.refNode[r] << 33) | ((uint64_t)float_order(t + 0.0f) << 1) | e;   // Start (0) before End (1)
        vals[i] = r;
    }
};

struct EdgeFlagsF {
    const uint64_t *skey;
    uint32_t *isEnd, *lead;
    WF_HD void operator()(size_t i) const {
        isEnd[i] = (uint32_t)(skey[i] & 1);
        lead[i] = (i == 0 || (skey[i - 1] >> 1) != (skey[i] >> 1)) ? (uint32_t)i : 0u;   // first edge of a run of equal (node, t)
    }
};

struct EvalAreasF {   // rbsp.cpp:301-311 for one distinct plane
    Level L;
    const float *primB;
    BuildParams P;
    uint32_t d;
    const uint64_t *skey;
    const uint32_t *sref, *leader;
    float *areaB, *areaA;
    uint32_t *err;
    WF_HD void operator()(size_t i) const {
        if (leader[i] != (uint32_t)i) return;
        const uint32_t k = (uint32_t)(skey[i] >> 33);
        if (L.kind[k] != kActive) return;
        const float t = primB[((size_t)L.refPrim[sref[i]] * P.M + d) * 2 + (skey[i] & 1)];
        const float *nb = L.bounds + ((size_t)k * P.M + d) * 2;
        if (!(t > nb[0] && t < nb[1])) { areaB[i] = -1.0f; return; }
        KdopMesh left, right;
        FacePoints fp;
        if (!kdop_cut(L.mesh[k], P.M, t, P.dirs + 3 * d, d, left, right, fp)) *err = kErrMeshOverflow;
        areaB[i] = kdop_area(left, P.dirs, P.M);
        areaA[i] = kdop_area(right, P.dirs, P.M);
    }
};

struct EvalCostF {   // rbsp.cpp:296-330 / kdtreeaccel.cpp:290-322 for one bound edge
    Level L;
    BuildParams P;
    uint32_t d;
    const uint64_t *skey;
    const uint32_t *sref, *leader, *endsBefore;
    const float *areaB, *areaA, *primB;
    WF_HD void operator()(size_t i) const {
        const uint32_t k = (uint32_t)(skey[i] >> 33);
        if (L.kind[k] != kActive) return;
        float aB, aA;
        if (P.kd) {
            const float t = primB[((size_t)L.refPrim[sref[i]] * 3 + d) * 2 + (skey[i] & 1)];
            const float *b = L.bounds + (size_t)k * 6;
            if (!(t > b[2 * d] && t < b[2 * d + 1])) return;
            const uint32_t o0 = (d + 1) % 3, o1 = (d + 2) % 3;
            const float d0 = b[2 * o0 + 1] - b[2 * o0], d1 = b[2 * o1 + 1] - b[2 * o1];
            aB = 2 * (d0 * d1 + (t - b[2 * d]) * (d0 + d1));
            aA = 2 * (d0 * d1 + (b[2 * d + 1] - t) * (d0 + d1));
        } else {
            const uint32_t ld = leader[i];
            aB = areaB[ld];
            if (aB == -1.0f) return;   // the plane is not strictly inside the node
            aA = areaA[ld];
        }
        const uint32_t s = 2 * L.refBase[k], il = (uint32_t)i - s, e = (uint32_t)(skey[i] & 1);
        const uint32_t eb = endsBefore[i] - endsBefore[s];
        const uint32_t nAbove = L.nPrims[k] - eb - e, nBelow = il - eb;
        const float invTotalSA = 1 / L.area[k];
        float cost;
        if (!P.rbspkd) {
            cost = sah_cost(P.traversalCost, P.isectCost, P.emptyBonus, aB, aA, invTotalSA, nBelow, nAbove);
        } else if (d < 3) {   // rbspKd.cpp:326-330
            cost = sah_cost(P.kdTraversalCost, P.isectCost, P.emptyBonus, aB, aA, invTotalSA, nBelow, nAbove);
        } else {              // rbspKd.cpp:379-385
            const float pBelow = aB * invTotalSA, pAbove = aA * invTotalSA;
            const float eb = (nAbove == 0 || nBelow == 0) ? P.emptyBonus : 0;
            const float costIntersection = P.isectCost * (1 - eb) * (pBelow * nBelow + pAbove * nAbove);
            const float costFixed = P.traversalCost + costIntersection;
            const float BSP_ALPHA = 0.1;
            cost = BSP_ALPHA * P.isectCost * (L.nPrims[k] - 1) + P.kdTraversalCost + costIntersection;
            if (costFixed < wf_inf())
                atomic_min_u64(L.bestFixed + k, ((uint64_t)float_order(costFixed + 0.0f) << 32) | ((uint64_t)d << 28) | il);
        }
        if (!(cost < wf_inf())) return;   // `cost < bestCost` is false for NaN and for +inf
        atomic_min_u64(L.best + k, ((uint64_t)float_order(cost + 0.0f) << 32) | ((uint64_t)d << 28) | il);
    }
};

struct DecideF {   // rbsp.cpp:335-344
    Level L;
    BuildParams P;
    uint32_t *interior;   // K + 1 flags (the last one 0) for the scan that numbers the children
    uint32_t *err;
    WF_HD void operator()(size_t k) const {
        interior[k] = 0;
        if (k >= L.K || L.kind[k] != kActive) return;
        const uint64_t b = L.best[k];
        const bool none = b == ~0ull;
        const uint32_t n = L.nPrims[k];
        const float oldCost = P.isectCost * float(n);
        const float bestCost = none ? wf_inf() : float_unorder((uint32_t)(b >> 32));
        uint32_t bad = L.bad[k];
        if (P.rbspkd) {   // rbspKd.cpp:403-411
            const uint64_t bf = L.bestFixed[k];
            const bool noneFixed = bf == ~0ull;
            const float bestCostFixed = noneFixed ? wf_inf() : float_unorder((uint32_t)(bf >> 32));
            if (bestCost > oldCost && bestCostFixed > oldCost) ++bad;
            if ((bestCost > 4 * oldCost && bestCostFixed > 4 * oldCost && n < 16) || (none && noneFixed) || bad == 3) { L.kind[k] = kLeaf; return; }
            if (none) { *err = kErrFixedOnly; L.kind[k] = kLeaf; return; }
        } else {
            if (bestCost > oldCost) ++bad;
            if ((bestCost > 4 * oldCost && n < 16) || none || bad == 3) { L.kind[k] = kLeaf; return; }
        }
        L.kind[k] = kInterior;
        L.bad[k] = bad;
        L.bestD[k] = (uint32_t)(b >> 28) & 15u;
        L.bestOff[k] = (uint32_t)b & 0x0fffffffu;
        interior[k] = 1;
    }
};

struct SplitF {   // the winning cut again: children's meshes, areas, extents (rbsp.cpp:358-395)
    Level L, Nx;
    const float *primB;
    BuildParams P;
    SortedDirs S;
    const uint32_t *rank;
    uint32_t *err;
    WF_HD void operator()(size_t k) const {
        if (L.kind[k] != kInterior) return;
        const uint32_t c = 2 * rank[k], d = L.bestD[k];
        const size_t pos = (size_t)2 * L.refBase[k] + L.bestOff[k];
        const float t = primB[((size_t)L.refPrim[S.ref[d][pos]] * P.M + d) * 2 + (S.key[d][pos] & 1)];
        Nx.depth[c] = Nx.depth[c + 1] = L.depth[k] - 1;
        Nx.bad[c] = Nx.bad[c + 1] = L.bad[k];
        L.split[k] = t;
        L.childBelow[k] = c;
        if (P.kd) {
            for (uint32_t j = 0; j < 6; ++j) Nx.bounds[(size_t)c * 6 + j] = Nx.bounds[(size_t)(c + 1) * 6 + j] = L.bounds[(size_t)k * 6 + j];
            Nx.bounds[(size_t)c * 6 + 2 * d + 1] = t;         // bounds0.pMax[axis] = tSplit
            Nx.bounds[(size_t)(c + 1) * 6 + 2 * d] = t;       // bounds1.pMin[axis] = tSplit
            return;
        }
        KdopMesh &left = Nx.mesh[c], &right = Nx.mesh[c + 1];
        FacePoints fp;
        if (!kdop_cut(L.mesh[k], P.M, t, P.dirs + 3 * d, d, left, right, fp)) *err = kErrMeshOverflow;
        Nx.area[c] = kdop_area(left, P.dirs, P.M);
        Nx.area[c + 1] = kdop_area(right, P.dirs, P.M);
        for (uint32_t dd = 0; dd < P.M; ++dd) {
            kdop_extent(left, P.dirs + 3 * dd, Nx.bounds[((size_t)c * P.M + dd) * 2], Nx.bounds[((size_t)c * P.M + dd) * 2 + 1]);
            kdop_extent(right, P.dirs + 3 * dd, Nx.bounds[((size_t)(c + 1) * P.M + dd) * 2], Nx.bounds[((size_t)(c + 1) * P.M + dd) * 2 + 1]);
        }
    }
};

struct RefFlagsF {   // rbsp.cpp:347-356 as two flag arrays over the winning direction's sorted edges (2N + 1 entries)
    Level L;
    SortedDirs S;
    uint32_t *fB, *fA;
    WF_HD void operator()(size_t p) const {
        uint32_t b = 0, a = 0;
        if (p < (size_t)2 * L.N) {
            const uint32_t k = L.refNode[p >> 1];
            if (L.kind[k] == kInterior) {
                const uint32_t e = (uint32_t)(S.key[L.bestD[k]][p] & 1), il = (uint32_t)p - 2 * L.refBase[k];
                b = (!e && il < L.bestOff[k]) ? 1u : 0u;
                a = (e && il > L.bestOff[k]) ? 1u : 0u;
            }
        }
        fB[p] = b;
        fA[p] = a;
    }
};

struct ChildCountsF {
    Level L, Nx;
    const uint32_t *rB, *rA;
    WF_HD void operator()(size_t k) const {
        if (L.kind[k] != kInterior) return;
        const size_t s = (size_t)2 * L.refBase[k], e = s + (size_t)2 * L.nPrims[k];
        const uint32_t n0 = rB[e] - rB[s], n1 = rA[e] - rA[s], base0 = rB[s] + rA[s], c = L.childBelow[k];
        Nx.nPrims[c] = n0;
        Nx.refBase[c] = base0;
        Nx.nPrims[c + 1] = n1;
        Nx.refBase[c + 1] = base0 + n0;
    }
};

struct ScatterRefsF {
    Level L, Nx;
    SortedDirs S;
    const uint32_t *fB, *fA, *rB, *rA;
    WF_HD void operator()(size_t p) const {
        if (!fB[p] && !fA[p]) return;
        const uint32_t k = L.refNode[p >> 1], c = L.childBelow[k];
        const size_t s = (size_t)2 * L.refBase[k];
        const uint32_t prim = L.refPrim[S.ref[L.bestD[k]][p]];
        if (fB[p]) {
            const uint32_t dst = Nx.refBase[c] + (rB[p] - rB[s]);
            Nx.refPrim[dst] = prim;
            Nx.refNode[dst] = c;
        } else {
            const uint32_t dst = Nx.refBase[c + 1] + (rA[p] - rA[s]);
            Nx.refPrim[dst] = prim;
            Nx.refNode[dst] = c + 1;
        }
    }
};

struct SubtreeSizeF {   // bottom-up: nodes and primitiveIndices entries of each subtree
    Level L, Nx;
    WF_HD void operator()(size_t k) const {
        if (L.kind[k] == kInterior) {
            const uint32_t c = L.childBelow[k];
            L.size[k] = 1 + Nx.size[c] + Nx.size[c + 1];
            L.idxCount[k] = Nx.idxCount[c] + Nx.idxCount[c + 1];
        } else {
            L.size[k] = 1;
            L.idxCount[k] = L.nPrims[k] > 1 ? L.nPrims[k] : 0;   // InitLeaf pushes indices only for np > 1 (rbsp.cpp:137-150)
        }
    }
};

struct AssignF {   // top-down: the reference's depth-first numbering — below child right after its parent, above child after
                   // the whole below subtree (rbsp.cpp:239-240,390-397)
    Level L, Nx;
    WF_HD void operator()(size_t k) const {
        if (L.kind[k] != kInterior) return;
        const uint32_t c = L.childBelow[k];
        Nx.pre[c] = L.pre[k] + 1;
        Nx.off[c] = L.off[k];
        Nx.pre[c + 1] = L.pre[k] + 1 + Nx.size[c];
        Nx.off[c + 1] = L.off[k] + Nx.idxCount[c];
    }
};

struct EmitF {   // RBSPNode::InitInterior / setAboveChild / InitLeaf, rbsp.cpp:19-27,137-150
    Level L, Nx;
    uint32_t M;
    uint32_t *nodes;     // 2 words per node
    uint32_t *indices;   // primitiveIndices
    WF_HD void operator()(size_t k) const {
        const uint32_t bo = bit_offset(M);
        uint32_t w0, w1;
        if (L.kind[k] == kInterior) {
            w0 = f2u(L.split[k]);
            w1 = L.bestD[k] | (Nx.pre[L.childBelow[k] + 1] << bo);
        } else {
            const uint32_t n = L.nPrims[k], base = L.refBase[k];
            w1 = M | (n << bo);
            if (n == 0) w0 = 0;
            else if (n == 1) w0 = L.refPrim[base];
            else {
                w0 = L.off[k];
                for (uint32_t i = 0; i < n; ++i) indices[w0 + i] = L.refPrim[base + i];
            }
        }
        nodes[2 * (size_t)L.pre[k]] = w0;
        nodes[2 * (size_t)L.pre[k] + 1] = w1;
    }
};

}  // namespace rbsp
}  // namespace b200rt
