// Level loop of the RBSP builder over a backend that provides memory, for_each (one kernel launch per functor), a stable
// radix sort of (u64 key, u32 value) pairs and two scans.  build_rbsp.cu instantiates it with the CUDA backend (the product);
// tests/hostbuild instantiates it with plain loops (test infrastructure).  See rbsp_build.cuh for the algorithm.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "rbsp_build.cuh"

namespace b200rt {
namespace rbsp {

// getDirections, RBSPShared.h:30-76 (Normalize = v * (1 / Length), geometry.h:281-286,1323-1325).  Returns M or 0.
inline uint32_t make_directions(uint32_t N, float *dirs) {
    if (N != 3 && N != 7 && N != 9 && N != 13) return 0;
    uint32_t n = 0;
    auto push = [&](float x, float y, float z, bool normalize) {
        if (normalize) {
            const float len = std::sqrt(x * x + y * y + z * z);
            const float inv = (float)1 / len;
            x *= inv; y *= inv; z *= inv;
        }
        dirs[3 * n] = x; dirs[3 * n + 1] = y; dirs[3 * n + 2] = z;
        ++n;
    };
    push(1, 0, 0, false); push(0, 1, 0, false); push(0, 0, 1, false);
    if (N == 7 || N == 13) { push(1, 1, 1, true); push(1, -1, 1, true); push(1, 1, -1, true); push(1, -1, -1, true); }
    if (N == 9 || N == 13) {
        push(1, 1, 0, true); push(1, 0, 1, true); push(0, 1, 1, true);
        push(1, -1, 0, true); push(1, 0, -1, true); push(0, 1, -1, true);
    }
    return n;
}

// calculateMaxDepth, geometry.h:1845-1847: round(2 + 1.6f * Log2Int(N))
inline uint32_t default_max_depth(int64_t n) {
    int lg = 0;
    while ((int64_t(1) << (lg + 1)) <= n) ++lg;
    return (uint32_t)std::round(2 + 1.6f * lg);
}

struct BuildResult {
    std::vector<uint32_t> nodes;     // 2 words per node
    std::vector<uint32_t> indices;   // primitiveIndices
    uint32_t levels = 0;
    uint64_t candidates = 0;         // bound edges looked at (2 n M summed over the levels)
    uint32_t error = 0;
};

template <class B>
void alloc_level(B &be, Level &L, uint32_t K, uint32_t M, bool meshes = true) {
    L.K = K;
    L.nPrims = be.template alloc<uint32_t>(K); L.depth = be.template alloc<uint32_t>(K);
    L.bad = be.template alloc<uint32_t>(K); L.refBase = be.template alloc<uint32_t>(K);
    L.area = be.template alloc<float>(K); L.bounds = be.template alloc<float>((size_t)K * M * 2);
    L.mesh = be.template alloc<KdopMesh>(meshes ? K : 0);
    L.kind = be.template alloc<uint32_t>(K); L.bestD = be.template alloc<uint32_t>(K);
    L.bestOff = be.template alloc<uint32_t>(K); L.childBelow = be.template alloc<uint32_t>(K);
    L.split = be.template alloc<float>(K); L.best = be.template alloc<uint64_t>(K); L.bestFixed = be.template alloc<uint64_t>(K);
    L.size = be.template alloc<uint32_t>(K); L.idxCount = be.template alloc<uint32_t>(K);
    L.pre = be.template alloc<uint32_t>(K); L.off = be.template alloc<uint32_t>(K);
}
template <class B>
void free_level(B &be, Level &L) {
    void *ps[] = {L.nPrims, L.depth, L.bad, L.refBase, L.area, L.bounds, L.mesh, L.kind, L.bestD, L.bestOff, L.childBelow,
                  L.split, L.best, L.bestFixed, L.size, L.idxCount, L.pre, L.off, L.refPrim, L.refNode};
    for (void *p : ps) be.free(p);
    L = Level();
}

// tris: n x 9 floats in BACKEND memory; world: the union of the triangles' boxes (host).
template <class B>
BuildResult build(B &be, const float *tris, uint32_t n, const BuildParams &P, const float world[6]) {
    BuildResult out;
    const uint32_t M = P.M;
    float *primB = be.template alloc<float>((size_t)n * M * 2);
    be.for_each((size_t)n * M, PrimBoundsF{tris, primB, P});
    uint32_t *err = be.template alloc<uint32_t>(1);
    be.for_each(1, FillF{err, 0u});

    // root: extents of all primitives per direction, the world box as a k-DOP, its area (rbsp.cpp:183-229)
    std::vector<Level> levels(1);
    Level *cur = &levels[0];
    alloc_level(be, *cur, 1, M, !P.kd);
    {
        uint32_t *ord = be.template alloc<uint32_t>(2 * M);
        std::vector<uint32_t> init(2 * M);
        for (uint32_t d = 0; d < M; ++d) { init[2 * d] = 0xffffffffu; init[2 * d + 1] = 0u; }
        be.upload(ord, init.data(), 2 * M);
        be.for_each((size_t)n * M, RootBoundsF{primB, ord, M});
        be.download(init.data(), ord, 2 * M);
        be.free(ord);
        std::vector<float> rb(2 * M);
        for (uint32_t i = 0; i < 2 * M; ++i) rb[i] = float_unorder(init[i]);
        if (n == 0) for (uint32_t d = 0; d < M; ++d) { rb[2 * d] = 3.402823466e+38f; rb[2 * d + 1] = -3.402823466e+38f; }
        if (P.kd) for (uint32_t d = 0; d < 3; ++d) { rb[2 * d] = world[d]; rb[2 * d + 1] = world[3 + d]; }   // the root's box
        be.upload(cur->bounds, rb.data(), 2 * M);
        if (!P.kd) {
            KdopMesh root;
            kdop_from_box(world, root);
            const float area = kdop_area(root, P.dirs, M);   // re-orients the root's edges, and the build keeps them so
            be.upload(cur->mesh, &root, 1);
            be.upload(cur->area, &area, 1);
        }
        const uint32_t one[4] = {n, P.maxDepth, 0u, 0u};
        be.upload(cur->nPrims, one + 0, 1); be.upload(cur->depth, one + 1, 1);
        be.upload(cur->bad, one + 2, 1); be.upload(cur->refBase, one + 3, 1);
        cur->N = n;
        cur->refPrim = be.template alloc<uint32_t>(n);
        cur->refNode = be.template alloc<uint32_t>(n);
        be.for_each(n, IotaF{cur->refPrim});
        be.for_each(n, FillF{cur->refNode, 0u});
    }

    for (;;) {
        Level &L = levels.back();
        const size_t E = (size_t)2 * L.N;
        be.for_each(L.K, InitNodeF{L, P});
        int nodeBits = 1;
        while ((1ull << nodeBits) < (uint64_t)L.K) ++nodeBits;
        // per direction: sort the level's bound edges, evaluate the candidate planes
        std::vector<uint64_t *> skey(M, nullptr);
        std::vector<uint32_t *> sref(M, nullptr);
        uint64_t *keyTmp = be.template alloc<uint64_t>(E);
        uint32_t *valTmp = be.template alloc<uint32_t>(E);
        uint32_t *isEnd = be.template alloc<uint32_t>(E), *lead = be.template alloc<uint32_t>(E);
        uint32_t *endsBefore = be.template alloc<uint32_t>(E), *leader = be.template alloc<uint32_t>(E);
        float *areaB = be.template alloc<float>(E), *areaA = be.template alloc<float>(E);
        for (uint32_t d = 0; d < M && E; ++d) {
            skey[d] = be.template alloc<uint64_t>(E);
            sref[d] = be.template alloc<uint32_t>(E);
            be.for_each(E, MakeKeysF{L, primB, M, d, keyTmp, valTmp});
            be.sort_pairs(keyTmp, skey[d], valTmp, sref[d], E, 33 + nodeBits);
            be.for_each(E, EdgeFlagsF{skey[d], isEnd, lead});
            be.exclusive_sum(isEnd, endsBefore, E);
            if (!P.kd) {
                be.inclusive_max(lead, leader, E);
                be.for_each_heavy(E, EvalAreasF{L, primB, P, d, skey[d], sref[d], leader, areaB, areaA, err});
            }
            be.for_each(E, EvalCostF{L, P, d, skey[d], sref[d], leader, endsBefore, areaB, areaA, primB});
            out.candidates += E;
        }
        void *tmp[] = {keyTmp, valTmp, isEnd, lead, endsBefore, leader, areaB, areaA};
        for (void *p : tmp) be.free(p);

        // leaf / interior, children numbered by a scan over the interior flags
        uint32_t *interior = be.template alloc<uint32_t>((size_t)L.K + 1), *rank = be.template alloc<uint32_t>((size_t)L.K + 1);
        be.for_each((size_t)L.K + 1, DecideF{L, P, interior, err});
        be.exclusive_sum(interior, rank, (size_t)L.K + 1);
        uint32_t nInterior = 0;
        be.download(&nInterior, rank + L.K, 1);
        be.download(&out.error, err, 1);
        be.free(interior);
        if (nInterior == 0 || out.error) {
            be.free(rank);
            for (uint32_t d = 0; d < M; ++d) { be.free(skey[d]); be.free(sref[d]); }
            be.free(L.mesh); L.mesh = nullptr;
            break;
        }
        levels.emplace_back();
        Level &Lc = levels[levels.size() - 2];   // the vector may have moved
        Level &Nx = levels.back();
        alloc_level(be, Nx, 2 * nInterior, M, !P.kd);
        SortedDirs S;
        for (uint32_t d = 0; d < kMaxDirs; ++d) { S.key[d] = d < M ? skey[d] : nullptr; S.ref[d] = d < M ? sref[d] : nullptr; }
        be.for_each_heavy(Lc.K, SplitF{Lc, Nx, primB, P, S, rank, err});
        be.free(rank);
        // primitive references of the children
        uint32_t *fB = be.template alloc<uint32_t>(E + 1), *fA = be.template alloc<uint32_t>(E + 1);
        uint32_t *rB = be.template alloc<uint32_t>(E + 1), *rA = be.template alloc<uint32_t>(E + 1);
        be.for_each(E + 1, RefFlagsF{Lc, S, fB, fA});
        be.exclusive_sum(fB, rB, E + 1);
        be.exclusive_sum(fA, rA, E + 1);
        be.for_each(Lc.K, ChildCountsF{Lc, Nx, rB, rA});
        uint32_t tot[2] = {0, 0};
        be.download(tot + 0, rB + E, 1);
        be.download(tot + 1, rA + E, 1);
        Nx.N = tot[0] + tot[1];
        Nx.refPrim = be.template alloc<uint32_t>(Nx.N);
        Nx.refNode = be.template alloc<uint32_t>(Nx.N);
        be.for_each(E, ScatterRefsF{Lc, Nx, S, fB, fA, rB, rA});
        void *tmp2[] = {fB, fA, rB, rA};
        for (void *p : tmp2) be.free(p);
        for (uint32_t d = 0; d < M; ++d) { be.free(skey[d]); be.free(sref[d]); }
        be.free(Lc.mesh); Lc.mesh = nullptr;
    }
    out.levels = (uint32_t)levels.size();
    if (!out.error) {
        // number the nodes in the reference's depth-first order and emit them
        Level none;
        for (size_t l = levels.size(); l-- > 0;) be.for_each(levels[l].K, SubtreeSizeF{levels[l], l + 1 < levels.size() ? levels[l + 1] : none});
        const uint32_t zero = 0;
        be.upload(levels[0].pre, &zero, 1);
        be.upload(levels[0].off, &zero, 1);
        for (size_t l = 0; l + 1 < levels.size(); ++l) be.for_each(levels[l].K, AssignF{levels[l], levels[l + 1]});
        uint32_t nNodes = 0, nIdx = 0;
        be.download(&nNodes, levels[0].size, 1);
        be.download(&nIdx, levels[0].idxCount, 1);
        uint32_t *dNodes = be.template alloc<uint32_t>((size_t)2 * nNodes), *dIdx = be.template alloc<uint32_t>((size_t)nIdx + 1);
        for (size_t l = 0; l < levels.size(); ++l) be.for_each(levels[l].K, EmitF{levels[l], l + 1 < levels.size() ? levels[l + 1] : none, M, dNodes, dIdx});
        out.nodes.resize((size_t)2 * nNodes);
        out.indices.resize(nIdx);
        be.download(out.nodes.data(), dNodes, (size_t)2 * nNodes);
        if (nIdx) be.download(out.indices.data(), dIdx, nIdx);
        be.free(dNodes);
        be.free(dIdx);
    }
    for (Level &L : levels) free_level(be, L);
    be.free(primB);
    be.free(err);
    return out;
}

}  // namespace rbsp
}  // namespace b200rt
