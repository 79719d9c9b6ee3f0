// Level loop of the sweep-SAH BVH builder (see bvh_build.cuh) over the same backend interface as rbsp_driver.h plus one
// collective: an inclusive box-union scan by segment key, forward or backward.
#pragma once
#include <cstdint>
#include <vector>

#include "bvh_build.cuh"

namespace b200rt {
namespace bvhb {

struct BvhResult {
    std::vector<uint32_t> nodes;   // 8 words per LinearBVHNode
    std::vector<uint32_t> order;   // primNumMapping
    uint32_t levels = 0;
    uint64_t candidates = 0;
    uint32_t error = 0;
};

template <class B>
void alloc_blevel(B &be, BLevel &L, uint32_t K) {
    L.K = K;
    L.start = be.template alloc<uint32_t>(K); L.end = be.template alloc<uint32_t>(K);
    L.kind = be.template alloc<uint32_t>(K); L.axis = be.template alloc<uint32_t>(K);
    L.bestI = be.template alloc<uint32_t>(K); L.child = be.template alloc<uint32_t>(K);
    L.box = be.template alloc<Box6>(K); L.best = be.template alloc<uint64_t>(K);
    L.size = be.template alloc<uint32_t>(K); L.pre = be.template alloc<uint32_t>(K); L.off = be.template alloc<uint32_t>(K);
}
template <class B>
void free_blevel(B &be, BLevel &L) {
    void *ps[] = {L.start, L.end, L.kind, L.axis, L.bestI, L.child, L.box, L.best, L.size, L.pre, L.off};
    for (void *p : ps) be.free(p);
    L = BLevel();
}

// tris: n x 9 floats in backend memory, n >= 1
template <class B>
BvhResult build_bvh(B &be, const float *tris, uint32_t n, const BvhParams &P, uint32_t maxLevels = 4096) {
    BvhResult out;
    Box6 *box = be.template alloc<Box6>(n);
    float *centroid = be.template alloc<float>((size_t)3 * n);
    be.for_each(n, PrimInfoF{tris, box, centroid});
    // one sort per axis for the whole scene
    uint32_t *sorted[3], *sortedNext[3];
    {
        uint64_t *kin = be.template alloc<uint64_t>(n), *kout = be.template alloc<uint64_t>(n);
        uint32_t *vin = be.template alloc<uint32_t>(n);
        for (uint32_t d = 0; d < 3; ++d) {
            sorted[d] = be.template alloc<uint32_t>(n);
            sortedNext[d] = be.template alloc<uint32_t>(n);
            be.for_each(n, CentroidKeyF{centroid, d, kin, vin});
            be.sort_pairs(kin, kout, vin, sorted[d], n, 64);
        }
        be.free(kin); be.free(kout); be.free(vin);
    }
    uint32_t *info = be.template alloc<uint32_t>(n), *infoNext = be.template alloc<uint32_t>(n);
    uint32_t *posNode = be.template alloc<uint32_t>(n), *posNodeNext = be.template alloc<uint32_t>(n);
    uint8_t *side = be.template alloc<uint8_t>(n);
    be.for_each(n, rbsp::IotaF{info});
    be.for_each(n, rbsp::FillF{posNode, 0u});
    Box6 *gathered = be.template alloc<Box6>(n), *l2r = be.template alloc<Box6>(n), *r2l = be.template alloc<Box6>((size_t)n + 1);
    uint32_t *flagA = be.template alloc<uint32_t>((size_t)n + 1), *flagB = be.template alloc<uint32_t>((size_t)n + 1);
    uint32_t *scanA = be.template alloc<uint32_t>((size_t)n + 1), *scanB = be.template alloc<uint32_t>((size_t)n + 1);
    uint32_t *pairF = be.template alloc<uint32_t>(n), *pairT = be.template alloc<uint32_t>(n);

    std::vector<BLevel> levels(1);
    alloc_blevel(be, levels[0], 1);
    {
        const uint32_t se[2] = {0u, n};
        be.upload(levels[0].start, se, 1);
        be.upload(levels[0].end, se + 1, 1);
    }
    for (;;) {
        BLevel L = levels.back();
        be.for_each(L.K, InitNodeF{L});
        for (uint32_t d = 0; d < 3; ++d) {
            be.for_each(n, GatherBoxF{sorted[d], box, gathered});
            be.scan_boxes(posNode, gathered, l2r, n, false);
            be.scan_boxes(posNode, gathered, r2l, n, true);
            if (d == 0) be.for_each(L.K, NodeBoxF{L, r2l});
            be.for_each(n, SweepCostF{L, P, d, posNode, l2r, r2l});
            out.candidates += n;
        }
        uint32_t *interior = be.template alloc<uint32_t>((size_t)L.K + 1), *rank = be.template alloc<uint32_t>((size_t)L.K + 1);
        be.for_each((size_t)L.K + 1, DecideF{L, P, interior});
        be.exclusive_sum(interior, rank, (size_t)L.K + 1);
        uint32_t nInterior = 0;
        be.download(&nInterior, rank + L.K, 1);
        be.free(interior);
        if (nInterior == 0 || levels.size() >= maxLevels) {
            be.free(rank);
            if (nInterior) out.error = 1;
            break;
        }
        levels.emplace_back();
        BLevel &Nx = levels.back();
        alloc_blevel(be, Nx, 2 * nInterior);
        be.for_each(L.K, ChildrenF{L, Nx, rank});
        be.free(rank);
        be.for_each(n, SideF{L, posNode, sorted[0], sorted[1], sorted[2], side});
        // each axis' sorted runs split into the children's runs
        for (uint32_t d = 0; d < 3; ++d) {
            be.for_each((size_t)n + 1, LeftFlagF{L, posNode, sorted[d], side, n, flagA});
            be.exclusive_sum(flagA, scanA, (size_t)n + 1);
            be.for_each(n, StablePartitionF{L, posNode, sorted[d], scanA, side, sortedNext[d]});
            std::swap(sorted[d], sortedNext[d]);
        }
        // primitiveInfo re-ordered as std::partition leaves it
        be.for_each((size_t)n + 1, MisplacedFlagF{L, posNode, info, side, n, flagA, flagB});
        be.exclusive_sum(flagA, scanA, (size_t)n + 1);
        be.exclusive_sum(flagB, scanB, (size_t)n + 1);
        uint32_t nPairs = 0;
        be.download(&nPairs, scanA + n, 1);
        be.for_each(n, PairUpF{L, posNode, flagA, flagB, scanA, scanB, pairF, pairT});
        be.for_each(n, CopyU32F{info, infoNext});
        be.for_each(nPairs, SwapPairsF{pairF, pairT, info, infoNext});
        std::swap(info, infoNext);
        be.for_each(n, NextPosNodeF{L, posNode, posNodeNext});
        std::swap(posNode, posNodeNext);
    }
    out.levels = (uint32_t)levels.size();
    if (!out.error) {
        BLevel none;
        for (size_t l = levels.size(); l-- > 0;) be.for_each(levels[l].K, SubtreeSizeF{levels[l], l + 1 < levels.size() ? levels[l + 1] : none});
        const uint32_t zero = 0;
        be.upload(levels[0].pre, &zero, 1);
        be.upload(levels[0].off, &zero, 1);
        for (size_t l = 0; l + 1 < levels.size(); ++l) be.for_each(levels[l].K, AssignF{levels[l], levels[l + 1]});
        uint32_t nNodes = 0;
        be.download(&nNodes, levels[0].size, 1);
        uint32_t *dNodes = be.template alloc<uint32_t>((size_t)8 * nNodes), *dOrder = be.template alloc<uint32_t>(n);
        for (size_t l = 0; l < levels.size(); ++l) be.for_each(levels[l].K, EmitF{levels[l], l + 1 < levels.size() ? levels[l + 1] : none, info, dNodes, dOrder});
        out.nodes.resize((size_t)8 * nNodes);
        out.order.resize(n);
        be.download(out.nodes.data(), dNodes, (size_t)8 * nNodes);
        be.download(out.order.data(), dOrder, n);
        be.free(dNodes);
        be.free(dOrder);
    }
    for (BLevel &L : levels) free_blevel(be, L);
    void *ps[] = {box, centroid, sorted[0], sorted[1], sorted[2], sortedNext[0], sortedNext[1], sortedNext[2], info, infoNext, posNode,
                  posNodeNext, side, gathered, l2r, r2l, flagA, flagB, scanA, scanB, pairF, pairT};
    for (void *p : ps) be.free(p);
    return out;
}

}  // namespace bvhb
}  // namespace b200rt
