// TEST INFRASTRUCTURE ONLY.  The RBSP builder's kernels (thesis-pbrt-v3_b200/csrc/rbsp_build.cuh: the functors the CUDA
// backend launches as grids) run here in plain host loops, over the same level loop (rbsp_driver.h), so that the builder is
// pinned against the reference's own trees on the CPU container.  The product (libb200rt.so) has no such host path.
#include <algorithm>
#include <cstdlib>
#include <numeric>

#include "rbsp_driver.h"
#include "bvh_driver.h"

using namespace b200rt::rbsp;

namespace {
struct HostBackend {
    template <class T> T *alloc(size_t n) { return (T *)std::malloc(std::max<size_t>(n, 1) * sizeof(T)); }
    void free(void *p) { std::free(p); }
    template <class T> void upload(T *dst, const T *src, size_t n) { if (n) std::memcpy(dst, src, n * sizeof(T)); }
    template <class T> void download(T *dst, const T *src, size_t n) { if (n) std::memcpy(dst, src, n * sizeof(T)); }
    template <class F> void for_each(size_t n, F f) { for (size_t i = 0; i < n; ++i) f(i); }
    template <class F> void for_each_heavy(size_t n, F f) { for (size_t i = 0; i < n; ++i) f(i); }
    void sort_pairs(const uint64_t *kin, uint64_t *kout, const uint32_t *vin, uint32_t *vout, size_t n, int bits) {
        std::vector<uint32_t> idx(n);
        std::iota(idx.begin(), idx.end(), 0u);
        const uint64_t mask = bits >= 64 ? ~0ull : ((1ull << bits) - 1);
        std::stable_sort(idx.begin(), idx.end(), [&](uint32_t a, uint32_t b) { return (kin[a] & mask) < (kin[b] & mask); });
        for (size_t i = 0; i < n; ++i) { kout[i] = kin[idx[i]]; vout[i] = vin[idx[i]]; }
    }
    void exclusive_sum(const uint32_t *in, uint32_t *out, size_t n) {
        uint32_t s = 0;
        for (size_t i = 0; i < n; ++i) { const uint32_t v = in[i]; out[i] = s; s += v; }
    }
    void scan_boxes(const uint32_t *keys, const b200rt::bvhb::Box6 *in, b200rt::bvhb::Box6 *out, size_t n, bool reverse) {
        b200rt::bvhb::BoxUnion op;
        for (size_t j = 0; j < n; ++j) {
            const size_t i = reverse ? n - 1 - j : j, prev = reverse ? i + 1 : i - 1;
            out[i] = (j > 0 && keys[prev] == keys[i]) ? op(out[prev], in[i]) : in[i];
        }
    }
    void inclusive_max(const uint32_t *in, uint32_t *out, size_t n) {
        uint32_t m = 0;
        for (size_t i = 0; i < n; ++i) { m = std::max(m, in[i]); out[i] = m; }
    }
};
}  // namespace

// nbDirections = 0 selects the kd-tree mode (KdTreeAccel::buildTree); nbDirections + 100 the RBSPKd mode (kdtraversalcost 1)
extern "C" int hb_build_rbsp(const float *tris, uint32_t n, uint32_t nbDirections, int isectCost, int traversalCost,
                             float emptyBonus, int maxPrims, int maxDepth, uint32_t **nodes, uint32_t *nNodes,
                             uint32_t **indices, uint32_t *nIndices, float *dirsOut, float *boundsOut, uint64_t *candidates) {
    BuildParams P{};
    P.kd = nbDirections == 0;
    if (nbDirections >= 100) { P.rbspkd = 1; P.kdTraversalCost = 1; nbDirections -= 100; }
    P.M = make_directions(P.kd ? 3 : nbDirections, P.dirs);
    if (!P.M) return 1;
    P.isectCost = (uint32_t)isectCost; P.traversalCost = (uint32_t)traversalCost; P.emptyBonus = emptyBonus;
    P.maxPrims = (uint32_t)maxPrims;
    P.maxDepth = maxDepth < 0 ? default_max_depth(n) : (uint32_t)maxDepth;
    float world[6] = {3.402823466e+38f, 3.402823466e+38f, 3.402823466e+38f, -3.402823466e+38f, -3.402823466e+38f, -3.402823466e+38f};
    for (uint32_t p = 0; p < n; ++p)
        for (int v = 0; v < 3; ++v)
            for (int c = 0; c < 3; ++c) {
                world[c] = std::min(world[c], tris[9 * p + 3 * v + c]);
                world[3 + c] = std::max(world[3 + c], tris[9 * p + 3 * v + c]);
            }
    HostBackend be;
    BuildResult r = build(be, tris, n, P, world);
    if (r.error) return 2;
    *nNodes = (uint32_t)(r.nodes.size() / 2);
    *nIndices = (uint32_t)r.indices.size();
    *nodes = (uint32_t *)std::malloc(std::max<size_t>(r.nodes.size(), 1) * 4);
    *indices = (uint32_t *)std::malloc(std::max<size_t>(r.indices.size(), 1) * 4);
    std::memcpy(*nodes, r.nodes.data(), r.nodes.size() * 4);
    std::memcpy(*indices, r.indices.data(), r.indices.size() * 4);
    std::memcpy(dirsOut, P.dirs, P.M * 12);
    std::memcpy(boundsOut, world, 24);
    if (candidates) *candidates = r.candidates;
    return 0;
}
extern "C" void hb_free(void *p) { std::free(p); }

extern "C" int hb_build_bvh(const float *tris, uint32_t n, int maxNodePrims, int isectCost, int traversalCost,
                            uint32_t **nodes, uint32_t *nNodes, uint32_t **order) {
    using namespace b200rt::bvhb;
    HostBackend be;
    BvhParams P{std::min(255, maxNodePrims), isectCost, traversalCost};
    BvhResult r = build_bvh(be, tris, n, P);
    if (r.error) return 2;
    *nNodes = (uint32_t)(r.nodes.size() / 8);
    *nodes = (uint32_t *)std::malloc(std::max<size_t>(r.nodes.size(), 1) * 4);
    *order = (uint32_t *)std::malloc(std::max<size_t>(r.order.size(), 1) * 4);
    std::memcpy(*nodes, r.nodes.data(), r.nodes.size() * 4);
    std::memcpy(*order, r.order.data(), r.order.size() * 4);
    return 0;
}
