// b200rt_build_rbsp: the RBSP tree builder on the B200 (include/b200rt.h).  The algorithm and its kernels (functors) are in
// rbsp_build.cuh, the level loop in rbsp_driver.h; this file is the CUDA backend — one grid per functor, CUB's radix sort and
// scans for the two collective primitives, stream-ordered allocation — and the C entry point.  There is no host fallback:
// without a device the call fails with B200RT_ENODEVICE.
#include <cuda_runtime.h>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <algorithm>
#include <atomic>
#include <cstdlib>

#include "../../include/b200rt.h"
#include <thrust/iterator/reverse_iterator.h>

#include "bvh_driver.h"
#include "rbsp_driver.h"
#include "scene_internal.h"

extern std::atomic<uint64_t> g_launches;   // b200rt.cu

namespace {
using namespace b200rt::rbsp;

template <class F>
__global__ void k_for_each(F f, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) f(i);
}
// the k-DOP functors keep two meshes and the per-face cut points in local memory (~9 KB per thread): small blocks
template <class F>
__global__ void __launch_bounds__(64) k_for_each_heavy(F f, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) f(i);
}

struct MaxU32 {
    __host__ __device__ uint32_t operator()(uint32_t a, uint32_t b) const { return a > b ? a : b; }
};

struct CudaBackend {
    cudaStream_t st = nullptr;
    cudaError_t status = cudaSuccess;
    uint64_t launches = 0;
    void check(cudaError_t e) { if (e != cudaSuccess && status == cudaSuccess) status = e; }
    template <class T> T *alloc(size_t n) {
        void *p = nullptr;
        check(cudaMallocAsync(&p, std::max<size_t>(n, 1) * sizeof(T), st));
        return (T *)p;
    }
    void free(void *p) { if (p) check(cudaFreeAsync(p, st)); }
    template <class T> void upload(T *dst, const T *src, size_t n) {
        if (!n || status != cudaSuccess) return;
        check(cudaMemcpyAsync(dst, src, n * sizeof(T), cudaMemcpyHostToDevice, st));
        check(cudaStreamSynchronize(st));   // src may be a temporary of the caller
    }
    template <class T> void download(T *dst, const T *src, size_t n) {
        if (!n || status != cudaSuccess) return;
        check(cudaMemcpyAsync(dst, src, n * sizeof(T), cudaMemcpyDeviceToHost, st));
        check(cudaStreamSynchronize(st));
    }
    template <class F> void for_each(size_t n, F f) {
        if (!n || status != cudaSuccess) return;
        k_for_each<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(f, n);
        check(cudaGetLastError());
        ++launches;
    }
    template <class F> void for_each_heavy(size_t n, F f) {
        if (!n || status != cudaSuccess) return;
        k_for_each_heavy<<<(unsigned)((n + 63) / 64), 64, 0, st>>>(f, n);
        check(cudaGetLastError());
        ++launches;
    }
    void sort_pairs(const uint64_t *kin, uint64_t *kout, const uint32_t *vin, uint32_t *vout, size_t n, int bits) {
        if (!n || status != cudaSuccess) return;
        size_t bytes = 0;
        check(cub::DeviceRadixSort::SortPairs(nullptr, bytes, kin, kout, vin, vout, (int64_t)n, 0, bits, st));
        void *tmp = alloc<uint8_t>(bytes);
        check(cub::DeviceRadixSort::SortPairs(tmp, bytes, kin, kout, vin, vout, (int64_t)n, 0, bits, st));
        free(tmp);
        ++launches;
    }
    void exclusive_sum(const uint32_t *in, uint32_t *out, size_t n) {
        if (!n || status != cudaSuccess) return;
        size_t bytes = 0;
        check(cub::DeviceScan::ExclusiveSum(nullptr, bytes, in, out, (int64_t)n, st));
        void *tmp = alloc<uint8_t>(bytes);
        check(cub::DeviceScan::ExclusiveSum(tmp, bytes, in, out, (int64_t)n, st));
        free(tmp);
        ++launches;
    }
    void scan_boxes(const uint32_t *keys, const b200rt::bvhb::Box6 *in, b200rt::bvhb::Box6 *out, size_t n, bool reverse) {
        if (!n || status != cudaSuccess) return;
        using b200rt::bvhb::BoxUnion;
        size_t bytes = 0;
        if (!reverse) {
            check(cub::DeviceScan::InclusiveScanByKey(nullptr, bytes, keys, in, out, BoxUnion(), (int64_t)n, cub::Equality(), st));
            void *tmp = alloc<uint8_t>(bytes);
            check(cub::DeviceScan::InclusiveScanByKey(tmp, bytes, keys, in, out, BoxUnion(), (int64_t)n, cub::Equality(), st));
            free(tmp);
        } else {
            auto rk = thrust::make_reverse_iterator(keys + n);
            auto ri = thrust::make_reverse_iterator(in + n);
            auto ro = thrust::make_reverse_iterator(out + n);
            check(cub::DeviceScan::InclusiveScanByKey(nullptr, bytes, rk, ri, ro, BoxUnion(), (int64_t)n, cub::Equality(), st));
            void *tmp = alloc<uint8_t>(bytes);
            check(cub::DeviceScan::InclusiveScanByKey(tmp, bytes, rk, ri, ro, BoxUnion(), (int64_t)n, cub::Equality(), st));
            free(tmp);
        }
        ++launches;
    }
    void inclusive_max(const uint32_t *in, uint32_t *out, size_t n) {
        if (!n || status != cudaSuccess) return;
        size_t bytes = 0;
        check(cub::DeviceScan::InclusiveScan(nullptr, bytes, in, out, MaxU32(), (int64_t)n, st));
        void *tmp = alloc<uint8_t>(bytes);
        check(cub::DeviceScan::InclusiveScan(tmp, bytes, in, out, MaxU32(), (int64_t)n, st));
        free(tmp);
        ++launches;
    }
};

}  // namespace

static bool have_device() {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return false; }
    return true;
}
static void open_backend(CudaBackend &be, cudaEvent_t *e0, cudaEvent_t *e1) {
    be.check(cudaStreamCreateWithFlags(&be.st, cudaStreamNonBlocking));
    int dev = 0;   // keep freed blocks in the pool between levels
    cudaGetDevice(&dev);
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
        uint64_t thr = ~0ull;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
    }
    be.check(cudaEventCreate(e0));
    be.check(cudaEventCreate(e1));
    be.check(cudaEventRecord(*e0, be.st));
}
static float close_backend(CudaBackend &be, cudaEvent_t e0, cudaEvent_t e1) {
    be.check(cudaEventRecord(e1, be.st));
    be.check(cudaStreamSynchronize(be.st));
    float ms = 0;
    if (be.status == cudaSuccess) cudaEventElapsedTime(&ms, e0, e1);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    if (be.st) cudaStreamDestroy(be.st);
    g_launches += be.launches;
    return ms;
}

// BVHAccel's sweep-SAH build on the device (bvh_build.cuh); outputs as b200rt_build_bvh
extern "C" int b200rt_build_bvh_device(const float *tri_verts, uint32_t n_prims, int max_node_prims, int isect_cost,
                                       int traversal_cost, void **out_nodes, uint32_t *out_node_count,
                                       uint32_t **out_prim_order, b200rt_build_stats *stats) {
    using b200rt::set_error;
    using namespace b200rt::bvhb;
    if (!tri_verts || !out_nodes || !out_node_count || !out_prim_order) return set_error(B200RT_EINVAL, "b200rt_build_bvh_device: null argument");
    *out_nodes = nullptr; *out_node_count = 0; *out_prim_order = nullptr;
    if (n_prims == 0) return B200RT_OK;
    if (n_prims >= (1u << 28)) return set_error(B200RT_EUNSUPPORTED, "b200rt_build_bvh_device: more than 2^28 primitives");
    if (!have_device()) return set_error(B200RT_ENODEVICE, "b200rt_build_bvh_device: no CUDA device (the builder has no host path)");
    CudaBackend be;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    open_backend(be, &e0, &e1);
    float *dTris = be.alloc<float>((size_t)n_prims * 9);
    be.upload(dTris, tri_verts, (size_t)n_prims * 9);
    BvhResult r;
    const BvhParams P{std::min(255, max_node_prims), isect_cost, traversal_cost};   // bvh.cpp:157
    if (be.status == cudaSuccess) r = build_bvh(be, dTris, n_prims, P);
    be.free(dTris);
    const float ms = close_backend(be, e0, e1);
    if (be.status != cudaSuccess) return set_error(B200RT_ECUDA, "b200rt_build_bvh_device: %s", cudaGetErrorString(be.status));
    if (r.error) return set_error(B200RT_EUNSUPPORTED, "b200rt_build_bvh_device: the tree is deeper than 4096 levels");
    uint32_t *nodes = (uint32_t *)std::malloc(std::max<size_t>(r.nodes.size(), 1) * 4);
    uint32_t *order = (uint32_t *)std::malloc(std::max<size_t>(r.order.size(), 1) * 4);
    if (!nodes || !order) { std::free(nodes); std::free(order); return set_error(B200RT_ENOMEM, "b200rt_build_bvh_device: out of host memory"); }
    std::memcpy(nodes, r.nodes.data(), r.nodes.size() * 4);
    std::memcpy(order, r.order.data(), r.order.size() * 4);
    *out_nodes = nodes;
    *out_node_count = (uint32_t)(r.nodes.size() / 8);
    *out_prim_order = order;
    if (stats) { stats->levels = r.levels; stats->candidates = r.candidates; stats->kernel_launches = be.launches; stats->build_ms = ms; }
    return B200RT_OK;
}

// nb_directions = 0: kd-tree mode; kd_traversal_cost >= 0: RBSPKd mode
static int build_on_device(const float *tri_verts, uint32_t n_prims, uint32_t nb_directions, int kd_traversal_cost, int isect_cost,
                           int traversal_cost, float empty_bonus, int max_prims, int max_depth, void **out_nodes,
                           uint32_t *out_node_count, uint32_t **out_prim_indices, uint32_t *out_n_indices,
                           float out_bounds[6], float *out_directions, b200rt_build_stats *stats) {
    using b200rt::set_error;
    if (!tri_verts || !out_nodes || !out_node_count || !out_prim_indices || !out_n_indices || !out_bounds)
        return set_error(B200RT_EINVAL, "b200rt_build_rbsp: null argument");
    if (n_prims >= (1u << 27)) return set_error(B200RT_EUNSUPPORTED, "b200rt_build_rbsp: more than 2^27 primitives");
    BuildParams P{};
    P.kd = nb_directions == 0;
    if (kd_traversal_cost >= 0) { P.rbspkd = 1; P.kdTraversalCost = (uint32_t)kd_traversal_cost; }
    P.M = make_directions(P.kd ? 3 : nb_directions, P.dirs);
    if (!P.M) return set_error(B200RT_EINVAL, "b200rt_build_rbsp: nbDirections must be 3, 7, 9 or 13 (RBSPShared.h:50-71), got %u", nb_directions);
    P.isectCost = (uint32_t)isect_cost;
    P.traversalCost = (uint32_t)traversal_cost;
    P.emptyBonus = empty_bonus;
    P.maxPrims = (uint32_t)max_prims;
    P.maxDepth = max_depth < 0 ? default_max_depth(n_prims) : (uint32_t)max_depth;
    if (P.maxDepth > 63) return set_error(B200RT_EUNSUPPORTED, "b200rt_build_rbsp: maxdepth %u exceeds the traversal's 64-entry stack", P.maxDepth);
    if (!have_device()) return set_error(B200RT_ENODEVICE, "b200rt_build_rbsp: no CUDA device (the builder has no host path)");
    // `bounds = Union(bounds, prim->WorldBound())` (rbsp.cpp:192): plain min / max of the coordinates
    float world[6] = {3.402823466e+38f, 3.402823466e+38f, 3.402823466e+38f, -3.402823466e+38f, -3.402823466e+38f, -3.402823466e+38f};
    for (uint32_t p = 0; p < n_prims; ++p)
        for (int v = 0; v < 3; ++v)
            for (int c = 0; c < 3; ++c) {
                world[c] = std::min(world[c], tri_verts[9 * (size_t)p + 3 * v + c]);
                world[3 + c] = std::max(world[3 + c], tri_verts[9 * (size_t)p + 3 * v + c]);
            }
    CudaBackend be;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    open_backend(be, &e0, &e1);
    float *dTris = be.alloc<float>((size_t)n_prims * 9);
    be.upload(dTris, tri_verts, (size_t)n_prims * 9);
    BuildResult r;
    if (be.status == cudaSuccess) r = build(be, dTris, n_prims, P, world);
    be.free(dTris);
    const float ms = close_backend(be, e0, e1);
    if (be.status != cudaSuccess) return set_error(B200RT_ECUDA, "b200rt_build_rbsp: %s", cudaGetErrorString(be.status));
    if (r.error == kErrFixedOnly)
        return set_error(B200RT_EUNSUPPORTED, "b200rt_build_rbspkd: a node has candidate planes only in the \"fixed\" search; the reference indexes edges[-1] there (rbspKd.cpp:417-428)");
    if (r.error) return set_error(B200RT_EUNSUPPORTED, "b200rt_build_rbsp: a node's k-DOP outgrew %d edges / %d cut points per face", kMaxEdges, kMaxFacePoints);
    const size_t nNodes = r.nodes.size() / 2;
    uint32_t *nodes = (uint32_t *)std::malloc(std::max<size_t>(r.nodes.size(), 1) * 4);
    uint32_t *idx = (uint32_t *)std::malloc(std::max<size_t>(r.indices.size(), 1) * 4);
    if (!nodes || !idx) { std::free(nodes); std::free(idx); return set_error(B200RT_ENOMEM, "b200rt_build_rbsp: out of host memory"); }
    std::memcpy(nodes, r.nodes.data(), r.nodes.size() * 4);
    std::memcpy(idx, r.indices.data(), r.indices.size() * 4);
    *out_nodes = nodes;
    *out_node_count = (uint32_t)nNodes;
    *out_prim_indices = idx;
    *out_n_indices = (uint32_t)r.indices.size();
    std::memcpy(out_bounds, world, 24);
    if (out_directions) std::memcpy(out_directions, P.dirs, P.M * 12);
    if (stats) {
        stats->levels = r.levels;
        stats->candidates = r.candidates;
        stats->kernel_launches = be.launches;
        stats->build_ms = ms;
    }
    return B200RT_OK;
}

extern "C" int b200rt_build_rbsp(const float *tri_verts, uint32_t n_prims, uint32_t nb_directions, int isect_cost,
                                 int traversal_cost, float empty_bonus, int max_prims, int max_depth, void **out_nodes,
                                 uint32_t *out_node_count, uint32_t **out_prim_indices, uint32_t *out_n_indices,
                                 float out_bounds[6], float *out_directions, b200rt_build_stats *stats) {
    if (!out_directions || nb_directions == 0) return b200rt::set_error(B200RT_EINVAL, "b200rt_build_rbsp: null argument / nbDirections 0");
    return build_on_device(tri_verts, n_prims, nb_directions, -1, isect_cost, traversal_cost, empty_bonus, max_prims, max_depth,
                           out_nodes, out_node_count, out_prim_indices, out_n_indices, out_bounds, out_directions, stats);
}

extern "C" int b200rt_build_rbspkd(const float *tri_verts, uint32_t n_prims, uint32_t nb_directions, int isect_cost,
                                   int traversal_cost, int kd_traversal_cost, float empty_bonus, int max_prims, int max_depth,
                                   void **out_nodes, uint32_t *out_node_count, uint32_t **out_prim_indices,
                                   uint32_t *out_n_indices, float out_bounds[6], float *out_directions, b200rt_build_stats *stats) {
    if (!out_directions || nb_directions == 0 || kd_traversal_cost < 0) return b200rt::set_error(B200RT_EINVAL, "b200rt_build_rbspkd: bad argument");
    return build_on_device(tri_verts, n_prims, nb_directions, kd_traversal_cost, isect_cost, traversal_cost, empty_bonus, max_prims,
                           max_depth, out_nodes, out_node_count, out_prim_indices, out_n_indices, out_bounds, out_directions, stats);
}

extern "C" int b200rt_build_kdtree_device(const float *tri_verts, uint32_t n_prims, int isect_cost, int traversal_cost,
                                          float empty_bonus, int max_prims, int max_depth, void **out_nodes,
                                          uint32_t *out_node_count, uint32_t **out_prim_indices, uint32_t *out_n_indices,
                                          float out_bounds[6], b200rt_build_stats *stats) {
    return build_on_device(tri_verts, n_prims, 0, -1, isect_cost, traversal_cost, empty_bonus, max_prims, max_depth, out_nodes,
                           out_node_count, out_prim_indices, out_n_indices, out_bounds, nullptr, stats);
}
