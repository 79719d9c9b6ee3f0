// libb200rt.so — C ABI (include/b200rt.h): scene upload + batched Intersect / IntersectP.
//
// Replaces, for ray batches, Scene::Intersect / IntersectP -> Aggregate::Intersect / IntersectP
// (reference src/core/scene.cpp:45-55 -> src/accelerators/{bvh,kdtreeaccel,rbsp,rbspKd}.cpp).
// There is no CPU path in this file: every compute entry point needs a CUDA device.
#include <cuda_runtime.h>
#include <limits>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>
#include "rt_kernels.cuh"
#include "scene_internal.h"

using namespace b200rt;

// ---- error plumbing ---------------------------------------------------------------------------
static thread_local char g_err[512] = "";
static int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver ? B200RT_ENODEVICE : B200RT_ECUDA, \
                        "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

int b200rt::set_error(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

std::atomic<uint64_t> g_launches{0};
int g_sm_count = 148;
#define B200RT_PC_SLOTS 256
// tuning knobs (read once in b200rt_init): B200RT_KERNEL=baseline selects the one-ray-per-thread kernel
// for A/B measurements; B200RT_THRESH = busy-lane count below which a persistent warp refills.
static int g_use_baseline = 0;
static int g_coherent_baseline = 1;   // B200RT_COHERENT_BASELINE=0: persistent kernel on every queue
static uint32_t g_thresh = 14;
static int g_pblocks[6][2];  // resident persistent blocks per SM, [kind][any]
static int g_persist_bsp = 1;
static int g_kd_prefetch = 0;      // B200RT_KD_PREFETCH=1: kd persistent kernel fetches the below child together with each node
static uint32_t g_bvh_steps = 4;
static uint32_t g_bvh_vote[2] = {1, 1}, g_bsp_vote[2] = {1, 1};   // the vote's weights (interior, leaf): B200RT_BVH_VOTE=wI,wL / B200RT_BSP_VOTE=wI,wL
static uint32_t g_sstack = 10;     // BVH stack entries kept in shared memory (B200RT_SSTACK; 0 = all in local memory)
static uint32_t g_toplevels = 0;   // BVH levels staged in shared memory per block (B200RT_TOPLEVELS; 0 = none, the default):
                                   // north-star item (3), built and measured — neutral at 3-6 levels, -1 % at 8
                                   // (profiles/r02_ab_bvh_toplevels*.log): those records are L1-resident anyway
static size_t g_bvh_attr[2] = {0, 0};   // MaxDynamicSharedMemorySize currently set on the two BVH kernels
static uint32_t g_bsp_steps = B200RT_BSP_STEPS;  // B200RT_BSP_STEPS_N
static uint32_t g_sstack_bsp = 0;
static int g_minb = 9;             // __launch_bounds__ min blocks/SM variant of the BVH kernels (B200RT_MINB: 8 | 9; 9 = 56 registers, 36 warps/SM: +2 % over 8, 10 spills — profiles/r01_ab_bvh_minblocks.log)
typedef void (*bvh_kernel_t)(const BvhView, const float4 *, uint64_t, const uint32_t *, b200rt_hit *, uint8_t *, PersistCounters *,
                             uint32_t, uint32_t, uint32_t, uint32_t, uint32_t);
static int g_bvh4 = 0;             // B200RT_BVH4=1: the two-levels-per-record kernel (BvhNode4)
static const void *bvh_fn(bool any) {
    bvh_kernel_t f;
    if (g_bvh4) {
        if (g_minb == 8) f = any ? bvh4_persistent_kernel<true, 8> : bvh4_persistent_kernel<false, 8>;
        else f = any ? bvh4_persistent_kernel<true, 9> : bvh4_persistent_kernel<false, 9>;
        return (const void *)f;
    }
    if (g_toplevels) {   // the instantiation that stages the top levels in shared memory
        if (g_minb == 8) f = any ? bvh_persistent_kernel<true, 8, true> : bvh_persistent_kernel<false, 8, true>;
        else f = any ? bvh_persistent_kernel<true, 9, true> : bvh_persistent_kernel<false, 9, true>;
        return (const void *)f;
    }
    if (g_minb == 8) f = any ? bvh_persistent_kernel<true, 8> : bvh_persistent_kernel<false, 8>;
    else f = any ? bvh_persistent_kernel<true, 9> : bvh_persistent_kernel<false, 9>;
    return (const void *)f;
}  // kd/RBSP/BSP todo entries kept in shared memory (B200RT_SSTACK_BSP)
#define B200RT_PIPE_CHUNK_MAX (1024u * 1024u)
static uint32_t g_pipe_kstreams = 2;          // B200RT_PIPE_KSTREAMS: kernel streams of the host-buffer pipeline (1 .. 4)
static int g_pipe_nokernel = 0;               // B200RT_PIPE_NOKERNEL=1 (measurement only): the pipeline's copies and events without the kernels
static int g_pipe_zerocopy = 0;               // B200RT_PIPE_ZEROCOPY=1 (experiment): kernels read PINNED host rays over PCIe themselves, no H2D stage
static int g_pipe_ramp = 1;                   // B200RT_PIPE_RAMP=0: uniform chunks (A/B)
static uint32_t g_pipe_chunk = 256u * 1024u;  // rays per pipeline stage (B200RT_PIPE_CHUNK overrides, <= 1 Mi)

extern "C" const char *b200rt_last_error(void) { return g_err; }
extern "C" int b200rt_version(void) { return B200RT_VERSION; }
extern "C" uint64_t b200rt_launch_count(void) { return g_launches.load(); }

extern "C" int b200rt_init(int device) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(B200RT_ENODEVICE, "no CUDA device (%s); b200rt has no CPU fallback",
                    e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    if (device < 0 || device >= n) return fail(B200RT_EINVAL, "device %d out of range [0,%d)", device, n);
    CU(cudaSetDevice(device));
    CU(cudaFree(0));
    cudaDeviceProp p;
    CU(cudaGetDeviceProperties(&p, device));
    g_sm_count = p.multiProcessorCount;
    if (const char *e = getenv("B200RT_COHERENT_BASELINE")) g_coherent_baseline = atoi(e);
    if (const char *e = getenv("B200RT_KERNEL")) g_use_baseline = std::strcmp(e, "baseline") == 0;
    if (const char *e = getenv("B200RT_THRESH")) g_thresh = (uint32_t)atoi(e);
    if (const char *e = getenv("B200RT_PIPE_CHUNK")) g_pipe_chunk = std::min<uint32_t>(B200RT_PIPE_CHUNK_MAX, std::max(1024, atoi(e)));
    if (const char *e = getenv("B200RT_PIPE_RAMP")) g_pipe_ramp = atoi(e);
    if (const char *e = getenv("B200RT_PIPE_ZEROCOPY")) g_pipe_zerocopy = atoi(e);
    if (const char *e = getenv("B200RT_PIPE_NOKERNEL")) g_pipe_nokernel = atoi(e);
    if (const char *e = getenv("B200RT_PIPE_KSTREAMS")) g_pipe_kstreams = (uint32_t)std::min(4, std::max(1, atoi(e)));
    if (const char *e = getenv("B200RT_BVH_STEPS")) g_bvh_steps = std::max(1, atoi(e));
    if (const char *e = getenv("B200RT_BVH_VOTE")) { unsigned a = 1, b = 1; if (sscanf(e, "%u,%u", &a, &b) >= 1) { g_bvh_vote[0] = std::max(1u, a); g_bvh_vote[1] = std::max(1u, b); } }
    if (const char *e = getenv("B200RT_BSP_VOTE")) { unsigned a = 1, b = 1; if (sscanf(e, "%u,%u", &a, &b) >= 1) { g_bsp_vote[0] = std::max(1u, a); g_bsp_vote[1] = std::max(1u, b); } }
    if (const char *e = getenv("B200RT_PERSIST_BSP")) g_persist_bsp = atoi(e);
    if (const char *e = getenv("B200RT_BSP_STEPS_N")) g_bsp_steps = (uint32_t)std::max(1, atoi(e));
    if (const char *e = getenv("B200RT_SSTACK_BSP")) g_sstack_bsp = (uint32_t)std::min(64, std::max(0, atoi(e)));
    if (const char *e = getenv("B200RT_SSTACK")) g_sstack = (uint32_t)std::min(64, std::max(0, atoi(e)));
    if (const char *e = getenv("B200RT_TOPLEVELS")) g_toplevels = (uint32_t)std::min(B200RT_TOP_MAXLEVELS, std::max(0, atoi(e)));
    // shared memory of the persistent kernels: the bottom of the traversal stack (+ the staged top levels of the BVH); the
    // carve-out asks for just that much, the rest of the 256 KB array stays L1 (node working set)
    int carve = 0;
    if (const char *e = getenv("B200RT_CARVEOUT")) carve = atoi(e);
    const size_t smB = (size_t)g_sstack_bsp * 12 * B200RT_PBLOCK;
    const int carveB = g_sstack_bsp ? (int)std::min<size_t>(100, (8 * smB * 100) / (228 * 1024) + 1) : carve;
#define OCC(k, a, fn)                                                                                        \
    CU(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smB));                      \
    if (carveB >= 0) CU(cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, carveB)); \
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_pblocks[k][a], fn, B200RT_PBLOCK, smB))
    {   // BVH kernels: dynamic shared memory = the shared part of the stack
        if (const char *e = getenv("B200RT_MINB")) g_minb = atoi(e);
        if (const char *e = getenv("B200RT_BVH4")) g_bvh4 = atoi(e);
        if (g_minb != 8) g_minb = 9;
        // the shared part of the stack + room for the staged top levels (at most 2^levels - 1 records; the launch passes the
        // scene's actual count): attribute, carve-out and residency are all taken for this size
        const size_t sm = (size_t)g_sstack * sizeof(uint2) * B200RT_PBLOCK +
                          ((g_toplevels && !g_bvh4) ? (((size_t)1 << g_toplevels) - 1) * sizeof(BvhNode2) : 0);
        const int bc = sm ? (int)std::min<size_t>(100, ((size_t)g_minb * (sm + (g_toplevels ? 1024 : 0)) * 100) / (228 * 1024) + 1) : carve;
        const void *fns[2] = {bvh_fn(false), bvh_fn(true)};
        for (int a = 0; a < 2; ++a) {
            CU(cudaFuncSetAttribute(fns[a], cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
            g_bvh_attr[a] = sm;
            if (bc >= 0) CU(cudaFuncSetAttribute(fns[a], cudaFuncAttributePreferredSharedMemoryCarveout, bc));
            CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_pblocks[B200RT_BVH][a], fns[a], B200RT_PBLOCK, sm));
        }
    }
    if (const char *e = getenv("B200RT_KD_PREFETCH")) g_kd_prefetch = atoi(e);
    if (g_kd_prefetch) {
        OCC(B200RT_KDTREE, 0, (bsp_persistent_kernel<B200RT_KDTREE, false, true>)); OCC(B200RT_KDTREE, 1, (bsp_persistent_kernel<B200RT_KDTREE, true, true>));
    } else {
        OCC(B200RT_KDTREE, 0, (bsp_persistent_kernel<B200RT_KDTREE, false>)); OCC(B200RT_KDTREE, 1, (bsp_persistent_kernel<B200RT_KDTREE, true>));
    }
    OCC(B200RT_RBSP, 0, (bsp_persistent_kernel<B200RT_RBSP, false>)); OCC(B200RT_RBSP, 1, (bsp_persistent_kernel<B200RT_RBSP, true>));
    OCC(B200RT_RBSPKD, 0, (bsp_persistent_kernel<B200RT_RBSPKD, false>)); OCC(B200RT_RBSPKD, 1, (bsp_persistent_kernel<B200RT_RBSPKD, true>));
    OCC(B200RT_BSP, 0, (bsp_persistent_kernel<B200RT_BSP, false>)); OCC(B200RT_BSP, 1, (bsp_persistent_kernel<B200RT_BSP, true>));
    OCC(B200RT_BSPKD, 0, (bsp_persistent_kernel<B200RT_BSPKD, false>)); OCC(B200RT_BSPKD, 1, (bsp_persistent_kernel<B200RT_BSPKD, true>));
#undef OCC
    return B200RT_OK;
}

// Triangle::Intersect's tail (triangle.cpp:294-315) rejects EVERY hit on a triangle whose UV
// parameterisation is degenerate (or dpdu x dpdv == 0) and whose geometric normal is zero.  It
// depends on the triangle only, so it becomes a per-triangle flag computed once at upload, in the
// same float/double arithmetic (host compiler: no contraction for these expressions, see build.py).
static bool tri_tail_rejects(const float *v, const float *uv) {
    static const float kDefault[6] = {0, 0, 1, 0, 1, 1};  // Triangle::GetUVs, triangle.h:109-121
    if (!uv) uv = kDefault;
    const float *p0 = v, *p1 = v + 3, *p2 = v + 6;
    volatile float duv02x = uv[0] - uv[4], duv02y = uv[1] - uv[5];
    volatile float duv12x = uv[2] - uv[4], duv12y = uv[3] - uv[5];
    float dp02[3], dp12[3];
    for (int i = 0; i < 3; ++i) { dp02[i] = p0[i] - p2[i]; dp12[i] = p1[i] - p2[i]; }
    volatile float m0 = duv02x * duv12y, m1 = duv02y * duv12x;
    const float determinant = m0 - m1;
    const bool degenerateUV = std::fabs(determinant) < 1e-8;
    bool need = degenerateUV;
    if (!need) {
        const float invdet = 1 / determinant;
        float dpdu[3], dpdv[3];
        for (int i = 0; i < 3; ++i) {
            volatile float a = duv12y * dp02[i], b = duv02y * dp12[i];
            volatile float du = a - b;
            dpdu[i] = du * invdet;
            volatile float c = -duv12x * dp02[i], d = duv02x * dp12[i];
            volatile float dv = c + d;
            dpdv[i] = dv * invdet;
        }
        // Cross() in double then narrowed (geometry.h:1287-1293); products of two floats are exact in double
        const double ax = dpdu[0], ay = dpdu[1], az = dpdu[2], bx = dpdv[0], by = dpdv[1], bz = dpdv[2];
        const float cx = (float)((ay * bz) - (az * by)), cy = (float)((az * bx) - (ax * bz)),
                    cz = (float)((ax * by) - (ay * bx));
        volatile float xx = cx * cx, yy = cy * cy, zz = cz * cz;
        volatile float s = xx + yy;
        need = (s + zz) == 0;
    }
    if (!need) return false;
    float e0[3], e1[3];
    for (int i = 0; i < 3; ++i) { e0[i] = p2[i] - p0[i]; e1[i] = p1[i] - p0[i]; }
    const double ax = e0[0], ay = e0[1], az = e0[2], bx = e1[0], by = e1[1], bz = e1[2];
    const float nx = (float)((ay * bz) - (az * by)), ny = (float)((az * bx) - (ax * bz)),
                nz = (float)((ax * by) - (ay * bx));
    volatile float xx = nx * nx, yy = ny * ny, zz = nz * nz;
    volatile float s = xx + yy;
    return (s + zz) == 0;
}

static void fill_record(float4 *rec, const float *v, uint32_t prim, uint32_t flags) {
    rec[0] = make_float4(v[0], v[1], v[2], 0.f);
    rec[1] = make_float4(v[3], v[4], v[5], 0.f);
    rec[2] = make_float4(v[6], v[7], v[8], 0.f);
    std::memcpy(&rec[0].w, &prim, 4);
    std::memcpy(&rec[1].w, &flags, 4);
}

// prim -> index of its sphere parameter row (valid where prim_kind[prim] != 0); validates the counts
static int sphere_table(const b200rt_scene_desc *d, std::vector<uint32_t> &sphereOf) {
    sphereOf.assign(d->n_prims, 0);
    uint32_t k = 0;
    if (d->prim_kind)
        for (uint32_t i = 0; i < d->n_prims; ++i) {
            if (d->prim_kind[i] > 1) return fail(B200RT_EUNSUPPORTED, "prim_kind[%u] = %u: only triangles (0) and full spheres (1)", i, d->prim_kind[i]);
            if (d->prim_kind[i]) sphereOf[i] = k++;
        }
    if (k != d->n_spheres) return fail(B200RT_EINVAL, "prim_kind marks %u spheres but n_spheres = %u", k, d->n_spheres);
    if (k && !d->sphere_params) return fail(B200RT_EINVAL, "sphere_params is null");
    return B200RT_OK;
}
static int upload_spheres(const b200rt_scene_desc *d, b200rt_scene *s) {
    if (!d->n_spheres) return B200RT_OK;
    CU(cudaMalloc(&s->d_spheres, (size_t)d->n_spheres * 64));
    CU(cudaMemcpy(s->d_spheres, d->sphere_params, (size_t)d->n_spheres * 64, cudaMemcpyHostToDevice));
    return B200RT_OK;
}
static void fill_sphere_record(float4 *rec, uint32_t prim, uint32_t flags, uint32_t sphereIdx) {
    rec[0] = rec[1] = rec[2] = make_float4(0.f, 0.f, 0.f, 0.f);
    flags |= B200RT_PRIM_SPHERE;
    std::memcpy(&rec[0].w, &prim, 4);
    std::memcpy(&rec[1].w, &flags, 4);
    std::memcpy(&rec[2].w, &sphereIdx, 4);
}

struct RefBvhNode {  // LinearBVHNode, bvh.cpp:123-152
    float lo[3], hi[3];
    int32_t offset;
    uint32_t word;
};

static int build_bvh_scene(const b200rt_scene_desc *d, b200rt_scene *s) {
    const RefBvhNode *ref = (const RefBvhNode *)d->nodes;
    const uint32_t N = d->node_count;
    if (d->n_indices != d->n_prims) return fail(B200RT_EINVAL, "BVH: n_indices (%u) must equal n_prims (%u)", d->n_indices, d->n_prims);
    // pass 1: validate + number the interior nodes (depth-first order = array order)
    std::vector<uint32_t> wide(N, 0xFFFFFFFFu);
    uint32_t nInterior = 0;
    for (uint32_t i = 0; i < N; ++i) {
        const bool leaf = (ref[i].word & 3u) == 3u;
        if (leaf) {
            const uint64_t np = ref[i].word >> 2, off = (uint32_t)ref[i].offset;
            if (np == 0 || ref[i].offset < 0 || off + np > d->n_prims)
                return fail(B200RT_EINVAL, "BVH leaf %u: primitive range [%lld,+%llu) outside %u prims", i, (long long)ref[i].offset, (unsigned long long)np, d->n_prims);
        } else {
            if (i + 1 >= N || ref[i].offset <= (int32_t)i + 1 || (uint32_t)ref[i].offset >= N)
                return fail(B200RT_EINVAL, "BVH interior node %u: bad child offsets (second=%d, nodes=%u)", i, ref[i].offset, N);
            wide[i] = nInterior++;
        }
    }
    // Renumber: interior nodes of depth < B200RT_TOP_MAXLEVELS first, level by level (breadth-first), the others after
    // them in the reference's depth-first order.  Only the device-side record order changes, never the traversal order.
    uint32_t topCount[12] = {0};
    {
        std::vector<uint32_t> depthOf(N, 0);
        for (uint32_t i = 0; i < N; ++i)
            if ((ref[i].word & 3u) != 3u) { depthOf[i + 1] = depthOf[i] + 1; depthOf[ref[i].offset] = depthOf[i] + 1; }
        std::vector<uint32_t> level{0};   // reference node numbers of the current level, in left-to-right order
        uint32_t next = 0;
        std::vector<uint8_t> placed(N, 0);
        // (only the levels that are staged are renumbered: the depth-first order keeps a first child next to its parent)
        for (int lv = 0; lv < (int)g_toplevels; ++lv) {
            std::vector<uint32_t> nxt;
            for (uint32_t i : level) {
                if ((ref[i].word & 3u) == 3u) continue;
                wide[i] = next++; placed[i] = 1;
                nxt.push_back(i + 1); nxt.push_back((uint32_t)ref[i].offset);
            }
            topCount[lv + 1] = next;
            level.swap(nxt);
        }
        for (uint32_t i = 0; i < N; ++i)
            if ((ref[i].word & 3u) != 3u && !placed[i]) wide[i] = next++;
        if (next != nInterior) return fail(B200RT_EINVAL, "BVH: inconsistent interior-node numbering");
    }
    // depth check (the reference's traversal stack holds 64 entries, bvh.cpp:362)
    {
        std::vector<uint32_t> depth(N, 0);
        uint32_t maxDepth = 0;
        depth[0] = 0;
        for (uint32_t i = 0; i < N; ++i) {
            if ((ref[i].word & 3u) != 3u) {
                depth[i + 1] = depth[i] + 1;
                depth[ref[i].offset] = depth[i] + 1;
                if (depth[i] + 1 > maxDepth) maxDepth = depth[i] + 1;
            }
        }
        if (maxDepth > 64) return fail(B200RT_EUNSUPPORTED, "BVH depth %u exceeds the reference's 64-entry traversal stack", maxDepth);
    }
    std::vector<uint32_t> sphereOf;
    { int rc = sphere_table(d, sphereOf); if (rc) return rc; }
    // triangles in slot order, flagged at leaf ends
    std::vector<float4> recs((size_t)3 * d->n_prims);
    std::vector<uint8_t> last(d->n_prims, 0);
    for (uint32_t i = 0; i < N; ++i)
        if ((ref[i].word & 3u) == 3u) last[(uint32_t)ref[i].offset + (ref[i].word >> 2) - 1] = 1;
    last[d->n_prims - 1] = 1;  // a leaf walk can never run past the array, whatever the input tree says
    for (uint32_t slot = 0; slot < d->n_prims; ++slot) {
        const uint32_t prim = d->prim_indices[slot];
        if (prim >= d->n_prims) return fail(B200RT_EINVAL, "BVH prim_indices[%u] = %u out of range", slot, prim);
        const float *v = d->tri_verts + 9 * (size_t)prim;
        uint32_t flags = last[slot] ? B200RT_TRI_LAST : 0u;
        if (d->prim_kind && d->prim_kind[prim]) { fill_sphere_record(&recs[3 * (size_t)slot], prim, flags, sphereOf[prim]); continue; }
        if (tri_tail_rejects(v, d->tri_uvs ? d->tri_uvs + 6 * (size_t)prim : nullptr)) flags |= B200RT_TRI_DEGENERATE;
        fill_record(&recs[3 * (size_t)slot], v, prim, flags);
    }
    auto ref_of = [&](uint32_t i) -> uint32_t {
        return (ref[i].word & 3u) == 3u ? (B200RT_LEAF_BIT | (uint32_t)ref[i].offset) : wide[i];
    };
    std::vector<BvhNode2> nodes(nInterior ? nInterior : 1);
    std::memset(nodes.data(), 0, nodes.size() * sizeof(BvhNode2));
    for (uint32_t i = 0; i < N; ++i) {
        if ((ref[i].word & 3u) == 3u) continue;
        const RefBvhNode &c0 = ref[i + 1], &c1 = ref[ref[i].offset];
        BvhNode2 &w = nodes[wide[i]];
        w.a = make_float4(c0.lo[0], c0.lo[1], c0.lo[2], c0.hi[0]);
        w.b = make_float4(c0.hi[1], c0.hi[2], c1.lo[0], c1.lo[1]);
        w.c = make_float4(c1.lo[2], c1.hi[0], c1.hi[1], c1.hi[2]);
        w.m = make_uint4(ref_of(i + 1), ref_of((uint32_t)ref[i].offset), ref[i].word & 3u, ref[i].word >> 2);
    }
    // ---- BvhNode4: two levels per record (rt_kernels.cuh) ----
    {
        std::vector<uint32_t> w4(N, 0xFFFFFFFFu);
        std::vector<uint32_t> order;          // reference node numbers of the wide records, depth-first
        if ((ref[0].word & 3u) != 3u) {
            std::vector<uint32_t> st{0};
            while (!st.empty()) {
                const uint32_t i = st.back(); st.pop_back();
                w4[i] = (uint32_t)order.size(); order.push_back(i);
                const uint32_t ch[2] = {i + 1, (uint32_t)ref[i].offset};
                for (int c = 1; c >= 0; --c) {
                    if ((ref[ch[c]].word & 3u) == 3u) continue;
                    const uint32_t g[2] = {ch[c] + 1, (uint32_t)ref[ch[c]].offset};
                    for (int k = 1; k >= 0; --k)
                        if ((ref[g[k]].word & 3u) != 3u) st.push_back(g[k]);
                }
            }
        }
        auto ref4_of = [&](uint32_t j) -> uint32_t {
            return (ref[j].word & 3u) == 3u ? (B200RT_LEAF_BIT | (uint32_t)ref[j].offset) : w4[j];
        };
        std::vector<BvhNode4> n4(order.size() ? order.size() : 1);
        std::memset(n4.data(), 0, n4.size() * sizeof(BvhNode4));
        const float inf = std::numeric_limits<float>::infinity();
        for (size_t w = 0; w < order.size(); ++w) {
            const uint32_t i = order[w];
            const uint32_t ch[2] = {i + 1, (uint32_t)ref[i].offset};
            uint32_t meta = ref[i].word & 3u;
            for (int c = 0; c < 2; ++c) {
                const RefBvhNode *slot[2] = {nullptr, nullptr};
                uint32_t sref[2] = {B200RT_EMPTY_SLOT, B200RT_EMPTY_SLOT};
                if ((ref[ch[c]].word & 3u) == 3u) { slot[0] = &ref[ch[c]]; sref[0] = ref4_of(ch[c]); }
                else {
                    const uint32_t g[2] = {ch[c] + 1, (uint32_t)ref[ch[c]].offset};
                    for (int k = 0; k < 2; ++k) { slot[k] = &ref[g[k]]; sref[k] = ref4_of(g[k]); }
                    meta |= (ref[ch[c]].word & 3u) << (2 + 2 * c);
                }
                float lo[2][3], hi[2][3];
                for (int k = 0; k < 2; ++k)
                    for (int a = 0; a < 3; ++a) { lo[k][a] = slot[k] ? slot[k]->lo[a] : inf; hi[k][a] = slot[k] ? slot[k]->hi[a] : -inf; }
                float4 *q = n4[w].q + 4 * c;
                q[0] = make_float4(lo[0][0], lo[0][1], lo[0][2], hi[0][0]);
                q[1] = make_float4(hi[0][1], hi[0][2], lo[1][0], lo[1][1]);
                q[2] = make_float4(lo[1][2], hi[1][0], hi[1][1], hi[1][2]);
                float r0, r1;
                std::memcpy(&r0, &sref[0], 4); std::memcpy(&r1, &sref[1], 4);
                q[3] = make_float4(r0, r1, 0.f, 0.f);
            }
            std::memcpy(&n4[w].q[3].z, &meta, 4);
        }
        CU(cudaMalloc(&s->d_nodes4, n4.size() * sizeof(BvhNode4)));
        CU(cudaMemcpy(s->d_nodes4, n4.data(), n4.size() * sizeof(BvhNode4), cudaMemcpyHostToDevice));
        s->bvh.nodes4 = (const BvhNode4 *)s->d_nodes4;
        s->bvh.rootRef4 = ref4_of(0);
    }
    CU(cudaMalloc(&s->d_nodes, nodes.size() * sizeof(BvhNode2)));
    CU(cudaMemcpy(s->d_nodes, nodes.data(), nodes.size() * sizeof(BvhNode2), cudaMemcpyHostToDevice));
    CU(cudaMalloc(&s->d_tris, recs.size() * sizeof(float4)));
    CU(cudaMemcpy(s->d_tris, recs.data(), recs.size() * sizeof(float4), cudaMemcpyHostToDevice));
    {   // original-primitive-order copy for the shading stage (hit.prim indexes this)
        std::vector<float4> byPrim((size_t)3 * d->n_prims);
        for (uint32_t slot = 0; slot < d->n_prims; ++slot)
            std::memcpy(&byPrim[3 * (size_t)d->prim_indices[slot]], &recs[3 * (size_t)slot], 3 * sizeof(float4));
        CU(cudaMalloc(&s->d_prim_tris, byPrim.size() * sizeof(float4)));
        CU(cudaMemcpy(s->d_prim_tris, byPrim.data(), byPrim.size() * sizeof(float4), cudaMemcpyHostToDevice));
    }
    { int rc = upload_spheres(d, s); if (rc) return rc; }
    s->bvh.nodes = (const BvhNode2 *)s->d_nodes;
    s->bvh.tris = s->d_tris;
    s->bvh.spheres = s->d_spheres;
    for (int k = 0; k < 3; ++k) { s->bvh.root[k] = ref[0].lo[k]; s->bvh.root[3 + k] = ref[0].hi[k]; }
    s->bvh.rootRef = ref_of(0);
    std::memcpy(s->bvh.topCount, topCount, sizeof(topCount));
    std::memcpy(s->bounds, s->bvh.root, sizeof(s->bounds));
    return B200RT_OK;
}

struct RefWideNode {  // BSPNode (BSP.h:170-183) / BSPKdNode (BSPKd.h:161-173): 20 bytes
    uint32_t w0, w1;
    float axis[3];
};

static int build_bsp_scene(const b200rt_scene_desc *d, b200rt_scene *s) {
    const int kind = (int)d->kind;
    const bool wideKind = kind == B200RT_BSP || kind == B200RT_BSPKD;
    const uint2 *ref = (const uint2 *)d->nodes;
    const RefWideNode *refW = (const RefWideNode *)d->nodes;
    const uint32_t N = d->node_count;
    uint32_t M = 3;
    if (kind == B200RT_RBSP || kind == B200RT_RBSPKD) {
        M = d->n_directions;
        if (!d->directions || M < 1 || M > 13) return fail(B200RT_EINVAL, "RBSP: need 1..13 directions (got %u)", M);
    }
    const uint32_t off = kind == B200RT_KDTREE ? 2u : kind == B200RT_BSP ? 1u : kind == B200RT_BSPKD ? 3u
                         : 32u - (uint32_t)__builtin_clz(M);  // genericBSP.cpp:11-22
    const uint32_t mask = (1u << off) - 1u, leafTag = kind == B200RT_KDTREE ? 3u : kind == B200RT_BSP ? 1u : kind == B200RT_BSPKD ? 3u : M;
    auto w0 = [&](uint32_t i) { return wideKind ? refW[i].w0 : ref[i].x; };
    auto w1 = [&](uint32_t i) { return wideKind ? refW[i].w1 : ref[i].y; };
    // validate + depth (todo stack of 64, kdtreeaccel.cpp:391)
    std::vector<uint32_t> depth(N, 0);
    uint32_t maxDepth = 0;
    for (uint32_t i = 0; i < N; ++i) {
        const uint32_t axis = w1(i) & mask;
        if (axis == leafTag) {
            const uint64_t np = w1(i) >> off;
            if (np == 1) { if (w0(i) >= d->n_prims) return fail(B200RT_EINVAL, "leaf %u: primitive %u out of range", i, w0(i)); }
            else if (np > 1 && (uint64_t)w0(i) + np > d->n_indices) return fail(B200RT_EINVAL, "leaf %u: index range out of bounds", i);
        } else {
            const uint32_t above = w1(i) >> off;
            // BSP: the tag is one bit; BSPKd: 0..2 = kd axis, 4 = general node (BSPKd.h:13,36-45)
            const bool axisOk = kind == B200RT_BSP ? true : kind == B200RT_BSPKD ? (axis < 3 || axis == 4) : axis < M;
            if (!axisOk || i + 1 >= N || above <= i + 1 || above >= N)
                return fail(B200RT_EINVAL, "interior node %u: bad axis/children (axis=%u above=%u nodes=%u)", i, axis, above, N);
            depth[i + 1] = depth[above] = depth[i] + 1;
            if (depth[i] + 1 > maxDepth) maxDepth = depth[i] + 1;
        }
    }
    if (maxDepth > 64) return fail(B200RT_EUNSUPPORTED, "tree depth %u exceeds the reference's 64-entry todo stack", maxDepth);
    for (uint32_t i = 0; i < d->n_indices; ++i)
        if (d->prim_indices[i] >= d->n_prims) return fail(B200RT_EINVAL, "prim_indices[%u] out of range", i);
    std::vector<uint32_t> sphereOf;
    { int rc = sphere_table(d, sphereOf); if (rc) return rc; }
    std::vector<float4> recs((size_t)3 * d->n_prims);
    for (uint32_t p = 0; p < d->n_prims; ++p) {
        if (d->prim_kind && d->prim_kind[p]) { fill_sphere_record(&recs[3 * (size_t)p], p, 0u, sphereOf[p]); continue; }
        const float *v = d->tri_verts + 9 * (size_t)p;
        uint32_t flags = tri_tail_rejects(v, d->tri_uvs ? d->tri_uvs + 6 * (size_t)p : nullptr) ? B200RT_TRI_DEGENERATE : 0u;
        fill_record(&recs[3 * (size_t)p], v, p, flags);
    }
    if (wideKind) {  // 20-byte reference nodes -> aligned 32-byte records (rt_kernels.cuh bsp_fetch)
        std::vector<float4> wide((size_t)2 * N);
        for (uint32_t i = 0; i < N; ++i) {
            float4 lo = make_float4(0.f, 0.f, refW[i].axis[0], refW[i].axis[1]);
            std::memcpy(&lo.x, &refW[i].w0, 4);
            std::memcpy(&lo.y, &refW[i].w1, 4);
            wide[2 * (size_t)i] = lo;
            wide[2 * (size_t)i + 1] = make_float4(refW[i].axis[2], 0.f, 0.f, 0.f);
        }
        CU(cudaMalloc(&s->d_nodes, wide.size() * sizeof(float4)));
        CU(cudaMemcpy(s->d_nodes, wide.data(), wide.size() * sizeof(float4), cudaMemcpyHostToDevice));
    } else {
        CU(cudaMalloc(&s->d_nodes, ((size_t)N + 1) * 8));   // + 1 record: the kd kernel's below-child prefetch may read nodes[N]
        CU(cudaMemsetAsync((char *)s->d_nodes + (size_t)N * 8, 0, 8));
        CU(cudaMemcpy(s->d_nodes, ref, (size_t)N * 8, cudaMemcpyHostToDevice));
    }
    CU(cudaMalloc(&s->d_indices, std::max<size_t>(1, d->n_indices) * 4));
    if (d->n_indices) CU(cudaMemcpy(s->d_indices, d->prim_indices, (size_t)d->n_indices * 4, cudaMemcpyHostToDevice));
    CU(cudaMalloc(&s->d_tris, std::max<size_t>(1, recs.size()) * sizeof(float4)));
    if (!recs.empty()) CU(cudaMemcpy(s->d_tris, recs.data(), recs.size() * sizeof(float4), cudaMemcpyHostToDevice));
    { int rc = upload_spheres(d, s); if (rc) return rc; }
    s->bsp.nodes = (const uint2 *)s->d_nodes;
    s->bsp.indices = s->d_indices;
    s->bsp.tris = s->d_tris;
    s->bsp.spheres = s->d_spheres;
    s->d_prim_tris = s->d_tris;
    std::memcpy(s->bsp.bounds, d->bounds, sizeof(s->bsp.bounds));
    std::memset(s->bsp.dirs, 0, sizeof(s->bsp.dirs));
    if (kind == B200RT_KDTREE || wideKind) { s->bsp.dirs[0] = s->bsp.dirs[4] = s->bsp.dirs[8] = 1.f; }
    else std::memcpy(s->bsp.dirs, d->directions, sizeof(float) * 3 * M);
    s->bsp.M = M; s->bsp.bitOffset = off; s->bsp.bitMask = mask; s->bsp.leafTag = leafTag; s->bsp.kind = kind;
    std::memcpy(s->bounds, d->bounds, sizeof(s->bounds));
    return B200RT_OK;
}

extern "C" int b200rt_scene_create(const b200rt_scene_desc *d, b200rt_scene **out) {
    if (!d || !out) return fail(B200RT_EINVAL, "null argument");
    *out = nullptr;
    if (d->kind > B200RT_BSPKD) return fail(B200RT_EINVAL, "unknown accelerator kind %u", d->kind);
    if (d->n_prims && !d->tri_verts) return fail(B200RT_EINVAL, "tri_verts is null");
    if (d->node_count && !d->nodes) return fail(B200RT_EINVAL, "nodes is null");
    if (d->n_indices && !d->prim_indices) return fail(B200RT_EINVAL, "prim_indices is null");
    int dev = 0;
    CU(cudaGetDevice(&dev));
    b200rt_scene *s = new (std::nothrow) b200rt_scene;
    if (!s) return fail(B200RT_ENOMEM, "out of host memory");
    s->kind = (int)d->kind;
    s->device = dev;
    s->n_prims = d->n_prims;
    int rc = B200RT_OK;
    if (d->node_count == 0 || d->n_prims == 0) s->empty = true;
    else rc = d->kind == B200RT_BVH ? build_bvh_scene(d, s) : build_bsp_scene(d, s);
    if (rc == B200RT_OK) {
        cudaError_t e = cudaMalloc(&s->d_counters, sizeof(Counters));
        if (e == cudaSuccess) e = cudaMalloc(&s->d_pc, B200RT_PC_SLOTS * sizeof(PersistCounters));
        if (e == cudaSuccess) e = cudaMemset(s->d_pc, 0, B200RT_PC_SLOTS * sizeof(PersistCounters));
        if (e != cudaSuccess) rc = fail(B200RT_ECUDA, "cudaMalloc(counters): %s", cudaGetErrorString(e));
    }
    // shading attributes of the wavefront stage (ABI v8): per-primitive UVs, vertex normals and orientation flags
    if (rc == B200RT_OK && !s->empty) {
        cudaError_t e = cudaSuccess;
        if (d->tri_uvs) {
            e = cudaMalloc(&s->d_prim_uv, (size_t)d->n_prims * 6 * sizeof(float));
            if (e == cudaSuccess) e = cudaMemcpy(s->d_prim_uv, d->tri_uvs, (size_t)d->n_prims * 6 * sizeof(float), cudaMemcpyHostToDevice);
        }
        if (e == cudaSuccess && d->tri_normals) {
            e = cudaMalloc(&s->d_prim_n, (size_t)d->n_prims * 9 * sizeof(float));
            if (e == cudaSuccess) e = cudaMemcpy(s->d_prim_n, d->tri_normals, (size_t)d->n_prims * 9 * sizeof(float), cudaMemcpyHostToDevice);
        }
        if (e == cudaSuccess && d->prim_flags) {
            for (uint32_t i = 0; i < d->n_prims && rc == B200RT_OK; ++i)
                if ((d->prim_flags[i] & 2u) && !d->tri_normals) rc = fail(B200RT_EINVAL, "prim_flags[%u] says vertex normals but tri_normals is null", i);
            if (rc == B200RT_OK) {
                e = cudaMalloc(&s->d_prim_flags, d->n_prims);
                if (e == cudaSuccess) e = cudaMemcpy(s->d_prim_flags, d->prim_flags, d->n_prims, cudaMemcpyHostToDevice);
            }
        }
        if (e != cudaSuccess) rc = fail(B200RT_ECUDA, "cudaMalloc(shading attributes): %s", cudaGetErrorString(e));
    }
    if (rc != B200RT_OK) { b200rt_scene_destroy(s); return rc; }
    *out = s;
    return B200RT_OK;
}

extern "C" void b200rt_scene_destroy(b200rt_scene *s) {
    if (!s) return;
    b200rt_wavefront_release(s);
    if (s->d_prim_tris != s->d_tris) cudaFree(s->d_prim_tris);
    cudaFree(s->d_nodes); cudaFree(s->d_indices); cudaFree(s->d_tris); cudaFree(s->d_spheres);
    cudaFree(s->d_prim_uv); cudaFree(s->d_prim_n); cudaFree(s->d_prim_flags); cudaFree(s->d_nodes4);
    cudaFree(s->d_stage_in); cudaFree(s->d_stage_out); cudaFree(s->d_counters); cudaFree(s->d_pc);
    for (auto &q : s->slots) {
        cudaFree(q.d_in); cudaFree(q.d_out);
        for (cudaEvent_t e : {q.evIn, q.evK, q.evOut}) if (e) cudaEventDestroy(e);
    }
    for (cudaStream_t st : {s->sIn, s->sK[0], s->sK[1], s->sK[2], s->sK[3], s->sOut}) if (st) cudaStreamDestroy(st);
    if (s->ev_end) cudaEventDestroy(s->ev_end);
    if (s->ev_start) cudaEventDestroy(s->ev_start);
    delete s;
}

extern "C" int b200rt_scene_bounds(const b200rt_scene *s, float out[6]) {
    if (!s || !out) return fail(B200RT_EINVAL, "null argument");
    std::memcpy(out, s->bounds, sizeof(s->bounds));
    return B200RT_OK;
}

// ---- launches -----------------------------------------------------------------------------------
static bool is_device_ptr(const void *p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

template <bool ANY, bool COUNT>
static void launch_kind(const b200rt_scene *s, const float4 *rays, uint64_t n, const uint32_t *n_dev, b200rt_hit *hits,
                        uint8_t *occ, Counters *cnt, cudaStream_t st, uint4 *per_ray = nullptr) {
    const uint64_t blocks64 = (n + B200RT_BLOCK - 1) / B200RT_BLOCK;
    const unsigned blocks = (unsigned)std::min<uint64_t>(blocks64, (uint64_t)g_sm_count * 64);
    switch (s->kind) {
        case B200RT_BVH: trace_kernel<B200RT_BVH, ANY, COUNT><<<blocks, B200RT_BLOCK, 0, st>>>(s->bvh, s->bsp, rays, n, n_dev, hits, occ, cnt, per_ray); break;
        case B200RT_KDTREE: trace_kernel<B200RT_KDTREE, ANY, COUNT><<<blocks, B200RT_BLOCK, 0, st>>>(s->bvh, s->bsp, rays, n, n_dev, hits, occ, cnt, per_ray); break;
        case B200RT_RBSP: trace_kernel<B200RT_RBSP, ANY, COUNT><<<blocks, B200RT_BLOCK, 0, st>>>(s->bvh, s->bsp, rays, n, n_dev, hits, occ, cnt, per_ray); break;
        case B200RT_RBSPKD: trace_kernel<B200RT_RBSPKD, ANY, COUNT><<<blocks, B200RT_BLOCK, 0, st>>>(s->bvh, s->bsp, rays, n, n_dev, hits, occ, cnt, per_ray); break;
        case B200RT_BSP: trace_kernel<B200RT_BSP, ANY, COUNT><<<blocks, B200RT_BLOCK, 0, st>>>(s->bvh, s->bsp, rays, n, n_dev, hits, occ, cnt, per_ray); break;
        default: trace_kernel<B200RT_BSPKD, ANY, COUNT><<<blocks, B200RT_BLOCK, 0, st>>>(s->bvh, s->bsp, rays, n, n_dev, hits, occ, cnt, per_ray); break;
    }
    g_launches.fetch_add(1);
}

// Fills the result for an empty scene (every ray misses) without touching a tree.
__global__ void miss_kernel(const float4 *__restrict__ rays, uint64_t n, b200rt_hit *hits, uint8_t *occ) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (occ) occ[i] = 0;
    if (hits) {
        b200rt_hit h;
        h.prim = B200RT_MISS; h.t = rays[2 * i].w; h.b0 = h.b1 = h.b2 = 0.f;
        hits[i] = h;
    }
}

int b200rt::launch_trace(const b200rt_scene *s, const float4 *rays, uint64_t n, const uint32_t *n_dev, b200rt_hit *hits,
                         uint8_t *occ, bool any, Counters *counters, cudaStream_t st, uint4 *per_ray, bool coherent) {
    if (n == 0) return B200RT_OK;
    if (s->empty) {
        miss_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(rays, n, any ? nullptr : hits, any ? occ : nullptr);
        g_launches.fetch_add(1);
    } else if (!counters && !g_use_baseline && (s->kind == B200RT_BVH || (g_persist_bsp && !(coherent && g_coherent_baseline)))) {
        // (kd / RBSP / BSP trees on a COHERENT queue — the wavefront's camera rays and their shadow rays — go to the
        // one-ray-per-thread kernel below: every lane is busy anyway, so the persistent kernel's votes are pure overhead there:
        // museum kd camera 1.80 vs 2.46 ms, first shadow queue 1.17 vs 1.47 ms, conference kd camera +40 %
        // (profiles/r02_ab_kd_queues.log); the BVH's persistent kernel is as fast on them, 4.31 vs 4.25 Grays/s)
        // production path: persistent warps (rt_persist.cuh); <= 2^30 rays per launch
        const uint64_t kMax = 1ull << 30;
        for (uint64_t off = 0; off < n; off += kMax) {
            const uint64_t m = std::min<uint64_t>(kMax, n - off);
            if (n_dev && off) return fail(B200RT_EINVAL, "device-side ray counts need n <= 2^30");
            PersistCounters *pc = s->d_pc + (s->pc_next.fetch_add(1) % B200RT_PC_SLOTS);
            const int perSM = std::max(1, g_pblocks[s->kind][any ? 1 : 0]);
            const uint64_t maxBlocks = (uint64_t)g_sm_count * perSM, wpb = B200RT_PBLOCK / 32;
            // big queues: 128-ray chunks (one atomic per 128 rays); small ones: 32-ray chunks so that
            // every resident warp gets work
            const uint32_t chunk = m >= maxBlocks * wpb * 128 * 4 ? 128u : 32u;
            const unsigned blocks = (unsigned)std::min<uint64_t>((m + chunk * wpb - 1) / (chunk * wpb), maxBlocks);
            b200rt_hit *ho = any ? nullptr : hits + off;
            uint8_t *oo = any ? occ + off : nullptr;
            const float4 *ri = rays + 2 * off;
            const uint32_t bspStepsW = pack_steps(g_bsp_steps, g_bsp_vote[0], g_bsp_vote[1]);
            uint32_t bvhStepsW = pack_steps(g_bvh_steps, g_bvh_vote[0], g_bvh_vote[1]);
#define PL(fn, view) fn<<<blocks, B200RT_PBLOCK, (size_t)g_sstack_bsp * 12 * B200RT_PBLOCK, st>>>(view, ri, m, n_dev, ho, oo, pc, g_thresh, chunk, g_sstack_bsp, bspStepsW)
            switch (s->kind * 2 + (any ? 1 : 0)) {
                case B200RT_BVH * 2: case B200RT_BVH * 2 + 1: {
                    uint32_t topN = (g_toplevels && !g_bvh4) ? s->bvh.topCount[g_toplevels] : 0u;
                    const size_t smem = (size_t)g_sstack * sizeof(uint2) * B200RT_PBLOCK + (size_t)topN * sizeof(BvhNode2);
                    unsigned blk = blocks;
                    if (topN) {   // experiment path: shared memory per block depends on the scene, so does the residency
                        // (b200rt_init sets the attribute for the stack alone; the staged records come on top, per scene)
                        if (g_bvh_attr[any ? 1 : 0] < smem) {
                            CU(cudaFuncSetAttribute(bvh_fn(any), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                            g_bvh_attr[any ? 1 : 0] = smem;
                        }
                        int per = 0;
                        CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, bvh_fn(any), B200RT_PBLOCK, smem));
                        blk = (unsigned)std::min<uint64_t>((m + chunk * wpb - 1) / (chunk * wpb), (uint64_t)g_sm_count * std::max(1, per));
                    }
                    void *args[] = {(void *)&s->bvh, (void *)&ri, (void *)&m, (void *)&n_dev, (void *)&ho, (void *)&oo, (void *)&pc,
                                    (void *)&g_thresh, (void *)&chunk, (void *)&bvhStepsW, (void *)&g_sstack, (void *)&topN};
                    CU(cudaLaunchKernel(bvh_fn(any), dim3(blk), dim3(B200RT_PBLOCK), args, smem, st));
                    break;
                }
                case B200RT_KDTREE * 2:
                    if (g_kd_prefetch) PL((bsp_persistent_kernel<B200RT_KDTREE, false, true>), s->bsp);
                    else PL((bsp_persistent_kernel<B200RT_KDTREE, false>), s->bsp);
                    break;
                case B200RT_KDTREE * 2 + 1:
                    if (g_kd_prefetch) PL((bsp_persistent_kernel<B200RT_KDTREE, true, true>), s->bsp);
                    else PL((bsp_persistent_kernel<B200RT_KDTREE, true>), s->bsp);
                    break;
                case B200RT_RBSP * 2: PL((bsp_persistent_kernel<B200RT_RBSP, false>), s->bsp); break;
                case B200RT_RBSP * 2 + 1: PL((bsp_persistent_kernel<B200RT_RBSP, true>), s->bsp); break;
                case B200RT_RBSPKD * 2: PL((bsp_persistent_kernel<B200RT_RBSPKD, false>), s->bsp); break;
                case B200RT_RBSPKD * 2 + 1: PL((bsp_persistent_kernel<B200RT_RBSPKD, true>), s->bsp); break;
                case B200RT_BSP * 2: PL((bsp_persistent_kernel<B200RT_BSP, false>), s->bsp); break;
                case B200RT_BSP * 2 + 1: PL((bsp_persistent_kernel<B200RT_BSP, true>), s->bsp); break;
                case B200RT_BSPKD * 2: PL((bsp_persistent_kernel<B200RT_BSPKD, false>), s->bsp); break;
                default: PL((bsp_persistent_kernel<B200RT_BSPKD, true>), s->bsp); break;
            }
#undef PL
#undef PLB
            g_launches.fetch_add(1);
        }
    } else if (any) {
        if (counters) launch_kind<true, true>(s, rays, n, n_dev, nullptr, occ, counters, st, per_ray);
        else launch_kind<true, false>(s, rays, n, n_dev, nullptr, occ, nullptr, st);
    } else {
        if (counters) launch_kind<false, true>(s, rays, n, n_dev, hits, nullptr, counters, st, per_ray);
        else launch_kind<false, false>(s, rays, n, n_dev, hits, nullptr, nullptr, st);
    }
    CU(cudaGetLastError());
    return B200RT_OK;
}

static int ensure(void **p, size_t *have, size_t want) {
    if (*have >= want) return B200RT_OK;
    cudaFree(*p);
    *p = nullptr; *have = 0;
    CU(cudaMalloc(p, want));
    *have = want;
    return B200RT_OK;
}

// Host-buffer batches: chunked 3-stage pipeline (H2D | trace | D2H) over three streams.

// Host rays in, host results out, in chunks through the scene's three-stream pipeline (scene_internal.h).  `wait`: block until
// the results are in `out` (the synchronous entry points); otherwise `user` is made to wait for them and the call returns.
static int run_pipelined(b200rt_scene *s, const b200rt_ray *rays, uint64_t n, void *out, bool any, cudaStream_t user, bool wait) {
    const size_t out_elem = any ? 1 : sizeof(b200rt_hit);
    if (!s->lanes_ready) {
        for (cudaStream_t *st : {&s->sIn, &s->sK[0], &s->sK[1], &s->sK[2], &s->sK[3], &s->sOut}) CU(cudaStreamCreateWithFlags(st, cudaStreamNonBlocking));
        for (auto &q : s->slots) {
            for (cudaEvent_t *e : {&q.evIn, &q.evK, &q.evOut}) CU(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
            CU(cudaMalloc(&q.d_in, (size_t)B200RT_PIPE_CHUNK_MAX * sizeof(b200rt_ray)));
            CU(cudaMalloc(&q.d_out, (size_t)B200RT_PIPE_CHUNK_MAX * sizeof(b200rt_hit)));
        }
        CU(cudaEventCreateWithFlags(&s->ev_start, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&s->ev_end, cudaEventDisableTiming));
        s->lanes_ready = true;
    }
    if (wait) {   // the blocking calls start after whatever the caller queued on its stream
        CU(cudaEventRecord(s->ev_start, user));
        CU(cudaStreamWaitEvent(s->sIn, s->ev_start, 0));
    }   // the enqueue calls read `rays` from the moment they are called: waiting for `user` here would chain every call behind
        // the previous one's results (the stream is made to wait for those below) and the queues would not overlap
    const b200rt_ray *mapped = nullptr;   // zero-copy experiment: the device's view of pinned host rays
    if (g_pipe_zerocopy) {
        cudaPointerAttributes a;
        if (cudaPointerGetAttributes(&a, rays) == cudaSuccess && a.type == cudaMemoryTypeHost && a.devicePointer) mapped = (const b200rt_ray *)a.devicePointer;
        else cudaGetLastError();
    }
    // Chunk schedule: a short head (c/4, c/2) so the first kernel starts after 1/4 of a chunk's copy time, full chunks
    // in the middle, and the last chunk split in two so the un-overlapped tail (last kernel + last D2H) is short too.
    int idx = 0;
    uint64_t m = 0;
    for (uint64_t off = 0; off < n; off += m, ++idx) {
        b200rt_scene::Slot &q = s->slots[s->slot_next++ % b200rt_scene::kSlots];
        const uint64_t rem = n - off, c = g_pipe_chunk;
        m = std::min<uint64_t>(c, rem);
        if (g_pipe_ramp) {
            if (idx == 0) m = std::min<uint64_t>(rem, std::max<uint64_t>(c / 4, 1024));
            else if (idx == 1) m = std::min<uint64_t>(rem, std::max<uint64_t>(c / 2, 1024));
            else if (rem <= c && rem > c / 2) m = rem / 2;
        }
        cudaStream_t sK = s->sK[s->chunk_next++ % g_pipe_kstreams];
        if (q.used) {   // the slot's previous chunk: its kernel has read d_in, its D2H has drained d_out
            CU(cudaStreamWaitEvent(s->sIn, q.evK, 0));
            CU(cudaStreamWaitEvent(sK, q.evOut, 0));
        }
        if (!mapped) {
            CU(cudaMemcpyAsync(q.d_in, rays + off, m * sizeof(b200rt_ray), cudaMemcpyHostToDevice, s->sIn));
            CU(cudaEventRecord(q.evIn, s->sIn));
            CU(cudaStreamWaitEvent(sK, q.evIn, 0));
        }
        int rc = g_pipe_nokernel ? 0 : launch_trace(s, mapped ? (const float4 *)(mapped + off) : (const float4 *)q.d_in, m, nullptr, any ? nullptr : (b200rt_hit *)q.d_out,
                                                    any ? (uint8_t *)q.d_out : nullptr, any, nullptr, sK);
        if (rc) return rc;
        CU(cudaEventRecord(q.evK, sK));
        CU(cudaStreamWaitEvent(s->sOut, q.evK, 0));
        CU(cudaMemcpyAsync((char *)out + off * out_elem, q.d_out, m * out_elem, cudaMemcpyDeviceToHost, s->sOut));
        CU(cudaEventRecord(q.evOut, s->sOut));
        q.used = true;
    }
    CU(cudaEventRecord(s->ev_end, s->sOut));   // D2H copies complete in order: the last one done = all done
    CU(cudaStreamWaitEvent(user, s->ev_end, 0));
    if (wait) CU(cudaEventSynchronize(s->ev_end));
    return B200RT_OK;
}

static int run(b200rt_scene *s, const b200rt_ray *rays, uint64_t n, b200rt_hit *hits, uint8_t *occ, bool any,
               b200rt_counters *counters, void *stream, b200rt_ray_stats *per_ray = nullptr, bool async = false) {
    if (!s) return fail(B200RT_EINVAL, "null scene");
    b200rt_counters dummy;
    if (per_ray && !counters) counters = &dummy;  // per-ray stats come from the instrumented kernels
    uint4 *d_per_ray = nullptr;
    bool per_ray_tmp = false;
    if (per_ray && n) {
        if (is_device_ptr(per_ray)) d_per_ray = (uint4 *)per_ray;
        else { CU(cudaMalloc(&d_per_ray, n * sizeof(uint4))); per_ray_tmp = true; }
        CU(cudaMemsetAsync(d_per_ray, 0, n * sizeof(uint4), (cudaStream_t)stream));
    }
    struct Guard { uint4 *p; bool own; ~Guard() { if (own) cudaFree(p); } } guard{d_per_ray, per_ray_tmp};
    if (n == 0) { if (counters) std::memset(counters, 0, sizeof(*counters)); return B200RT_OK; }
    if (!rays || (any ? !occ : !hits)) return fail(B200RT_EINVAL, "null ray/result buffer");
    cudaStream_t st = (cudaStream_t)stream;
    const bool rays_dev = is_device_ptr(rays);
    const void *outp = any ? (const void *)occ : (const void *)hits;
    const bool out_dev = is_device_ptr(outp);
    const size_t out_elem = any ? 1 : sizeof(b200rt_hit);
    std::unique_lock<std::mutex> lock(s->mu, std::defer_lock);
    const float4 *d_rays = (const float4 *)rays;
    void *d_out = (void *)outp;
    if (!rays_dev || !out_dev || counters) lock.lock();
    if (!rays_dev && !out_dev && !counters && (async || n > g_pipe_chunk))
        return run_pipelined(s, rays, n, (void *)outp, any, st, !async);
    if (async) return fail(B200RT_EINVAL, "the asynchronous calls take HOST rays and HOST results (device buffers are asynchronous through the plain calls)");
    if (!rays_dev) {
        int rc = ensure(&s->d_stage_in, &s->stage_in_bytes, n * sizeof(b200rt_ray));
        if (rc) return rc;
        CU(cudaMemcpyAsync(s->d_stage_in, rays, n * sizeof(b200rt_ray), cudaMemcpyHostToDevice, st));
        d_rays = (const float4 *)s->d_stage_in;
    }
    if (!out_dev) {
        int rc = ensure(&s->d_stage_out, &s->stage_out_bytes, n * out_elem);
        if (rc) return rc;
        d_out = s->d_stage_out;
    }
    if (counters) CU(cudaMemsetAsync(s->d_counters, 0, sizeof(Counters), st));
    {
        int rc = launch_trace(s, d_rays, n, nullptr, any ? nullptr : (b200rt_hit *)d_out, any ? (uint8_t *)d_out : nullptr, any,
                              counters ? s->d_counters : nullptr, st, s->empty ? nullptr : d_per_ray);
        if (rc) return rc;
    }
    CU(cudaGetLastError());
    if (per_ray_tmp) CU(cudaMemcpyAsync(per_ray, d_per_ray, n * sizeof(uint4), cudaMemcpyDeviceToHost, st));
    if (!out_dev) CU(cudaMemcpyAsync((void *)outp, d_out, n * out_elem, cudaMemcpyDeviceToHost, st));
    if (counters) {
        Counters c;
        CU(cudaMemcpyAsync(&c, s->d_counters, sizeof(c), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        counters->rays = s->empty ? n : c.rays; counters->nodes_fetched = c.nodes_fetched; counters->nodes_passed = c.nodes_passed;
        counters->leaf_visits = c.leaf_visits; counters->prim_tests = c.prim_tests; counters->index_reads = c.index_reads;
    } else if (!rays_dev || !out_dev) {
        CU(cudaStreamSynchronize(st));
    }
    return B200RT_OK;
}

extern "C" int b200rt_intersect_closest(b200rt_scene *s, const b200rt_ray *rays, uint64_t n, b200rt_hit *hits, void *stream) {
    return run(s, rays, n, hits, nullptr, false, nullptr, stream);
}
extern "C" int b200rt_intersect_any(b200rt_scene *s, const b200rt_ray *rays, uint64_t n, uint8_t *occ, void *stream) {
    return run(s, rays, n, nullptr, occ, true, nullptr, stream);
}
extern "C" int b200rt_intersect_closest_async(b200rt_scene *s, const b200rt_ray *rays, uint64_t n, b200rt_hit *hits, void *stream) {
    return run(s, rays, n, hits, nullptr, false, nullptr, stream, nullptr, true);
}
extern "C" int b200rt_intersect_any_async(b200rt_scene *s, const b200rt_ray *rays, uint64_t n, uint8_t *occ, void *stream) {
    return run(s, rays, n, nullptr, occ, true, nullptr, stream, nullptr, true);
}
extern "C" int b200rt_intersect_wait(b200rt_scene *s) {
    if (!s) return fail(B200RT_EINVAL, "null scene");
    std::lock_guard<std::mutex> lock(s->mu);
    if (s->lanes_ready) CU(cudaEventSynchronize(s->ev_end));
    return B200RT_OK;
}
extern "C" int b200rt_intersect_closest_counted(b200rt_scene *s, const b200rt_ray *rays, uint64_t n, b200rt_hit *hits,
                                                b200rt_counters *c, void *stream) {
    if (!c) return fail(B200RT_EINVAL, "null counters");
    return run(s, rays, n, hits, nullptr, false, c, stream);
}
extern "C" int b200rt_intersect_any_counted(b200rt_scene *s, const b200rt_ray *rays, uint64_t n, uint8_t *occ,
                                            b200rt_counters *c, void *stream) {
    if (!c) return fail(B200RT_EINVAL, "null counters");
    return run(s, rays, n, nullptr, occ, true, c, stream);
}
static_assert(sizeof(b200rt_ray_stats) == sizeof(uint4), "b200rt_ray_stats is four 32-bit words");
extern "C" int b200rt_intersect_closest_stats(b200rt_scene *s, const b200rt_ray *rays, uint64_t n, b200rt_hit *hits,
                                              b200rt_ray_stats *per_ray, void *stream) {
    if (!per_ray) return fail(B200RT_EINVAL, "null per-ray stats buffer");
    return run(s, rays, n, hits, nullptr, false, nullptr, stream, per_ray);
}
extern "C" int b200rt_intersect_any_stats(b200rt_scene *s, const b200rt_ray *rays, uint64_t n, uint8_t *occ,
                                          b200rt_ray_stats *per_ray, void *stream) {
    if (!per_ray) return fail(B200RT_EINVAL, "null per-ray stats buffer");
    return run(s, rays, n, nullptr, occ, true, nullptr, stream, per_ray);
}
