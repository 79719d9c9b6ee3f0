// Internal: the scene object behind the opaque b200rt_scene handle, shared by b200rt.cu and wavefront.cu.
#pragma once
#include <mutex>
#include <atomic>
#include "rt_persist.cuh"

struct b200rt_scene {
    int kind = 0;
    int device = 0;
    b200rt::BvhView bvh{};
    b200rt::BspView bsp{};
    void *d_nodes = nullptr;
    void *d_nodes4 = nullptr;       // BVH: two-levels-per-record layout (BvhNode4)
    uint32_t *d_indices = nullptr;
    float4 *d_tris = nullptr;       // traversal order (BVH: leaf slots; kd/RBSP: original primitive order)
    float4 *d_spheres = nullptr;    // 4 x float4 per sphere primitive (rt_math.cuh sphere_test)
    float4 *d_prim_tris = nullptr;  // original primitive order, for the shading stage (== d_tris for kd/RBSP)
    float *d_prim_uv = nullptr;     // 6 floats per primitive (ABI v8; nullptr = the default UVs)
    float *d_prim_n = nullptr;      // 9 floats per primitive: vertex normals (nullptr = none)
    uint8_t *d_prim_flags = nullptr;  // bit 0 orientation flip, bit 1 has vertex normals
    float bounds[6] = {0, 0, 0, 0, 0, 0};
    uint32_t n_prims = 0;
    bool empty = false;  // zero primitives: every ray misses (bvh.cpp:355 `if (!nodes) return false`)
    // staging for host-pointer calls
    std::mutex mu;
    void *d_stage_in = nullptr, *d_stage_out = nullptr;
    size_t stage_in_bytes = 0, stage_out_bytes = 0;
    b200rt::Counters *d_counters = nullptr;
    // host-pointer calls are pipelined in chunks over three streams BY FUNCTION — host-to-device copies, kernels, device-to-host
    // copies — and a ring of staging slots: the H2D engine runs ahead of the kernels by up to the ring's depth and never waits
    // for a D2H of its own chunk; consecutive (asynchronous) calls continue the same pipeline without draining it
    struct Slot {
        void *d_in = nullptr, *d_out = nullptr;
        cudaEvent_t evIn = nullptr, evK = nullptr, evOut = nullptr;   // H2D done, kernel done (d_in free), D2H done (d_out free)
        bool used = false;
    };
    static constexpr int kSlots = 6, kKernelStreams = 4;
    Slot slots[kSlots];
    uint32_t slot_next = 0;
    cudaStream_t sIn = nullptr, sK[kKernelStreams] = {nullptr, nullptr, nullptr, nullptr}, sOut = nullptr;   // kernels of consecutive chunks alternate over g_pipe_kstreams streams: their tails overlap
    uint32_t chunk_next = 0;
    cudaEvent_t ev_start = nullptr, ev_end = nullptr;
    bool lanes_ready = false;
    // persistent kernels: ring of work-counter slots (one per in-flight launch; each slot is re-armed by
    // the last warp of the launch that used it)
    b200rt::PersistCounters *d_pc = nullptr;
    mutable std::atomic<uint32_t> pc_next{0};
};


// frees the wavefront workspace attached to a scene (wavefront.cu)
void b200rt_wavefront_release(const b200rt_scene *s);

namespace b200rt {
// Launches the traversal kernel on device-resident buffers.  n_dev (nullable): device-side ray count
// (queue sizes produced by the wavefront stages) — the kernel traces min(n, *n_dev) rays.
int launch_trace(const b200rt_scene *s, const float4 *rays, uint64_t n, const uint32_t *n_dev, b200rt_hit *hits,
                 uint8_t *occ, bool any, Counters *counters, cudaStream_t st, uint4 *per_ray = nullptr, bool coherent = false);
int set_error(int code, const char *fmt, ...);
}  // namespace b200rt
