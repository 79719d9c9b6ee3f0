// b200rt_intersect_closest_noalloc: Aggregate::Intersect over a batch of HOST rays without any allocation inside the call.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The existing host-buffer entry point
// (b200rt_intersect_closest, b200rt.cu `run` / `run_pipelined`) stages rays and hits in device buffers it cudaMalloc's on
// first use and grows on demand; here the same traversal kernels (launch_trace: the persistent-warp kernels of
// rt_persist.cuh, every tree kind) read their ray queue from and write their hit records to a file-scope
// `static __device__` scratch array of fixed size.  Every GPU that loads the library holds the array (3.25 MiB) until exit.
//
// The array's NAME is a host-side stand-in in host code, not a device address: the device address is fetched once per
// device with cudaGetSymbolAddress (that device current, return code checked) and that pointer is what the kernels get.
//
// All calls share the array and it keeps the last call's contents: the hit records of this call's n rays are cleared before
// the launch (the kernels overwrite every one of them anyway), nothing else has to start at zero.  Calls that overlap on the
// GPU (two host threads, two processes' worth of streams on one context) would race on the array; inside one process the
// entry point serialises callers on a mutex and returns only after its stream has drained, so overlap cannot happen through
// this entry point itself.
#include <cuda_runtime.h>

#include <mutex>

#include "../../include/b200rt.h"
#include "scene_internal.h"

namespace {

constexpr uint64_t kMaxRays = B200RT_NOALLOC_MAX_RAYS;
constexpr size_t kRayBytes = kMaxRays * sizeof(b200rt_ray);       // 2 MiB
constexpr size_t kHitBytes = kMaxRays * sizeof(b200rt_hit);       // 1.25 MiB
static_assert(sizeof(b200rt_ray) == 32 && sizeof(b200rt_hit) == 20, "record sizes of include/b200rt.h");
static_assert(kRayBytes % 256 == 0, "the hit region starts 256-byte aligned");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kRayBytes + kHitBytes];

constexpr int kMaxDevices = 64;
unsigned char *g_scratch_ptr[kMaxDevices] = {};   // device address of g_noalloc_scratch, per device
std::mutex g_mu;

}  // namespace

extern "C" int b200rt_intersect_closest_noalloc(b200rt_scene *s, const b200rt_ray *rays, uint64_t n, b200rt_hit *hits,
                                                void *stream) {
    using b200rt::set_error;
    if (!s || (n && (!rays || !hits))) return set_error(B200RT_EINVAL, "b200rt_intersect_closest_noalloc: null argument");
    if (n > kMaxRays)
        return set_error(B200RT_EINVAL, "b200rt_intersect_closest_noalloc: %llu rays exceed the static scratch (B200RT_NOALLOC_MAX_RAYS = %llu); "
                         "split the batch or use b200rt_intersect_closest", (unsigned long long)n, (unsigned long long)kMaxRays);
    if (n == 0) return B200RT_OK;
    std::lock_guard<std::mutex> lock(g_mu);
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return set_error(B200RT_ENODEVICE, "b200rt_intersect_closest_noalloc: cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev != s->device) return set_error(B200RT_EINVAL, "b200rt_intersect_closest_noalloc: the scene lives on device %d, the current device is %d", s->device, dev);
    if (dev < 0 || dev >= kMaxDevices) return set_error(B200RT_EUNSUPPORTED, "b200rt_intersect_closest_noalloc: device index %d", dev);
    if (!g_scratch_ptr[dev]) {   // set-up, once per device, with that device current
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) return set_error(B200RT_ECUDA, "b200rt_intersect_closest_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        g_scratch_ptr[dev] = (unsigned char *)p;
    }
    unsigned char *base = g_scratch_ptr[dev];
    float4 *dRays = (float4 *)base;
    b200rt_hit *dHits = (b200rt_hit *)(base + kRayBytes);
    cudaStream_t st = (cudaStream_t)stream;
    e = cudaMemcpyAsync(dRays, rays, n * sizeof(b200rt_ray), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemsetAsync(dHits, 0, n * sizeof(b200rt_hit), st);   // the last call's records
    if (e != cudaSuccess) return set_error(B200RT_ECUDA, "b200rt_intersect_closest_noalloc: upload: %s", cudaGetErrorString(e));
    const int rc = b200rt::launch_trace(s, dRays, n, nullptr, dHits, nullptr, false, nullptr, st);
    if (rc) return rc;
    e = cudaMemcpyAsync(hits, dHits, n * sizeof(b200rt_hit), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return set_error(B200RT_ECUDA, "b200rt_intersect_closest_noalloc: %s", cudaGetErrorString(e));
    return B200RT_OK;
}
