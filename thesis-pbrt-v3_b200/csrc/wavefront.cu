// Wavefront path-tracing driver (include/b200rt_wavefront.h).
//
// SamplerIntegrator::Render's tile/pixel/sample loops (reference src/core/integrator.cpp:230-360)
// and PathIntegrator::Li's bounce loop (src/integrators/path.cpp:64-204) restructured into
// device-resident ray queues.  Per wavefront step:
//   closest-hit queue --trace--> hits --shade--> { shadow queue --any-hit--> resolve , next closest queue }
// Queue sizes live on the device (no host round trip inside a batch); queues are compacted with
// warp ballots + one atomic per warp.  The arithmetic of a path vertex is wf_shade.cuh (host/device, pinned against
// the compiled reference on the CPU by tests/hostshade); this file holds the kernels around it and the host driver.
#include <cuda_runtime.h>
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <vector>
#include "../../include/b200rt_wavefront.h"
#include "scene_internal.h"
#include "wf_shade.cuh"
#include "wf_sampler_host.h"
#include "wf_setup_host.h"

using namespace b200rt;

extern std::atomic<uint64_t> g_launches;
extern int g_sm_count;

#define WCU(call)                                                                                   \
    do {                                                                                            \
        cudaError_t e_ = (call);                                                                    \
        if (e_ != cudaSuccess)                                                                      \
            return set_error(B200RT_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

namespace {

typedef WfTiling Tiling;

// warp-aggregated queue append: one atomic per warp (ballot + popc + shuffle)
__device__ __forceinline__ uint32_t warp_append(bool pred, uint32_t *counter) {
    const unsigned m = __ballot_sync(0xFFFFFFFFu, pred);
    if (m == 0) return 0;
    const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(counter, (uint32_t)__popc(m));
    base = __shfl_sync(0xFFFFFFFFu, base, leader);
    return base + __popc(m & ((1u << lane) - 1u));
}

// Block-aggregated append to up to three queues at once: one atomic per queue per BLOCK.  Every warp of a launch appends to the
// same two or three counters; with one atomic per warp the shade kernel spent 18 % of its stall samples waiting for them
// (same-address atomics serialise in L2).  All threads of the block must call it.  slot[q] is only meaningful where pred[q].
#define B200RT_WF_TB 256
#ifndef B200RT_SHADE_PHASES
#define B200RT_SHADE_PHASES 1   // k_shade<true>: a block-wide barrier between the two halves of the vertex (0: one call, A/B)
#endif
__device__ __forceinline__ void block_append3(bool p0, bool p1, bool p2, uint32_t *c0, uint32_t *c1, uint32_t *c2,
                                              uint32_t &s0, uint32_t &s1, uint32_t &s2) {
    __shared__ uint32_t wcount[3][32], wbase[3][32];   // up to 1024 threads a block
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    const unsigned m0 = __ballot_sync(0xFFFFFFFFu, p0), m1 = __ballot_sync(0xFFFFFFFFu, p1), m2 = __ballot_sync(0xFFFFFFFFu, p2);
    if (lane == 0) { wcount[0][warp] = __popc(m0); wcount[1][warp] = __popc(m1); wcount[2][warp] = __popc(m2); }
    __syncthreads();
    if (threadIdx.x < 3) {
        const uint32_t q = threadIdx.x;
        uint32_t total = 0;
        for (uint32_t w = 0; w < nw; ++w) { wbase[q][w] = total; total += wcount[q][w]; }
        uint32_t *c = q == 0 ? c0 : (q == 1 ? c1 : c2);
        const uint32_t base = total ? atomicAdd(c, total) : 0u;
        for (uint32_t w = 0; w < nw; ++w) wbase[q][w] += base;
    }
    __syncthreads();
    const uint32_t below = (1u << lane) - 1u;
    s0 = wbase[0][warp] + __popc(m0 & below);
    s1 = wbase[1][warp] + __popc(m1 & below);
    s2 = wbase[2][warp] + __popc(m2 & below);
}

// Sampler "sobol": SobolSampleFloat XORs one matrix column per set bit of the sample index (lowdiscrepancy.h:252-266).  The same
// XOR by bytes: T[dim][b][x] = XOR of the columns 8 b + i over the set bits i of x, so that a 32-bit index costs four table reads
// instead of a loop over its set bits (a quarter of k_shade<true>'s instructions on scenes/living-room).
__global__ void k_sobol_bytes(const uint32_t *__restrict__ sobol32, uint32_t *__restrict__ T) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (uint32_t)kSobolDims * 1024u) return;
    const uint32_t dim = i >> 10, b = (i >> 8) & 3u, x = i & 255u;
    const uint32_t *col = sobol32 + dim * kSobolMatrixSize + 8u * b;
    uint32_t v = 0;
    for (uint32_t k = 0; k < 8; ++k)
        if ((x >> k) & 1u) v ^= col[k];
    T[i] = v;
}

// SpatialLightDistribution: one thread computes one voxel's Distribution1D (wf_shade.cuh spatial_voxel_distribution)
__global__ void k_light_voxels(const WfLight *__restrict__ lights, uint32_t nLights, const float *__restrict__ wb6, const uint32_t *__restrict__ nVox3,
                               uint32_t nVoxels, const float *__restrict__ halton5, float *cdfHi, float *pickPdf) {
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nVoxels) return;
    float contrib[B200RT_MAX_LIGHTS];
    const float wb[6] = {wb6[0], wb6[1], wb6[2], wb6[3], wb6[4], wb6[5]};
    const uint32_t nVox[3] = {nVox3[0], nVox3[1], nVox3[2]};
    spatial_voxel_distribution(lights, nLights, wb, nVox, v, halton5, contrib, cdfHi + (size_t)v * nLights, pickPdf + (size_t)v * nLights);
}

// ---- stage kernels ------------------------------------------------------------------------------
// Camera rays for path slots [batchStart, batchStart + nSlots) (Sampler::GetCameraSample + PerspectiveCamera::
// GenerateRayDifferential, src/cameras/perspective.cpp:97-143, pinhole; wf_shade.cuh camera_ray).
__global__ void k_camera(const __grid_constant__ WfParams P, uint64_t batchStart, uint32_t nSlots, float4 *rays,
                         uint32_t *qPath, uint32_t *count, float4 *beta, float4 *L, uint32_t *pdim, uint4 *pinfo) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    bool active = false;
    uint32_t x = 0, y = 0, s = 0;
    if (i < nSlots) {
        active = path_pixel(P.tl, batchStart + i, x, y, s);
        beta[i] = make_float4(1.f, 1.f, 1.f, 0.f);
        L[i] = make_float4(0.f, 0.f, 0.f, active ? 1.f : 0.f);
        pdim[i] = kCameraDims;
        // the path's pixel, sample number and global sample index: the same at every bounce, so k_shade reads them back
        // instead of redoing path_pixel's divisions and GetIndexForSample
        const uint64_t idx = active ? path_sample_index(P, x, y, s) : 0ull;
        pinfo[i] = make_uint4(x | (y << 16), s, (uint32_t)idx, (uint32_t)(idx >> 32));
    }
    const uint32_t slot = warp_append(active, count);
    if (!active) return;
    V3 o, d;
    camera_ray(P, x, y, s, o, d);
    rays[2 * (size_t)slot] = make_float4(o.x, o.y, o.z, __int_as_float(0x7f800000));
    rays[2 * (size_t)slot + 1] = make_float4(d.x, d.y, d.z, 0.f);
    qPath[slot] = i;
}

// Shade one wavefront: PathIntegrator::Li's body for one bounce (wf_shade.cuh shade_vertex), then the queue appends.
// FULL = false: the lean instantiation for one point light + one matte Kd (the bench workload); FULL = true adds the
// area / distant / infinite lights with MIS, the per-primitive material table and the specular bookkeeping.
// No __launch_bounds__ on purpose.  k_shade<true> waits for instructions and for its slowest warps (ncu on scenes/living-room:
// 15 no_instruction and 6.6 barrier stalls per issue at 16 resident warps), so 3 and 4 resident blocks per SM were tried: 80 / 64
// registers with spills, no faster (741 / 751 ms against 720 in a run whose frames still included the buffer allocation,
// profiles/r02_ab_shade_full.log); and the mere bound (256, 1) on the lean
// instantiation changed its code enough (84 registers) to cost the headline frame 5 %, 106.3 against 101.3 ms on the same box.
template <bool FULL>
__global__ void k_shade(const __grid_constant__ WfParams P, uint32_t depth, uint64_t batchStart,
                        const uint32_t *__restrict__ nCur, const float4 *__restrict__ rays,
                        const uint32_t *__restrict__ qPath, const b200rt_hit *__restrict__ hits,
                        const float4 *__restrict__ primTris, float4 *beta, uint32_t *pdim, const uint4 *__restrict__ pinfo,
                        float4 *nextRays, uint32_t *nextPath,
                        uint32_t *nNext, float4 *shRays, uint32_t *shPath, float4 *shContrib, uint32_t *nShadow,
                        float4 *L, float4 *misRays, uint32_t *misPath, float4 *misContrib, uint32_t *nMis) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t n = *nCur;
    if (blockIdx.x * blockDim.x >= n) return;   // the launch covers the batch's capacity, the queue is usually shorter (block-uniform)
    const bool mis = FULL && P.anyMis;
    ShadeOut o;
    o.wantShadow = o.wantNext = o.wantMis = false;
    uint32_t path = 0;
    if (!FULL || !B200RT_SHADE_PHASES) {
        if (i < n) {
            const b200rt_hit h = hits[i];
            path = qPath[i];
            const float4 ro = rays[2 * (size_t)i], rd = rays[2 * (size_t)i + 1];
            const float4 bt = beta[path];
            V3 p0 = v3(0.f, 0.f, 0.f), p1 = p0, p2 = p0;
            bool isSphere = false;
            if (h.prim != B200RT_MISS) {
                const float4 a = __ldg(primTris + 3 * (size_t)h.prim), b = __ldg(primTris + 3 * (size_t)h.prim + 1),
                             c = __ldg(primTris + 3 * (size_t)h.prim + 2);
                p0 = v3(a.x, a.y, a.z); p1 = v3(b.x, b.y, b.z); p2 = v3(c.x, c.y, c.z);
                isSphere = (__float_as_uint(b.w) & B200RT_PRIM_SPHERE) != 0;
            }
            const uint4 pi = pinfo[path];
            const uint32_t x = pi.x & 0xffffu, y = pi.x >> 16, s = pi.y;
            shade_vertex<FULL>(P, depth, x, y, s, (uint64_t)pi.z | ((uint64_t)pi.w << 32), pdim[path], v3(ro.x, ro.y, ro.z), v3(rd.x, rd.y, rd.z), h.prim, h.t, h.b0, h.b1, h.b2,
                               p0, p1, p2, isSphere, v3(bt.x, bt.y, bt.z), bt.w, o);
        }
    } else {
        // the four quarters of the vertex with block-wide barriers between them: the warps of the block then run the same quarter of
        // this long kernel at the same time and share the instruction cache (see the block size of its launches)
        ShadeMid mid;
        bool more = false;
        V3 rayD = v3(0.f, 0.f, 0.f), bta = rayD;
        if (i < n) {
            const b200rt_hit h = hits[i];
            path = qPath[i];
            const float4 ro = rays[2 * (size_t)i], rd = rays[2 * (size_t)i + 1];
            const float4 bt = beta[path];
            V3 p0 = v3(0.f, 0.f, 0.f), p1 = p0, p2 = p0;
            bool isSphere = false;
            if (h.prim != B200RT_MISS) {
                const float4 a = __ldg(primTris + 3 * (size_t)h.prim), b = __ldg(primTris + 3 * (size_t)h.prim + 1),
                             c = __ldg(primTris + 3 * (size_t)h.prim + 2);
                p0 = v3(a.x, a.y, a.z); p1 = v3(b.x, b.y, b.z); p2 = v3(c.x, c.y, c.z);
                isSphere = (__float_as_uint(b.w) & B200RT_PRIM_SPHERE) != 0;
            }
            const uint4 pi = pinfo[path];
            const uint32_t x = pi.x & 0xffffu, y = pi.x >> 16, s = pi.y;
            rayD = v3(rd.x, rd.y, rd.z); bta = v3(bt.x, bt.y, bt.z);
            more = shade_vertex_setup<FULL>(P, depth, x, y, s, (uint64_t)pi.z | ((uint64_t)pi.w << 32), pdim[path], v3(ro.x, ro.y, ro.z), rayD, h.prim, h.t,
                                            h.b0, h.b1, h.b2, p0, p1, p2, isSphere, bta, bt.w, o, mid);
        }
        __syncthreads();
        if (more) shade_vertex_light<FULL>(P, bta, mid, o);
        __syncthreads();
        if (more) shade_vertex_mis<FULL>(P, bta, mid, o);
        __syncthreads();
        if (more) shade_vertex_bounce<FULL>(P, depth, rayD, bta, mid, o);
    }
    if (i < n) {
        if (FULL && (o.Le.x != 0.f || o.Le.y != 0.f || o.Le.z != 0.f)) {
            float4 l = L[path];
            l.x += o.Le.x; l.y += o.Le.y; l.z += o.Le.z;
            L[path] = l;
        }
        if (o.wantNext) {
            beta[path] = make_float4(o.beta.x, o.beta.y, o.beta.z, o.betaW);
            pdim[path] = o.dim;
        }
    }
    uint32_t sj, nj, mj;
    block_append3(o.wantShadow, o.wantNext, mis && o.wantMis, nShadow, nNext, nMis, sj, nj, mj);
    if (o.wantShadow) {
        shRays[2 * (size_t)sj] = make_float4(o.so.x, o.so.y, o.so.z, 1.f - kShadowEps);
        shRays[2 * (size_t)sj + 1] = make_float4(o.sd.x, o.sd.y, o.sd.z, 0.f);
        shPath[sj] = path;
        shContrib[sj] = make_float4(o.contrib.x, o.contrib.y, o.contrib.z, 0.f);
    }
    if (o.wantNext) {
        nextRays[2 * (size_t)nj] = make_float4(o.no.x, o.no.y, o.no.z, __int_as_float(0x7f800000));
        nextRays[2 * (size_t)nj + 1] = make_float4(o.nd.x, o.nd.y, o.nd.z, 0.f);
        nextPath[nj] = path;
    }
    if (mis) {  // launch-uniform
        if (o.wantMis) {
            misRays[2 * (size_t)mj] = make_float4(o.mo.x, o.mo.y, o.mo.z, __int_as_float(0x7f800000));
            misRays[2 * (size_t)mj + 1] = make_float4(o.md.x, o.md.y, o.md.z, 0.f);
            misPath[mj] = path;
            misContrib[mj] = make_float4(o.mcontrib.x, o.mcontrib.y, o.mcontrib.z, __uint_as_float(o.misPrim));
        }
    }
}

// The BSDF-sampled ray of EstimateDirect (integrator.cpp:196-212): it contributes when its closest hit is the emitter
// (`lightIsect.primitive->GetAreaLight() == &light`, Li = lightIsect.Le(-wi); seen from outside the sphere is front-facing),
// or, for an infinite light, when it escapes (`light.Le(ray)`): misContrib.w holds the primitive number the chosen light asks for.
__global__ void k_mis_resolve(const uint32_t *__restrict__ nMis, const b200rt_hit *__restrict__ hits,
                              const uint32_t *__restrict__ misPath, const float4 *__restrict__ misContrib, float4 *L) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= *nMis) return;
    const float4 c = misContrib[j];
    if (hits[j].prim != __float_as_uint(c.w)) return;
    float4 l = L[misPath[j]];   // one MIS ray per path and depth: no two threads share a path
    l.x += c.x; l.y += c.y; l.z += c.z;
    L[misPath[j]] = l;
}

// VisibilityTester::Unoccluded (light.cpp:59-64): add the light sample when the shadow ray is free.
__global__ void k_shadow_resolve(const uint32_t *__restrict__ nShadow, const uint8_t *__restrict__ occluded,
                                 const uint32_t *__restrict__ shPath, const float4 *__restrict__ shContrib, float4 *L) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= *nShadow || occluded[j]) return;
    const float4 c = shContrib[j];
    float4 l = L[shPath[j]];  // one shadow ray per path per bounce: single writer
    l.x += c.x; l.y += c.y; l.z += c.z;
    L[shPath[j]] = l;
}

// FilmTile::AddSample with the box filter (film.h:130-169).  A sample lands in the pixels [p0, p1) with
// p0 = ceil(pFilm - 0.5 - radius), p1 = floor(pFilm - 0.5 + radius) + 1, clipped to the tile's pixel bounds (the tile plus
// a one-pixel halo, Film::GetFilmTile, film.cpp:96-107): its own pixel — and, when pFilm sits exactly on a pixel boundary
// (Halton's first samples have offset 0; pixel + offset can round up to the next integer), the neighbour across that
// boundary too.  Each pixel sums its own samples in sample order into (sum R, sum G, sum B, sum weight); contributions to
// neighbours go, summed per direction, to eight per-direction lists that k_splat applies after the last batch (within one
// direction every target pixel appears at most once, so the result does not depend on the execution order).
__global__ void k_film(const __grid_constant__ WfParams P, uint64_t batchStart, uint32_t nPixels, const float4 *__restrict__ L,
                       float4 *film, uint32_t *splatCount, uint32_t *splatPix, float4 *splatVal, uint32_t splatCap) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nPixels) return;
    uint32_t x, y, s;
    if (!path_pixel(P.tl, batchStart + (uint64_t)i * P.tl.spp, x, y, s)) return;
    const int tile = (int)P.tl.tile;
    const int tx0 = ((int)x / tile) * tile, ty0 = ((int)y / tile) * tile;
    const int bx0 = max(tx0 - 1, 0), by0 = max(ty0 - 1, 0);
    const int bx1 = min(min(tx0 + tile, (int)P.tl.xres) + 1, (int)P.tl.xres), by1 = min(min(ty0 + tile, (int)P.tl.yres) + 1, (int)P.tl.yres);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 nb[8];
    bool anyNb = false;
    PathSampler ps;
    for (uint32_t k = 0; k < P.tl.spp; ++k) {
        const float4 l = L[(size_t)i * P.tl.spp + k];
        path_sampler_init(P, x, y, k, 0, ps);
        const float u0 = get1d(P, ps), u1 = get1d(P, ps);
        // pFilm in window coordinates differs from the full-frame one by integers: the rounding of px + u is the full frame's
        const float fx = ((float)ps.px + u0) - 0.5f, fy = ((float)ps.py + u1) - 0.5f;
        const int p0x = max((int)ceilf(fx - 0.5f) - (int)P.tl.x0, bx0), p1x = min((int)floorf(fx + 0.5f) + 1 - (int)P.tl.x0, bx1);
        const int p0y = max((int)ceilf(fy - 0.5f) - (int)P.tl.y0, by0), p1y = min((int)floorf(fy + 0.5f) + 1 - (int)P.tl.y0, by1);
        if (p0x == (int)x && p1x == (int)x + 1 && p0y == (int)y && p1y == (int)y + 1) {
            acc.x += l.x; acc.y += l.y; acc.z += l.z; acc.w += l.w;
            continue;
        }
        if (!anyNb) { for (int q = 0; q < 8; ++q) nb[q] = make_float4(0.f, 0.f, 0.f, 0.f); anyNb = true; }
        for (int yy = p0y; yy < p1y; ++yy)
            for (int xx = p0x; xx < p1x; ++xx) {
                const int dx = xx - (int)x, dy = yy - (int)y;
                if (dx == 0 && dy == 0) { acc.x += l.x; acc.y += l.y; acc.z += l.z; acc.w += l.w; continue; }
                if (dx < -1 || dx > 1 || dy < -1 || dy > 1) continue;
                const int dir = (dy + 1) * 3 + (dx + 1), q = dir < 4 ? dir : dir - 1;
                nb[q].x += l.x; nb[q].y += l.y; nb[q].z += l.z; nb[q].w += l.w;
            }
    }
    float4 f = film[(size_t)y * P.tl.xres + x];
    f.x += acc.x; f.y += acc.y; f.z += acc.z; f.w += acc.w;
    film[(size_t)y * P.tl.xres + x] = f;
    if (anyNb)
        for (int q = 0; q < 8; ++q) {
            if (nb[q].w == 0.f) continue;
            const int dir = q < 4 ? q : q + 1, dx = dir % 3 - 1, dy = dir / 3 - 1;
            const uint32_t slot = atomicAdd(splatCount + q, 1u);
            if (slot < splatCap) {
                splatPix[(size_t)q * splatCap + slot] = (uint32_t)((int)y + dy) * P.tl.xres + (uint32_t)((int)x + dx);
                splatVal[(size_t)q * splatCap + slot] = nb[q];
            }
        }
}
// Applies one direction's list of neighbour contributions (see k_film): distinct target pixels, plain read-modify-write.
__global__ void k_splat(const uint32_t *__restrict__ splatCount, const uint32_t *__restrict__ splatPix, const float4 *__restrict__ splatVal,
                        uint32_t splatCap, uint32_t q, float4 *film) {
    const uint32_t n = min(splatCount[q], splatCap);
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
        const uint32_t pix = splatPix[(size_t)q * splatCap + j];
        const float4 v = splatVal[(size_t)q * splatCap + j];
        float4 f = film[pix];
        f.x += v.x; f.y += v.y; f.z += v.z; f.w += v.w;
        film[pix] = f;
    }
}

// Per-pixel traversal statistics (b200rt_render_heatmaps): the rays of queue entries [0, *nQ) add their per-ray
// stats to the pixel of their path.  plane = 0 for closest-hit queues, 1 for shadow queues.
__global__ void k_heat(const __grid_constant__ WfParams P, uint64_t batchStart, const uint32_t *__restrict__ nQ,
                       const uint32_t *__restrict__ qPath, const uint4 *__restrict__ perRay, uint32_t *heat, uint32_t plane) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= *nQ) return;
    uint32_t x, y, s;
    path_pixel(P.tl, batchStart + qPath[i], x, y, s);
    const size_t npx = (size_t)P.tl.xres * P.tl.yres, pix = (size_t)y * P.tl.xres + x;
    const uint4 v = perRay[i];
    if (v.x) atomicAdd(heat + (B200RT_HEAT_PRIM + plane) * npx + pix, v.x);
    if (v.y) atomicAdd(heat + (B200RT_HEAT_NODE + plane) * npx + pix, v.y);
    if (v.z) atomicAdd(heat + (B200RT_HEAT_LEAF + plane) * npx + pix, v.z);
}

// The per-batch buffers of one wavefront in flight.
struct QSet {
    float4 *rays[2], *shRays, *shContrib, *beta, *L, *misRays, *misContrib;
    uint32_t *qPath[2], *shPath, *counts, *pdim, *misPath;
    uint4 *pinfo;   // per path: pixel, sample number, global sample index (k_camera)
    b200rt_hit *hits;
    uint8_t *occ;
};

struct Workspace {
    std::mutex mu;   // one render at a time per scene: the queues below are shared state
    // A second set of batch buffers + two side streams: consecutive batches are independent, so batch k+1's camera / shade /
    // film kernels and the head of its traversal run in the gaps batch k's traversal tails leave (B200RT_WF_STREAMS=1: off)
    static constexpr int kMaxSets = 4;
    QSet alt[kMaxSets - 1] = {};
    size_t altCap[kMaxSets - 1] = {0, 0, 0};
    bool altMis[kMaxSets - 1] = {false, false, false};
    cudaStream_t side[kMaxSets] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t evFork = nullptr, evJoin[kMaxSets] = {nullptr, nullptr, nullptr, nullptr};
    size_t cap = 0;
    float4 *misRays = nullptr, *misContrib = nullptr;   // lights with area only (EstimateDirect's BSDF-sampled rays)
    uint32_t *misPath = nullptr;
    float *mats = nullptr;      // kMatFloats per material
    WfTexture *tex = nullptr;   // image textures (ABI v9) and their pyramids
    float *texData = nullptr;
    const void *texHost = nullptr, *texFirst = nullptr;
    uint32_t texN = 0;
    uint8_t *primMat = nullptr;
    size_t mats_n = 0, primMat_n = 0;
    WfLight *lights = nullptr;   // B200RT_MAX_LIGHTS records (ABI v10), allocated on the first render with more than one light
    float *voxTable = nullptr;   // "spatial": cdfHi then pickPdf of every voxel's distribution, + 640 Halton floats + bound + counts
    size_t voxFloats = 0;
    uint4 *perRay = nullptr;   // heat-map mode only
    uint32_t *heat = nullptr;
    size_t heat_px = 0;
    float4 *rays[2] = {nullptr, nullptr}, *shRays = nullptr, *shContrib = nullptr, *beta = nullptr, *L = nullptr;
    uint32_t *qPath[2] = {nullptr, nullptr}, *shPath = nullptr, *counts = nullptr, *pdim = nullptr;
    uint4 *pinfo = nullptr;
    b200rt_hit *hits = nullptr;
    uint8_t *occ = nullptr;
    float4 *film = nullptr;
    size_t film_px = 0;
    // contributions of boundary samples to neighbouring pixels (k_film / k_splat): 8 lists of splatCap entries
    uint32_t *splatCount = nullptr, *splatPix = nullptr;
    float4 *splatVal = nullptr;
    uint32_t splatCap = 0;
    // sample-stream tables on the device
    uint16_t *haltonPerms = nullptr;
    HaltonDim *haltonDims = nullptr;
    uint32_t *sobol32 = nullptr;
    uint32_t *sobolBytes = nullptr;   // per dimension, 4 x 256 words: the XOR of the matrix columns a byte of the index selects
    uint64_t *sobolVdc = nullptr;   // 2 x 52: the VdC row and its inverse for the current resolution
    const void *sobolHost = nullptr;
    uint32_t sobolRow = ~0u;
    // pinned host mirror of the per-batch queue counters (read once, after the last batch)
    uint32_t *hCounts = nullptr;
    size_t hCountsCap = 0;
    // timing events, created once and reused (only recorded when the caller asks for statistics)
    std::vector<cudaEvent_t> events;
    // recorded queues of the last batch (b200rt_render_keep_queues)
    bool keep = false;
    std::vector<float4 *> kept;
    std::vector<uint32_t *> keptPath;
    std::vector<float4 *> keptContrib;
    std::vector<uint32_t> keptCount;
    size_t keptCap = 0;
};

void free_kept(Workspace *w) {
    for (float4 *p : w->kept) cudaFree(p);
    for (uint32_t *p : w->keptPath) cudaFree(p);
    for (float4 *p : w->keptContrib) cudaFree(p);
    w->kept.clear(); w->keptPath.clear(); w->keptContrib.clear(); w->keptCount.clear();
    w->keptCap = 0;
}
void free_queues(Workspace *w) {
    for (int k = 0; k < 2; ++k) { cudaFree(w->rays[k]); cudaFree(w->qPath[k]); w->rays[k] = nullptr; w->qPath[k] = nullptr; }
    cudaFree(w->shRays); cudaFree(w->shContrib); cudaFree(w->beta); cudaFree(w->L); cudaFree(w->shPath); cudaFree(w->pdim); cudaFree(w->pinfo); w->pinfo = nullptr;
    cudaFree(w->counts); cudaFree(w->hits); cudaFree(w->occ); cudaFree(w->perRay);
    cudaFree(w->misRays); cudaFree(w->misContrib); cudaFree(w->misPath);
    w->shRays = w->shContrib = w->beta = w->L = w->misRays = w->misContrib = nullptr;
    w->shPath = w->pdim = w->counts = w->misPath = nullptr;
    w->hits = nullptr; w->occ = nullptr; w->perRay = nullptr;
    w->cap = 0;
    free_kept(w);
}
void free_alt(Workspace *w, int i) {
    QSet &a = w->alt[i];
    for (int k = 0; k < 2; ++k) { cudaFree(a.rays[k]); cudaFree(a.qPath[k]); }
    cudaFree(a.shRays); cudaFree(a.shContrib); cudaFree(a.beta); cudaFree(a.L); cudaFree(a.misRays); cudaFree(a.misContrib);
    cudaFree(a.shPath); cudaFree(a.counts); cudaFree(a.pdim); cudaFree(a.pinfo); cudaFree(a.misPath); cudaFree(a.hits); cudaFree(a.occ);
    a = QSet{};
    w->altCap[i] = 0; w->altMis[i] = false;
}
void free_ws(Workspace *w) {
    free_queues(w);
    for (int i = 0; i < Workspace::kMaxSets - 1; ++i) free_alt(w, i);
    for (int k = 0; k < Workspace::kMaxSets; ++k) { if (w->side[k]) cudaStreamDestroy(w->side[k]); if (w->evJoin[k]) cudaEventDestroy(w->evJoin[k]); }
    if (w->evFork) cudaEventDestroy(w->evFork);
    cudaFree(w->film); cudaFree(w->heat); cudaFree(w->mats); cudaFree(w->primMat); cudaFree(w->tex); cudaFree(w->texData); cudaFree(w->lights); cudaFree(w->voxTable);
    cudaFree(w->splatCount); cudaFree(w->splatPix); cudaFree(w->splatVal);
    cudaFree(w->haltonPerms); cudaFree(w->haltonDims); cudaFree(w->sobol32); cudaFree(w->sobolBytes); cudaFree(w->sobolVdc);
    if (w->hCounts) cudaFreeHost(w->hCounts);
    for (cudaEvent_t e : w->events) cudaEventDestroy(e);
}

}  // namespace

// the workspace hangs off the scene handle
static std::mutex g_ws_mu;
static std::vector<std::pair<const b200rt_scene *, Workspace *>> g_ws;
static Workspace *ws_of(const b200rt_scene *s) {
    std::lock_guard<std::mutex> g(g_ws_mu);
    for (auto &p : g_ws) if (p.first == s) return p.second;
    g_ws.emplace_back(s, new Workspace());
    return g_ws.back().second;
}
void b200rt_wavefront_release(const b200rt_scene *s) {
    std::lock_guard<std::mutex> g(g_ws_mu);
    for (size_t i = 0; i < g_ws.size(); ++i)
        if (g_ws[i].first == s) { free_ws(g_ws[i].second); delete g_ws[i].second; g_ws.erase(g_ws.begin() + i); return; }
}

extern "C" int b200rt_render_tiles(const b200rt_render_desc *d, uint32_t *n_tiles, uint64_t *n_pixels) {
    Tiling t;
    int rc = tiling_of(d, &t);
    if (rc) return rc;
    if (n_tiles) *n_tiles = t.n_tiles_rank;
    if (n_pixels) {
        uint64_t px = 0;
        for (uint32_t k = 0; k < t.n_tiles_rank; ++k) {
            const uint32_t idx = t.rank + k * t.world, tx = idx % t.ntx, ty = idx / t.ntx;
            px += (uint64_t)std::min(t.tile, t.xres - tx * t.tile) * std::min(t.tile, t.yres - ty * t.tile);
        }
        *n_pixels = px;
    }
    return B200RT_OK;
}

extern "C" int b200rt_render_keep_queues(b200rt_scene *s, int keep) {
    if (!s) return set_error(B200RT_EINVAL, "null scene");
    Workspace *w = ws_of(s);
    std::lock_guard<std::mutex> g(w->mu);
    w->keep = keep != 0;
    return B200RT_OK;
}

extern "C" int b200rt_render_queue(b200rt_scene *s, uint32_t q, b200rt_ray *rays, uint64_t *n) {
    if (!s || !n) return set_error(B200RT_EINVAL, "null argument");
    Workspace *w = ws_of(s);
    std::lock_guard<std::mutex> g(w->mu);
    if (q >= w->kept.size()) { *n = 0; return B200RT_OK; }
    const uint64_t have = w->keptCount[q];
    if (!rays) { *n = have; return B200RT_OK; }
    const uint64_t m = std::min<uint64_t>(have, *n);
    WCU(cudaMemcpy(rays, w->kept[q], m * sizeof(b200rt_ray), cudaMemcpyDefault));
    *n = m;
    return B200RT_OK;
}

extern "C" int b200rt_render_queue_paths(b200rt_scene *s, uint32_t q, uint32_t *path, float *contrib, uint64_t *n) {
    if (!s || !n) return set_error(B200RT_EINVAL, "null argument");
    Workspace *w = ws_of(s);
    std::lock_guard<std::mutex> g(w->mu);
    if (q >= w->kept.size()) { *n = 0; return B200RT_OK; }
    const uint64_t m = std::min<uint64_t>(w->keptCount[q], *n);
    if (path) WCU(cudaMemcpy(path, w->keptPath[q], m * 4, cudaMemcpyDeviceToHost));
    if (contrib) {
        if (w->keptContrib[q]) WCU(cudaMemcpy(contrib, w->keptContrib[q], m * 16, cudaMemcpyDeviceToHost));
        else std::memset(contrib, 0, m * 16);
    }
    *n = m;
    return B200RT_OK;
}

// wavefronts in flight (B200RT_WF_STREAMS in the environment, or b200rt_render_set_streams): 1 .. 4, default 2
static std::atomic<int> g_wf_sets{[] { const char *e = getenv("B200RT_WF_STREAMS"); const int v = e ? atoi(e) : 2; return v < 1 ? 1 : (v > 4 ? 4 : v); }()};
extern "C" int b200rt_render_set_streams(int n) {
    if (n < 1 || n > 4) return set_error(B200RT_EINVAL, "b200rt_render_set_streams: 1 .. 4");
    g_wf_sets.store(n);
    return B200RT_OK;
}
static int render_impl(b200rt_scene *s, const b200rt_render_desc *d, float *film, uint32_t *heat, b200rt_render_stats *stats,
                       void *stream);
extern "C" int b200rt_render(b200rt_scene *s, const b200rt_render_desc *d, float *film, b200rt_render_stats *stats,
                             void *stream) {
    return render_impl(s, d, film, nullptr, stats, stream);
}
extern "C" int b200rt_render_heatmaps(b200rt_scene *s, const b200rt_render_desc *d, float *film, uint32_t *heat,
                                      b200rt_render_stats *stats, void *stream) {
    if (!heat) return set_error(B200RT_EINVAL, "null heat-map buffer");
    return render_impl(s, d, film, heat, stats, stream);
}

// Uploads the sample-stream tables `d->sampler` needs and fills P.smp.  maxDims: an upper bound of the dimensions a path can draw.
static int setup_sampler(Workspace *w, const b200rt_render_desc *d, WfParams &P, uint32_t maxDims, cudaStream_t st) {
    std::memset(&P.smp, 0, sizeof(P.smp));
    P.smp.kind = d->sampler; P.smp.seed = d->seed;
    if (d->sampler == B200RT_SAMPLER_HASH) return B200RT_OK;
    if (d->sampler == B200RT_SAMPLER_HALTON) {
        if (maxDims > (uint32_t)kHaltonDims)
            return set_error(B200RT_EUNSUPPORTED, "Sampler \"halton\" has %d dimensions; max_depth %u can draw %u", kHaltonDims, d->max_depth, maxDims);
        const HaltonTables &T = halton_tables();
        if (!w->haltonPerms) {
            WCU(cudaMalloc(&w->haltonPerms, T.perms.size() * sizeof(uint16_t)));
            WCU(cudaMalloc(&w->haltonDims, T.dims.size() * sizeof(HaltonDim)));
            WCU(cudaMemcpyAsync(w->haltonPerms, T.perms.data(), T.perms.size() * sizeof(uint16_t), cudaMemcpyHostToDevice, st));
            WCU(cudaMemcpyAsync(w->haltonDims, T.dims.data(), T.dims.size() * sizeof(HaltonDim), cudaMemcpyHostToDevice, st));
        }
        // sampleBounds = Film::GetSampleBounds() = the crop window for a box filter of radius 0.5 (film.cpp:91-99)
        halton_setup((int)P.tl.xres, (int)P.tl.yres, &P.smp);
        P.smp.perms = w->haltonPerms; P.smp.dimInfo = w->haltonDims; P.smp.nDims = (uint32_t)T.dims.size();
        return B200RT_OK;
    }
    if (d->sampler == B200RT_SAMPLER_SOBOL) {
        if (!d->sobol_matrices32 || !d->sobol_vdc || !d->sobol_vdc_inv)
            return set_error(B200RT_EINVAL, "Sampler \"sobol\": the generator matrices (sobol_matrices32, sobol_vdc, sobol_vdc_inv) are required");
        if (maxDims > (uint32_t)kSobolDims)
            return set_error(B200RT_EUNSUPPORTED, "Sampler \"sobol\" has %d dimensions; max_depth %u can draw %u", kSobolDims, d->max_depth, maxDims);
        if (d->spp & (d->spp - 1))
            return set_error(B200RT_EINVAL, "Sampler \"sobol\": pixelsamples must be a power of two (the reference rounds it up, samplers/sobol.h; do that in the caller)");
        sobol_setup((int)P.tl.xres, (int)P.tl.yres, (int)P.tl.x0, (int)P.tl.y0, &P.smp);
        if (P.smp.log2Res > 0 && P.smp.log2Res > d->sobol_vdc_rows) return set_error(B200RT_EINVAL, "sobol_vdc_rows %u < log2(resolution) %u", d->sobol_vdc_rows, P.smp.log2Res);
        if (!w->sobol32) {
            WCU(cudaMalloc(&w->sobol32, (size_t)kSobolDims * kSobolMatrixSize * 4)); WCU(cudaMalloc(&w->sobolVdc, 2 * kSobolMatrixSize * 8));
            WCU(cudaMalloc(&w->sobolBytes, (size_t)kSobolDims * 1024 * 4));
        }
        if (w->sobolHost != d->sobol_matrices32) {
            WCU(cudaMemcpyAsync(w->sobol32, d->sobol_matrices32, (size_t)kSobolDims * kSobolMatrixSize * 4, cudaMemcpyHostToDevice, st));
            k_sobol_bytes<<<(kSobolDims * 1024 + 255) / 256, 256, 0, st>>>(w->sobol32, w->sobolBytes);
            g_launches.fetch_add(1);
            w->sobolHost = d->sobol_matrices32; w->sobolRow = ~0u;
        }
        if (P.smp.log2Res > 0 && w->sobolRow != P.smp.log2Res - 1) {
            const size_t row = (size_t)(P.smp.log2Res - 1) * kSobolMatrixSize;
            WCU(cudaMemcpyAsync(w->sobolVdc, d->sobol_vdc + row, kSobolMatrixSize * 8, cudaMemcpyHostToDevice, st));
            WCU(cudaMemcpyAsync(w->sobolVdc + kSobolMatrixSize, d->sobol_vdc_inv + row, kSobolMatrixSize * 8, cudaMemcpyHostToDevice, st));
            w->sobolRow = P.smp.log2Res - 1;
        }
        P.smp.sobolBytes = w->sobolBytes;
        P.smp.sobol32 = w->sobol32; P.smp.vdc = w->sobolVdc; P.smp.vdcInv = w->sobolVdc + kSobolMatrixSize;
        return B200RT_OK;
    }
    return set_error(B200RT_EINVAL, "unknown sampler %u", d->sampler);
}

static int render_impl(b200rt_scene *s, const b200rt_render_desc *d, float *film, uint32_t *heat, b200rt_render_stats *stats,
                       void *stream) {
    if (!s || !d || !film) return set_error(B200RT_EINVAL, "null argument");
    WfParams P;
    std::memset(&P, 0, sizeof(P));
    int rc = tiling_of(d, &P.tl);
    if (rc) return rc;
    P.cam = d->camera;
    P.cam_exact = d->camera_exact;
    if (d->camera_exact) { std::memcpy(P.cm.r2c, d->raster_to_camera, 64); std::memcpy(P.cm.c2w, d->camera_to_world, 64); }
    std::memcpy(P.kd, d->kd, 12);
    P.max_depth = d->max_depth;
    if (d->max_depth > 68) return set_error(B200RT_EINVAL, "max_depth %u too large", d->max_depth);
    if (P.tl.xres > 65000u || P.tl.yres > 65000u) return set_error(B200RT_EUNSUPPORTED, "film window %u x %u: at most 65000 pixels a side (a path's pixel is kept in 16 + 16 bits)", P.tl.xres, P.tl.yres);
    std::vector<WfLight> lights;
    rc = lights_of(d, s->n_prims, lights);
    if (rc) return rc;
    P.light = lights[0];
    P.nLights = (uint32_t)lights.size();
    P.lights = nullptr;
    bool mis = false, notPoint = false;   // mis = lights with area: EstimateDirect traces a BSDF-sampled ray too
    for (const WfLight &l : lights) {
        mis = mis || l.kind == B200RT_LIGHT_SPHERE || l.kind == B200RT_LIGHT_INFINITE;
        notPoint = notPoint || l.kind != B200RT_LIGHT_POINT;
    }
    P.anyMis = mis ? 1u : 0u;
    if (d->materials && (!d->prim_material || d->n_materials == 0 || d->n_materials > 256))
        return set_error(B200RT_EINVAL, "materials need prim_material and 1..256 records");
    const uint32_t mstride = d->material_stride ? d->material_stride : 8;
    if (mstride != 8 && mstride != 20) return set_error(B200RT_EINVAL, "material_stride must be 8 or 20");
    cudaStream_t st = (cudaStream_t)stream;
    Workspace *w = ws_of(s);
    std::lock_guard<std::mutex> wsLock(w->mu);
    const uint32_t tp = P.tl.tile * P.tl.tile;
    const uint64_t unit = (uint64_t)tp * P.tl.spp;  // batches hold whole tiles so a pixel's samples never split
    static const uint64_t envBatch = [] { const char *e = getenv("B200RT_WF_BATCH"); return e ? (uint64_t)atoll(e) : 0ull; }();
    // 2^25 paths per batch (about 8 GB of queue buffers per batch in flight: the card has 180): the deep, thin bounces of a batch
    // cost launches whatever their size, so fewer, larger batches pay — scenes/living-room (maxdepth 65) 1 312 ms at 2^22,
    // 1 022 at 2^23, 873 at 2^24, 789 at 2^25; the headline frame 104.6 -> 101.1 ms (profiles/r02_ab_wf_batch.log)
    uint64_t want = d->max_batch ? d->max_batch : (envBatch ? envBatch : (1u << 25));
    uint64_t cap = std::max<uint64_t>(unit, want / unit * unit);
    cap = std::min<uint64_t>(cap, std::max<uint64_t>(unit, P.tl.n_paths_rank));
    // a share that fits one batch is split in two so that the second set of queue buffers has something to overlap with (one
    // rank of an 8-rank headline frame: 14.42 -> 14.18 ms, profiles/r02_ab_rank_share_batches.log); batches smaller than ~2 M
    // paths lose more to the thin deep-bounce launches than the overlap returns
    if (!d->max_batch && !envBatch && g_wf_sets.load() >= 2 && !heat && !w->keep && P.tl.n_paths_rank <= cap && P.tl.n_paths_rank >= (3ull << 20))
        cap = std::max<uint64_t>(unit, ((P.tl.n_paths_rank + 1) / 2 + unit - 1) / unit * unit);
    if (cap > 0x7fffffffull) return set_error(B200RT_EINVAL, "batch too large");
    const uint32_t misBase = 2 * (d->max_depth + 2);          // counts[misBase + depth] = MIS queue size at depth
    const uint32_t nCounts = misBase + d->max_depth + 1;
    if (w->cap < cap && !d->max_batch) {
        // the default batch is sized for a B200's 180 GB; on a card (or a shared one) with less free memory the batch shrinks to
        // what fits rather than failing: about 200 bytes of queue buffers per path (+ 52 with MIS rays) and batch in flight
        size_t freeB = 0, totalB = 0;
        if (cudaMemGetInfo(&freeB, &totalB) == cudaSuccess) {
            const size_t perPath = 2 * 32 + 2 * 4 + 32 + 16 + 4 + 16 + 16 + 4 + 16 + sizeof(b200rt_hit) + 1 + (mis ? 52 : 0);
            const size_t sets = (heat || w->keep) ? 1 : (size_t)g_wf_sets.load();
            while (cap > unit && (double)cap * perPath * sets > 0.8 * (double)freeB) cap = std::max<uint64_t>(unit, cap / 2 / unit * unit);
            if (cap < w->cap) cap = w->cap / unit * unit;   // never below what an earlier frame already holds (its buffers are part of what is not free)
        } else cudaGetLastError();
    }
    if (w->cap < cap) {
        free_queues(w);
        for (int k = 0; k < 2; ++k) { WCU(cudaMalloc(&w->rays[k], cap * 32)); WCU(cudaMalloc(&w->qPath[k], cap * 4)); }
        WCU(cudaMalloc(&w->shRays, cap * 32)); WCU(cudaMalloc(&w->shContrib, cap * 16)); WCU(cudaMalloc(&w->shPath, cap * 4));
        WCU(cudaMalloc(&w->beta, cap * 16)); WCU(cudaMalloc(&w->L, cap * 16)); WCU(cudaMalloc(&w->pdim, cap * 4)); WCU(cudaMalloc(&w->pinfo, cap * 16));
        WCU(cudaMalloc(&w->hits, cap * sizeof(b200rt_hit))); WCU(cudaMalloc(&w->occ, cap));
        WCU(cudaMalloc(&w->counts, 4 * 3 * 72));
        w->cap = cap;
    }
    if (mis && !w->misRays) {
        WCU(cudaMalloc(&w->misRays, w->cap * 32)); WCU(cudaMalloc(&w->misContrib, w->cap * 16)); WCU(cudaMalloc(&w->misPath, w->cap * 4));
    }
    if (d->materials) {
        std::vector<float> dev((size_t)d->n_materials * kMatFloats);
        bool hasSpec = false;
        for (uint32_t k = 0; k < d->n_materials; ++k)
            if (!material_lobes(d->materials + (size_t)mstride * k, mstride, k, dev.data() + (size_t)k * kMatFloats, &hasSpec, d->textures ? d->n_textures : 0u)) return B200RT_EINVAL;
        P.has_specular = hasSpec ? 1u : 0u;
        if (w->mats_n < d->n_materials) { cudaFree(w->mats); w->mats = nullptr; WCU(cudaMalloc(&w->mats, (size_t)d->n_materials * kMatFloats * 4)); w->mats_n = d->n_materials; }
        if (w->primMat_n < s->n_prims) { cudaFree(w->primMat); w->primMat = nullptr; WCU(cudaMalloc(&w->primMat, s->n_prims)); w->primMat_n = s->n_prims; }
        for (uint32_t k = 0; k < s->n_prims; ++k)
            if (d->prim_material[k] >= d->n_materials) return set_error(B200RT_EINVAL, "prim_material[%u] out of range", k);
        WCU(cudaMemcpy(w->mats, dev.data(), dev.size() * 4, cudaMemcpyHostToDevice));
        WCU(cudaMemcpy(w->primMat, d->prim_material, s->n_prims, cudaMemcpyHostToDevice));
        P.mats = w->mats; P.primMat = w->primMat;
    }
    if (P.nLights > 1) {
        if (!w->lights) WCU(cudaMalloc(&w->lights, B200RT_MAX_LIGHTS * sizeof(WfLight)));
        WCU(cudaMemcpy(w->lights, lights.data(), lights.size() * sizeof(WfLight), cudaMemcpyHostToDevice));
        P.lights = w->lights;
        if (d->light_strategy == B200RT_LIGHTS_SPATIAL) {
            // every voxel's distribution, before the frame (the reference computes a voxel's the first time a path lands in it;
            // the values are the same): nv threads x 128 points x nLights light samples
            std::memcpy(P.wb, s->bounds, sizeof(P.wb));
            spatial_voxel_counts(P.wb, 64, P.nVox);
            const size_t nv = (size_t)P.nVox[0] * P.nVox[1] * P.nVox[2], per = nv * P.nLights;
            if (per > (48u << 20)) return set_error(B200RT_EUNSUPPORTED, "spatial light distribution: %zu voxels x %u lights is too large a table", nv, P.nLights);
            const size_t need = 2 * per + 640 + 16;
            if (w->voxFloats < need) {
                cudaFree(w->voxTable); w->voxTable = nullptr; w->voxFloats = 0;
                WCU(cudaMalloc(&w->voxTable, need * sizeof(float)));
                w->voxFloats = need;
            }
            float host[640 + 16];
            spatial_halton_table(host);
            std::memcpy(host + 640, P.wb, 24);
            std::memcpy(host + 640 + 6, P.nVox, 12);
            float *aux = w->voxTable + 2 * per;
            WCU(cudaMemcpy(aux, host, sizeof(host), cudaMemcpyHostToDevice));
            k_light_voxels<<<(unsigned)((nv + 63) / 64), 64, 0, st>>>(w->lights, P.nLights, aux + 640, (const uint32_t *)(aux + 640 + 6), (uint32_t)nv,
                                                                  aux, w->voxTable, w->voxTable + per);
            g_launches.fetch_add(1);
            P.voxCdf = w->voxTable; P.voxPdf = w->voxTable + per;
        } else if (d->light_strategy != B200RT_LIGHTS_WEIGHTS) return set_error(B200RT_EINVAL, "unknown light_strategy %u", d->light_strategy);
    }
    P.tex = nullptr; P.nTex = 0;
    if (d->materials && d->textures && d->n_textures) {
        // image textures: the pyramids are uploaded when the host array (pointer, count, first texel pointer) changes
        if (!P.cam_exact) return set_error(B200RT_EINVAL, "image textures need the exact camera (camera_exact): the ray differentials come from its matrices");
        std::vector<WfTexture> T(d->n_textures);
        size_t total = 0;
        for (uint32_t k = 0; k < d->n_textures; ++k) {
            const b200rt_texture &t = d->textures[k];
            if (!t.texels || !t.width || !t.height || (t.width & (t.width - 1)) || (t.height & (t.height - 1)) || t.n_levels == 0 || t.n_levels > 16)
                return set_error(B200RT_EINVAL, "texture %u: a pyramid with power-of-two level 0 and 1..16 levels is required (b200rt_mipmap_build)", k);
            if (t.filter > B200RT_TEXFILTER_EWA) return set_error(B200RT_EINVAL, "texture %u: unknown filter %u", k, t.filter);
            T[k].nLevels = t.n_levels; T[k].w0 = t.width; T[k].h0 = t.height;
            T[k].su = t.su; T[k].sv = t.sv; T[k].du = t.du; T[k].dv = t.dv;
            T[k].ewa = t.filter == B200RT_TEXFILTER_EWA ? 1u : 0u;
            T[k].maxAniso = t.max_anisotropy > 0.f ? t.max_anisotropy : 8.f;
            size_t off = 0;
            for (uint32_t l = 0; l < 16; ++l) {
                T[k].levelOffset[l] = (uint32_t)off;
                if (l < t.n_levels) off += (size_t)std::max(1u, t.width >> l) * std::max(1u, t.height >> l);
            }
            total += off;
        }
        if (w->texHost != (const void *)d->textures || w->texN != d->n_textures || w->texFirst != (const void *)d->textures[0].texels) {
            cudaFree(w->texData); cudaFree(w->tex); w->texData = nullptr; w->tex = nullptr;
            // the texels of every pyramid, then MIPMap::weightLut (128 floats, shared by the EWA lookups of every texture)
            WCU(cudaMalloc(&w->texData, (std::max<size_t>(total, 1) * 3 + 128) * sizeof(float)));
            size_t base = 0;
            for (uint32_t k = 0; k < d->n_textures; ++k) {
                const b200rt_texture &t = d->textures[k];
                size_t n = 0;
                for (uint32_t l = 0; l < t.n_levels; ++l) n += (size_t)std::max(1u, t.width >> l) * std::max(1u, t.height >> l);
                WCU(cudaMemcpy(w->texData + base * 3, t.texels, n * 3 * sizeof(float), cudaMemcpyHostToDevice));
                base += n;
            }
            float lut[128];
            b200rt_ewa_weights(lut);
            WCU(cudaMemcpy(w->texData + std::max<size_t>(total, 1) * 3, lut, sizeof(lut), cudaMemcpyHostToDevice));
            WCU(cudaMalloc(&w->tex, T.size() * sizeof(WfTexture)));
            w->texHost = d->textures; w->texN = d->n_textures; w->texFirst = d->textures[0].texels;
        }
        {   // the records (filter, mapping) are refreshed every render; the texels above only when the host array changed
            size_t base = 0;
            for (uint32_t k = 0; k < d->n_textures; ++k) {
                T[k].texels = w->texData + base * 3;
                T[k].lut = w->texData + std::max<size_t>(total, 1) * 3;
                for (uint32_t l = 0; l < T[k].nLevels; ++l) base += (size_t)std::max(1u, T[k].w0 >> l) * std::max(1u, T[k].h0 >> l);
            }
            WCU(cudaMemcpy(w->tex, T.data(), T.size() * sizeof(WfTexture), cudaMemcpyHostToDevice));
        }
        P.tex = w->tex; P.nTex = d->n_textures;
    }
    P.primUV = s->d_prim_uv; P.primN = s->d_prim_n; P.primFlags = s->d_prim_flags;
    // camera 5 + per bounce: light choice 1 + uLight 2 + uScattering 2 + BSDF 2 + Russian roulette 1
    rc = setup_sampler(w, d, P, 5 + 8 * (d->max_depth + 1), st);
    if (rc) return rc;
    // the lean instantiation: one point light, one Lambertian Kd, no vertex normals / UVs / orientation flags
    const bool full = mis || notPoint || P.nLights > 1 || P.mats || P.primN || P.primUV || P.primFlags;
    const size_t npx = (size_t)P.tl.xres * P.tl.yres;
    cudaPointerAttributes pa;
    bool film_dev = cudaPointerGetAttributes(&pa, film) == cudaSuccess &&
                    (pa.type == cudaMemoryTypeDevice || pa.type == cudaMemoryTypeManaged);
    cudaGetLastError();
    float4 *d_film = (float4 *)film;
    if (!film_dev) {
        if (w->film_px < npx) { cudaFree(w->film); w->film = nullptr; WCU(cudaMalloc(&w->film, npx * 16)); w->film_px = npx; }
        WCU(cudaMemcpyAsync(w->film, film, npx * 16, cudaMemcpyHostToDevice, st));
        d_film = w->film;
    }
    uint32_t *d_heat = heat;
    bool heat_dev = true;
    if (heat) {
        if (!w->perRay) WCU(cudaMalloc(&w->perRay, w->cap * sizeof(uint4)));
        heat_dev = cudaPointerGetAttributes(&pa, heat) == cudaSuccess &&
                   (pa.type == cudaMemoryTypeDevice || pa.type == cudaMemoryTypeManaged);
        cudaGetLastError();
        if (!heat_dev) {
            if (w->heat_px < npx) { cudaFree(w->heat); w->heat = nullptr; WCU(cudaMalloc(&w->heat, npx * 4 * B200RT_HEAT_PLANES)); w->heat_px = npx; }
            WCU(cudaMemcpyAsync(w->heat, heat, npx * 4 * B200RT_HEAT_PLANES, cudaMemcpyHostToDevice, st));
            d_heat = w->heat;
        }
    }
    // kept queues: closest q = 2*depth, shadow q = 2*depth+1, MIS q = 2*(max_depth+1) + depth
    const size_t nq = 2 * (size_t)(d->max_depth + 1) + (mis ? d->max_depth + 1 : 0);
    if (w->keep) {
        if (w->kept.size() != nq || w->keptCap < cap) {
            free_kept(w);
            w->kept.assign(nq, nullptr); w->keptPath.assign(nq, nullptr); w->keptContrib.assign(nq, nullptr);
            for (size_t q = 0; q < nq; ++q) {
                WCU(cudaMalloc(&w->kept[q], cap * 32)); WCU(cudaMalloc(&w->keptPath[q], cap * 4));
                if ((q & 1) || q >= 2 * (size_t)(d->max_depth + 1)) WCU(cudaMalloc(&w->keptContrib[q], cap * 16));
            }
            w->keptCap = cap;
        }
        w->keptCount.assign(nq, 0);
    }
    {
        const uint32_t wantCap = (uint32_t)std::min<size_t>(npx, npx / 4 + 4096);
        if (w->splatCap < wantCap) {
            cudaFree(w->splatPix); cudaFree(w->splatVal); w->splatPix = nullptr; w->splatVal = nullptr;
            if (!w->splatCount) WCU(cudaMalloc(&w->splatCount, 16 * 4));
            WCU(cudaMalloc(&w->splatPix, (size_t)8 * wantCap * 4));
            WCU(cudaMalloc(&w->splatVal, (size_t)8 * wantCap * 16));
            w->splatCap = wantCap;
        }
        WCU(cudaMemsetAsync(w->splatCount, 0, 16 * 4, st));
    }
    const uint64_t nBatches = (P.tl.n_paths_rank + cap - 1) / cap;
    if (w->hCountsCap < nBatches * nCounts + 8) {
        if (w->hCounts) cudaFreeHost(w->hCounts);
        w->hCounts = nullptr; w->hCountsCap = 0;
        WCU(cudaMallocHost(&w->hCounts, (nBatches * nCounts + 8) * 4));
        w->hCountsCap = nBatches * nCounts + 8;
    }
    // timing events come from the workspace's pool and are only recorded when statistics are asked for
    size_t evUsed = 0;
    bool evFail = false;
    auto ev = [&](cudaStream_t on) -> cudaEvent_t {
        if (evUsed == w->events.size()) {
            cudaEvent_t e = nullptr;
            if (cudaEventCreate(&e) != cudaSuccess) { evFail = true; return nullptr; }
            w->events.push_back(e);
        }
        cudaEvent_t e = w->events[evUsed++];
        cudaEventRecord(e, on);
        return e;
    };
    struct Span { cudaEvent_t a, b; bool any; };
    std::vector<Span> spans;
    const bool timed = stats != nullptr;
    const uint64_t launches0 = g_launches.load();
    uint64_t nClosest = 0, nShadow = 0, nCamera = 0, nMis = 0;
    // ---- one or two wavefronts in flight ----
    const int envSets = g_wf_sets.load();
    const int nSets = (heat || w->keep) ? 1 : (int)std::min<uint64_t>((uint64_t)envSets, nBatches);
    const bool two = nSets >= 2;
    if (two) {
        if (!w->evFork) WCU(cudaEventCreateWithFlags(&w->evFork, cudaEventDisableTiming));
        for (int k = 0; k < nSets; ++k)
            if (!w->side[k]) { WCU(cudaStreamCreateWithFlags(&w->side[k], cudaStreamNonBlocking)); WCU(cudaEventCreateWithFlags(&w->evJoin[k], cudaEventDisableTiming)); }
        for (int i = 0; i < nSets - 1; ++i) {
            if (w->altCap[i] >= w->cap && (!mis || w->altMis[i])) continue;
            free_alt(w, i);
            QSet &q = w->alt[i];
            const size_t c = w->cap;
            for (int k = 0; k < 2; ++k) { WCU(cudaMalloc(&q.rays[k], c * 32)); WCU(cudaMalloc(&q.qPath[k], c * 4)); }
            WCU(cudaMalloc(&q.shRays, c * 32)); WCU(cudaMalloc(&q.shContrib, c * 16)); WCU(cudaMalloc(&q.shPath, c * 4));
            WCU(cudaMalloc(&q.beta, c * 16)); WCU(cudaMalloc(&q.L, c * 16)); WCU(cudaMalloc(&q.pdim, c * 4)); WCU(cudaMalloc(&q.pinfo, c * 16));
            WCU(cudaMalloc(&q.hits, c * sizeof(b200rt_hit))); WCU(cudaMalloc(&q.occ, c));
            WCU(cudaMalloc(&q.counts, 4 * 3 * 72));
            if (mis) { WCU(cudaMalloc(&q.misRays, c * 32)); WCU(cudaMalloc(&q.misContrib, c * 16)); WCU(cudaMalloc(&q.misPath, c * 4)); w->altMis[i] = true; }
            w->altCap[i] = c;
        }
    }
    QSet prim;
    prim.rays[0] = w->rays[0]; prim.rays[1] = w->rays[1]; prim.qPath[0] = w->qPath[0]; prim.qPath[1] = w->qPath[1];
    prim.shRays = w->shRays; prim.shContrib = w->shContrib; prim.shPath = w->shPath; prim.beta = w->beta; prim.L = w->L; prim.pdim = w->pdim; prim.pinfo = w->pinfo;
    prim.hits = w->hits; prim.occ = w->occ; prim.counts = w->counts;
    prim.misRays = w->misRays; prim.misContrib = w->misContrib; prim.misPath = w->misPath;
    cudaEvent_t t0 = timed ? ev(st) : nullptr;
    if (two) {   // the side streams start after everything queued on the caller's stream so far (uploads, film copy, memsets)
        WCU(cudaEventRecord(w->evFork, st));
        for (int k = 0; k < nSets; ++k) WCU(cudaStreamWaitEvent(w->side[k], w->evFork, 0));
    }
    const unsigned TB = B200RT_WF_TB;
    // k_shade<true>'s block size (B200RT_WF_TB_FULL: 64 | 128 | 256 | 512).  The kernel is 23 k instructions long and its warps wait
    // for instructions more than for anything else (ncu: 15 no_instruction stalls per issue); the warps of ONE block start together
    // and run the same stretch of code at about the same time, warps of different blocks do not — so the larger the block, the
    // better the instruction cache is shared: living-room 1280x720 @ 64 spp 938 ms at 64 threads, 871 at 128, 699 at 256,
    // 557 at 512 (one block of 16 warps per SM at 114 registers; profiles/r02_ab_shade_tb.log)
    static const unsigned envTbFull = [] { const char *e = getenv("B200RT_WF_TB_FULL"); const int v = e ? atoi(e) : 512; return (unsigned)(v == 64 || v == 128 || v == 256 || v == 512 ? v : 512); }();
    const unsigned tbFull = envTbFull;
    static const unsigned tbLean = [] { const char *e = getenv("B200RT_WF_TB_LEAN"); const int v = e ? atoi(e) : 64; return (unsigned)(v == 32 || v == 64 || v == 128 || v == 256 || v == 512 ? v : 64); }();
    auto shadeFull = k_shade<true>;
    uint64_t batch = 0;
    for (uint64_t start = 0; start < P.tl.n_paths_rank; start += cap, ++batch) {
        const uint32_t nSlots = (uint32_t)std::min<uint64_t>(cap, P.tl.n_paths_rank - start);
        const int set = two ? (int)(batch % (uint64_t)nSets) : 0;
        const QSet &Q = set ? w->alt[set - 1] : prim;
        cudaStream_t bs = two ? w->side[set] : st;
        const unsigned gridSlots = (nSlots + TB - 1) / TB;
        WCU(cudaMemsetAsync(Q.counts, 0, 4 * nCounts, bs));
        // counts[2*depth] = closest queue size at depth, counts[2*depth+1] = shadow queue size at depth
        k_camera<<<gridSlots, TB, 0, bs>>>(P, start, nSlots, Q.rays[0], Q.qPath[0], Q.counts, Q.beta, Q.L, Q.pdim, Q.pinfo);
        g_launches.fetch_add(1);
        int cur = 0;
        for (uint32_t depth = 0; depth <= d->max_depth; ++depth) {
            if (w->keep) {
                WCU(cudaMemcpyAsync(w->kept[2 * depth], Q.rays[cur], (size_t)nSlots * 32, cudaMemcpyDeviceToDevice, bs));
                WCU(cudaMemcpyAsync(w->keptPath[2 * depth], Q.qPath[cur], (size_t)nSlots * 4, cudaMemcpyDeviceToDevice, bs));
            }
            cudaEvent_t a = timed ? ev(bs) : nullptr;
            rc = launch_trace(s, Q.rays[cur], nSlots, Q.counts + 2 * depth, Q.hits, nullptr, false,
                              heat ? s->d_counters : nullptr, bs, heat ? w->perRay : nullptr, depth == 0);
            if (rc) return rc;
            if (timed) spans.push_back({a, ev(bs), false});
            if (heat) {
                k_heat<<<gridSlots, TB, 0, bs>>>(P, start, Q.counts + 2 * depth, Q.qPath[cur], w->perRay, d_heat, 0u);
                g_launches.fetch_add(1);
            }
            // path.cpp:110: the last intersection is found, then the path ends — after the emitted radiance a specular
            // bounce may still see there (path.cpp:97-107): one more shade launch, which only adds Le
            if (depth == d->max_depth && !(P.has_specular && mis)) break;
            (full ? shadeFull : k_shade<false>)<<<full ? (nSlots + tbFull - 1) / tbFull : (nSlots + tbLean - 1) / tbLean, full ? tbFull : tbLean, 0, bs>>>(P, depth, start, Q.counts + 2 * depth, Q.rays[cur], Q.qPath[cur], Q.hits,
                                              s->d_prim_tris, Q.beta, Q.pdim, Q.pinfo, Q.rays[cur ^ 1], Q.qPath[cur ^ 1],
                                              Q.counts + 2 * (depth + 1), Q.shRays, Q.shPath, Q.shContrib,
                                              Q.counts + 2 * depth + 1, Q.L, Q.misRays, Q.misPath, Q.misContrib,
                                              Q.counts + misBase + depth);
            g_launches.fetch_add(1);
            if (depth == d->max_depth) break;
            if (w->keep) {
                WCU(cudaMemcpyAsync(w->kept[2 * depth + 1], Q.shRays, (size_t)nSlots * 32, cudaMemcpyDeviceToDevice, bs));
                WCU(cudaMemcpyAsync(w->keptPath[2 * depth + 1], Q.shPath, (size_t)nSlots * 4, cudaMemcpyDeviceToDevice, bs));
                WCU(cudaMemcpyAsync(w->keptContrib[2 * depth + 1], Q.shContrib, (size_t)nSlots * 16, cudaMemcpyDeviceToDevice, bs));
                if (mis) {
                    const size_t q = 2 * (size_t)(d->max_depth + 1) + depth;
                    WCU(cudaMemcpyAsync(w->kept[q], Q.misRays, (size_t)nSlots * 32, cudaMemcpyDeviceToDevice, bs));
                    WCU(cudaMemcpyAsync(w->keptPath[q], Q.misPath, (size_t)nSlots * 4, cudaMemcpyDeviceToDevice, bs));
                    WCU(cudaMemcpyAsync(w->keptContrib[q], Q.misContrib, (size_t)nSlots * 16, cudaMemcpyDeviceToDevice, bs));
                }
            }
            a = timed ? ev(bs) : nullptr;
            rc = launch_trace(s, Q.shRays, nSlots, Q.counts + 2 * depth + 1, nullptr, Q.occ, true,
                              heat ? s->d_counters : nullptr, bs, heat ? w->perRay : nullptr, depth == 0);
            if (rc) return rc;
            if (timed) spans.push_back({a, ev(bs), true});
            if (heat) {
                k_heat<<<gridSlots, TB, 0, bs>>>(P, start, Q.counts + 2 * depth + 1, Q.shPath, w->perRay, d_heat, 1u);
                g_launches.fetch_add(1);
            }
            k_shadow_resolve<<<gridSlots, TB, 0, bs>>>(Q.counts + 2 * depth + 1, Q.occ, Q.shPath, Q.shContrib, Q.L);
            g_launches.fetch_add(1);
            if (mis) {  // EstimateDirect's BSDF-sampled ray: one more closest-hit queue per bounce (Q.hits is free again)
                a = timed ? ev(bs) : nullptr;
                rc = launch_trace(s, Q.misRays, nSlots, Q.counts + misBase + depth, Q.hits, nullptr, false,
                                  heat ? s->d_counters : nullptr, bs, heat ? w->perRay : nullptr);
                if (rc) return rc;
                if (timed) spans.push_back({a, ev(bs), false});
                if (heat) {
                    k_heat<<<gridSlots, TB, 0, bs>>>(P, start, Q.counts + misBase + depth, Q.misPath, w->perRay, d_heat, 0u);
                    g_launches.fetch_add(1);
                }
                k_mis_resolve<<<gridSlots, TB, 0, bs>>>(Q.counts + misBase + depth, Q.hits, Q.misPath, Q.misContrib, Q.L);
                g_launches.fetch_add(1);
            }
            cur ^= 1;
            // deep paths (scenes/living-room: maxdepth 65): Russian roulette empties the queues long before maxdepth, so
            // look at the next queue's size every 4th depth and stop launching once no path is left
            if (depth >= 7 && (depth & 3) == 3 && depth + 1 < d->max_depth) {
                uint32_t *left = w->hCounts + batch * nCounts + 2 * (depth + 1);
                WCU(cudaMemcpyAsync(left, Q.counts + 2 * (depth + 1), 4, cudaMemcpyDeviceToHost, bs));
                WCU(cudaStreamSynchronize(bs));
                if (*left == 0) break;
            }
        }
        const uint32_t nPix = nSlots / P.tl.spp;
        k_film<<<(nPix + TB - 1) / TB, TB, 0, bs>>>(P, start, nPix, Q.L, d_film, w->splatCount, w->splatPix, w->splatVal, w->splatCap);
        g_launches.fetch_add(1);
        // the batch's queue sizes go to pinned host memory without stopping the stream; they are read after the last batch
        WCU(cudaMemcpyAsync(w->hCounts + batch * nCounts, Q.counts, 4 * nCounts, cudaMemcpyDeviceToHost, bs));
    }
    if (two)
        for (int k = 0; k < nSets; ++k) { WCU(cudaEventRecord(w->evJoin[k], w->side[k])); WCU(cudaStreamWaitEvent(st, w->evJoin[k], 0)); }
    // boundary samples' contributions to neighbouring pixels, one direction after the other (grid-stride over the
    // device-side list sizes: no host round trip; an overflow of the lists is reported after the frame)
    uint32_t *hSplat = w->hCounts + nBatches * nCounts;
    for (uint32_t q = 0; q < 8; ++q) {
        k_splat<<<32, TB, 0, st>>>(w->splatCount, w->splatPix, w->splatVal, w->splatCap, q, d_film);
        g_launches.fetch_add(1);
    }
    WCU(cudaMemcpyAsync(hSplat, w->splatCount, 8 * 4, cudaMemcpyDeviceToHost, st));
    if (!film_dev) WCU(cudaMemcpyAsync(film, d_film, npx * 16, cudaMemcpyDeviceToHost, st));
    if (heat && !heat_dev) WCU(cudaMemcpyAsync(heat, d_heat, npx * 4 * B200RT_HEAT_PLANES, cudaMemcpyDeviceToHost, st));
    cudaEvent_t t1 = timed ? ev(st) : nullptr;
    WCU(cudaStreamSynchronize(st));
    WCU(cudaGetLastError());
    if (evFail) return set_error(B200RT_ECUDA, "cudaEventCreate failed");
    for (uint32_t q = 0; q < 8; ++q)
        if (hSplat[q] > w->splatCap) return set_error(B200RT_EUNSUPPORTED, "more than %u samples on pixel boundaries in one direction", w->splatCap);
    for (uint64_t b = 0; b < nBatches; ++b) {
        const uint32_t *hc = w->hCounts + b * nCounts;
        nCamera += hc[0];
        for (uint32_t depth = 0; depth <= d->max_depth; ++depth) {
            nClosest += hc[2 * depth];
            if (depth < d->max_depth) nShadow += hc[2 * depth + 1];
            if (mis && depth < d->max_depth) { nClosest += hc[misBase + depth]; nMis += hc[misBase + depth]; }
            if (w->keep && b + 1 == nBatches) {
                w->keptCount[2 * depth] = hc[2 * depth];
                if (depth < d->max_depth) {
                    w->keptCount[2 * depth + 1] = hc[2 * depth + 1];
                    if (mis) w->keptCount[2 * (size_t)(d->max_depth + 1) + depth] = hc[misBase + depth];
                }
            }
        }
    }
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        stats->camera_rays = nCamera; stats->closest_rays = nClosest; stats->shadow_rays = nShadow; stats->mis_rays = nMis;
        stats->launches = g_launches.load() - launches0;
        stats->n_tiles = P.tl.n_tiles_rank;
        for (const Span &sp : spans) (sp.any ? stats->launches_any : stats->launches_closest) += 1;
        for (const Span &sp : spans) {
            float ms = 0;
            cudaEventElapsedTime(&ms, sp.a, sp.b);
            (sp.any ? stats->ms_any : stats->ms_closest) += ms;
        }
        cudaEventElapsedTime(&stats->ms_total, t0, t1);
    }
    return B200RT_OK;
}
