// Wavefront path-tracing driver (include/b200rt_wavefront.h).
//
// SamplerIntegrator::Render's tile/pixel/sample loops (reference src/core/integrator.cpp:230-360)
// and PathIntegrator::Li's bounce loop (src/integrators/path.cpp:64-204) restructured into
// device-resident ray queues.  Per wavefront step:
//   closest-hit queue --trace--> hits --shade--> { shadow queue --any-hit--> resolve , next closest queue }
// Queue sizes live on the device (no host round trip inside a batch); queues are compacted with
// warp ballots + one atomic per warp.
#include <cuda_runtime.h>
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <vector>
#include "../../include/b200rt_wavefront.h"
#include "scene_internal.h"
#include "wf_bsdf.cuh"

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

constexpr float kPi = 3.14159265358979323846f;
constexpr float kInvPi = 0.31830988618379067154f;
constexpr float kShadowEpsilon = 0.0001f;  // pbrt.h:204
constexpr float kGamma7 = rt_gamma(7);

struct Tiling {
    uint32_t xres, yres, tile, ntx, nty, rank, world, spp;   // xres, yres = extent of the film's crop window
    uint32_t x0, y0;            // its first pixel inside the camera's full frame (Film::croppedPixelBounds.pMin)
    uint32_t n_tiles_rank;      // tiles owned by this rank
    uint64_t n_paths_rank;      // n_tiles_rank * tile^2 * spp (incl. slots outside the image on edge tiles)
};

struct WfParams {
    b200rt_camera cam;
    Tiling tl;
    float light_pos[3], light_I[3], kd[3];
    uint32_t max_depth, seed;
    uint32_t light_kind, light_prim;
    float light_radius;
    float l2w[9];             // B200RT_LIGHT_INFINITE: rotation of LightToWorld, row-major
    const float4 *mats;       // 2 x float4 per material: (kind, Kd.rgb) (Ks.rgb, roughness); nullptr = one matte kd
    const uint8_t *primMat;   // material index per primitive
    uint32_t has_specular;    // some material is "mirror" / "glass": beta.w then carries the path's specularBounce flag and
                              // etaScale (path.cpp:76-83,153-160): 0 = (false, 1); else |w| = etaScale, w < 0 = specular bounce
};
constexpr float kGamma5 = rt_gamma(5);

// ---- counter-based sampler: value = hash(seed, pixel, sample, dimension) ------------------------
__device__ __forceinline__ uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}
__device__ __forceinline__ float rnd(uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t dim) {
    uint32_t h = mix32(seed ^ 0x9e3779b9U);
    h = mix32(h ^ pixel);
    h = mix32(h + sample * 0x85ebca6bU);
    h = mix32(h ^ (dim * 0xc2b2ae35U + 0x27d4eb2fU));
    return (h >> 8) * 5.9604644775390625e-08f;  // [0,1)
}

// path slot -> (pixel x, y, sample); false when the slot lies outside the image (edge tiles)
__device__ __forceinline__ bool path_pixel(const Tiling &t, uint64_t gpath, uint32_t &x, uint32_t &y, uint32_t &s) {
    const uint32_t tp = t.tile * t.tile;
    const uint64_t pixOrd = gpath / t.spp;
    s = (uint32_t)(gpath % t.spp);
    const uint32_t tileOrd = (uint32_t)(pixOrd / tp), inTile = (uint32_t)(pixOrd % tp);
    const uint32_t tileIdx = t.rank + tileOrd * t.world;  // tiles t with t % world == rank
    const uint32_t tx = tileIdx % t.ntx, ty = tileIdx / t.ntx;
    x = tx * t.tile + inTile % t.tile;
    y = ty * t.tile + inTile / t.tile;
    return ty < t.nty && x < t.xres && y < t.yres;
}

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

// OffsetRayOrigin, src/core/geometry.h:1782-1802
__device__ __forceinline__ float3 offset_ray_origin(float3 p, float3 pError, float3 n, float3 w) {
    const float d = fabsf(n.x) * pError.x + fabsf(n.y) * pError.y + fabsf(n.z) * pError.z;
    float3 off = make_float3(d * n.x, d * n.y, d * n.z);
    if (w.x * n.x + w.y * n.y + w.z * n.z < 0) { off.x = -off.x; off.y = -off.y; off.z = -off.z; }
    float3 po = make_float3(p.x + off.x, p.y + off.y, p.z + off.z);
    if (off.x > 0) po.x = next_up(po.x); else if (off.x < 0) po.x = next_down(po.x);
    if (off.y > 0) po.y = next_up(po.y); else if (off.y < 0) po.y = next_down(po.y);
    if (off.z > 0) po.z = next_up(po.z); else if (off.z < 0) po.z = next_down(po.z);
    return po;
}

// ---- stage kernels ------------------------------------------------------------------------------
// Camera rays for path slots [batchStart, batchStart + nSlots) (PerspectiveCamera::GenerateRay,
// src/cameras/perspective.cpp:97-143, pinhole; pixel sample = box filter footprint).
__global__ void k_camera(const __grid_constant__ WfParams P, uint64_t batchStart, uint32_t nSlots, float4 *rays,
                         uint32_t *qPath, uint32_t *count, float4 *beta, float4 *L) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    bool active = false;
    uint32_t x = 0, y = 0, s = 0;
    if (i < nSlots) {
        active = path_pixel(P.tl, batchStart + i, x, y, s);
        beta[i] = make_float4(1.f, 1.f, 1.f, 0.f);
        L[i] = make_float4(0.f, 0.f, 0.f, active ? 1.f : 0.f);
    }
    const uint32_t slot = warp_append(active, count);
    if (!active) return;
    const uint32_t gx = x + P.tl.x0, gy = y + P.tl.y0;   // pixel in the full frame
    const uint32_t pix = gy * P.cam.xres + gx;
    const float fx = (gx + rnd(P.seed, pix, s, 0)) / P.cam.xres, fy = (gy + rnd(P.seed, pix, s, 1)) / P.cam.yres;
    const float sx = 2.f * fx - 1.f, sy = 1.f - 2.f * fy;  // raster x grows along camera +x, raster y downwards (camera.h ScreenToRaster)
    float dx = P.cam.dir[0] + sx * P.cam.right[0] + sy * P.cam.up[0];
    float dy = P.cam.dir[1] + sx * P.cam.right[1] + sy * P.cam.up[1];
    float dz = P.cam.dir[2] + sx * P.cam.right[2] + sy * P.cam.up[2];
    const float inv = 1.f / sqrtf(dx * dx + dy * dy + dz * dz);
    rays[2 * (size_t)slot] = make_float4(P.cam.eye[0], P.cam.eye[1], P.cam.eye[2], __int_as_float(0x7f800000));
    rays[2 * (size_t)slot + 1] = make_float4(dx * inv, dy * inv, dz * inv, 0.f);
    qPath[slot] = i;
}


// ---- shading of one hit for materials with a glossy lobe (Material "plastic") --------------------------------
// The same steps as the matte code in k_shade — direct lighting by UniformSampleOneLight / EstimateDirect
// (integrator.cpp:86-217) and the path's next direction (path.cpp:144-164) — written against the general BSDF
// of wf_bsdf.cuh (f, Pdf, Sample_f over two lobes).
struct ShadeOut {
    bool wantShadow, wantNext, wantMis;
    float3 so, sd, contrib, no, nd, mo, md, mcontrib, beta;
    float betaW;   // the new beta.w (see WfParams::has_specular)
};
struct Frame { float3 s, t, n; };
__device__ __forceinline__ float3 to_local(const Frame &f, float3 v) {
    return make_float3(v.x * f.s.x + v.y * f.s.y + v.z * f.s.z, v.x * f.t.x + v.y * f.t.y + v.z * f.t.z, v.x * f.n.x + v.y * f.n.y + v.z * f.n.z);
}
__device__ __forceinline__ float3 to_world(const Frame &f, float3 v) {
    return make_float3(f.s.x * v.x + f.t.x * v.y + f.n.x * v.z, f.s.y * v.x + f.t.y * v.y + f.n.y * v.z, f.s.z * v.x + f.t.z * v.y + f.n.z * v.z);
}
// Sphere::Sample(ref, u, &pdf) for a reference point outside the sphere (sphere.cpp:254-291): point + normal on the
// sphere inside the subtended cone; returns the uniform cone pdf.
__device__ __forceinline__ float sphere_cone_sample(const WfParams &P, float3 wl /* centre - p */, float d2, float u0, float u1,
                                                    float3 &pW, float3 &nW) {
    const float r2 = P.light_radius * P.light_radius;
    const float dc = sqrtf(d2);
    const float3 wc = make_float3(wl.x / dc, wl.y / dc, wl.z / dc);
    float3 wcX;
    if (fabsf(wc.x) > fabsf(wc.y)) { const float il = 1.f / sqrtf(wc.x * wc.x + wc.z * wc.z); wcX = make_float3(-wc.z * il, 0.f, wc.x * il); }
    else { const float il = 1.f / sqrtf(wc.y * wc.y + wc.z * wc.z); wcX = make_float3(0.f, wc.z * il, -wc.y * il); }
    const float3 wcY = make_float3(wc.y * wcX.z - wc.z * wcX.y, wc.z * wcX.x - wc.x * wcX.z, wc.x * wcX.y - wc.y * wcX.x);
    const float cosThetaMax = sqrtf(fmaxf(0.f, 1.f - r2 / d2));
    const float cosTheta = (1.f - u0) + u0 * cosThetaMax;
    const float sinTheta = sqrtf(fmaxf(0.f, 1.f - cosTheta * cosTheta));
    const float phi = u1 * 2.f * kPi;
    const float ds = dc * cosTheta - sqrtf(fmaxf(0.f, r2 - dc * dc * sinTheta * sinTheta));
    const float cosAlpha = (dc * dc + r2 - ds * ds) / (2.f * dc * P.light_radius);
    const float sinAlpha = sqrtf(fmaxf(0.f, 1.f - cosAlpha * cosAlpha));
    const float cx = -sinAlpha * cosf(phi), cy = -sinAlpha * sinf(phi), cz = -cosAlpha;
    nW = make_float3(cx * wcX.x + cy * wcY.x + cz * wc.x, cx * wcX.y + cy * wcY.y + cz * wc.y, cx * wcX.z + cy * wcY.z + cz * wc.z);
    pW = make_float3(P.light_pos[0] + P.light_radius * nW.x, P.light_pos[1] + P.light_radius * nW.y, P.light_pos[2] + P.light_radius * nW.z);
    return 1.f / (2.f * kPi * (1.f - cosThetaMax));
}

__device__ void shade_general(const WfParams &P, const WfMat &m, uint32_t depth, uint32_t pix, uint32_t s, float3 p, float3 pe,
                              float3 ng, float3 nf, float3 wo, float4 bt, ShadeOut &o) {
    const float etaScale = bt.w == 0.f ? 1.f : fabsf(bt.w);
    o.betaW = etaScale;
    const uint32_t dim = 2 + 3 * depth;
    if (m.kind == B200RT_MAT_MIRROR || m.kind == B200RT_MAT_GLASS) {
        // A perfectly specular BSDF: no direct lighting (path.cpp:130-141 skips UniformSampleOneLight when there is no
        // non-specular component), only the next direction.  The frame's z is the shape's TRUE normal (isect.n =
        // isect.shading.n for meshes without "N", triangle.cpp:346,421-424), not the one on wo's side.
        const float3 ns = m.flip ? make_float3(-ng.x, -ng.y, -ng.z) : ng;
        Frame fs;
        fs.n = ns;
        if (fabsf(ns.x) > fabsf(ns.y)) { const float il = 1.f / sqrtf(ns.x * ns.x + ns.z * ns.z); fs.s = make_float3(-ns.z * il, 0.f, ns.x * il); }
        else { const float il = 1.f / sqrtf(ns.y * ns.y + ns.z * ns.z); fs.s = make_float3(0.f, ns.z * il, -ns.y * il); }
        fs.t = make_float3(ns.y * fs.s.z - ns.z * fs.s.y, ns.z * fs.s.x - ns.x * fs.s.z, ns.x * fs.s.y - ns.y * fs.s.x);
        const float3 woS = to_local(fs, wo);
        float3 wiS, wgt;
        bool transmitted;
        if (!specular_sample(m, woS, rnd(P.seed, pix, s, dim), wiS, wgt, transmitted)) return;
        float3 nb = make_float3(bt.x * wgt.x, bt.y * wgt.y, bt.z * wgt.z);
        float es = etaScale;
        if (transmitted) es *= woS.z > 0.f ? m.alpha * m.alpha : 1.f / (m.alpha * m.alpha);   // path.cpp:154-160
        bool alive = nb.x > 0.f || nb.y > 0.f || nb.z > 0.f;
        const float mx = fmaxf(nb.x * es, fmaxf(nb.y * es, nb.z * es));   // rrBeta = beta * etaScale (path.cpp:192)
        if (alive && mx < 1.f && depth > 3) {
            const float q = fmaxf(.05f, 1.f - mx);
            if (rnd(P.seed, pix, s, dim + 2) < q) alive = false;
            else { const float k = 1.f / (1.f - q); nb.x *= k; nb.y *= k; nb.z *= k; }
        }
        if (alive) {
            o.beta = nb;
            o.betaW = -es;
            o.nd = to_world(fs, wiS);
            o.no = offset_ray_origin(p, pe, ng, o.nd);
            o.wantNext = true;
        }
        return;
    }
    Frame fr;
    fr.n = nf;
    if (fabsf(nf.x) > fabsf(nf.y)) { const float il = 1.f / sqrtf(nf.x * nf.x + nf.z * nf.z); fr.s = make_float3(-nf.z * il, 0.f, nf.x * il); }
    else { const float il = 1.f / sqrtf(nf.y * nf.y + nf.z * nf.z); fr.s = make_float3(0.f, nf.z * il, -nf.y * il); }
    fr.t = make_float3(nf.y * fr.s.z - nf.z * fr.s.y, nf.z * fr.s.x - nf.x * fr.s.z, nf.x * fr.s.y - nf.y * fr.s.x);
    const float3 woL = to_local(fr, wo);
    const float3 wl = make_float3(P.light_pos[0] - p.x, P.light_pos[1] - p.y, P.light_pos[2] - p.z);
    const float d2 = wl.x * wl.x + wl.y * wl.y + wl.z * wl.z;
    if (P.light_kind == B200RT_LIGHT_SPHERE) {
        const float r2 = P.light_radius * P.light_radius;
        const float3 pO = offset_ray_origin(p, pe, ng, wl);
        const float dO2 = (pO.x - P.light_pos[0]) * (pO.x - P.light_pos[0]) + (pO.y - P.light_pos[1]) * (pO.y - P.light_pos[1]) +
                          (pO.z - P.light_pos[2]) * (pO.z - P.light_pos[2]);
        if (dO2 > r2 && d2 > 0.f) {
            float3 pW, nW;
            const float lightPdf = sphere_cone_sample(P, wl, d2, rnd(P.seed, pix, s, 1000 + 4 * depth), rnd(P.seed, pix, s, 1001 + 4 * depth), pW, nW);
            float3 wi = make_float3(pW.x - p.x, pW.y - p.y, pW.z - p.z);
            const float wl2 = wi.x * wi.x + wi.y * wi.y + wi.z * wi.z;
            if (wl2 > 0.f) {
                const float il = 1.f / sqrtf(wl2);
                wi.x *= il; wi.y *= il; wi.z *= il;
                if ((nW.x * -wi.x + nW.y * -wi.y + nW.z * -wi.z) > 0.f) {
                    const float3 wiL = to_local(fr, wi);
                    float3 f; float scatPdf;
                    bsdf_eval(m, woL, wiL, f, scatPdf);
                    const float c = fabsf(wiL.z);
                    if (f.x * c > 0.f || f.y * c > 0.f || f.z * c > 0.f) {
                        const float g = c * ((lightPdf * lightPdf) / (lightPdf * lightPdf + scatPdf * scatPdf)) / lightPdf;
                        o.contrib = make_float3(bt.x * f.x * P.light_I[0] * g, bt.y * f.y * P.light_I[1] * g, bt.z * f.z * P.light_I[2] * g);
                        const float3 peL = make_float3(kGamma5 * fabsf(pW.x), kGamma5 * fabsf(pW.y), kGamma5 * fabsf(pW.z));
                        o.so = offset_ray_origin(p, pe, ng, make_float3(pW.x - p.x, pW.y - p.y, pW.z - p.z));
                        const float3 tg = offset_ray_origin(pW, peL, nW, make_float3(o.so.x - pW.x, o.so.y - pW.y, o.so.z - pW.z));
                        o.sd = make_float3(tg.x - o.so.x, tg.y - o.so.y, tg.z - o.so.z);
                        o.wantShadow = true;
                    }
                }
            }
            float3 wiL, f; float scatPdf;
            if (bsdf_sample(m, woL, rnd(P.seed, pix, s, 1002 + 4 * depth), rnd(P.seed, pix, s, 1003 + 4 * depth), wiL, f, scatPdf)) {
                const float c = fabsf(wiL.z);
                const float g = c * ((scatPdf * scatPdf) / (scatPdf * scatPdf + lightPdf * lightPdf)) / scatPdf;
                o.mcontrib = make_float3(bt.x * f.x * P.light_I[0] * g, bt.y * f.y * P.light_I[1] * g, bt.z * f.z * P.light_I[2] * g);
                o.md = to_world(fr, wiL);
                o.mo = offset_ray_origin(p, pe, ng, o.md);
                o.wantMis = true;
            }
        }
    } else if (P.light_kind == B200RT_LIGHT_INFINITE) {
        // InfiniteAreaLight with a 1x1 (constant) map: Distribution2D::SampleContinuous(u) = u with mapPdf 1
        // (sampling.h:95-137), so Sample_Li (infinite.cpp:104-131) is theta = u[1] pi, phi = u[0] 2 pi,
        // wi = LightToWorld(sin t cos p, sin t sin p, cos t), pdf = 1 / (2 pi^2 sin t); Li = L.
        {
            const float theta = rnd(P.seed, pix, s, 1001 + 4 * depth) * kPi, phi = rnd(P.seed, pix, s, 1000 + 4 * depth) * 2.f * kPi;
            const float cosT = cosf(theta), sinT = sinf(theta);
            if (sinT != 0.f) {
                const float lightPdf = 1.f / (2.f * kPi * kPi * sinT);
                const float lx = sinT * cosf(phi), ly = sinT * sinf(phi), lz = cosT;
                const float3 wi = make_float3(P.l2w[0] * lx + P.l2w[1] * ly + P.l2w[2] * lz, P.l2w[3] * lx + P.l2w[4] * ly + P.l2w[5] * lz,
                                              P.l2w[6] * lx + P.l2w[7] * ly + P.l2w[8] * lz);
                const float3 wiL = to_local(fr, wi);
                float3 f; float scatPdf;
                bsdf_eval(m, woL, wiL, f, scatPdf);
                const float c = fabsf(wiL.z);
                if (lightPdf > 0.f && (f.x * c > 0.f || f.y * c > 0.f || f.z * c > 0.f)) {
                    const float g = c * ((lightPdf * lightPdf) / (lightPdf * lightPdf + scatPdf * scatPdf)) / lightPdf;
                    o.contrib = make_float3(bt.x * f.x * P.light_I[0] * g, bt.y * f.y * P.light_I[1] * g, bt.z * f.z * P.light_I[2] * g);
                    if (o.contrib.x > 0.f || o.contrib.y > 0.f || o.contrib.z > 0.f) {
                        const float r2x = 2.f * P.light_radius;   // vis = (ref, ref.p + wi * (2 * worldRadius)) (infinite.cpp:127-128)
                        const float3 pOut = make_float3(p.x + wi.x * r2x, p.y + wi.y * r2x, p.z + wi.z * r2x);
                        o.so = offset_ray_origin(p, pe, ng, make_float3(pOut.x - p.x, pOut.y - p.y, pOut.z - p.z));
                        o.sd = make_float3(pOut.x - o.so.x, pOut.y - o.so.y, pOut.z - o.so.z);
                        o.wantShadow = true;
                    }
                }
            }
        }
        // the BSDF-sampled ray (integrator.cpp:176-212): lightPdf = Pdf_Li(wi) (infinite.cpp:133-142); it carries Le when it escapes
        float3 wiL, f; float scatPdf;
        if (bsdf_sample(m, woL, rnd(P.seed, pix, s, 1002 + 4 * depth), rnd(P.seed, pix, s, 1003 + 4 * depth), wiL, f, scatPdf)) {
            const float3 wiW = to_world(fr, wiL);
            const float lz = P.l2w[2] * wiW.x + P.l2w[5] * wiW.y + P.l2w[8] * wiW.z;   // WorldToLight = transpose
            const float sinT = sinf(acosf(fminf(fmaxf(lz, -1.f), 1.f)));              // SphericalTheta (geometry.h)
            if (sinT != 0.f) {
                const float lightPdf = 1.f / (2.f * kPi * kPi * sinT);
                const float c = fabsf(wiL.z);
                const float g = c * ((scatPdf * scatPdf) / (scatPdf * scatPdf + lightPdf * lightPdf)) / scatPdf;
                o.mcontrib = make_float3(bt.x * f.x * P.light_I[0] * g, bt.y * f.y * P.light_I[1] * g, bt.z * f.z * P.light_I[2] * g);
                o.md = wiW;
                o.mo = offset_ray_origin(p, pe, ng, o.md);
                o.wantMis = true;
            }
        }
    } else if (P.light_kind == B200RT_LIGHT_DISTANT) {   // DistantLight::Sample_Li (lights/distant.cpp:49-59)
        const float3 wiL = to_local(fr, make_float3(P.light_pos[0], P.light_pos[1], P.light_pos[2]));
        float3 f; float scatPdf;
        bsdf_eval(m, woL, wiL, f, scatPdf);
        const float g = fabsf(wiL.z);
        o.contrib = make_float3(bt.x * f.x * P.light_I[0] * g, bt.y * f.y * P.light_I[1] * g, bt.z * f.z * P.light_I[2] * g);
        if (o.contrib.x > 0.f || o.contrib.y > 0.f || o.contrib.z > 0.f) {
            const float r2x = 2.f * P.light_radius;
            const float3 pOut = make_float3(p.x + P.light_pos[0] * r2x, p.y + P.light_pos[1] * r2x, p.z + P.light_pos[2] * r2x);
            o.so = offset_ray_origin(p, pe, ng, make_float3(pOut.x - p.x, pOut.y - p.y, pOut.z - p.z));
            o.sd = make_float3(pOut.x - o.so.x, pOut.y - o.so.y, pOut.z - o.so.z);
            o.wantShadow = true;
        }
    } else if (d2 > 0.f) {   // PointLight::Sample_Li (point.cpp:44-53): Li = I / d^2, pdf = 1
        const float invd = 1.f / sqrtf(d2);
        const float3 wiL = to_local(fr, make_float3(wl.x * invd, wl.y * invd, wl.z * invd));
        float3 f; float scatPdf;
        bsdf_eval(m, woL, wiL, f, scatPdf);
        const float g = fabsf(wiL.z) / d2;
        o.contrib = make_float3(bt.x * f.x * P.light_I[0] * g, bt.y * f.y * P.light_I[1] * g, bt.z * f.z * P.light_I[2] * g);
        if (o.contrib.x > 0.f || o.contrib.y > 0.f || o.contrib.z > 0.f) {
            o.so = offset_ray_origin(p, pe, ng, wl);
            o.sd = make_float3(P.light_pos[0] - o.so.x, P.light_pos[1] - o.so.y, P.light_pos[2] - o.so.z);
            o.wantShadow = true;
        }
    }
    // the path's next direction: bsdf->Sample_f, beta *= f |cos| / pdf (path.cpp:144-151), Russian roulette (:190-199)
    float3 wiL, f; float pdf;
    if (bsdf_sample(m, woL, rnd(P.seed, pix, s, dim), rnd(P.seed, pix, s, dim + 1), wiL, f, pdf)) {
        const float c = fabsf(wiL.z) / pdf;
        float3 nb = make_float3(bt.x * f.x * c, bt.y * f.y * c, bt.z * f.z * c);
        bool alive = nb.x > 0.f || nb.y > 0.f || nb.z > 0.f;
        const float mx = fmaxf(nb.x * etaScale, fmaxf(nb.y * etaScale, nb.z * etaScale));   // etaScale == 1 without glass: exact
        if (alive && mx < 1.f && depth > 3) {
            const float q = fmaxf(.05f, 1.f - mx);
            if (rnd(P.seed, pix, s, dim + 2) < q) alive = false;
            else { const float k = 1.f / (1.f - q); nb.x *= k; nb.y *= k; nb.z *= k; }
        }
        if (alive) {
            o.beta = nb;
            o.nd = to_world(fr, wiL);
            o.no = offset_ray_origin(p, pe, ng, o.nd);
            o.wantNext = true;
        }
    }
}

// Shade one wavefront (PathIntegrator::Li body for one bounce, path.cpp:89-199; direct lighting =
// UniformSampleOneLight with one point light, integrator.cpp:86-217; matte BSDF, reflection.h
// LambertianReflection; CosineSampleHemisphere, sampling.h).
// FULL = false: the lean instantiation for one point light + one matte Kd (the bench workload); FULL = true adds the
// sphere area light with MIS, the per-primitive material table and the plastic BSDF.
template <bool FULL>
__global__ void k_shade(const __grid_constant__ WfParams P, uint32_t depth, uint64_t batchStart,
                        const uint32_t *__restrict__ nCur, const float4 *__restrict__ rays,
                        const uint32_t *__restrict__ qPath, const b200rt_hit *__restrict__ hits,
                        const float4 *__restrict__ primTris, float4 *beta, float4 *nextRays, uint32_t *nextPath,
                        uint32_t *nNext, float4 *shRays, uint32_t *shPath, float4 *shContrib, uint32_t *nShadow,
                        float4 *L, float4 *misRays, uint32_t *misPath, float4 *misContrib, uint32_t *nMis) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t n = *nCur;
    const bool area = FULL && P.light_kind == B200RT_LIGHT_SPHERE;
    const bool infinite = FULL && P.light_kind == B200RT_LIGHT_INFINITE;
    bool wantShadow = false, wantNext = false, wantMis = false;
    float3 so = make_float3(0, 0, 0), sd = so, no = so, nd = so, contrib = so, mo = so, md = so, mcontrib = so;
    uint32_t path = 0;
    if (i < n) {
        const b200rt_hit h = hits[i];
        // path.cpp:97-107: emitted radiance is added for camera rays and after a specular bounce (every other vertex gets
        // it through EstimateDirect's MIS pair)
        bool seesLe = depth == 0;
        if (FULL && P.has_specular && depth > 0 && (area || infinite)) seesLe = beta[qPath[i]].w < 0.f;
        if (infinite && seesLe && h.prim == B200RT_MISS) {
            // a camera ray that escapes sees the environment: L += beta * light->Le(ray) (path.cpp:97-107, bounces == 0;
            // infinite.cpp:95-101 with the constant map).  Later escapes are accounted for by the MIS ray.
            const uint32_t pth = qPath[i];
            const float4 bt = beta[pth];
            float4 l = L[pth];
            l.x += bt.x * P.light_I[0]; l.y += bt.y * P.light_I[1]; l.z += bt.z * P.light_I[2];
            L[pth] = l;
        }
        // path.cpp:110: the path ends at maxdepth AFTER the emitted radiance was added; the launch at depth == max_depth
        // exists only for that (scenes with specular materials)
        if (h.prim != B200RT_MISS && (depth < P.max_depth || (area && seesLe))) {
            path = qPath[i];
            const float4 rd = rays[2 * (size_t)i + 1];
            const float4 a = __ldg(primTris + 3 * (size_t)h.prim), b = __ldg(primTris + 3 * (size_t)h.prim + 1),
                         c = __ldg(primTris + 3 * (size_t)h.prim + 2);
            // pHit, pError (triangle.cpp:318-326)
            const float3 p = make_float3(h.b0 * a.x + h.b1 * b.x + h.b2 * c.x, h.b0 * a.y + h.b1 * b.y + h.b2 * c.y,
                                         h.b0 * a.z + h.b1 * b.z + h.b2 * c.z);
            const float3 pe = make_float3(kGamma7 * (fabsf(h.b0 * a.x) + fabsf(h.b1 * b.x) + fabsf(h.b2 * c.x)),
                                          kGamma7 * (fabsf(h.b0 * a.y) + fabsf(h.b1 * b.y) + fabsf(h.b2 * c.y)),
                                          kGamma7 * (fabsf(h.b0 * a.z) + fabsf(h.b1 * b.z) + fabsf(h.b2 * c.z)));
            // geometric normal = Normalize(Cross(dp02, dp12)) (triangle.cpp:346)
            const float3 e0 = make_float3(a.x - c.x, a.y - c.y, a.z - c.z), e1 = make_float3(b.x - c.x, b.y - c.y, b.z - c.z);
            float3 ng = make_float3(e0.y * e1.z - e0.z * e1.y, e0.z * e1.x - e0.x * e1.z, e0.x * e1.y - e0.y * e1.x);
            const float nl = sqrtf(ng.x * ng.x + ng.y * ng.y + ng.z * ng.z);
            // a sphere primitive (killeroo-simple's emitter: matte Kd = 0) ends the path: nothing to reflect.  Seen by a
            // camera ray it adds its emitted radiance (path.cpp:97-107: bounces == 0; later hits are covered by MIS).
            if (area && seesLe && h.prim == P.light_prim) {
                const float4 ro = rays[2 * (size_t)i];
                const float3 ph = make_float3(ro.x + h.t * rd.x, ro.y + h.t * rd.y, ro.z + h.t * rd.z);
                const float cosl = (ph.x - P.light_pos[0]) * -rd.x + (ph.y - P.light_pos[1]) * -rd.y + (ph.z - P.light_pos[2]) * -rd.z;
                if (cosl > 0.f) {  // DiffuseAreaLight::L, one-sided (diffuse.h:56-58)
                    const float4 bt = beta[path];
                    float4 l = L[path];
                    l.x += bt.x * P.light_I[0]; l.y += bt.y * P.light_I[1]; l.z += bt.z * P.light_I[2];
                    L[path] = l;
                }
            }
            if (depth < P.max_depth && nl > 0.f && !(__float_as_uint(b.w) & B200RT_PRIM_SPHERE)) {
                ng.x /= nl; ng.y /= nl; ng.z /= nl;
                float kdr = P.kd[0], kdg = P.kd[1], kdb = P.kd[2];
                WfMat gm;
                gm.glossy = false;
                gm.kind = B200RT_MAT_MATTE; gm.flip = false; gm.alpha = 0.f;
                if (FULL && P.mats) {
                    const uint32_t mi = __ldg(P.primMat + h.prim);
                    const float4 m0 = __ldg(P.mats + 2 * (size_t)mi), m1 = __ldg(P.mats + 2 * (size_t)mi + 1);
                    kdr = m0.y; kdg = m0.z; kdb = m0.w;
                    const uint32_t kf = (uint32_t)m0.x;
                    gm.kind = kf & 15u; gm.flip = (kf & B200RT_MAT_REVERSE_ORIENTATION) != 0;
                    gm.glossy = (gm.kind == B200RT_MAT_PLASTIC || gm.kind == B200RT_MAT_SUBSTRATE) && (m1.x > 0.f || m1.y > 0.f || m1.z > 0.f);
                    gm.ks[0] = m1.x; gm.ks[1] = m1.y; gm.ks[2] = m1.z; gm.alpha = m1.w;
                    if (gm.kind == B200RT_MAT_MATTE) gm.alpha = fminf(fmaxf(m1.x, 0.f), 90.f);   // "sigma" (matte.cpp:55), > 0 = Oren-Nayar
                }
                const float3 wo = make_float3(-rd.x, -rd.y, -rd.z);
                const float side = (ng.x * wo.x + ng.y * wo.y + ng.z * wo.z) < 0.f ? -1.f : 1.f;
                const float3 nf = make_float3(side * ng.x, side * ng.y, side * ng.z);  // wo's hemisphere
                const float4 bt = beta[path];
                if (FULL && (gm.glossy || infinite || P.has_specular || gm.kind >= B200RT_MAT_MIRROR || gm.alpha > 0.f)) {   // Material "plastic" / the infinite light / scenes with specular materials: the general BSDF + light code
                    gm.kd[0] = kdr; gm.kd[1] = kdg; gm.kd[2] = kdb;
                    gm.diffuse = kdr > 0.f || kdg > 0.f || kdb > 0.f;
                    uint32_t gx, gy, gs;
                    path_pixel(P.tl, batchStart + path, gx, gy, gs);
                    ShadeOut o;
                    o.wantShadow = o.wantNext = o.wantMis = false;
                    shade_general(P, gm, depth, (gy + P.tl.y0) * P.cam.xres + gx + P.tl.x0, gs, p, pe, ng, nf, wo, bt, o);
                    wantShadow = o.wantShadow; so = o.so; sd = o.sd; contrib = o.contrib;
                    wantMis = o.wantMis; mo = o.mo; md = o.md; mcontrib = o.mcontrib;
                    wantNext = o.wantNext; no = o.no; nd = o.nd;
                    if (o.wantNext) beta[path] = make_float4(o.beta.x, o.beta.y, o.beta.z, P.has_specular ? o.betaW : 0.f);
                } else {
                // -- direct lighting: PointLight::Sample_Li (point.cpp:44-53) + matte f * |cos|
                const float3 wl = make_float3(P.light_pos[0] - p.x, P.light_pos[1] - p.y, P.light_pos[2] - p.z);
                const float d2 = wl.x * wl.x + wl.y * wl.y + wl.z * wl.z;
                uint32_t x, y, s;
                path_pixel(P.tl, batchStart + path, x, y, s);
                const uint32_t pix = (y + P.tl.y0) * P.cam.xres + x + P.tl.x0, dim = 2 + 3 * depth;
                if (area) {
                    // ---- EstimateDirect with MIS (integrator.cpp:109-217) against the sphere light ----
                    const float r2 = P.light_radius * P.light_radius;
                    const float3 pO = offset_ray_origin(p, pe, ng, wl);  // Sphere::Sample's inside test (sphere.cpp:237-239)
                    const float dO2 = (pO.x - P.light_pos[0]) * (pO.x - P.light_pos[0]) + (pO.y - P.light_pos[1]) * (pO.y - P.light_pos[1]) +
                                      (pO.z - P.light_pos[2]) * (pO.z - P.light_pos[2]);
                    if (dO2 > r2 && d2 > 0.f) {  // reference points outside the sphere (the only case the config scenes have)
                        const float dc = sqrtf(d2);
                        const float3 wc = make_float3(wl.x / dc, wl.y / dc, wl.z / dc);
                        float3 wcX;
                        if (fabsf(wc.x) > fabsf(wc.y)) { const float il = 1.f / sqrtf(wc.x * wc.x + wc.z * wc.z); wcX = make_float3(-wc.z * il, 0.f, wc.x * il); }
                        else { const float il = 1.f / sqrtf(wc.y * wc.y + wc.z * wc.z); wcX = make_float3(0.f, wc.z * il, -wc.y * il); }
                        const float3 wcY = make_float3(wc.y * wcX.z - wc.z * wcX.y, wc.z * wcX.x - wc.x * wcX.z, wc.x * wcX.y - wc.y * wcX.x);
                        const float sinThetaMax2 = r2 / d2;
                        const float cosThetaMax = sqrtf(fmaxf(0.f, 1.f - sinThetaMax2));
                        const float lightPdf = 1.f / (2.f * kPi * (1.f - cosThetaMax));  // UniformConePdf
                        // (a) light sample: Sphere::Sample(ref, u) cone sampling (sphere.cpp:254-291)
                        {
                            const float u0 = rnd(P.seed, pix, s, 1000 + 4 * depth), u1 = rnd(P.seed, pix, s, 1001 + 4 * depth);
                            const float cosTheta = (1.f - u0) + u0 * cosThetaMax;
                            const float sinTheta = sqrtf(fmaxf(0.f, 1.f - cosTheta * cosTheta));
                            const float phi = u1 * 2.f * kPi;
                            const float ds = dc * cosTheta - sqrtf(fmaxf(0.f, r2 - dc * dc * sinTheta * sinTheta));
                            const float cosAlpha = (dc * dc + r2 - ds * ds) / (2.f * dc * P.light_radius);
                            const float sinAlpha = sqrtf(fmaxf(0.f, 1.f - cosAlpha * cosAlpha));
                            const float cx = -sinAlpha * cosf(phi), cy = -sinAlpha * sinf(phi), cz = -cosAlpha;
                            const float3 nW = make_float3(cx * wcX.x + cy * wcY.x + cz * wc.x, cx * wcX.y + cy * wcY.y + cz * wc.y,
                                                          cx * wcX.z + cy * wcY.z + cz * wc.z);
                            const float3 pW = make_float3(P.light_pos[0] + P.light_radius * nW.x, P.light_pos[1] + P.light_radius * nW.y,
                                                          P.light_pos[2] + P.light_radius * nW.z);
                            float3 wi = make_float3(pW.x - p.x, pW.y - p.y, pW.z - p.z);
                            const float wl2 = wi.x * wi.x + wi.y * wi.y + wi.z * wi.z;
                            if (wl2 > 0.f) {
                                const float il = 1.f / sqrtf(wl2);
                                wi.x *= il; wi.y *= il; wi.z *= il;
                                const bool emits = (nW.x * -wi.x + nW.y * -wi.y + nW.z * -wi.z) > 0.f;
                                const float cosi = nf.x * wi.x + nf.y * wi.y + nf.z * wi.z;
                                if (emits && cosi > 0.f) {  // reflective Lambertian lobe: wi in wo's hemisphere
                                    const float scatPdf = cosi * kInvPi;
                                    const float wgt = (lightPdf * lightPdf) / (lightPdf * lightPdf + scatPdf * scatPdf);  // PowerHeuristic
                                    const float g = cosi * kInvPi * wgt / lightPdf;
                                    contrib = make_float3(bt.x * kdr * P.light_I[0] * g, bt.y * kdg * P.light_I[1] * g, bt.z * kdb * P.light_I[2] * g);
                                    if (contrib.x > 0.f || contrib.y > 0.f || contrib.z > 0.f) {
                                        // Interaction::SpawnRayTo(const Interaction &) (interaction.h:73-78)
                                        const float3 peL = make_float3(kGamma5 * fabsf(pW.x), kGamma5 * fabsf(pW.y), kGamma5 * fabsf(pW.z));
                                        so = offset_ray_origin(p, pe, ng, make_float3(pW.x - p.x, pW.y - p.y, pW.z - p.z));
                                        const float3 tg = offset_ray_origin(pW, peL, nW, make_float3(so.x - pW.x, so.y - pW.y, so.z - pW.z));
                                        sd = make_float3(tg.x - so.x, tg.y - so.y, tg.z - so.z);
                                        wantShadow = true;
                                    }
                                }
                            }
                        }
                        // (b) BSDF sample: CosineSampleHemisphere; the ray is traced to see whether it reaches the light
                        {
                            const float u0 = 2.f * rnd(P.seed, pix, s, 1002 + 4 * depth) - 1.f, u1 = 2.f * rnd(P.seed, pix, s, 1003 + 4 * depth) - 1.f;
                            float dxl = 0.f, dyl = 0.f;
                            if (u0 != 0.f || u1 != 0.f) {
                                float r, th;
                                if (fabsf(u0) > fabsf(u1)) { r = u0; th = (kPi / 4.f) * (u1 / u0); }
                                else { r = u1; th = (kPi / 2.f) - (kPi / 4.f) * (u0 / u1); }
                                dxl = r * cosf(th); dyl = r * sinf(th);
                            }
                            const float dzl = sqrtf(fmaxf(0.f, 1.f - dxl * dxl - dyl * dyl));
                            float3 ss;
                            if (fabsf(nf.x) > fabsf(nf.y)) { const float il = 1.f / sqrtf(nf.x * nf.x + nf.z * nf.z); ss = make_float3(-nf.z * il, 0.f, nf.x * il); }
                            else { const float il = 1.f / sqrtf(nf.y * nf.y + nf.z * nf.z); ss = make_float3(0.f, nf.z * il, -nf.y * il); }
                            const float3 ts = make_float3(nf.y * ss.z - nf.z * ss.y, nf.z * ss.x - nf.x * ss.z, nf.x * ss.y - nf.y * ss.x);
                            const float3 wi = make_float3(ss.x * dxl + ts.x * dyl + nf.x * dzl, ss.y * dxl + ts.y * dyl + nf.y * dzl,
                                                          ss.z * dxl + ts.z * dyl + nf.z * dzl);
                            const float scatPdf = dzl * kInvPi;
                            if (scatPdf > 0.f && (kdr > 0.f || kdg > 0.f || kdb > 0.f)) {  // !f.IsBlack() && scatteringPdf > 0
                                const float wgt = (scatPdf * scatPdf) / (scatPdf * scatPdf + lightPdf * lightPdf);
                                // f * |cos| / pdf = Kd; Li = L when the ray's closest hit is the emitter
                                mcontrib = make_float3(bt.x * kdr * P.light_I[0] * wgt, bt.y * kdg * P.light_I[1] * wgt, bt.z * kdb * P.light_I[2] * wgt);
                                mo = offset_ray_origin(p, pe, ng, wi);
                                md = wi;
                                wantMis = true;
                            }
                        }
                    }
                } else if (FULL && P.light_kind == B200RT_LIGHT_DISTANT) {
                    // DistantLight::Sample_Li (lights/distant.cpp:49-59): wi = wLight, Li = L, pdf = 1
                    const float cosl = nf.x * P.light_pos[0] + nf.y * P.light_pos[1] + nf.z * P.light_pos[2];
                    if (cosl > 0.f) {
                        const float g = cosl * kInvPi;
                        contrib = make_float3(bt.x * kdr * P.light_I[0] * g, bt.y * kdg * P.light_I[1] * g,
                                              bt.z * kdb * P.light_I[2] * g);
                        if (contrib.x > 0.f || contrib.y > 0.f || contrib.z > 0.f) {
                            const float r2x = 2.f * P.light_radius;   // pOutside = ref.p + wLight * (2 * worldRadius)
                            const float3 pOut = make_float3(p.x + P.light_pos[0] * r2x, p.y + P.light_pos[1] * r2x, p.z + P.light_pos[2] * r2x);
                            so = offset_ray_origin(p, pe, ng, make_float3(pOut.x - p.x, pOut.y - p.y, pOut.z - p.z));
                            sd = make_float3(pOut.x - so.x, pOut.y - so.y, pOut.z - so.z);
                            wantShadow = true;
                        }
                    }
                } else if (d2 > 0.f) {
                    const float invd = 1.f / sqrtf(d2);
                    const float cosl = (nf.x * wl.x + nf.y * wl.y + nf.z * wl.z) * invd;
                    if (cosl > 0.f) {  // SameHemisphere(wo, wi) for a reflective Lambertian lobe
                        const float g = cosl * kInvPi / d2;
                        contrib = make_float3(bt.x * kdr * P.light_I[0] * g, bt.y * kdg * P.light_I[1] * g,
                                              bt.z * kdb * P.light_I[2] * g);
                        if (contrib.x > 0.f || contrib.y > 0.f || contrib.z > 0.f) {
                            // Interaction::SpawnRayTo (interaction.h:68-72)
                            so = offset_ray_origin(p, pe, ng, wl);
                            sd = make_float3(P.light_pos[0] - so.x, P.light_pos[1] - so.y, P.light_pos[2] - so.z);
                            wantShadow = true;
                        }
                    }
                }
                // -- BSDF sample: CosineSampleHemisphere via ConcentricSampleDisk (sampling.cpp)
                const float u0 = 2.f * rnd(P.seed, pix, s, dim) - 1.f, u1 = 2.f * rnd(P.seed, pix, s, dim + 1) - 1.f;
                float dxl = 0.f, dyl = 0.f;
                if (u0 != 0.f || u1 != 0.f) {
                    float r, th;
                    if (fabsf(u0) > fabsf(u1)) { r = u0; th = (kPi / 4.f) * (u1 / u0); }
                    else { r = u1; th = (kPi / 2.f) - (kPi / 4.f) * (u0 / u1); }
                    dxl = r * cosf(th); dyl = r * sinf(th);
                }
                const float dzl = sqrtf(fmaxf(0.f, 1.f - dxl * dxl - dyl * dyl));
                // CoordinateSystem(nf, &s, &t) (geometry.h)
                float3 ss;
                if (fabsf(nf.x) > fabsf(nf.y)) { const float il = 1.f / sqrtf(nf.x * nf.x + nf.z * nf.z); ss = make_float3(-nf.z * il, 0.f, nf.x * il); }
                else { const float il = 1.f / sqrtf(nf.y * nf.y + nf.z * nf.z); ss = make_float3(0.f, nf.z * il, -nf.y * il); }
                const float3 ts = make_float3(nf.y * ss.z - nf.z * ss.y, nf.z * ss.x - nf.x * ss.z, nf.x * ss.y - nf.y * ss.x);
                const float3 wi = make_float3(ss.x * dxl + ts.x * dyl + nf.x * dzl, ss.y * dxl + ts.y * dyl + nf.y * dzl,
                                              ss.z * dxl + ts.z * dyl + nf.z * dzl);
                if (dzl > 0.f) {
                    // beta *= f * |cos| / pdf = Kd (path.cpp:151)
                    float3 nb = make_float3(bt.x * kdr, bt.y * kdg, bt.z * kdb);
                    bool alive = nb.x > 0.f || nb.y > 0.f || nb.z > 0.f;
                    // Russian roulette (path.cpp:190-199): bounces > 3 and max(beta) < rrThreshold = 1
                    const float mx = fmaxf(nb.x, fmaxf(nb.y, nb.z));
                    if (alive && mx < 1.f && depth > 3) {
                        const float q = fmaxf(.05f, 1.f - mx);
                        if (rnd(P.seed, pix, s, dim + 2) < q) alive = false;
                        else { const float k = 1.f / (1.f - q); nb.x *= k; nb.y *= k; nb.z *= k; }
                    }
                    if (alive) {
                        beta[path] = make_float4(nb.x, nb.y, nb.z, 0.f);
                        no = offset_ray_origin(p, pe, ng, wi);  // Interaction::SpawnRay (interaction.h:64-67)
                        nd = wi;
                        wantNext = true;
                    }
                }
                }  // matte
            }
        }
    }
    const uint32_t sj = warp_append(wantShadow, nShadow);
    if (wantShadow) {
        shRays[2 * (size_t)sj] = make_float4(so.x, so.y, so.z, 1.f - kShadowEpsilon);
        shRays[2 * (size_t)sj + 1] = make_float4(sd.x, sd.y, sd.z, 0.f);
        shPath[sj] = path;
        shContrib[sj] = make_float4(contrib.x, contrib.y, contrib.z, 0.f);
    }
    const uint32_t nj = warp_append(wantNext, nNext);
    if (wantNext) {
        nextRays[2 * (size_t)nj] = make_float4(no.x, no.y, no.z, __int_as_float(0x7f800000));
        nextRays[2 * (size_t)nj + 1] = make_float4(nd.x, nd.y, nd.z, 0.f);
        nextPath[nj] = path;
    }
    if (area || infinite) {  // warp-uniform
        const uint32_t mj = warp_append(wantMis, nMis);
        if (wantMis) {
            misRays[2 * (size_t)mj] = make_float4(mo.x, mo.y, mo.z, __int_as_float(0x7f800000));
            misRays[2 * (size_t)mj + 1] = make_float4(md.x, md.y, md.z, 0.f);
            misPath[mj] = path;
            misContrib[mj] = make_float4(mcontrib.x, mcontrib.y, mcontrib.z, 0.f);
        }
    }
}

// The BSDF-sampled ray of EstimateDirect (integrator.cpp:196-212): it contributes when its closest hit is the emitter
// (`lightIsect.primitive->GetAreaLight() == &light`, Li = lightIsect.Le(-wi); seen from outside the sphere is front-facing).
__global__ void k_mis_resolve(const uint32_t *__restrict__ nMis, const b200rt_hit *__restrict__ hits, uint32_t lightPrim,
                              const uint32_t *__restrict__ misPath, const float4 *__restrict__ misContrib, float4 *L) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= *nMis || hits[j].prim != lightPrim) return;
    const float4 c = misContrib[j];
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

// FilmTile::AddSample with the box filter (film.h:130-169): each pixel sums its own samples, in
// sample order (deterministic), into (sum R, sum G, sum B, sum weight).
__global__ void k_film(const __grid_constant__ WfParams P, uint64_t batchStart, uint32_t nPixels, const float4 *__restrict__ L,
                       float4 *film) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nPixels) return;
    uint32_t x, y, s;
    if (!path_pixel(P.tl, batchStart + (uint64_t)i * P.tl.spp, x, y, s)) return;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (uint32_t k = 0; k < P.tl.spp; ++k) {
        const float4 l = L[(size_t)i * P.tl.spp + k];
        acc.x += l.x; acc.y += l.y; acc.z += l.z; acc.w += l.w;
    }
    float4 f = film[(size_t)y * P.tl.xres + x];
    f.x += acc.x; f.y += acc.y; f.z += acc.z; f.w += acc.w;
    film[(size_t)y * P.tl.xres + x] = f;
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

struct Workspace {
    size_t cap = 0;
    float4 *misRays = nullptr, *misContrib = nullptr;   // area-light mode only (EstimateDirect's BSDF-sampled rays)
    uint32_t *misPath = nullptr;
    float4 *mats = nullptr;
    uint8_t *primMat = nullptr;
    size_t mats_n = 0, primMat_n = 0;
    uint4 *perRay = nullptr;   // heat-map mode only
    uint32_t *heat = nullptr;
    size_t heat_px = 0;
    float4 *rays[2] = {nullptr, nullptr}, *shRays = nullptr, *shContrib = nullptr, *beta = nullptr, *L = nullptr;
    uint32_t *qPath[2] = {nullptr, nullptr}, *shPath = nullptr, *counts = nullptr;
    b200rt_hit *hits = nullptr;
    uint8_t *occ = nullptr;
    float4 *film = nullptr;
    size_t film_px = 0;
    // recorded queues of the last batch (b200rt_render_keep_queues)
    bool keep = false;
    std::vector<float4 *> kept;
    std::vector<uint32_t> keptCount;
    size_t keptCap = 0;
};

void free_ws(Workspace *w) {
    for (int k = 0; k < 2; ++k) { cudaFree(w->rays[k]); cudaFree(w->qPath[k]); }
    cudaFree(w->shRays); cudaFree(w->shContrib); cudaFree(w->beta); cudaFree(w->L); cudaFree(w->shPath);
    cudaFree(w->counts); cudaFree(w->hits); cudaFree(w->occ); cudaFree(w->film); cudaFree(w->perRay); cudaFree(w->heat);
    cudaFree(w->misRays); cudaFree(w->misContrib); cudaFree(w->misPath); cudaFree(w->mats); cudaFree(w->primMat);
    for (float4 *p : w->kept) cudaFree(p);
    *w = Workspace();
}

int tiling_of(const b200rt_render_desc *d, Tiling *t) {
    if (!d || d->camera.xres == 0 || d->camera.yres == 0 || d->spp == 0 || d->world == 0 || d->rank >= d->world)
        return set_error(B200RT_EINVAL, "bad render description (resolution/spp/rank/world)");
    t->xres = d->camera.xres; t->yres = d->camera.yres;
    t->x0 = t->y0 = 0;
    if (d->crop[0] | d->crop[1] | d->crop[2] | d->crop[3]) {   // Film "cropwindow" in pixels: [x0, x1) x [y0, y1)
        if (d->crop[0] >= d->crop[1] || d->crop[2] >= d->crop[3] || d->crop[1] > d->camera.xres || d->crop[3] > d->camera.yres)
            return set_error(B200RT_EINVAL, "bad crop window [%u, %u) x [%u, %u)", d->crop[0], d->crop[1], d->crop[2], d->crop[3]);
        t->x0 = d->crop[0]; t->y0 = d->crop[2];
        t->xres = d->crop[1] - d->crop[0]; t->yres = d->crop[3] - d->crop[2];
    }
    t->tile = d->tile_size ? d->tile_size : 16;
    t->ntx = (t->xres + t->tile - 1) / t->tile;
    t->nty = (t->yres + t->tile - 1) / t->tile;
    t->rank = d->rank; t->world = d->world; t->spp = d->spp;
    const uint32_t total = t->ntx * t->nty;
    t->n_tiles_rank = total > t->rank ? (total - t->rank + t->world - 1) / t->world : 0;
    t->n_paths_rank = (uint64_t)t->n_tiles_rank * t->tile * t->tile * t->spp;
    return B200RT_OK;
}

}  // namespace

// the workspace hangs off the scene handle
struct b200rt_wavefront_ws : Workspace {};
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
    ws_of(s)->keep = keep != 0;
    return B200RT_OK;
}

extern "C" int b200rt_render_queue(b200rt_scene *s, uint32_t q, b200rt_ray *rays, uint64_t *n) {
    if (!s || !n) return set_error(B200RT_EINVAL, "null argument");
    Workspace *w = ws_of(s);
    if (q >= w->kept.size()) { *n = 0; return B200RT_OK; }
    const uint64_t have = w->keptCount[q];
    if (!rays) { *n = have; return B200RT_OK; }
    const uint64_t m = std::min<uint64_t>(have, *n);
    WCU(cudaMemcpy(rays, w->kept[q], m * sizeof(b200rt_ray), cudaMemcpyDefault));
    *n = m;
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
static int render_impl(b200rt_scene *s, const b200rt_render_desc *d, float *film, uint32_t *heat, b200rt_render_stats *stats,
                       void *stream) {
    if (!s || !d || !film) return set_error(B200RT_EINVAL, "null argument");
    WfParams P;
    std::memset(&P, 0, sizeof(P));
    int rc = tiling_of(d, &P.tl);
    if (rc) return rc;
    P.cam = d->camera;
    std::memcpy(P.light_pos, d->light_pos, 12); std::memcpy(P.light_I, d->light_I, 12); std::memcpy(P.kd, d->kd, 12);
    P.max_depth = d->max_depth; P.seed = d->seed;
    if (d->max_depth > 68) return set_error(B200RT_EINVAL, "max_depth %u too large", d->max_depth);
    P.light_kind = d->light_kind; P.light_prim = d->light_prim; P.light_radius = d->light_radius;
    const bool area = d->light_kind == B200RT_LIGHT_SPHERE;
    const bool infinite = d->light_kind == B200RT_LIGHT_INFINITE;
    const bool mis = area || infinite;   // lights with area: EstimateDirect traces a BSDF-sampled ray too
    if (infinite) {
        bool zero = true;
        for (int k = 0; k < 9; ++k) zero = zero && d->light_to_world[k] == 0.f;
        for (int k = 0; k < 9; ++k) P.l2w[k] = zero ? (k % 4 == 0 ? 1.f : 0.f) : d->light_to_world[k];
        if (!(d->light_radius > 0.f)) return set_error(B200RT_EINVAL, "infinite light: light_radius must be the scene's bounding-sphere radius");
    }
    if (d->light_kind > B200RT_LIGHT_INFINITE) return set_error(B200RT_EINVAL, "unknown light_kind %u", d->light_kind);
    const bool distant = d->light_kind == B200RT_LIGHT_DISTANT;
    if (distant) {
        const float l2 = d->light_pos[0] * d->light_pos[0] + d->light_pos[1] * d->light_pos[1] + d->light_pos[2] * d->light_pos[2];
        if (!(d->light_radius > 0.f) || !(std::fabs(l2 - 1.f) < 1e-3f))
            return set_error(B200RT_EINVAL, "distant light: light_pos must be the unit direction to the light and light_radius the scene's bounding-sphere radius");
    }
    if (area && (!(d->light_radius > 0.f) || d->light_prim >= s->n_prims))
        return set_error(B200RT_EINVAL, "sphere light: radius must be > 0 and light_prim a primitive of the scene");
    if (d->materials && (!d->prim_material || d->n_materials == 0 || d->n_materials > 256))
        return set_error(B200RT_EINVAL, "materials need prim_material and 1..256 records");
    cudaStream_t st = (cudaStream_t)stream;
    Workspace *w = ws_of(s);
    const uint32_t tp = P.tl.tile * P.tl.tile;
    const uint64_t unit = (uint64_t)tp * P.tl.spp;  // batches hold whole tiles so a pixel's samples never split
    uint64_t want = d->max_batch ? d->max_batch : (1u << 22);
    uint64_t cap = std::max<uint64_t>(unit, want / unit * unit);
    cap = std::min<uint64_t>(cap, std::max<uint64_t>(unit, P.tl.n_paths_rank));
    if (cap > 0x7fffffffull) return set_error(B200RT_EINVAL, "batch too large");
    const uint32_t misBase = 2 * (d->max_depth + 2);          // counts[misBase + depth] = MIS queue size at depth
    const uint32_t nCounts = misBase + d->max_depth + 1;
    if (w->cap < cap) {
        const bool keep = w->keep;
        free_ws(w);
        w->keep = keep;
        for (int k = 0; k < 2; ++k) { WCU(cudaMalloc(&w->rays[k], cap * 32)); WCU(cudaMalloc(&w->qPath[k], cap * 4)); }
        WCU(cudaMalloc(&w->shRays, cap * 32)); WCU(cudaMalloc(&w->shContrib, cap * 16)); WCU(cudaMalloc(&w->shPath, cap * 4));
        WCU(cudaMalloc(&w->beta, cap * 16)); WCU(cudaMalloc(&w->L, cap * 16));
        WCU(cudaMalloc(&w->hits, cap * sizeof(b200rt_hit))); WCU(cudaMalloc(&w->occ, cap));
        WCU(cudaMalloc(&w->counts, 4 * 3 * 72));
        w->cap = cap;
    }
    if (mis && !w->misRays) {
        WCU(cudaMalloc(&w->misRays, w->cap * 32)); WCU(cudaMalloc(&w->misContrib, w->cap * 16)); WCU(cudaMalloc(&w->misPath, w->cap * 4));
    }
    if (d->materials) {
        for (uint32_t k = 0; k < d->n_materials; ++k)
        {
            const float kf = d->materials[8 * k];
            const uint32_t kind = (uint32_t)kf & 15u;
            if (!(kf >= 0.f && kf < 64.f) || kf != (float)(uint32_t)kf || kind > B200RT_MAT_SUBSTRATE)
                return set_error(B200RT_EUNSUPPORTED, "material %u: only matte, plastic, mirror, glass, metal and substrate are supported", k);
            if ((kind == B200RT_MAT_METAL || kind == B200RT_MAT_SUBSTRATE) && !(d->materials[8 * k + 7] > 0.f))
                return set_error(B200RT_EINVAL, "material %u: roughness must be > 0 (perfectly smooth microfacet lobes are not supported)", k);
            if (kind == B200RT_MAT_GLASS && !(d->materials[8 * k + 7] > 0.f))
                return set_error(B200RT_EINVAL, "material %u: glass needs its \"index\" of refraction in the roughness slot", k);
            if (kind == B200RT_MAT_MIRROR || kind == B200RT_MAT_GLASS) P.has_specular = 1;
        }
        if (w->mats_n < d->n_materials) { cudaFree(w->mats); w->mats = nullptr; WCU(cudaMalloc(&w->mats, (size_t)d->n_materials * 32)); w->mats_n = d->n_materials; }
        if (w->primMat_n < s->n_prims) { cudaFree(w->primMat); w->primMat = nullptr; WCU(cudaMalloc(&w->primMat, s->n_prims)); w->primMat_n = s->n_prims; }
        for (uint32_t k = 0; k < s->n_prims; ++k)
            if (d->prim_material[k] >= d->n_materials) return set_error(B200RT_EINVAL, "prim_material[%u] out of range", k);
        std::vector<float> mt(d->materials, d->materials + 8 * (size_t)d->n_materials);
        for (uint32_t k = 0; k < d->n_materials; ++k)
            if (const uint32_t kf = (uint32_t)mt[8 * k]; !(kf & B200RT_MAT_NO_REMAP) && ((kf & 15u) == B200RT_MAT_PLASTIC ||
                                                                                            (kf & 15u) == B200RT_MAT_METAL || (kf & 15u) == B200RT_MAT_SUBSTRATE)) {  // TrowbridgeReitzDistribution::RoughnessToAlpha (microfacet.h:123-128), remaproughness true
                const float x = std::log(std::max(mt[8 * k + 7], 1e-3f));
                mt[8 * k + 7] = 1.62142f + 0.819955f * x + 0.1734f * x * x + 0.0171201f * x * x * x + 0.000640711f * x * x * x * x;
            }
        WCU(cudaMemcpy(w->mats, mt.data(), mt.size() * 4, cudaMemcpyHostToDevice));
        WCU(cudaMemcpyAsync(w->primMat, d->prim_material, s->n_prims, cudaMemcpyHostToDevice, st));
        P.mats = w->mats; P.primMat = w->primMat;
    }
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
    if (w->keep) {
        const size_t nq = 2 * (size_t)(d->max_depth + 1);
        if (w->kept.size() != nq || w->keptCap < cap) {
            for (float4 *p : w->kept) cudaFree(p);
            w->kept.assign(nq, nullptr);
            for (size_t q = 0; q < nq; ++q) WCU(cudaMalloc(&w->kept[q], cap * 32));
            w->keptCap = cap;
        }
        w->keptCount.assign(nq, 0);
    }
    std::vector<cudaEvent_t> evs;
    auto ev = [&]() { cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, st); evs.push_back(e); return e; };
    struct Span { cudaEvent_t a, b; bool any; };
    std::vector<Span> spans;
    const uint64_t launches0 = g_launches.load();
    uint64_t nClosest = 0, nShadow = 0, nCamera = 0, nMis = 0;
    std::vector<uint32_t> hostCounts(nCounts);
    cudaEvent_t t0 = ev();
    const unsigned TB = 256;
    for (uint64_t start = 0; start < P.tl.n_paths_rank; start += cap) {
        const uint32_t nSlots = (uint32_t)std::min<uint64_t>(cap, P.tl.n_paths_rank - start);
        const unsigned gridSlots = (nSlots + TB - 1) / TB;
        WCU(cudaMemsetAsync(w->counts, 0, 4 * nCounts, st));
        // counts[2*depth] = closest queue size at depth, counts[2*depth+1] = shadow queue size at depth
        k_camera<<<gridSlots, TB, 0, st>>>(P, start, nSlots, w->rays[0], w->qPath[0], w->counts, w->beta, w->L);
        g_launches.fetch_add(1);
        int cur = 0;
        for (uint32_t depth = 0; depth <= d->max_depth; ++depth) {
            if (w->keep) WCU(cudaMemcpyAsync(w->kept[2 * depth], w->rays[cur], (size_t)nSlots * 32, cudaMemcpyDeviceToDevice, st));
            cudaEvent_t a = ev();
            rc = launch_trace(s, w->rays[cur], nSlots, w->counts + 2 * depth, w->hits, nullptr, false,
                              heat ? s->d_counters : nullptr, st, heat ? w->perRay : nullptr);
            if (rc) return rc;
            spans.push_back({a, ev(), false});
            if (heat) {
                k_heat<<<gridSlots, TB, 0, st>>>(P, start, w->counts + 2 * depth, w->qPath[cur], w->perRay, d_heat, 0u);
                g_launches.fetch_add(1);
            }
            // path.cpp:110: the last intersection is found, then the path ends — after the emitted radiance a specular
            // bounce may still see there (path.cpp:97-107): one more shade launch, which only adds Le
            if (depth == d->max_depth && !(P.has_specular && mis)) break;
            (mis || distant || P.mats ? k_shade<true> : k_shade<false>)<<<gridSlots, TB, 0, st>>>(P, depth, start, w->counts + 2 * depth, w->rays[cur], w->qPath[cur], w->hits,
                                              s->d_prim_tris, w->beta, w->rays[cur ^ 1], w->qPath[cur ^ 1],
                                              w->counts + 2 * (depth + 1), w->shRays, w->shPath, w->shContrib,
                                              w->counts + 2 * depth + 1, w->L, w->misRays, w->misPath, w->misContrib,
                                              w->counts + misBase + depth);
            g_launches.fetch_add(1);
            if (depth == d->max_depth) break;
            if (w->keep) WCU(cudaMemcpyAsync(w->kept[2 * depth + 1], w->shRays, (size_t)nSlots * 32, cudaMemcpyDeviceToDevice, st));
            a = ev();
            rc = launch_trace(s, w->shRays, nSlots, w->counts + 2 * depth + 1, nullptr, w->occ, true,
                              heat ? s->d_counters : nullptr, st, heat ? w->perRay : nullptr);
            if (rc) return rc;
            spans.push_back({a, ev(), true});
            if (heat) {
                k_heat<<<gridSlots, TB, 0, st>>>(P, start, w->counts + 2 * depth + 1, w->shPath, w->perRay, d_heat, 1u);
                g_launches.fetch_add(1);
            }
            k_shadow_resolve<<<gridSlots, TB, 0, st>>>(w->counts + 2 * depth + 1, w->occ, w->shPath, w->shContrib, w->L);
            g_launches.fetch_add(1);
            if (mis) {  // EstimateDirect's BSDF-sampled ray: one more closest-hit queue per bounce (w->hits is free again)
                a = ev();
                rc = launch_trace(s, w->misRays, nSlots, w->counts + misBase + depth, w->hits, nullptr, false,
                                  heat ? s->d_counters : nullptr, st, heat ? w->perRay : nullptr);
                if (rc) return rc;
                spans.push_back({a, ev(), false});
                if (heat) {
                    k_heat<<<gridSlots, TB, 0, st>>>(P, start, w->counts + misBase + depth, w->misPath, w->perRay, d_heat, 0u);
                    g_launches.fetch_add(1);
                }
                k_mis_resolve<<<gridSlots, TB, 0, st>>>(w->counts + misBase + depth, w->hits, infinite ? (uint32_t)B200RT_MISS : d->light_prim, w->misPath, w->misContrib, w->L);
                g_launches.fetch_add(1);
            }
            cur ^= 1;
            // deep paths (scenes/living-room: maxdepth 65): Russian roulette empties the queues long before maxdepth, so
            // look at the next queue's size every 4th depth and stop launching once no path is left
            if (depth >= 7 && (depth & 3) == 3 && depth + 1 < d->max_depth) {
                uint32_t left = 0;
                WCU(cudaMemcpyAsync(&left, w->counts + 2 * (depth + 1), 4, cudaMemcpyDeviceToHost, st));
                WCU(cudaStreamSynchronize(st));
                if (left == 0) break;
            }
        }
        const uint32_t nPix = nSlots / P.tl.spp;
        k_film<<<(nPix + TB - 1) / TB, TB, 0, st>>>(P, start, nPix, w->L, d_film);
        g_launches.fetch_add(1);
        WCU(cudaMemcpyAsync(hostCounts.data(), w->counts, 4 * nCounts, cudaMemcpyDeviceToHost, st));
        WCU(cudaStreamSynchronize(st));
        nCamera += hostCounts[0];
        for (uint32_t depth = 0; depth <= d->max_depth; ++depth) {
            nClosest += hostCounts[2 * depth];
            if (depth < d->max_depth) nShadow += hostCounts[2 * depth + 1];
            if (mis && depth < d->max_depth) { nClosest += hostCounts[misBase + depth]; nMis += hostCounts[misBase + depth]; }
            if (w->keep) {
                w->keptCount[2 * depth] = hostCounts[2 * depth];
                if (depth < d->max_depth) w->keptCount[2 * depth + 1] = hostCounts[2 * depth + 1];
            }
        }
    }
    if (!film_dev) WCU(cudaMemcpyAsync(film, d_film, npx * 16, cudaMemcpyDeviceToHost, st));
    if (heat && !heat_dev) WCU(cudaMemcpyAsync(heat, d_heat, npx * 4 * B200RT_HEAT_PLANES, cudaMemcpyDeviceToHost, st));
    cudaEvent_t t1 = ev();
    WCU(cudaStreamSynchronize(st));
    WCU(cudaGetLastError());
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
    for (cudaEvent_t e : evs) cudaEventDestroy(e);
    return B200RT_OK;
}
