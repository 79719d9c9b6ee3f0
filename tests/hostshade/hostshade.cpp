// TEST INFRASTRUCTURE ONLY.  The shade stage of the wavefront driver (thesis-pbrt-v3_b200/csrc/wf_shade.cuh and the
// headers under it: sampler, camera, triangle tail, BSDFs, lights, one path vertex) compiled for the HOST with g++, so that
// the restatement can be pinned against the compiled reference (oracle/_ref: sample values, camera rays, every spawned ray
// of a traced render) in the CPU test suite, without a GPU.  Nothing in the product library links or calls this; the device
// build of the same text is what b200rt_render runs.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <algorithm>
#include <thread>
#include <vector>
#include "../../thesis-pbrt-v3_b200/csrc/wf_setup_host.h"
#include "../../thesis-pbrt-v3_b200/csrc/wf_sampler_host.h"

namespace b200rt {
static char g_err[512];
int set_error(int code, const char *fmt, ...) {
    va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
    return code;
}
}  // namespace b200rt
using namespace b200rt;

struct hs_ctx {
    WfParams P;
    bool full;
    std::vector<float> tris, uv, normals, mats;
    std::vector<uint8_t> flags, isSphere, primMat;
    std::vector<uint32_t> sobol32;
    std::vector<uint64_t> vdc;
    std::vector<WfTexture> tex;
    std::vector<WfLight> lights;
    std::vector<float> vox;   // "spatial": cdfHi then pickPdf of every voxel
};

extern "C" const char *hs_last_error() { return g_err; }

extern "C" hs_ctx *hs_create(const b200rt_render_desc *d, uint32_t n_prims, const float *tri_verts, const float *tri_uvs,
                             const float *tri_normals, const uint8_t *prim_flags, const uint8_t *prim_is_sphere,
                             const float *scene_bounds /* 6 floats: Scene::WorldBound(), needed by the spatial light distribution */) {
    hs_ctx *c = new hs_ctx;
    WfParams &P = c->P;
    std::memset(&P, 0, sizeof(P));
    if (tiling_of(d, &P.tl)) { delete c; return nullptr; }
    P.cam = d->camera; P.cam_exact = d->camera_exact;
    std::memcpy(P.cm.r2c, d->raster_to_camera, 64); std::memcpy(P.cm.c2w, d->camera_to_world, 64);
    std::memcpy(P.kd, d->kd, 12);
    P.max_depth = d->max_depth;
    if (lights_of(d, n_prims, c->lights)) { delete c; return nullptr; }
    P.light = c->lights[0];
    P.nLights = (uint32_t)c->lights.size();
    P.lights = P.nLights > 1 ? c->lights.data() : nullptr;
    if (P.nLights > 1 && d->light_strategy == B200RT_LIGHTS_SPATIAL) {
        // every voxel's distribution (the device computes them with k_light_voxels), here on a few host threads
        std::memcpy(P.wb, scene_bounds, sizeof(P.wb));
        spatial_voxel_counts(P.wb, 64, P.nVox);
        const size_t nv = (size_t)P.nVox[0] * P.nVox[1] * P.nVox[2], per = nv * P.nLights;
        c->vox.resize(2 * per);
        float halton[640];
        spatial_halton_table(halton);
        const unsigned nt = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
        std::vector<std::thread> th;
        for (unsigned t = 0; t < nt; ++t)
            th.emplace_back([&, t] {
                std::vector<float> contrib(P.nLights);
                for (size_t v = t; v < nv; v += nt)
                    spatial_voxel_distribution(c->lights.data(), P.nLights, P.wb, P.nVox, (uint32_t)v, halton, contrib.data(),
                                               c->vox.data() + v * P.nLights, c->vox.data() + per + v * P.nLights);
            });
        for (auto &x : th) x.join();
        P.voxCdf = c->vox.data(); P.voxPdf = c->vox.data() + per;
    }
    bool notPoint = false;
    for (const WfLight &l : c->lights) {
        if (l.kind == B200RT_LIGHT_SPHERE || l.kind == B200RT_LIGHT_INFINITE) P.anyMis = 1;
        notPoint = notPoint || l.kind != B200RT_LIGHT_POINT;
    }
    c->tris.assign(tri_verts, tri_verts + 9 * (size_t)n_prims);
    if (tri_uvs) { c->uv.assign(tri_uvs, tri_uvs + 6 * (size_t)n_prims); P.primUV = c->uv.data(); }
    if (tri_normals) { c->normals.assign(tri_normals, tri_normals + 9 * (size_t)n_prims); P.primN = c->normals.data(); }
    if (prim_flags) { c->flags.assign(prim_flags, prim_flags + n_prims); P.primFlags = c->flags.data(); }
    c->isSphere.assign(n_prims, 0);
    if (prim_is_sphere) c->isSphere.assign(prim_is_sphere, prim_is_sphere + n_prims);
    if (d->materials) {
        const uint32_t stride = d->material_stride ? d->material_stride : 8;
        c->mats.resize((size_t)d->n_materials * kMatFloats);
        bool spec = false;
        for (uint32_t k = 0; k < d->n_materials; ++k)
            if (!material_lobes(d->materials + (size_t)stride * k, stride, k, c->mats.data() + (size_t)k * kMatFloats, &spec, d->textures ? d->n_textures : 0u)) { delete c; return nullptr; }
        P.has_specular = spec;
        c->primMat.assign(d->prim_material, d->prim_material + n_prims);
        P.mats = c->mats.data(); P.primMat = c->primMat.data();
        if (d->textures && d->n_textures) {   // the caller's pyramids stay alive as long as the context
            c->tex.resize(d->n_textures);
            for (uint32_t k = 0; k < d->n_textures; ++k) {
                const b200rt_texture &t = d->textures[k];
                WfTexture &T = c->tex[k];
                T.texels = t.texels; T.nLevels = t.n_levels; T.w0 = t.width; T.h0 = t.height;
                T.su = t.su; T.sv = t.sv; T.du = t.du; T.dv = t.dv;
                static float lut[128];
                b200rt_ewa_weights(lut);
                T.ewa = t.filter == B200RT_TEXFILTER_EWA ? 1u : 0u; T.maxAniso = t.max_anisotropy > 0.f ? t.max_anisotropy : 8.f; T.lut = lut;
                size_t off = 0;
                for (uint32_t l = 0; l < 16; ++l) {
                    T.levelOffset[l] = (uint32_t)off;
                    if (l < t.n_levels) off += (size_t)std::max(1u, t.width >> l) * std::max(1u, t.height >> l);
                }
            }
            P.tex = c->tex.data(); P.nTex = d->n_textures;
        }
    }
    P.smp.kind = d->sampler; P.smp.seed = d->seed;
    if (d->sampler == B200RT_SAMPLER_HALTON) {
        const HaltonTables &T = halton_tables();
        halton_setup((int)P.tl.xres, (int)P.tl.yres, &P.smp);
        P.smp.perms = T.perms.data(); P.smp.dimInfo = T.dims.data(); P.smp.nDims = (uint32_t)T.dims.size();
    } else if (d->sampler == B200RT_SAMPLER_SOBOL) {
        sobol_setup((int)P.tl.xres, (int)P.tl.yres, (int)P.tl.x0, (int)P.tl.y0, &P.smp);
        c->sobol32.assign(d->sobol_matrices32, d->sobol_matrices32 + (size_t)kSobolDims * kSobolMatrixSize);
        c->vdc.resize(2 * kSobolMatrixSize);
        if (P.smp.log2Res > 0) {
            const size_t row = (size_t)(P.smp.log2Res - 1) * kSobolMatrixSize;
            std::memcpy(c->vdc.data(), d->sobol_vdc + row, kSobolMatrixSize * 8);
            std::memcpy(c->vdc.data() + kSobolMatrixSize, d->sobol_vdc_inv + row, kSobolMatrixSize * 8);
        }
        P.smp.sobol32 = c->sobol32.data(); P.smp.vdc = c->vdc.data(); P.smp.vdcInv = c->vdc.data() + kSobolMatrixSize;
    }
    c->full = notPoint || P.nLights > 1 || P.mats || P.primN || P.primUV || P.primFlags;
    return c;
}
extern "C" void hs_destroy(hs_ctx *c) { delete c; }
extern "C" uint64_t hs_n_paths(const hs_ctx *c) { return c->P.tl.n_paths_rank; }

// GlobalSampler::SampleDimension(GetIndexForSample(s), dim) at full-frame pixel (px, py)
extern "C" void hs_sample_values(const hs_ctx *c, uint64_t n, const int32_t *px, const int32_t *py, const uint32_t *s,
                                 const uint32_t *dim, float *out) {
    for (uint64_t i = 0; i < n; ++i) {
        const uint64_t idx = sampler_index(c->P.smp, px[i], py[i], s[i]);
        out[i] = sampler_dim(c->P.smp, idx, dim[i], px[i], py[i], (uint32_t)py[i] * c->P.cam.xres + (uint32_t)px[i], s[i]);
    }
}

// k_camera on the host: path slots [start, start + nSlots)
extern "C" uint32_t hs_camera(const hs_ctx *c, uint64_t start, uint32_t nSlots, float *rays, uint32_t *qPath, float *beta, float *L,
                              uint32_t *pdim) {
    const WfParams &P = c->P;
    uint32_t count = 0;
    for (uint32_t i = 0; i < nSlots; ++i) {
        uint32_t x, y, s;
        const bool active = path_pixel(P.tl, start + i, x, y, s);
        beta[4 * i] = beta[4 * i + 1] = beta[4 * i + 2] = 1.f; beta[4 * i + 3] = 0.f;
        L[4 * i] = L[4 * i + 1] = L[4 * i + 2] = 0.f; L[4 * i + 3] = active ? 1.f : 0.f;
        pdim[i] = kCameraDims;
        if (!active) continue;
        V3 o, d;
        camera_ray(P, x, y, s, o, d);
        float *r = rays + 8 * (size_t)count;
        r[0] = o.x; r[1] = o.y; r[2] = o.z; r[3] = wf_inf(); r[4] = d.x; r[5] = d.y; r[6] = d.z; r[7] = 0.f;
        qPath[count++] = i;
    }
    return count;
}

// k_shade on the host.  hits: n x (prim u32, t, b0, b1, b2).  Queue outputs are appended in order; counts returned in nOut[3]
// = (next, shadow, mis).
extern "C" void hs_shade(const hs_ctx *c, uint32_t depth, uint64_t start, uint32_t n, const float *rays, const uint32_t *qPath,
                         const uint32_t *hits, float *beta, uint32_t *pdim, float *L, float *nextRays, uint32_t *nextPath,
                         float *shRays, uint32_t *shPath, float *shContrib, float *misRays, uint32_t *misPath, float *misContrib,
                         uint32_t *nOut) {
    const WfParams &P = c->P;
    uint32_t nNext = 0, nSh = 0, nMis = 0;
    for (uint32_t i = 0; i < n; ++i) {
        const uint32_t path = qPath[i];
        const float *r = rays + 8 * (size_t)i;
        const uint32_t *h = hits + 5 * (size_t)i;
        const uint32_t prim = h[0];
        V3 p0 = v3(0, 0, 0), p1 = p0, p2 = p0;
        bool sph = false;
        if (prim != B200RT_MISS) {
            const float *t = c->tris.data() + 9 * (size_t)prim;
            p0 = v3(t[0], t[1], t[2]); p1 = v3(t[3], t[4], t[5]); p2 = v3(t[6], t[7], t[8]);
            sph = c->isSphere[prim] != 0;
        }
        uint32_t x, y, s;
        path_pixel(P.tl, start + path, x, y, s);
        ShadeOut o;
        const V3 bt = v3(beta[4 * path], beta[4 * path + 1], beta[4 * path + 2]);
        if (c->full) shade_vertex<true>(P, depth, x, y, s, path_sample_index(P, x, y, s), pdim[path], v3(r[0], r[1], r[2]), v3(r[4], r[5], r[6]), prim, u2f(h[1]), u2f(h[2]), u2f(h[3]), u2f(h[4]), p0, p1, p2, sph, bt, beta[4 * path + 3], o);
        else shade_vertex<false>(P, depth, x, y, s, path_sample_index(P, x, y, s), pdim[path], v3(r[0], r[1], r[2]), v3(r[4], r[5], r[6]), prim, u2f(h[1]), u2f(h[2]), u2f(h[3]), u2f(h[4]), p0, p1, p2, sph, bt, beta[4 * path + 3], o);
        L[4 * path] += o.Le.x; L[4 * path + 1] += o.Le.y; L[4 * path + 2] += o.Le.z;
        if (o.wantNext) {
            beta[4 * path] = o.beta.x; beta[4 * path + 1] = o.beta.y; beta[4 * path + 2] = o.beta.z; beta[4 * path + 3] = o.betaW;
            pdim[path] = o.dim;
            float *q = nextRays + 8 * (size_t)nNext;
            q[0] = o.no.x; q[1] = o.no.y; q[2] = o.no.z; q[3] = wf_inf(); q[4] = o.nd.x; q[5] = o.nd.y; q[6] = o.nd.z; q[7] = 0.f;
            nextPath[nNext++] = path;
        }
        if (o.wantShadow) {
            float *q = shRays + 8 * (size_t)nSh;
            q[0] = o.so.x; q[1] = o.so.y; q[2] = o.so.z; q[3] = 1.f - kShadowEps; q[4] = o.sd.x; q[5] = o.sd.y; q[6] = o.sd.z; q[7] = 0.f;
            shContrib[4 * (size_t)nSh] = o.contrib.x; shContrib[4 * (size_t)nSh + 1] = o.contrib.y; shContrib[4 * (size_t)nSh + 2] = o.contrib.z; shContrib[4 * (size_t)nSh + 3] = 0.f;
            shPath[nSh++] = path;
        }
        if (o.wantMis) {
            float *q = misRays + 8 * (size_t)nMis;
            q[0] = o.mo.x; q[1] = o.mo.y; q[2] = o.mo.z; q[3] = wf_inf(); q[4] = o.md.x; q[5] = o.md.y; q[6] = o.md.z; q[7] = 0.f;
            misContrib[4 * (size_t)nMis] = o.mcontrib.x; misContrib[4 * (size_t)nMis + 1] = o.mcontrib.y; misContrib[4 * (size_t)nMis + 2] = o.mcontrib.z; std::memcpy(misContrib + 4 * (size_t)nMis + 3, &o.misPrim, 4);
            misPath[nMis++] = path;
        }
    }
    nOut[0] = nNext; nOut[1] = nSh; nOut[2] = nMis;
}

// sin_ref / cos_ref (wf_core.cuh) against this machine's libm sinf / cosf over the float bit patterns [lo, hi) with the given
// stride, both signs; returns the number of mismatches.
#include <cmath>
extern "C" uint64_t hs_sincos_check(uint32_t lo, uint32_t hi, uint32_t stride) {
    uint64_t bad = 0;
    for (uint32_t u = lo; u < hi; u += stride)
        for (int sg = 0; sg < 2; ++sg) {
            const float x = u2f(u | (sg ? 0x80000000u : 0u));
            const float a = sinf(x), b = sin_ref(x), c = cosf(x), d = cos_ref(x);
            bad += (f2u(a) != f2u(b)) + (f2u(c) != f2u(d));
        }
    return bad;
}

// log_ref (wf_texture.cuh) against this machine's libm logf over the positive float bit patterns [lo, hi) with the given stride
extern "C" uint64_t hs_log_check(uint32_t lo, uint32_t hi, uint32_t stride) {
    uint64_t bad = 0;
    for (uint32_t u = lo; u < hi; u += stride) {
        const float x = u2f(u);
        bad += f2u(logf(x)) != f2u(log_ref(x));
    }
    return bad;
}
// MIPMap::Lookup of one texture (tex_lookup) for a list of (s, t, width) triples: 3 floats out per lookup
extern "C" void hs_tex_lookup(const b200rt_texture *t, const float *stw, uint32_t n, float *out) {
    WfTexture T;
    T.texels = t->texels; T.nLevels = t->n_levels; T.w0 = t->width; T.h0 = t->height; T.su = t->su; T.sv = t->sv; T.du = t->du; T.dv = t->dv;
    size_t off = 0;
    for (uint32_t l = 0; l < 16; ++l) { T.levelOffset[l] = (uint32_t)off; if (l < t->n_levels) off += (size_t)std::max(1u, t->width >> l) * std::max(1u, t->height >> l); }
    for (uint32_t i = 0; i < n; ++i) {
        const V3 v = tex_lookup(T, stw[3 * i], stw[3 * i + 1], stw[3 * i + 2]);
        out[3 * i] = v.x; out[3 * i + 1] = v.y; out[3 * i + 2] = v.z;
    }
}
// MIPMap::Lookup(st, dst0, dst1) of one texture (tex_evaluate's two filters with an identity mapping) for lookups of 6 floats:
// s, t, dst0.x, dst0.y, dst1.x, dst1.y
extern "C" void hs_tex_lookup_diff(const b200rt_texture *t, const float *q, uint32_t n, float *out) {
    WfTexture T;
    static float lut[128];
    b200rt_ewa_weights(lut);
    T.texels = t->texels; T.nLevels = t->n_levels; T.w0 = t->width; T.h0 = t->height; T.su = 1.f; T.sv = 1.f; T.du = 0.f; T.dv = 0.f;
    T.ewa = t->filter == B200RT_TEXFILTER_EWA ? 1u : 0u; T.maxAniso = t->max_anisotropy > 0.f ? t->max_anisotropy : 8.f; T.lut = lut;
    size_t off = 0;
    for (uint32_t l = 0; l < 16; ++l) { T.levelOffset[l] = (uint32_t)off; if (l < t->n_levels) off += (size_t)std::max(1u, t->width >> l) * std::max(1u, t->height >> l); }
    for (uint32_t i = 0; i < n; ++i) {
        const float *r = q + 6 * i;
        V3 v;
        if (T.ewa) v = tex_lookup_ewa(T, r[0], r[1], r[2], r[3], r[4], r[5]);
        else v = tex_lookup(T, r[0], r[1], 2 * smax(smax(fabsf(r[2]), fabsf(r[3])), smax(fabsf(r[4]), fabsf(r[5]))));
        out[3 * i] = v.x; out[3 * i + 1] = v.y; out[3 * i + 2] = v.z;
    }
}
