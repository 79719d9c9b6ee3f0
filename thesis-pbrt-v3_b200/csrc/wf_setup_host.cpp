// See wf_setup_host.h.
#include <algorithm>
#include <cmath>
#include <cstring>
#include "wf_setup_host.h"

namespace b200rt {

int tiling_of(const b200rt_render_desc *d, WfTiling *t) {
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

namespace {
// TrowbridgeReitzDistribution::RoughnessToAlpha (microfacet.h:123-128)
float roughness_to_alpha(float roughness) {
    roughness = std::max(roughness, 1e-3f);
    const float x = std::log(roughness);
    return 1.62142f + 0.819955f * x + 0.1734f * x * x + 0.0171201f * x * x * x + 0.000640711f * x * x * x * x;
}
float clamp01inf(float v) { return v < 0.f ? 0.f : v; }   // Spectrum::Clamp(0, Infinity)

}  // namespace

bool material_lobes(const float *rec, uint32_t stride, uint32_t index, float *out, bool *hasSpecular, uint32_t nTextures) {
    std::memset(out, 0, sizeof(float) * kMatFloats);
    const float kf = rec[0];
    if (!(kf >= 0.f && kf < 64.f) || kf != (float)(uint32_t)kf) { set_error(B200RT_EUNSUPPORTED, "material %u: bad kind", index); return false; }
    const uint32_t kind = (uint32_t)kf & 15u;
    const bool remap = !((uint32_t)kf & B200RT_MAT_NO_REMAP);
    uint32_t n = 0;
    float eta = 1.f;
    auto add = [&](uint32_t lk, const float *c0, const float *c1, float a, float e0, float e1) {
        float *l = out + 4 + n * kLobeFloats;
        std::memcpy(l, &lk, 4);
        for (int c = 0; c < 3; ++c) { l[1 + c] = c0 ? c0[c] : 0.f; l[4 + c] = c1 ? c1[c] : 0.f; }
        l[7] = a; l[8] = e0; l[9] = e1;
        ++n;
    };
    auto black = [](const float *c) { return c[0] == 0.f && c[1] == 0.f && c[2] == 0.f; };
    float c0[3] = {clamp01inf(rec[1]), clamp01inf(rec[2]), clamp01inf(rec[3])};
    float c1[3] = {clamp01inf(rec[4]), clamp01inf(rec[5]), clamp01inf(rec[6])};
    const float rough = rec[7];
    switch (kind) {
    case B200RT_MAT_MATTE: {   // matte.cpp:44-62
        const float sig = std::min(std::max(rec[4], 0.f), 90.f);
        if (!black(c0)) {
            if (sig == 0) add(LOBE_LAMBERT, c0, nullptr, 0, 0, 0);
            else {   // OrenNayar's constructor (reflection.h:414-420)
                const float sr = (3.14159265358979323846f / 180.f) * sig, s2 = sr * sr;
                add(LOBE_OREN, c0, nullptr, 0, 1.f - (s2 / (2.f * (s2 + 0.33f))), 0.45f * s2 / (s2 + 0.09f));
            }
        }
        break;
    }
    case B200RT_MAT_PLASTIC:   // plastic.cpp:44-70
        if (!black(c0)) add(LOBE_LAMBERT, c0, nullptr, 0, 0, 0);
        if (!black(c1)) {
            const float a = remap ? roughness_to_alpha(rough) : rough;
            if (!(a > 0.f)) { set_error(B200RT_EINVAL, "material %u: roughness must be > 0", index); return false; }
            add(LOBE_MICRO_DIEL, c1, nullptr, a, 1.5f, 1.f);
        }
        break;
    case B200RT_MAT_METAL: {   // metal.cpp:60-82: eta = rec[1..3], k = rec[4..6] (not clamped)
        const float a = remap ? roughness_to_alpha(rough) : rough;
        if (!(a > 0.f)) { set_error(B200RT_EINVAL, "material %u: roughness must be > 0 (perfectly smooth microfacet lobes are not supported)", index); return false; }
        add(LOBE_MICRO_COND, rec + 1, rec + 4, a, 0, 0);
        break;
    }
    case B200RT_MAT_SUBSTRATE:   // substrate.cpp:45-66
        if (!black(c0) || !black(c1)) {
            const float a = remap ? roughness_to_alpha(rough) : rough;
            if (!(a > 0.f)) { set_error(B200RT_EINVAL, "material %u: roughness must be > 0 (perfectly smooth microfacet lobes are not supported)", index); return false; }
            add(LOBE_BLEND, c0, c1, a, 0, 0);
        }
        break;
    case B200RT_MAT_MIRROR:   // mirror.cpp:44-56
        if (!black(c0)) add(LOBE_SPEC_REFL_NOOP, c0, nullptr, 0, 0, 0);
        *hasSpecular = true;
        break;
    case B200RT_MAT_GLASS:   // glass.cpp:44-63, roughness 0, allowMultipleLobes (the path integrator passes true)
        if (!(rough > 0.f)) { set_error(B200RT_EINVAL, "material %u: glass needs its \"index\" of refraction in the roughness slot", index); return false; }
        eta = rough;
        if (!(black(c0) && black(c1))) add(LOBE_FRESNEL_SPEC, c0, c1, 0, 1.f, eta);
        *hasSpecular = true;
        break;
    case B200RT_MAT_UBER: {   // uber.cpp:45-100
        if (stride < 20) { set_error(B200RT_EINVAL, "material %u: \"uber\" needs 20-float records (material_stride)", index); return false; }
        const float e = rec[17];
        float op[3], t[3], kd[3], ks[3], kr[3], kt[3];
        bool tBlack = true;
        for (int c = 0; c < 3; ++c) {
            op[c] = clamp01inf(rec[14 + c]);
            t[c] = clamp01inf(-op[c] + 1.f);
            tBlack = tBlack && t[c] == 0.f;
            kd[c] = op[c] * c0[c]; ks[c] = op[c] * c1[c];
            kr[c] = op[c] * clamp01inf(rec[8 + c]); kt[c] = op[c] * clamp01inf(rec[11 + c]);
        }
        if (!tBlack) { eta = 1.f; add(LOBE_SPEC_TRANS, t, nullptr, 0, 1.f, 1.f); *hasSpecular = true; }
        else eta = e;
        if (!black(kd)) add(LOBE_LAMBERT, kd, nullptr, 0, 0, 0);
        if (!black(ks)) {
            const float a = remap ? roughness_to_alpha(rough) : rough;
            if (!(a > 0.f)) { set_error(B200RT_EINVAL, "material %u: roughness must be > 0", index); return false; }
            add(LOBE_MICRO_DIEL, ks, nullptr, a, 1.f, e);
        }
        if (!black(kr)) { add(LOBE_SPEC_REFL_DIEL, kr, nullptr, 0, 1.f, e); *hasSpecular = true; }
        if (!black(kt)) { add(LOBE_SPEC_TRANS, kt, nullptr, 0, 1.f, e); *hasSpecular = true; }
        break;
    }
    default:
        set_error(B200RT_EUNSUPPORTED, "material %u: only matte, plastic, mirror, glass, metal, substrate and uber are supported", index);
        return false;
    }
    std::memcpy(out, &n, 4);
    out[1] = eta;
    // ABI v9: image textures on Kd (slot 18) and on uber's opacity (slot 19), texture index + 1.  The lobes of such a material
    // depend on the hit, so the record switches to the TEXTURED layout wf_shade.cuh's textured_lobes reads (the constant
    // record above has served as the validation of the kind, the roughness and the constants).
    if (stride >= 20 && (rec[18] != 0.f || rec[19] != 0.f)) {
        const float kdT = rec[18], opT = rec[19];
        if (!(kdT >= 0.f && opT >= 0.f && kdT == (float)(uint32_t)kdT && opT == (float)(uint32_t)opT && kdT <= (float)nTextures && opT <= (float)nTextures)) {
            set_error(B200RT_EINVAL, "material %u: texture indices (%g, %g) outside the %u textures of the description", index, kdT, opT, nTextures);
            return false;
        }
        if (!(kind == B200RT_MAT_MATTE || kind == B200RT_MAT_PLASTIC || kind == B200RT_MAT_SUBSTRATE || kind == B200RT_MAT_UBER) || (opT != 0.f && kind != B200RT_MAT_UBER)) {
            set_error(B200RT_EUNSUPPORTED, "material %u: image textures are supported on the Kd of matte / plastic / substrate / uber and on uber's opacity", index);
            return false;
        }
        const float alpha = (kind == B200RT_MAT_MATTE) ? 0.f : (remap ? roughness_to_alpha(rough) : rough);
        if (kind != B200RT_MAT_MATTE && !(alpha > 0.f)) { set_error(B200RT_EINVAL, "material %u: roughness must be > 0", index); return false; }
        std::memset(out, 0, sizeof(float) * kMatFloats);
        const uint32_t kdI = (uint32_t)kdT, opI = (uint32_t)opT, k32 = kind;
        std::memcpy(out + 2, &kdI, 4); std::memcpy(out + 3, &opI, 4);
        std::memcpy(out + 4, &k32, 4);
        for (int c = 0; c < 3; ++c) { out[5 + c] = c0[c]; out[8 + c] = c1[c]; }
        out[11] = alpha;
        if (kind == B200RT_MAT_UBER) {
            for (int c = 0; c < 3; ++c) { out[12 + c] = clamp01inf(rec[8 + c]); out[15 + c] = clamp01inf(rec[11 + c]); out[18 + c] = clamp01inf(rec[14 + c]); }
            out[21] = rec[17];
            *hasSpecular = true;   // 1 - opacity may be non-black at some hit
        } else {
            out[18] = out[19] = out[20] = 1.f;
        }
        if (kind == B200RT_MAT_MATTE) {
            const float sig = std::min(std::max(rec[4], 0.f), 90.f);
            out[8] = out[9] = out[10] = 0.f;   // slot 4 of a matte record is sigma, not Ks
            if (sig != 0) {
                const float sr = (3.14159265358979323846f / 180.f) * sig, s2 = sr * sr;
                out[22] = 1.f - (s2 / (2.f * (s2 + 0.33f))); out[23] = 0.45f * s2 / (s2 + 0.09f);
            }
        }
    }
    return true;
}


// One light record, validated (the kinds and their fields: include/b200rt_wavefront.h).
static int light_from(uint32_t kind, uint32_t prim, const float *pos, const float *I, float radius, const float *l2w, uint32_t n_prims, WfLight *Lt) {
    std::memset(Lt, 0, sizeof(*Lt));
    std::memcpy(Lt->pos, pos, 12); std::memcpy(Lt->I, I, 12);
    Lt->kind = kind; Lt->prim = prim; Lt->radius = radius;
    Lt->cdfHi = 1.f; Lt->pickPdf = 1.f;
    if (kind > B200RT_LIGHT_INFINITE) return set_error(B200RT_EINVAL, "unknown light_kind %u", kind);
    if (kind == B200RT_LIGHT_INFINITE) {
        bool zero = true;
        for (int k = 0; k < 9; ++k) zero = zero && l2w[k] == 0.f;
        for (int k = 0; k < 9; ++k) Lt->l2w[k] = zero ? (k % 4 == 0 ? 1.f : 0.f) : l2w[k];
        if (!(radius > 0.f)) return set_error(B200RT_EINVAL, "infinite light: light_radius must be the scene's bounding-sphere radius");
    }
    if (kind == B200RT_LIGHT_DISTANT) {
        const float l2 = pos[0] * pos[0] + pos[1] * pos[1] + pos[2] * pos[2];
        if (!(radius > 0.f) || !(std::fabs(l2 - 1.f) < 1e-3f))
            return set_error(B200RT_EINVAL, "distant light: light_pos must be the unit direction to the light and light_radius the scene's bounding-sphere radius");
    }
    if (kind == B200RT_LIGHT_SPHERE && (!(radius > 0.f) || prim >= n_prims))
        return set_error(B200RT_EINVAL, "sphere light: radius must be > 0 and light_prim a primitive of the scene");
    return B200RT_OK;
}

int light_of(const b200rt_render_desc *d, uint32_t n_prims, WfLight *Lt) {
    return light_from(d->light_kind, d->light_prim, d->light_pos, d->light_I, d->light_radius, d->light_to_world, n_prims, Lt);
}

// Scene::lights and the distribution UniformSampleOneLight draws from: Distribution1D(func, n) (sampling.h:57-69) over the
// records' weights, restated with its float operations so that the device's choice is SampleDiscrete's for every u.
int lights_of(const b200rt_render_desc *d, uint32_t n_prims, std::vector<WfLight> &out) {
    out.clear();
    if (d->n_lights == 0) {
        out.resize(1);
        return light_of(d, n_prims, &out[0]);
    }
    if (!d->lights) return set_error(B200RT_EINVAL, "n_lights = %u without a lights array", d->n_lights);
    if (d->n_lights > B200RT_MAX_LIGHTS) return set_error(B200RT_EUNSUPPORTED, "%u lights: at most %u", d->n_lights, B200RT_MAX_LIGHTS);
    const int n = (int)d->n_lights;
    out.resize((size_t)n);
    std::vector<float> func((size_t)n), cdf((size_t)n + 1);
    for (int k = 0; k < n; ++k) {
        const b200rt_light &l = d->lights[k];
        const int rc = light_from(l.kind, l.prim, l.pos, l.I, l.radius, l.to_world, n_prims, &out[(size_t)k]);
        if (rc) return rc;
        if (!(l.weight >= 0.f)) return set_error(B200RT_EINVAL, "light %d: weight must be >= 0", k);
        func[(size_t)k] = l.weight;
    }
    cdf[0] = 0;
    for (int i = 1; i < n + 1; ++i) cdf[(size_t)i] = cdf[(size_t)i - 1] + func[(size_t)i - 1] / n;
    const float funcInt = cdf[(size_t)n];
    if (funcInt == 0) {
        for (int i = 1; i < n + 1; ++i) cdf[(size_t)i] = float(i) / float(n);
    } else {
        for (int i = 1; i < n + 1; ++i) cdf[(size_t)i] /= funcInt;
    }
    for (int k = 0; k < n; ++k) {
        out[(size_t)k].cdfHi = cdf[(size_t)k + 1];
        out[(size_t)k].pickPdf = (funcInt > 0) ? func[(size_t)k] / (funcInt * n) : 0.f;
    }
    return B200RT_OK;
}

void spatial_halton_table(float *out) {
    static const uint32_t bases[5] = {2, 3, 5, 7, 11};
    for (uint32_t i = 0; i < 128; ++i) {
        out[5 * i] = (float)((double)reverse_bits64(i) * 0x1p-64);   // RadicalInverse(0, a) (lowdiscrepancy.cpp:430-435)
        for (int b = 1; b < 5; ++b) out[5 * i + b] = radical_inverse(bases[b], i);
    }
}

}  // namespace b200rt
