// BSDFs of the wavefront shade stage, in the local shading frame (z = the normal on wo's side, so wo.z >= 0).
//
//   Material "matte"   (materials/matte.cpp:44-62)            : LambertianReflection(Kd), or OrenNayar(Kd, sigma) for sigma > 0
//   Material "plastic" (materials/plastic.cpp:44-70)           : LambertianReflection(Kd) + MicrofacetReflection(Ks,
//        TrowbridgeReitzDistribution(alpha, alpha) with alpha = RoughnessToAlpha(roughness), FresnelDielectric(1.5, 1))
//   Material "metal"   (materials/metal.cpp:60-82)              : MicrofacetReflection(1, TrowbridgeReitz(alpha, alpha),
//                                                                FresnelConductor(1, eta, k))
//   Material "substrate" (materials/substrate.cpp:45-66)        : FresnelBlend(Kd, Ks, TrowbridgeReitz(alpha, alpha))
//   Material "mirror"  (materials/mirror.cpp:44-56)            : SpecularReflection(Kr, FresnelNoOp)
//   Material "glass"   (materials/glass.cpp:44-63, roughness 0; PathIntegrator asks for allowMultipleLobes)
//                                                              : FresnelSpecular(Kr, Kt, 1, index, Radiance)
// and the BSDF container's rules for several lobes (reflection.cpp BSDF::f / Pdf / Sample_f: lobe picked uniformly by
// u[0], pdf = mean of the lobes' pdfs, f = sum of the lobes).  Isotropic lobes only, so any orthonormal frame around
// the normal is equivalent to the reference's dpdu frame.
#pragma once
#include <cuda_runtime.h>

namespace b200rt {

struct WfMat {
    float kd[3], ks[3];
    float alpha;     // TrowbridgeReitz alpha (already remapped from "roughness" on the host); matte: "sigma" in degrees
    bool diffuse;    // !Kd.IsBlack()
    bool glossy;     // !Ks.IsBlack()
    // perfectly specular materials: kd = Kr, ks = Kt, alpha = "index"; `flip` = reverseOrientation ^ transformSwapsHandedness
    // of the shape (triangle.cpp:421-424), which decides on which side of a triangle the glass is
    uint32_t kind;   // B200RT_MAT_*
    bool flip;
};

constexpr float kBsdfPi = 3.14159265358979323846f;
constexpr float kBsdfInvPi = 0.31830988618379067154f;
constexpr float kOneMinusEpsilon = 0.99999994f;

// FrDielectric (reflection.cpp:47-68)
__device__ __forceinline__ float fr_dielectric(float cosThetaI, float etaI, float etaT) {
    cosThetaI = fminf(fmaxf(cosThetaI, -1.f), 1.f);
    if (!(cosThetaI > 0.f)) { const float t = etaI; etaI = etaT; etaT = t; cosThetaI = fabsf(cosThetaI); }
    const float sinThetaI = sqrtf(fmaxf(0.f, 1.f - cosThetaI * cosThetaI));
    const float sinThetaT = etaI / etaT * sinThetaI;
    if (sinThetaT >= 1.f) return 1.f;
    const float cosThetaT = sqrtf(fmaxf(0.f, 1.f - sinThetaT * sinThetaT));
    const float Rparl = ((etaT * cosThetaI) - (etaI * cosThetaT)) / ((etaT * cosThetaI) + (etaI * cosThetaT));
    const float Rperp = ((etaI * cosThetaI) - (etaT * cosThetaT)) / ((etaI * cosThetaI) + (etaT * cosThetaT));
    return (Rparl * Rparl + Rperp * Rperp) / 2.f;
}

// TrowbridgeReitzDistribution, alphax == alphay == a (microfacet.cpp:155-163, 176-184; microfacet.h:54-60)
__device__ __forceinline__ float tr_D(const float3 wh, float a) {
    const float cos2 = wh.z * wh.z, sin2 = fmaxf(0.f, 1.f - cos2);
    const float tan2 = sin2 / cos2;
    if (isinf(tan2)) return 0.f;
    const float e = tan2 / (a * a);   // (Cos2Phi + Sin2Phi) / a^2 * tan2Theta
    return 1.f / (kBsdfPi * a * a * cos2 * cos2 * (1.f + e) * (1.f + e));
}
__device__ __forceinline__ float tr_lambda(const float3 w, float a) {
    const float sin2 = fmaxf(0.f, 1.f - w.z * w.z);
    const float absTan = fabsf(sqrtf(sin2) / w.z);
    if (isinf(absTan)) return 0.f;
    const float at = (a * absTan) * (a * absTan);
    return (-1.f + sqrtf(1.f + at)) / 2.f;
}
// MicrofacetDistribution::Pdf with sampleVisibleArea (microfacet.cpp:338-344)
__device__ __forceinline__ float tr_pdf(const float3 wo, const float3 wh, float a) {
    const float g1 = 1.f / (1.f + tr_lambda(wo, a));
    return tr_D(wh, a) * g1 * fabsf(wo.x * wh.x + wo.y * wh.y + wo.z * wh.z) / fabsf(wo.z);
}
// TrowbridgeReitzSample11 + TrowbridgeReitzSample (microfacet.cpp:238-305), visible-normal sampling
__device__ __forceinline__ float3 tr_sample_wh(const float3 wo, float U1, float U2, float a) {
    float3 ws = make_float3(a * wo.x, a * wo.y, wo.z);
    const float il = 1.f / sqrtf(ws.x * ws.x + ws.y * ws.y + ws.z * ws.z);
    ws.x *= il; ws.y *= il; ws.z *= il;
    float slope_x, slope_y;
    const float cosTheta = ws.z;
    if (cosTheta > .9999f) {
        const float r = sqrtf(U1 / (1.f - U1)), phi = 6.28318530718f * U2;
        slope_x = r * cosf(phi); slope_y = r * sinf(phi);
    } else {
        const float sinTheta = sqrtf(fmaxf(0.f, 1.f - cosTheta * cosTheta));
        const float tanTheta = sinTheta / cosTheta;
        const float aa = 1.f / tanTheta;
        const float G1 = 2.f / (1.f + sqrtf(1.f + 1.f / (aa * aa)));
        const float A = 2.f * U1 / G1 - 1.f;
        float tmp = 1.f / (A * A - 1.f);
        if (tmp > 1e10f) tmp = 1e10f;
        const float B = tanTheta;
        const float D = sqrtf(fmaxf(B * B * tmp * tmp - (A * A - B * B) * tmp, 0.f));
        const float s1 = B * tmp - D, s2 = B * tmp + D;
        slope_x = (A < 0.f || s2 > 1.f / tanTheta) ? s1 : s2;
        float S;
        if (U2 > 0.5f) { S = 1.f; U2 = 2.f * (U2 - .5f); }
        else { S = -1.f; U2 = 2.f * (.5f - U2); }
        const float z = (U2 * (U2 * (U2 * 0.27385f - 0.73369f) + 0.46341f)) /
                        (U2 * (U2 * (U2 * 0.093073f + 0.309420f) - 1.000000f) + 0.597999f);
        slope_y = S * z * sqrtf(1.f + slope_x * slope_x);
    }
    const float sinT = sqrtf(fmaxf(0.f, 1.f - ws.z * ws.z));
    const float cosPhi = sinT == 0.f ? 1.f : fminf(fmaxf(ws.x / sinT, -1.f), 1.f);
    const float sinPhi = sinT == 0.f ? 0.f : fminf(fmaxf(ws.y / sinT, -1.f), 1.f);
    const float t = cosPhi * slope_x - sinPhi * slope_y;
    slope_y = sinPhi * slope_x + cosPhi * slope_y;
    slope_x = t;
    slope_x *= a; slope_y *= a;
    float3 wh = make_float3(-slope_x, -slope_y, 1.f);
    const float ih = 1.f / sqrtf(wh.x * wh.x + wh.y * wh.y + 1.f);
    wh.x *= ih; wh.y *= ih; wh.z *= ih;
    return wh;
}

// MicrofacetReflection::f (reflection.cpp:226-236), scalar part (times Ks)
__device__ __forceinline__ float glossy_f(const float3 wo, const float3 wi, float a) {
    const float cosO = fabsf(wo.z), cosI = fabsf(wi.z);
    float3 wh = make_float3(wi.x + wo.x, wi.y + wo.y, wi.z + wo.z);
    if (cosI == 0.f || cosO == 0.f) return 0.f;
    if (wh.x == 0.f && wh.y == 0.f && wh.z == 0.f) return 0.f;
    const float il = 1.f / sqrtf(wh.x * wh.x + wh.y * wh.y + wh.z * wh.z);
    wh.x *= il; wh.y *= il; wh.z *= il;
    const float F = fr_dielectric(wi.x * wh.x + wi.y * wh.y + wi.z * wh.z, 1.5f, 1.f);   // FresnelDielectric(1.5f, 1.f), plastic.cpp:59
    const float G = 1.f / (1.f + tr_lambda(wo, a) + tr_lambda(wi, a));
    return tr_D(wh, a) * G * F / (4.f * cosI * cosO);
}
__device__ __forceinline__ float glossy_pdf(const float3 wo, const float3 wi, float a) {
    if (!(wo.z * wi.z > 0.f)) return 0.f;
    float3 wh = make_float3(wo.x + wi.x, wo.y + wi.y, wo.z + wi.z);
    const float il = 1.f / sqrtf(wh.x * wh.x + wh.y * wh.y + wh.z * wh.z);
    wh.x *= il; wh.y *= il; wh.z *= il;
    return tr_pdf(wo, wh, a) / (4.f * (wo.x * wh.x + wo.y * wh.y + wo.z * wh.z));
}

// FrConductor (reflection.cpp:71-94) for one channel, etaI = 1 (FresnelConductor::Evaluate passes |cos|)
__device__ __forceinline__ float fr_conductor(float cosThetaI, float eta, float k) {
    cosThetaI = fminf(fmaxf(cosThetaI, -1.f), 1.f);
    const float c2 = cosThetaI * cosThetaI, s2 = 1.f - c2;
    const float eta2 = eta * eta, etak2 = k * k;
    const float t0 = eta2 - etak2 - s2;
    const float a2plusb2 = sqrtf(t0 * t0 + 4.f * eta2 * etak2);
    const float t1 = a2plusb2 + c2;
    const float a = sqrtf(0.5f * (a2plusb2 + t0));
    const float t2 = 2.f * cosThetaI * a;
    const float Rs = (t1 - t2) / (t1 + t2);
    const float t3 = c2 * a2plusb2 + s2 * s2;
    const float t4 = t2 * s2;
    const float Rp = Rs * (t3 - t4) / (t3 + t4);
    return 0.5f * (Rp + Rs);
}
// MicrofacetReflection::f (reflection.cpp:226-236) with R = 1 and FresnelConductor(1, eta = m.kd, k = m.ks)
__device__ __forceinline__ float3 metal_f(const WfMat &m, const float3 wo, const float3 wi) {
    const float cosO = fabsf(wo.z), cosI = fabsf(wi.z);
    float3 wh = make_float3(wi.x + wo.x, wi.y + wo.y, wi.z + wo.z);
    if (cosI == 0.f || cosO == 0.f) return make_float3(0.f, 0.f, 0.f);
    if (wh.x == 0.f && wh.y == 0.f && wh.z == 0.f) return make_float3(0.f, 0.f, 0.f);
    const float il = 1.f / sqrtf(wh.x * wh.x + wh.y * wh.y + wh.z * wh.z);
    wh.x *= il; wh.y *= il; wh.z *= il;
    const float c = fabsf(wi.x * wh.x + wi.y * wh.y + wi.z * wh.z);
    const float G = 1.f / (1.f + tr_lambda(wo, m.alpha) + tr_lambda(wi, m.alpha));
    const float s = tr_D(wh, m.alpha) * G / (4.f * cosI * cosO);
    return make_float3(s * fr_conductor(c, m.kd[0], m.ks[0]), s * fr_conductor(c, m.kd[1], m.ks[1]), s * fr_conductor(c, m.kd[2], m.ks[2]));
}
// FresnelBlend::f (reflection.cpp:285-298) and ::Pdf (:470-475), Rd = m.kd, Rs = m.ks
__device__ __forceinline__ float pow5f(float v) { return (v * v) * (v * v) * v; }
__device__ __forceinline__ float3 blend_f(const WfMat &m, const float3 wo, const float3 wi) {
    const float dterm = (28.f / (23.f * kBsdfPi)) * (1.f - pow5f(1.f - .5f * fabsf(wi.z))) * (1.f - pow5f(1.f - .5f * fabsf(wo.z)));
    float3 f = make_float3(m.kd[0] * (1.f - m.ks[0]) * dterm, m.kd[1] * (1.f - m.ks[1]) * dterm, m.kd[2] * (1.f - m.ks[2]) * dterm);
    float3 wh = make_float3(wi.x + wo.x, wi.y + wo.y, wi.z + wo.z);
    if (wh.x == 0.f && wh.y == 0.f && wh.z == 0.f) return make_float3(0.f, 0.f, 0.f);
    const float il = 1.f / sqrtf(wh.x * wh.x + wh.y * wh.y + wh.z * wh.z);
    wh.x *= il; wh.y *= il; wh.z *= il;
    const float d = wi.x * wh.x + wi.y * wh.y + wi.z * wh.z;
    const float s = tr_D(wh, m.alpha) / (4.f * fabsf(d) * fmaxf(fabsf(wi.z), fabsf(wo.z)));
    const float p = pow5f(1.f - d);
    f.x += s * (m.ks[0] + p * (1.f - m.ks[0])); f.y += s * (m.ks[1] + p * (1.f - m.ks[1])); f.z += s * (m.ks[2] + p * (1.f - m.ks[2]));
    return f;
}
__device__ __forceinline__ float blend_pdf(const WfMat &m, const float3 wo, const float3 wi) {
    if (!(wo.z * wi.z > 0.f)) return 0.f;
    float3 wh = make_float3(wo.x + wi.x, wo.y + wi.y, wo.z + wi.z);
    const float il = 1.f / sqrtf(wh.x * wh.x + wh.y * wh.y + wh.z * wh.z);
    wh.x *= il; wh.y *= il; wh.z *= il;
    return .5f * (fabsf(wi.z) * kBsdfInvPi + tr_pdf(wo, wh, m.alpha) / (4.f * (wo.x * wh.x + wo.y * wh.y + wo.z * wh.z)));
}

// BSDF::f and BSDF::Pdf (reflection.cpp): every lobe here is reflective, so f = 0 across the surface.
// (inlined on purpose: shade_general evaluates the BSDF at up to four places and samples it at three, which makes
// k_shade<true> 21.7 k SASS lines, and a SMALL launch shows instruction-fetch stalls in ncu (`no_instruction` 8-9 warps per
// issue, profiles/r01_ncu_full_k_shade_living_room_summary.txt).  Measured out of line (__noinline__: 10.1 k lines, 200-byte
// stack frame), same box, alternating runs: 1280x720 @ 64 spp living-room frames of 970 / 977 ms inline vs 980 / 981 ms out
// of line (profiles/r01_ab_shade_inline.log) — not adopted.)
#ifdef B200RT_BSDF_NOINLINE   // A/B build only (scripts/ab_shade_inline.sh)
#define B200RT_BSDF_FN __device__ __noinline__
#else
#define B200RT_BSDF_FN __device__ __forceinline__
#endif
B200RT_BSDF_FN void bsdf_eval(const WfMat &m, const float3 wo, const float3 wi, float3 &f, float &pdf) {
    f = make_float3(0.f, 0.f, 0.f);
    pdf = 0.f;
    if (m.kind == 4u /* B200RT_MAT_METAL */) {
        if (wo.z * wi.z > 0.f) f = metal_f(m, wo, wi);
        pdf = glossy_pdf(wo, wi, m.alpha);
        return;
    }
    if (m.kind == 5u /* B200RT_MAT_SUBSTRATE */) {
        if (!m.diffuse && !m.glossy) return;   // substrate.cpp:56: no BxDF when Kd and Ks are black
        if (wo.z * wi.z > 0.f) f = blend_f(m, wo, wi);
        pdf = blend_pdf(m, wo, wi);
        return;
    }
    const int n = (m.diffuse ? 1 : 0) + (m.glossy ? 1 : 0);
    if (n == 0) return;
    const bool same = wo.z * wi.z > 0.f;
    if (m.diffuse) {
        if (same) {
            float on = 1.f;
            if (m.kind == 0u && m.alpha > 0.f) {   // OrenNayar (reflection.h:414-420, reflection.cpp:197-219); sampled like any BxDF (cosine)
                const float sigma = (kBsdfPi / 180.f) * m.alpha, sigma2 = sigma * sigma;
                const float A = 1.f - (sigma2 / (2.f * (sigma2 + 0.33f))), B = 0.45f * sigma2 / (sigma2 + 0.09f);
                const float sinI = sqrtf(fmaxf(0.f, 1.f - wi.z * wi.z)), sinO = sqrtf(fmaxf(0.f, 1.f - wo.z * wo.z));
                float maxCos = 0.f;
                if (sinI > 1e-4f && sinO > 1e-4f) {
                    const float cpi = fminf(fmaxf(wi.x / sinI, -1.f), 1.f), spi = fminf(fmaxf(wi.y / sinI, -1.f), 1.f);
                    const float cpo = fminf(fmaxf(wo.x / sinO, -1.f), 1.f), spo = fminf(fmaxf(wo.y / sinO, -1.f), 1.f);
                    maxCos = fmaxf(0.f, cpi * cpo + spi * spo);
                }
                float sinAlpha, tanBeta;
                if (fabsf(wi.z) > fabsf(wo.z)) { sinAlpha = sinO; tanBeta = sinI / fabsf(wi.z); }
                else { sinAlpha = sinI; tanBeta = sinO / fabsf(wo.z); }
                on = A + B * maxCos * sinAlpha * tanBeta;
            }
            f.x += m.kd[0] * kBsdfInvPi * on; f.y += m.kd[1] * kBsdfInvPi * on; f.z += m.kd[2] * kBsdfInvPi * on;
            pdf += fabsf(wi.z) * kBsdfInvPi;
        }
    }
    if (m.glossy) {
        if (same) { const float g = glossy_f(wo, wi, m.alpha); f.x += m.ks[0] * g; f.y += m.ks[1] * g; f.z += m.ks[2] * g; }
        pdf += glossy_pdf(wo, wi, m.alpha);
    }
    pdf /= (float)n;
}

// BSDF::Sample_f (reflection.cpp): false = the sample carries nothing (pdf == 0 or f black).
B200RT_BSDF_FN bool bsdf_sample(const WfMat &m, const float3 wo, float u0, float u1, float3 &wi, float3 &f, float &pdf) {
    int n = (m.diffuse ? 1 : 0) + (m.glossy ? 1 : 0);
    bool pickGlossy;
    if (m.kind == 4u) { n = 1; pickGlossy = true; }   // metal: the one microfacet lobe
    else if (m.kind == 5u) {
        // FresnelBlend::Sample_f (reflection.cpp:450-468): u0 < .5 the cosine lobe, else the microfacet half vector
        if (n == 0 || wo.z == 0.f) return false;
        n = 1;
        pickGlossy = !(u0 < .5f);
        u0 = pickGlossy ? fminf(2.f * (u0 - .5f), kOneMinusEpsilon) : fminf(2.f * u0, kOneMinusEpsilon);
    } else {
        if (n == 0) return false;
        int comp = (int)floorf(u0 * n);
        if (comp > n - 1) comp = n - 1;
        u0 = fminf(u0 * n - comp, kOneMinusEpsilon);
        pickGlossy = m.glossy && (!m.diffuse || comp == 1);
    }
    if (!pickGlossy) {
        // BxDF::Sample_f: CosineSampleHemisphere via ConcentricSampleDisk (sampling.cpp)
        const float a0 = 2.f * u0 - 1.f, a1 = 2.f * u1 - 1.f;
        float dx = 0.f, dy = 0.f;
        if (a0 != 0.f || a1 != 0.f) {
            float r, th;
            if (fabsf(a0) > fabsf(a1)) { r = a0; th = (kBsdfPi / 4.f) * (a1 / a0); }
            else { r = a1; th = (kBsdfPi / 2.f) - (kBsdfPi / 4.f) * (a0 / a1); }
            dx = r * cosf(th); dy = r * sinf(th);
        }
        wi = make_float3(dx, dy, sqrtf(fmaxf(0.f, 1.f - dx * dx - dy * dy)));
        if (m.kind != 5u && wi.z * kBsdfInvPi == 0.f) return false;
    } else {
        if (wo.z == 0.f) return false;
        const float3 wh = tr_sample_wh(wo, u0, u1, m.alpha);
        const float d = wo.x * wh.x + wo.y * wh.y + wo.z * wh.z;
        wi = make_float3(-wo.x + 2.f * d * wh.x, -wo.y + 2.f * d * wh.y, -wo.z + 2.f * d * wh.z);   // Reflect(wo, wh)
        if (!(wo.z * wi.z > 0.f)) return false;
        if (m.kind != 5u && tr_pdf(wo, wh, m.alpha) / (4.f * d) == 0.f) return false;
    }
    bsdf_eval(m, wo, wi, f, pdf);
    return pdf > 0.f && (f.x > 0.f || f.y > 0.f || f.z > 0.f);
}

// The single specular lobe of "mirror" / "glass" through BSDF::Sample_f (reflection.cpp:703-770; one matching
// component, so u0 reaches the BxDF unchanged): SpecularReflection::Sample_f (reflection.cpp:136-143) with FresnelNoOp,
// or FresnelSpecular::Sample_f (reflection.cpp:477-511, TransportMode::Radiance).  wo in the frame of the TRUE shading
// normal (wo.z < 0 = seen from the back: for glass, from inside).  Returns false when the path ends (f black / pdf 0);
// else wi, w = f |cos wi| / pdf per channel (what path.cpp:151 multiplies beta by) and whether the lobe transmitted.
__device__ __forceinline__ bool specular_sample(const WfMat &m, const float3 wo, float u0, float3 &wi, float3 &w, bool &transmitted) {
    transmitted = false;
    if (wo.z == 0.f) return false;   // BSDF::Sample_f (reflection.cpp:735)
    if (m.kind == 2u /* B200RT_MAT_MIRROR */) {
        if (!(m.kd[0] > 0.f || m.kd[1] > 0.f || m.kd[2] > 0.f)) return false;   // R.IsBlack(): no BxDF at all
        wi = make_float3(-wo.x, -wo.y, wo.z);
        const float a = fabsf(wi.z);
        w = make_float3(m.kd[0] / a * a, m.kd[1] / a * a, m.kd[2] / a * a);
        return w.x > 0.f || w.y > 0.f || w.z > 0.f;
    }
    const bool rBlack = !(m.kd[0] > 0.f || m.kd[1] > 0.f || m.kd[2] > 0.f), tBlack = !(m.ks[0] > 0.f || m.ks[1] > 0.f || m.ks[2] > 0.f);
    if (rBlack && tBlack) return false;   // glass.cpp:58
    const float eta = m.alpha;
    const float F = fr_dielectric(wo.z, 1.f, eta);
    if (u0 < F) {
        wi = make_float3(-wo.x, -wo.y, wo.z);
        const float a = fabsf(wi.z);
        w = make_float3(F * m.kd[0] / a * a / F, F * m.kd[1] / a * a / F, F * m.kd[2] / a * a / F);
        return w.x > 0.f || w.y > 0.f || w.z > 0.f;
    }
    const bool entering = wo.z > 0.f;
    const float etaI = entering ? 1.f : eta, etaT = entering ? eta : 1.f;
    // Refract(wo, Faceforward(Normal3f(0, 0, 1), wo), etaI / etaT, wi) (reflection.h:96-108)
    const float nz = wo.z < 0.f ? -1.f : 1.f, e = etaI / etaT;
    const float cosThetaI = nz * wo.z;
    const float sin2ThetaI = fmaxf(0.f, 1.f - cosThetaI * cosThetaI);
    const float sin2ThetaT = e * e * sin2ThetaI;
    if (sin2ThetaT >= 1.f) return false;   // total internal reflection on the transmission branch: f = 0, the path ends
    const float cosThetaT = sqrtf(1.f - sin2ThetaT);
    wi = make_float3(e * -wo.x, e * -wo.y, e * -wo.z + (e * cosThetaI - cosThetaT) * nz);
    const float k = (1.f - F) * ((etaI * etaI) / (etaT * etaT)), a = fabsf(wi.z), pdf = 1.f - F;
    w = make_float3(m.ks[0] * k / a * a / pdf, m.ks[1] * k / a * a / pdf, m.ks[2] * k / a * a / pdf);
    transmitted = true;
    return pdf > 0.f && (w.x > 0.f || w.y > 0.f || w.z > 0.f);
}

}  // namespace b200rt
