// BSDFs of the wavefront shade stage: the reference's BSDF container (src/core/reflection.cpp:670-785, reflection.h:165-235)
// over a small list of lobes, in the reference's OWN shading frame (ss = Normalize(shading.dpdu), ts = Cross(ns, ss),
// ns = shading.n; BSDF::BSDF, reflection.h:170-175) — with matched sample streams the frame decides where a sampled
// direction points, so "any orthonormal frame" is not good enough.
//
// Lobes (one WfLobe each; the host turns a material record into its lobe list exactly as the material's
// ComputeScatteringFunctions adds BxDFs):
//   LambertianReflection (reflection.cpp:180-182)                    Material "matte" sigma 0, "plastic", "uber" Kd
//   OrenNayar (reflection.cpp:197-219, reflection.h:414-420)         Material "matte" sigma > 0
//   MicrofacetReflection + TrowbridgeReitzDistribution (reflection.cpp:226-236,405-423; microfacet.cpp:155-345)
//        with FresnelDielectric (plastic, uber Ks) or FresnelConductor (metal)
//   FresnelBlend (reflection.cpp:280-298,450-475)                    Material "substrate"
//   SpecularReflection (reflection.cpp:136-143) with FresnelNoOp (mirror) or FresnelDielectric (uber Kr)
//   SpecularTransmission (reflection.cpp:150-167)                    "uber" Kt / opacity
//   FresnelSpecular (reflection.cpp:477-511)                         smooth "glass" under the path integrator
#pragma once
#include "wf_core.cuh"

namespace b200rt {

enum { BSDF_REFLECTION = 1, BSDF_TRANSMISSION = 2, BSDF_DIFFUSE = 4, BSDF_GLOSSY = 8, BSDF_SPECULAR = 16, BSDF_ALL = 31 };
enum {
    LOBE_LAMBERT = 0,         // c0 = R
    LOBE_OREN = 1,            // c0 = R, e0 = A, e1 = B
    LOBE_MICRO_DIEL = 2,      // c0 = R, a = alpha, e0 = etaI, e1 = etaT
    LOBE_MICRO_COND = 3,      // c0 = eta, c1 = k, a = alpha  (R = 1, etaI = 1)
    LOBE_BLEND = 4,           // c0 = Rd, c1 = Rs, a = alpha
    LOBE_SPEC_REFL_NOOP = 5,  // c0 = R
    LOBE_SPEC_REFL_DIEL = 6,  // c0 = R, e0 = etaI, e1 = etaT
    LOBE_SPEC_TRANS = 7,      // c0 = T, e0 = etaA, e1 = etaB
    LOBE_FRESNEL_SPEC = 8     // c0 = R, c1 = T, e0 = etaA, e1 = etaB
};
constexpr int kMaxLobes = 5;
constexpr int kLobeFloats = 12;                       // 3 x float4
constexpr int kMatFloats = 4 + kMaxLobes * kLobeFloats;   // header (nLobes, eta, -, -) + lobes = 64 floats = 256 B

struct WfLobe {
    uint32_t kind;
    V3 c0, c1;
    float a, e0, e1;
};

WF_HD uint32_t lobe_type(uint32_t kind) {
    switch (kind) {
    case LOBE_LAMBERT: case LOBE_OREN: return BSDF_REFLECTION | BSDF_DIFFUSE;
    case LOBE_MICRO_DIEL: case LOBE_MICRO_COND: case LOBE_BLEND: return BSDF_REFLECTION | BSDF_GLOSSY;
    case LOBE_SPEC_REFL_NOOP: case LOBE_SPEC_REFL_DIEL: return BSDF_REFLECTION | BSDF_SPECULAR;
    case LOBE_SPEC_TRANS: return BSDF_TRANSMISSION | BSDF_SPECULAR;
    default: return BSDF_REFLECTION | BSDF_TRANSMISSION | BSDF_SPECULAR;
    }
}
WF_HD bool type_matches(uint32_t type, uint32_t flags) { return (type & flags) == type; }   // BxDF::MatchesFlags

// ---- reflection.h trigonometry of local directions -----------------------------------------------------------------
WF_HD float cos2_theta(V3 w) { return w.z * w.z; }
WF_HD float sin2_theta(V3 w) { return smax(0.f, 1.f - cos2_theta(w)); }
WF_HD float sin_theta(V3 w) { return sqrtf(sin2_theta(w)); }
WF_HD float tan_theta(V3 w) { return sin_theta(w) / w.z; }
WF_HD float tan2_theta(V3 w) { return sin2_theta(w) / cos2_theta(w); }
WF_HD float cos_phi(V3 w) { const float s = sin_theta(w); return (s == 0) ? 1.f : clampf(w.x / s, -1.f, 1.f); }
WF_HD float sin_phi(V3 w) { const float s = sin_theta(w); return (s == 0) ? 0.f : clampf(w.y / s, -1.f, 1.f); }
WF_HD bool same_hemisphere(V3 a, V3 b) { return a.z * b.z > 0; }
WF_HD V3 spec_mul(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
WF_HD V3 spec_div(V3 a, float f) { return v3(a.x / f, a.y / f, a.z / f); }   // Spectrum::operator/(Float): true divisions
WF_HD bool spec_black(V3 a) { return a.x == 0.f && a.y == 0.f && a.z == 0.f; }

// FrDielectric (reflection.cpp:47-68)
WF_HD float fr_dielectric(float cosThetaI, float etaI, float etaT) {
    cosThetaI = clampf(cosThetaI, -1.f, 1.f);
    const bool entering = cosThetaI > 0.f;
    if (!entering) { const float t = etaI; etaI = etaT; etaT = t; cosThetaI = fabsf(cosThetaI); }
    const float sinThetaI = sqrtf(smax(0.f, 1.f - cosThetaI * cosThetaI));
    const float sinThetaT = etaI / etaT * sinThetaI;
    if (sinThetaT >= 1.f) return 1.f;
    const float cosThetaT = sqrtf(smax(0.f, 1.f - sinThetaT * sinThetaT));
    const float Rparl = ((etaT * cosThetaI) - (etaI * cosThetaT)) / ((etaT * cosThetaI) + (etaI * cosThetaT));
    const float Rperp = ((etaI * cosThetaI) - (etaT * cosThetaT)) / ((etaI * cosThetaI) + (etaT * cosThetaT));
    return (Rparl * Rparl + Rperp * Rperp) / 2.f;
}
// FrConductor (reflection.cpp:71-94) for one channel with etai = 1; FresnelConductor::Evaluate passes |cos|.
// `1. - cosThetaI2` and `0.5 * (Rp + Rs)` are double expressions in the reference (rounded to float on assignment).
WF_HD float fr_conductor(float cosThetaI, float etat, float k) {
    cosThetaI = clampf(cosThetaI, -1.f, 1.f);
    const float eta = etat / 1.f, etak = k / 1.f;
    const float cosThetaI2 = cosThetaI * cosThetaI;
    const float sinThetaI2 = (float)(1. - (double)cosThetaI2);
    const float eta2 = eta * eta, etak2 = etak * etak;
    const float t0 = eta2 - etak2 - sinThetaI2;
    const float a2plusb2 = sqrtf(t0 * t0 + 4.f * eta2 * etak2);
    const float t1 = a2plusb2 + cosThetaI2;
    const float a = sqrtf(0.5f * (a2plusb2 + t0));
    const float t2 = 2.f * cosThetaI * a;
    const float Rs = (t1 - t2) / (t1 + t2);
    const float t3 = cosThetaI2 * a2plusb2 + sinThetaI2 * sinThetaI2;
    const float t4 = t2 * sinThetaI2;
    const float Rp = Rs * (t3 - t4) / (t3 + t4);
    return 0.5f * (Rp + Rs);
}

// ---- TrowbridgeReitzDistribution with alphax = alphay = a (microfacet.cpp) ------------------------------------------
WF_HD float tr_D(V3 wh, float a) {
    const float tan2 = tan2_theta(wh);
    if (isinf(tan2)) return 0.f;
    const float cos4 = cos2_theta(wh) * cos2_theta(wh);
    const float cp = cos_phi(wh), sp = sin_phi(wh);
    const float e = ((cp * cp) / (a * a) + (sp * sp) / (a * a)) * tan2;
    return 1.f / (kPiF * a * a * cos4 * (1.f + e) * (1.f + e));
}
WF_HD float tr_lambda(V3 w, float a) {
    const float absTan = fabsf(tan_theta(w));
    if (isinf(absTan)) return 0.f;
    const float cp = cos_phi(w), sp = sin_phi(w);
    const float alpha = sqrtf((cp * cp) * a * a + (sp * sp) * a * a);
    const float at = (alpha * absTan) * (alpha * absTan);
    return (-1.f + sqrtf(1.f + at)) / 2.f;
}
WF_HD float tr_G(V3 wo, V3 wi, float a) { return 1.f / (1.f + tr_lambda(wo, a) + tr_lambda(wi, a)); }
// MicrofacetDistribution::Pdf, sampleVisibleArea (microfacet.cpp:338-344)
WF_HD float tr_pdf(V3 wo, V3 wh, float a) {
    const float g1 = 1.f / (1.f + tr_lambda(wo, a));
    return tr_D(wh, a) * g1 * fabsf(vdot(wo, wh)) / fabsf(wo.z);
}
// TrowbridgeReitzSample11 (microfacet.cpp:238-281).  Its normal-incidence branch is written with unqualified sqrt / cos /
// sin and double literals in the reference: the arithmetic there is double.
WF_HD void tr_sample11(float cosTheta, float U1, float U2, float &slope_x, float &slope_y) {
    if ((double)cosTheta > .9999) {
        const float r = (float)sqrt((double)(U1 / (1.f - U1)));
        const float phi = (float)(6.28318530718 * (double)U2);
        slope_x = (float)((double)r * cos((double)phi));
        slope_y = (float)((double)r * sin((double)phi));
        return;
    }
    const float sinTheta = sqrtf(smax(0.f, 1.f - cosTheta * cosTheta));
    const float tanTheta = sinTheta / cosTheta;
    const float aa = 1.f / tanTheta;
    const float G1 = 2.f / (1.f + sqrtf(1.f + 1.f / (aa * aa)));
    const float A = 2.f * U1 / G1 - 1.f;
    float tmp = 1.f / (A * A - 1.f);
    if ((double)tmp > 1e10) tmp = 1e10f;
    const float B = tanTheta;
    const float D = sqrtf(smax(B * B * tmp * tmp - (A * A - B * B) * tmp, 0.f));
    const float s1 = B * tmp - D, s2 = B * tmp + D;
    slope_x = (A < 0.f || s2 > 1.f / tanTheta) ? s1 : s2;
    float S;
    if (U2 > 0.5f) { S = 1.f; U2 = 2.f * (U2 - .5f); }
    else { S = -1.f; U2 = 2.f * (.5f - U2); }
    const float z = (U2 * (U2 * (U2 * 0.27385f - 0.73369f) + 0.46341f)) /
                    (U2 * (U2 * (U2 * 0.093073f + 0.309420f) - 1.000000f) + 0.597999f);
    slope_y = S * z * sqrtf(1.f + slope_x * slope_x);
}
// TrowbridgeReitzSample + TrowbridgeReitzDistribution::Sample_wh, sampleVisibleArea (microfacet.cpp:283-336)
WF_HD V3 tr_sample_wh(V3 wo, float u0, float u1, float a) {
    const bool flip = wo.z < 0;
    const V3 wi = flip ? vneg(wo) : wo;
    const V3 ws = vnormalize(v3(a * wi.x, a * wi.y, wi.z));
    float slope_x, slope_y;
    tr_sample11(ws.z, u0, u1, slope_x, slope_y);
    const float cp = cos_phi(ws), sp = sin_phi(ws);
    const float t = cp * slope_x - sp * slope_y;
    slope_y = sp * slope_x + cp * slope_y;
    slope_x = t;
    slope_x = a * slope_x;
    slope_y = a * slope_y;
    V3 wh = vnormalize(v3(-slope_x, -slope_y, 1.f));
    if (flip) wh = vneg(wh);
    return wh;
}

// ConcentricSampleDisk + CosineSampleHemisphere (sampling.cpp:113-130, sampling.h:159-163)
WF_HD V3 cosine_sample_hemisphere(float u0, float u1) {
    const float ox = 2.f * u0 - 1.f, oy = 2.f * u1 - 1.f;
    float dx = 0.f, dy = 0.f;
    if (!(ox == 0 && oy == 0)) {
        float theta, r;
        if (fabsf(ox) > fabsf(oy)) { r = ox; theta = kPiOver4F * (oy / ox); }
        else { r = oy; theta = kPiOver2F - kPiOver4F * (ox / oy); }
        dx = r * cos_ref(theta); dy = r * sin_ref(theta);
    }
    const float z = sqrtf(smax(0.f, 1.f - dx * dx - dy * dy));
    return v3(dx, dy, z);
}

// ---- one lobe: BxDF::f / Pdf / Sample_f ----------------------------------------------------------------------------
WF_HD_NOINLINE V3 microfacet_f(const WfLobe &l, V3 wo, V3 wi) {   // MicrofacetReflection::f (reflection.cpp:226-236)
    const float cosO = fabsf(wo.z), cosI = fabsf(wi.z);
    V3 wh = vadd(wi, wo);
    if (cosI == 0 || cosO == 0) return v3(0.f, 0.f, 0.f);
    if (wh.x == 0 && wh.y == 0 && wh.z == 0) return v3(0.f, 0.f, 0.f);
    wh = vnormalize(wh);
    const float D = tr_D(wh, l.a), G = tr_G(wo, wi, l.a), den = 4.f * cosI * cosO;
    const float c = vdot(wi, wh);
    if (l.kind == LOBE_MICRO_DIEL) {
        const float F = fr_dielectric(c, l.e0, l.e1);
        return spec_div(v3(l.c0.x * D * G * F, l.c0.y * D * G * F, l.c0.z * D * G * F), den);
    }
    const V3 F = v3(fr_conductor(fabsf(c), l.c0.x, l.c1.x), fr_conductor(fabsf(c), l.c0.y, l.c1.y), fr_conductor(fabsf(c), l.c0.z, l.c1.z));
    return spec_div(v3(1.f * D * G * F.x, 1.f * D * G * F.y, 1.f * D * G * F.z), den);
}
WF_HD float pow5f(float v) { return (v * v) * (v * v) * v; }
WF_HD_NOINLINE V3 blend_f(const WfLobe &l, V3 wo, V3 wi) {   // FresnelBlend::f (reflection.cpp:285-298)
    const float k = 28.f / (23.f * kPiF);
    const float ti = 1.f - pow5f(1.f - .5f * fabsf(wi.z)), to = 1.f - pow5f(1.f - .5f * fabsf(wo.z));
    const V3 diffuse = v3(k * l.c0.x * (1.f - l.c1.x) * ti * to, k * l.c0.y * (1.f - l.c1.y) * ti * to, k * l.c0.z * (1.f - l.c1.z) * ti * to);
    V3 wh = vadd(wi, wo);
    if (wh.x == 0 && wh.y == 0 && wh.z == 0) return v3(0.f, 0.f, 0.f);
    wh = vnormalize(wh);
    const float d = vdot(wi, wh);
    const float s = tr_D(wh, l.a) / (4.f * fabsf(d) * smax(fabsf(wi.z), fabsf(wo.z)));
    const float p = pow5f(1.f - d);
    return v3(diffuse.x + s * (l.c1.x + p * (1.f - l.c1.x)), diffuse.y + s * (l.c1.y + p * (1.f - l.c1.y)),
              diffuse.z + s * (l.c1.z + p * (1.f - l.c1.z)));
}
WF_HD V3 lobe_f(const WfLobe &l, V3 wo, V3 wi) {
    switch (l.kind) {
    case LOBE_LAMBERT: return vmul(l.c0, kInvPiF);
    case LOBE_OREN: {   // reflection.cpp:197-219
        const float sinI = sin_theta(wi), sinO = sin_theta(wo);
        float maxCos = 0.f;
        if ((double)sinI > 1e-4 && (double)sinO > 1e-4) {
            const float dCos = cos_phi(wi) * cos_phi(wo) + sin_phi(wi) * sin_phi(wo);
            maxCos = smax(0.f, dCos);
        }
        float sinAlpha, tanBeta;
        if (fabsf(wi.z) > fabsf(wo.z)) { sinAlpha = sinO; tanBeta = sinI / fabsf(wi.z); }
        else { sinAlpha = sinI; tanBeta = sinO / fabsf(wo.z); }
        return vmul(vmul(l.c0, kInvPiF), l.e0 + l.e1 * maxCos * sinAlpha * tanBeta);
    }
    case LOBE_MICRO_DIEL: case LOBE_MICRO_COND: return microfacet_f(l, wo, wi);
    case LOBE_BLEND: return blend_f(l, wo, wi);
    default: return v3(0.f, 0.f, 0.f);   // specular lobes: f = 0 (reflection.h)
    }
}
WF_HD float lobe_pdf(const WfLobe &l, V3 wo, V3 wi) {
    switch (l.kind) {
    case LOBE_LAMBERT: case LOBE_OREN: return same_hemisphere(wo, wi) ? fabsf(wi.z) * kInvPiF : 0.f;   // BxDF::Pdf
    case LOBE_MICRO_DIEL: case LOBE_MICRO_COND: {   // MicrofacetReflection::Pdf (reflection.cpp:419-423)
        if (!same_hemisphere(wo, wi)) return 0.f;
        const V3 wh = vnormalize(vadd(wo, wi));
        return tr_pdf(wo, wh, l.a) / (4.f * vdot(wo, wh));
    }
    case LOBE_BLEND: {   // FresnelBlend::Pdf (reflection.cpp:470-475)
        if (!same_hemisphere(wo, wi)) return 0.f;
        const V3 wh = vnormalize(vadd(wo, wi));
        const float pdf_wh = tr_pdf(wo, wh, l.a);
        return .5f * (fabsf(wi.z) * kInvPiF + pdf_wh / (4.f * vdot(wo, wh)));
    }
    default: return 0.f;
    }
}
WF_HD bool refract_ref(V3 wi, V3 n, float eta, V3 &wt) {   // reflection.h:96-108
    const float cosThetaI = vdot(n, wi);
    const float sin2ThetaI = smax(0.f, 1.f - cosThetaI * cosThetaI);
    const float sin2ThetaT = eta * eta * sin2ThetaI;
    if (sin2ThetaT >= 1) return false;
    const float cosThetaT = sqrtf(1.f - sin2ThetaT);
    const float k = eta * cosThetaI - cosThetaT;
    wt = v3(eta * -wi.x + k * n.x, eta * -wi.y + k * n.y, eta * -wi.z + k * n.z);
    return true;
}
// BxDF::Sample_f of one lobe.  Returns f; pdf = 0 means "no sample".  sampledType: the lobe's type, narrowed by FresnelSpecular.
WF_HD V3 lobe_sample(const WfLobe &l, V3 wo, float u0, float u1, V3 &wi, float &pdf, uint32_t &sampledType) {
    const V3 zero = v3(0.f, 0.f, 0.f);
    pdf = 0.f;
    switch (l.kind) {
    case LOBE_LAMBERT: case LOBE_OREN:   // BxDF::Sample_f (reflection.cpp:378-385)
        wi = cosine_sample_hemisphere(u0, u1);
        if (wo.z < 0) wi.z *= -1.f;
        pdf = lobe_pdf(l, wo, wi);
        return lobe_f(l, wo, wi);
    case LOBE_MICRO_DIEL: case LOBE_MICRO_COND: {   // MicrofacetReflection::Sample_f (reflection.cpp:405-417)
        if (wo.z == 0) return zero;
        const V3 wh = tr_sample_wh(wo, u0, u1, l.a);
        const float d2 = 2.f * vdot(wo, wh);
        wi = v3(-wo.x + d2 * wh.x, -wo.y + d2 * wh.y, -wo.z + d2 * wh.z);   // Reflect(wo, wh)
        if (!same_hemisphere(wo, wi)) return zero;
        pdf = tr_pdf(wo, wh, l.a) / (4.f * vdot(wo, wh));
        return microfacet_f(l, wo, wi);
    }
    case LOBE_BLEND: {   // FresnelBlend::Sample_f (reflection.cpp:450-468)
        if ((double)u0 < .5) {
            u0 = smin(2.f * u0, kOneMinusEps);
            wi = cosine_sample_hemisphere(u0, u1);
            if (wo.z < 0) wi.z *= -1.f;
        } else {
            u0 = smin(2.f * (u0 - .5f), kOneMinusEps);
            const V3 wh = tr_sample_wh(wo, u0, u1, l.a);
            const float d2 = 2.f * vdot(wo, wh);
            wi = v3(-wo.x + d2 * wh.x, -wo.y + d2 * wh.y, -wo.z + d2 * wh.z);
            if (!same_hemisphere(wo, wi)) return zero;
        }
        pdf = lobe_pdf(l, wo, wi);
        return blend_f(l, wo, wi);
    }
    case LOBE_SPEC_REFL_NOOP: case LOBE_SPEC_REFL_DIEL: {   // SpecularReflection::Sample_f (reflection.cpp:136-143)
        wi = v3(-wo.x, -wo.y, wo.z);
        pdf = 1.f;
        const float F = l.kind == LOBE_SPEC_REFL_NOOP ? 1.f : fr_dielectric(wi.z, l.e0, l.e1);
        return spec_div(vmul(l.c0, F), fabsf(wi.z));   // fresnel->Evaluate(cos) * R / AbsCosTheta(wi)
    }
    case LOBE_SPEC_TRANS: {   // SpecularTransmission::Sample_f (reflection.cpp:150-167), TransportMode::Radiance
        const bool entering = wo.z > 0;
        const float etaI = entering ? l.e0 : l.e1, etaT = entering ? l.e1 : l.e0;
        if (!refract_ref(wo, faceforward(v3(0.f, 0.f, 1.f), wo), etaI / etaT, wi)) return zero;
        pdf = 1.f;
        const float F = fr_dielectric(wi.z, l.e0, l.e1);
        V3 ft = v3(l.c0.x * (1.f - F), l.c0.y * (1.f - F), l.c0.z * (1.f - F));
        ft = vmul(ft, (etaI * etaI) / (etaT * etaT));
        return spec_div(ft, fabsf(wi.z));
    }
    default: {   // FresnelSpecular::Sample_f (reflection.cpp:477-511)
        const float F = fr_dielectric(wo.z, l.e0, l.e1);
        if (u0 < F) {
            wi = v3(-wo.x, -wo.y, wo.z);
            sampledType = BSDF_SPECULAR | BSDF_REFLECTION;
            pdf = F;
            return spec_div(v3(F * l.c0.x, F * l.c0.y, F * l.c0.z), fabsf(wi.z));
        }
        const bool entering = wo.z > 0;
        const float etaI = entering ? l.e0 : l.e1, etaT = entering ? l.e1 : l.e0;
        if (!refract_ref(wo, faceforward(v3(0.f, 0.f, 1.f), wo), etaI / etaT, wi)) return zero;
        V3 ft = vmul(l.c1, 1.f - F);
        ft = vmul(ft, (etaI * etaI) / (etaT * etaT));
        sampledType = BSDF_SPECULAR | BSDF_TRANSMISSION;
        pdf = 1.f - F;
        return spec_div(ft, fabsf(wi.z));
    }
    }
}

// ---- the BSDF container --------------------------------------------------------------------------------------------
struct WfBsdf {
    V3 ns, ng, ss, ts;
    float eta;
    uint32_t nLobes;
    const float *lobes;   // nLobes x kLobeFloats in device memory (16-byte aligned), or nullptr = the one Lambertian lobe `lean`
    V3 leanKd;
};
WF_HD WfLobe bsdf_lobe(const WfBsdf &b, uint32_t i) {
    WfLobe l;
    if (!b.lobes) { l.kind = LOBE_LAMBERT; l.c0 = b.leanKd; l.c1 = v3(0.f, 0.f, 0.f); l.a = l.e0 = l.e1 = 0.f; return l; }
    const float *p = b.lobes + (size_t)i * kLobeFloats;
    l.kind = f2u(p[0]);
    l.c0 = v3(p[1], p[2], p[3]); l.c1 = v3(p[4], p[5], p[6]);
    l.a = p[7]; l.e0 = p[8]; l.e1 = p[9];
    return l;
}
WF_HD V3 bsdf_to_local(const WfBsdf &b, V3 v) { return v3(vdot(v, b.ss), vdot(v, b.ts), vdot(v, b.ns)); }
WF_HD V3 bsdf_to_world(const WfBsdf &b, V3 v) {
    return v3(b.ss.x * v.x + b.ts.x * v.y + b.ns.x * v.z, b.ss.y * v.x + b.ts.y * v.y + b.ns.y * v.z, b.ss.z * v.x + b.ts.z * v.y + b.ns.z * v.z);
}
WF_HD uint32_t bsdf_num_components(const WfBsdf &b, uint32_t flags) {
    uint32_t n = 0;
    for (uint32_t i = 0; i < b.nLobes; ++i) n += type_matches(lobe_type(bsdf_lobe(b, i).kind), flags) ? 1u : 0u;
    return n;
}
// The three per-lobe switches out of line: k_shade<true> evaluates BSDFs at three places (the light sample's f / Pdf, EstimateDirect's
// BSDF sample, the path's next direction); inlined, every place carries its own copy of every lobe's code, and a kernel of warps
// that diverge over materials then waits for instructions (ncu: 15 no_instruction stalls per issue on scenes/living-room).
// OUT = false keeps them inline (the lean kernel has one Lambert lobe and nothing to share).
WF_HD_NOINLINE V3 lobe_f_out(const WfLobe &l, V3 wo, V3 wi) { return lobe_f(l, wo, wi); }
WF_HD_NOINLINE float lobe_pdf_out(const WfLobe &l, V3 wo, V3 wi) { return lobe_pdf(l, wo, wi); }
WF_HD_NOINLINE V3 lobe_sample_out(const WfLobe &l, V3 wo, float u0, float u1, V3 &wi, float &pdf, uint32_t &sampledType) {
    return lobe_sample(l, wo, u0, u1, wi, pdf, sampledType);
}
// BSDF::f (reflection.cpp:670-683)
template <bool OUT = false>
WF_HD V3 bsdf_f(const WfBsdf &b, V3 woW, V3 wiW, uint32_t flags) {
    const V3 wi = bsdf_to_local(b, wiW), wo = bsdf_to_local(b, woW);
    V3 f = v3(0.f, 0.f, 0.f);
    if (wo.z == 0) return f;
    const bool reflect = vdot(wiW, b.ng) * vdot(woW, b.ng) > 0;
    for (uint32_t i = 0; i < b.nLobes; ++i) {
        const WfLobe l = bsdf_lobe(b, i);
        const uint32_t t = lobe_type(l.kind);
        if (type_matches(t, flags) && ((reflect && (t & BSDF_REFLECTION)) || (!reflect && (t & BSDF_TRANSMISSION)))) f = vadd(f, OUT ? lobe_f_out(l, wo, wi) : lobe_f(l, wo, wi));
    }
    return f;
}
// BSDF::Pdf (reflection.cpp:770-785)
template <bool OUT = false>
WF_HD float bsdf_pdf(const WfBsdf &b, V3 woW, V3 wiW, uint32_t flags) {
    if (b.nLobes == 0) return 0.f;
    const V3 wo = bsdf_to_local(b, woW), wi = bsdf_to_local(b, wiW);
    if (wo.z == 0) return 0.f;
    float pdf = 0.f;
    int matching = 0;
    for (uint32_t i = 0; i < b.nLobes; ++i) {
        const WfLobe l = bsdf_lobe(b, i);
        if (type_matches(lobe_type(l.kind), flags)) { ++matching; pdf += OUT ? lobe_pdf_out(l, wo, wi) : lobe_pdf(l, wo, wi); }
    }
    return matching > 0 ? pdf / matching : 0.f;
}
// BSDF::Sample_f (reflection.cpp:703-768).  Returns f (black with pdf = 0 when nothing was sampled).
template <bool OUT = false>
WF_HD V3 bsdf_sample(const WfBsdf &b, V3 woW, float u0, float u1, uint32_t flags, V3 &wiW, float &pdf, uint32_t &sampledType) {
    const V3 zero = v3(0.f, 0.f, 0.f);
    pdf = 0.f; sampledType = 0;
    const int matching = (int)bsdf_num_components(b, flags);
    if (matching == 0) return zero;
    int comp = (int)floorf(u0 * matching);
    if (matching - 1 < comp) comp = matching - 1;
    uint32_t chosen = 0;
    int count = comp;
    for (uint32_t i = 0; i < b.nLobes; ++i)
        if (type_matches(lobe_type(bsdf_lobe(b, i).kind), flags) && count-- == 0) { chosen = i; break; }
    const WfLobe l = bsdf_lobe(b, chosen);
    const uint32_t ltype = lobe_type(l.kind);
    const float ur = smin(u0 * matching - comp, kOneMinusEps);
    const V3 wo = bsdf_to_local(b, woW);
    if (wo.z == 0) return zero;
    V3 wi;
    sampledType = ltype;
    V3 f = OUT ? lobe_sample_out(l, wo, ur, u1, wi, pdf, sampledType) : lobe_sample(l, wo, ur, u1, wi, pdf, sampledType);
    if (pdf == 0) { sampledType = 0; return zero; }
    wiW = bsdf_to_world(b, wi);
    if (!(ltype & BSDF_SPECULAR) && matching > 1)
        for (uint32_t i = 0; i < b.nLobes; ++i) {
            if (i == chosen) continue;
            const WfLobe o = bsdf_lobe(b, i);
            if (type_matches(lobe_type(o.kind), flags)) pdf += OUT ? lobe_pdf_out(o, wo, wi) : lobe_pdf(o, wo, wi);
        }
    if (matching > 1) pdf /= matching;
    if (!(ltype & BSDF_SPECULAR)) {
        const bool reflect = vdot(wiW, b.ng) * vdot(woW, b.ng) > 0;
        f = zero;
        for (uint32_t i = 0; i < b.nLobes; ++i) {
            const WfLobe o = bsdf_lobe(b, i);
            const uint32_t t = lobe_type(o.kind);
            if (type_matches(t, flags) && ((reflect && (t & BSDF_REFLECTION)) || (!reflect && (t & BSDF_TRANSMISSION)))) f = vadd(f, OUT ? lobe_f_out(o, wo, wi) : lobe_f(o, wo, wi));
        }
    }
    return f;
}

}  // namespace b200rt
