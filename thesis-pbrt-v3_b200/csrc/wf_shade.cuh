// One path vertex of PathIntegrator::Li (src/integrators/path.cpp:84-200) as a pure host/device function: the triangle's
// SurfaceInteraction (Triangle::Intersect's tail, src/shapes/triangle.cpp:294-425), the material's BSDF, direct lighting by
// UniformSampleOneLight / EstimateDirect (src/core/integrator.cpp:86-217) and the next direction with Russian roulette.
// Sample dimensions are consumed in the reference's order, so with a reference sample stream
// (wf_sampler.cuh) every spawned ray is the reference's ray.
#pragma once
#include "../../include/b200rt_wavefront.h"
#include "wf_bsdf.cuh"
#include "wf_texture.cuh"
#include "wf_sampler.cuh"

#ifndef B200RT_SHADE_OUTLINE
#define B200RT_SHADE_OUTLINE 1   // 0: every BSDF evaluation of k_shade<true> inlines the lobe code (A/B)
#endif

namespace b200rt {

struct WfTiling {
    uint32_t xres, yres, tile, ntx, nty, rank, world, spp;   // xres, yres = extent of the film's crop window
    uint32_t x0, y0;            // its first pixel inside the camera's full frame (Film::croppedPixelBounds.pMin)
    uint32_t n_tiles_rank;      // tiles owned by this rank
    uint64_t n_paths_rank;      // n_tiles_rank * tile^2 * spp (incl. slots outside the image on edge tiles)
};

struct WfLight {
    uint32_t kind;     // B200RT_LIGHT_*
    uint32_t prim;     // sphere area light: its primitive number
    float pos[3];      // point: position; sphere: centre; distant: unit direction TOWARDS the light
    float I[3];        // point: intensity; others: radiance
    float radius;      // sphere: radius; distant / infinite: the scene's bounding-sphere radius
    float l2w[9];      // infinite: rotation of LightToWorld, row-major
    float cdfHi;       // cdf[k + 1] of the light distribution (Distribution1D, sampling.h:57-69) for light k
    float pickPdf;     // Distribution1D::SampleDiscrete's pdf of choosing this light: func[k] / (funcInt * n)
};

struct WfParams {
    b200rt_camera cam;
    CamMat cm;                // the reference camera's matrices (cam_exact != 0)
    uint32_t cam_exact;
    WfTiling tl;
    WfLight light;            // lights[0]: the scene's only light unless nLights > 1
    const WfLight *lights;    // nLights records (ABI v10) when nLights > 1, else unused
    uint32_t nLights;
    // "lightsamplestrategy" "spatial" (SpatialLightDistribution, lightdistrib.cpp:95-300): one Distribution1D per voxel of the scene's
    // bound, all of them computed before the frame (k_light_voxels); voxCdf / voxPdf hold cdfHi / pickPdf of light k of voxel v at
    // [v * nLights + k], v = (ix * nVox[1] + iy) * nVox[2] + iz.  nullptr = the lights' own cdfHi / pickPdf (uniform, power)
    const float *voxCdf, *voxPdf;
    uint32_t nVox[3];
    float wb[6];              // Scene::WorldBound(): pMin, pMax
    uint32_t anyMis;          // some light has area (sphere / infinite): EstimateDirect traces BSDF-sampled rays, emitters are visible
    float kd[3];
    uint32_t max_depth;
    const float *mats;        // kMatFloats per material (wf_bsdf.cuh); nullptr = one Lambertian lobe with R = kd
    const uint8_t *primMat;   // material index per primitive
    const float *primUV;      // 6 floats per primitive or nullptr = (0,0) (1,0) (1,1) (triangle.h:109-121)
    const float *primN;       // 9 floats per primitive (world-space vertex normals, TriangleMesh::n) or nullptr
    const uint8_t *primFlags; // bit 0: reverseOrientation ^ transformSwapsHandedness; bit 1: the primitive has vertex normals
    uint32_t has_specular;    // some material has a specular lobe (path.cpp's specularBounce / etaScale bookkeeping is live)
    WfSampler smp;
    const WfTexture *tex;     // imagemap textures (ABI v9) or nullptr; materials name them in their header words 2 / 3
    uint32_t nTex;
};

// path slot -> (pixel x, y inside the crop window, sample); false when the slot lies outside the image (edge tiles)
WF_HD bool path_pixel(const WfTiling &t, uint64_t gpath, uint32_t &x, uint32_t &y, uint32_t &s) {
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

// The sample stream of one path: pixel, sample number and the running dimension counter of GlobalSampler::Get1D / Get2D
// (sampler.cpp:179-195; the path integrator requests no sample arrays, so no dimension is skipped).
struct PathSampler {
    uint64_t index;
    int32_t px, py;
    uint32_t pixOrd, s, dim;
};
WF_HD void path_sampler_init(const WfParams &P, uint32_t x, uint32_t y, uint32_t s, uint32_t dim, PathSampler &ps) {
    ps.px = (int32_t)(x + P.tl.x0); ps.py = (int32_t)(y + P.tl.y0);
    ps.pixOrd = (uint32_t)ps.py * P.cam.xres + (uint32_t)ps.px;
    ps.s = s; ps.dim = dim;
    ps.index = sampler_index(P.smp, ps.px, ps.py, s);
}
// ... with the sample's global index already known (k_camera stores it per path: the index is the same at every bounce)
WF_HD void path_sampler_init(const WfParams &P, uint32_t x, uint32_t y, uint32_t s, uint32_t dim, uint64_t index, PathSampler &ps) {
    ps.px = (int32_t)(x + P.tl.x0); ps.py = (int32_t)(y + P.tl.y0);
    ps.pixOrd = (uint32_t)ps.py * P.cam.xres + (uint32_t)ps.px;
    ps.s = s; ps.dim = dim;
    ps.index = index;
}
WF_HD uint64_t path_sample_index(const WfParams &P, uint32_t x, uint32_t y, uint32_t s) {
    return sampler_index(P.smp, (int32_t)(x + P.tl.x0), (int32_t)(y + P.tl.y0), s);
}
WF_HD float get1d(const WfParams &P, PathSampler &ps) { return sampler_dim(P.smp, ps.index, ps.dim++, ps.px, ps.py, ps.pixOrd, ps.s); }

// ---- camera ray of one path slot -------------------------------------------------------------------------------------
// Sampler::GetCameraSample (sampler.cpp:46-52): pFilm = pixel + Get2D (dims 0, 1), time = Get1D (2), pLens = Get2D (3, 4).
WF_HD void camera_ray(const WfParams &P, uint32_t x, uint32_t y, uint32_t s, V3 &o, V3 &d) {
    PathSampler ps;
    path_sampler_init(P, x, y, s, 0, ps);
    const float u0 = get1d(P, ps), u1 = get1d(P, ps);
    if (P.cam_exact) {
        camera_ray_ref(P.cm, (float)ps.px + u0, (float)ps.py + u1, o, d);
        return;
    }
    // the round-1 pinhole frame (eye / dir / right / up): same image up to rounding, not the reference's float operations
    const float fx = ((float)ps.px + u0) / P.cam.xres, fy = ((float)ps.py + u1) / P.cam.yres;
    const float sx = 2.f * fx - 1.f, sy = 1.f - 2.f * fy;
    const float dx = P.cam.dir[0] + sx * P.cam.right[0] + sy * P.cam.up[0];
    const float dy = P.cam.dir[1] + sx * P.cam.right[1] + sy * P.cam.up[1];
    const float dz = P.cam.dir[2] + sx * P.cam.right[2] + sy * P.cam.up[2];
    const float inv = 1.f / sqrtf(dx * dx + dy * dy + dz * dz);
    o = v3(P.cam.eye[0], P.cam.eye[1], P.cam.eye[2]);
    d = v3(dx * inv, dy * inv, dz * inv);
}
constexpr uint32_t kCameraDims = 5;   // the first dimension PathIntegrator::Li draws

// ---- Triangle::Intersect's tail (triangle.cpp:294-425) for the accepted hit ---------------------------------------------
struct WfIsect {
    V3 p, pError, n, wo;   // Interaction::p, pError, n (geometric, oriented), wo = Normalize(-ray.d)
    V3 ns, dpdu;           // shading.n, shading.dpdu
    V3 gdpdu, gdpdv;       // SurfaceInteraction::dpdu / dpdv (what ComputeDifferentials solves with)
    float u, v;            // SurfaceInteraction::uv = b0 uv0 + b1 uv1 + b2 uv2 (triangle.cpp:331)
};
WF_HD void triangle_isect(const WfParams &P, uint32_t prim, V3 p0, V3 p1, V3 p2, float b0, float b1, float b2, V3 rayD, WfIsect &it) {
    float uv[6] = {0.f, 0.f, 1.f, 0.f, 1.f, 1.f};
    if (P.primUV) for (int k = 0; k < 6; ++k) uv[k] = P.primUV[6 * (size_t)prim + k];
    const uint32_t pf = P.primFlags ? P.primFlags[prim] : 0u;
    const bool flip = (pf & 1u) != 0, hasN = (pf & 2u) != 0 && P.primN;
    const float duv02x = uv[0] - uv[4], duv02y = uv[1] - uv[5], duv12x = uv[2] - uv[4], duv12y = uv[3] - uv[5];
    const V3 dp02 = vsub(p0, p2), dp12 = vsub(p1, p2);
    const float determinant = duv02x * duv12y - duv02y * duv12x;
    const bool degenerateUV = (double)fabsf(determinant) < 1e-8;
    V3 dpdu = v3(0.f, 0.f, 0.f), dpdv = v3(0.f, 0.f, 0.f);
    if (!degenerateUV) {
        const float invdet = 1.f / determinant;
        dpdu = vmul(vsub(vmul(dp02, duv12y), vmul(dp12, duv02y)), invdet);
        dpdv = vmul(vadd(vmul(dp02, -duv12x), vmul(dp12, duv02x)), invdet);
    }
    if (degenerateUV || vlen2(vcross(dpdu, dpdv)) == 0) {
        const V3 ng = vcross(vsub(p2, p0), vsub(p1, p0));   // zero only for triangles the upload flagged as rejected
        coordinate_system(vnormalize(ng), dpdu, dpdv);
    }
    const float g7 = wf_gamma(7);
    it.pError = v3(g7 * (fabsf(b0 * p0.x) + fabsf(b1 * p1.x) + fabsf(b2 * p2.x)), g7 * (fabsf(b0 * p0.y) + fabsf(b1 * p1.y) + fabsf(b2 * p2.y)),
                   g7 * (fabsf(b0 * p0.z) + fabsf(b1 * p1.z) + fabsf(b2 * p2.z)));
    it.p = v3(b0 * p0.x + b1 * p1.x + b2 * p2.x, b0 * p0.y + b1 * p1.y + b2 * p2.y, b0 * p0.z + b1 * p1.z + b2 * p2.z);
    it.wo = vnormalize(vneg(rayD));   // Interaction's constructor normalises wo (interaction.h:60)
    V3 n = vnormalize(vcross(dp02, dp12));   // isect->n = isect->shading.n (triangle.cpp:346)
    V3 ns = n;
    V3 sdpdu = dpdu;
    if (hasN) {
        const float *q = P.primN + 9 * (size_t)prim;
        const V3 n0 = v3(q[0], q[1], q[2]), n1 = v3(q[3], q[4], q[5]), n2 = v3(q[6], q[7], q[8]);
        V3 nsI = v3(b0 * n0.x + b1 * n1.x + b2 * n2.x, b0 * n0.y + b1 * n1.y + b2 * n2.y, b0 * n0.z + b1 * n1.z + b2 * n2.z);
        if (vlen2(nsI) > 0) nsI = vnormalize(nsI); else nsI = n;
        V3 ss = vnormalize(dpdu);
        V3 ts = vcross(ss, nsI);
        if (vlen2(ts) > 0.f) { ts = vnormalize(ts); ss = vcross(ts, nsI); }
        else coordinate_system(nsI, ss, ts);
        // SetShadingGeometry(ss, ts, dndu, dndv, true) (interaction.cpp:73-92)
        ns = vnormalize(vcross(ss, ts));
        if (flip) ns = vneg(ns);
        n = faceforward(n, ns);
        sdpdu = ss;
        n = faceforward(n, ns);   // triangle.cpp:419-420 (idempotent)
    } else if (flip) {
        n = vneg(n); ns = n;
    }
    it.n = n; it.ns = ns; it.dpdu = sdpdu;
    it.gdpdu = dpdu; it.gdpdv = dpdv;
    it.u = b0 * uv[0] + b1 * uv[2] + b2 * uv[4];
    it.v = b0 * uv[1] + b1 * uv[3] + b2 * uv[5];
}

// Lobes of a material whose Kd and / or opacity is an image texture, rebuilt per hit from the texture values (the constant-
// texture materials get theirs once on the host, wf_setup_host.cpp material_lobes — same rules: matte.cpp:44-62,
// plastic.cpp:44-70, substrate.cpp:45-66, uber.cpp:45-100).  m: the material's 64-float record in its TEXTURED layout:
//   m[4] kind, m[5..7] Kd, m[8..10] Ks, m[11] alpha (already remapped), m[12..14] Kr, m[15..17] Kt, m[18..20] opacity,
//   m[21] eta, m[22], m[23] Oren-Nayar A, B (sigma > 0) or 0, 0
WF_HD void put_lobe(float *L, uint32_t &n, uint32_t kind, V3 c0, V3 c1, float a, float e0, float e1) {
    float *l = L + n * kLobeFloats;
    l[0] = u2f(kind);
    l[1] = c0.x; l[2] = c0.y; l[3] = c0.z; l[4] = c1.x; l[5] = c1.y; l[6] = c1.z;
    l[7] = a; l[8] = e0; l[9] = e1;
    for (int i = 10; i < kLobeFloats; ++i) l[i] = 0.f;
    ++n;
}
WF_HD V3 spec_clamp0(V3 c) { return v3(c.x < 0.f ? 0.f : c.x, c.y < 0.f ? 0.f : c.y, c.z < 0.f ? 0.f : c.z); }   // Spectrum::Clamp()
WF_HD void textured_lobes(const float *m, bool haveKd, V3 kdTex, bool haveOp, V3 opTex, float *L, uint32_t &n, float &eta) {
    const uint32_t kind = f2u(m[4]);
    const V3 Kd = haveKd ? spec_clamp0(kdTex) : v3(m[5], m[6], m[7]);
    const V3 Ks = v3(m[8], m[9], m[10]), zero = v3(0.f, 0.f, 0.f);
    const float alpha = m[11];
    n = 0;
    eta = 1.f;
    if (kind == B200RT_MAT_MATTE) {
        if (!spec_black(Kd)) {
            if (m[22] == 0.f && m[23] == 0.f) put_lobe(L, n, LOBE_LAMBERT, Kd, zero, 0, 0, 0);
            else put_lobe(L, n, LOBE_OREN, Kd, zero, 0, m[22], m[23]);
        }
    } else if (kind == B200RT_MAT_PLASTIC) {
        if (!spec_black(Kd)) put_lobe(L, n, LOBE_LAMBERT, Kd, zero, 0, 0, 0);
        if (!spec_black(Ks)) put_lobe(L, n, LOBE_MICRO_DIEL, Ks, zero, alpha, 1.5f, 1.f);
    } else if (kind == B200RT_MAT_SUBSTRATE) {
        if (!spec_black(Kd) || !spec_black(Ks)) put_lobe(L, n, LOBE_BLEND, Kd, Ks, alpha, 0, 0);
    } else {   // B200RT_MAT_UBER
        const float e = m[21];
        const V3 op = haveOp ? spec_clamp0(opTex) : v3(m[18], m[19], m[20]);
        const V3 t = spec_clamp0(v3(-op.x + 1.f, -op.y + 1.f, -op.z + 1.f));
        if (!spec_black(t)) { eta = 1.f; put_lobe(L, n, LOBE_SPEC_TRANS, t, zero, 0, 1.f, 1.f); }
        else eta = e;
        const V3 kd = spec_mul(op, Kd), ks = spec_mul(op, Ks), kr = spec_mul(op, v3(m[12], m[13], m[14])), kt = spec_mul(op, v3(m[15], m[16], m[17]));
        if (!spec_black(kd)) put_lobe(L, n, LOBE_LAMBERT, kd, zero, 0, 0, 0);
        if (!spec_black(ks)) put_lobe(L, n, LOBE_MICRO_DIEL, ks, zero, alpha, 1.f, e);
        if (!spec_black(kr)) put_lobe(L, n, LOBE_SPEC_REFL_DIEL, kr, zero, 0, 1.f, e);
        if (!spec_black(kt)) put_lobe(L, n, LOBE_SPEC_TRANS, kt, zero, 0, 1.f, e);
    }
}

// ---- what one vertex produces ------------------------------------------------------------------------------------------
struct ShadeOut {
    bool wantShadow, wantNext, wantMis;
    V3 Le;                          // beta * emitted radiance seen at this vertex (already weighted)
    V3 so, sd, contrib;             // shadow ray (tMax = 1 - ShadowEpsilon) and the radiance it carries if unoccluded
    V3 mo, md, mcontrib;            // BSDF-sampled MIS ray of EstimateDirect and the radiance it carries if it reaches the light
    uint32_t misPrim;               // ... i.e. if its closest hit is this primitive (the chosen sphere light) / B200RT_MISS (infinite light)
    V3 no, nd, beta;                // next path ray and the new throughput
    float betaW;                    // new (specularBounce, etaScale) code: 0 = (false, 1); else |w| = etaScale, w < 0 = specular bounce
    uint32_t dim;                   // the path's next sample dimension
};

WF_HD V3 spec3(const float *c) { return v3(c[0], c[1], c[2]); }

// Light::Sample_Li + the VisibilityTester's ray, for the reference point `it`.  Returns Li; pdf = 0 means no sample.
// isDelta as IsDeltaLight(light.flags).
template <bool FULL>
WF_HD V3 light_sample(const WfLight &Lt, const WfIsect &it, float u0, float u1, V3 &wi, float &pdf, V3 &so, V3 &sd) {
    const V3 zero = v3(0.f, 0.f, 0.f);
    pdf = 0.f;
    const V3 pL = spec3(Lt.pos);
    if (!FULL || Lt.kind == B200RT_LIGHT_POINT) {
        // PointLight::Sample_Li (point.cpp:43-52); VisibilityTester -> Interaction::SpawnRayTo(const Interaction &) (interaction.h:73-78)
        // towards a bare point (n = 0, pError = 0: the target is the light's position itself)
        const V3 dl = vsub(pL, it.p);
        wi = vnormalize(dl);
        pdf = 1.f;
        so = offset_ray_origin_ref(it.p, it.pError, it.n, dl);
        sd = vsub(pL, so);
        const float d2 = vlen2(dl);   // DistanceSquared(pLight, ref.p) = (p1 - p2).LengthSquared()
        return spec_div(spec3(Lt.I), d2);
    }
    if (Lt.kind == B200RT_LIGHT_DISTANT) {   // DistantLight::Sample_Li (distant.cpp:49-59)
        wi = pL;
        pdf = 1.f;
        const V3 pOut = vadd(it.p, vmul(pL, 2.f * Lt.radius));
        so = offset_ray_origin_ref(it.p, it.pError, it.n, vsub(pOut, it.p));
        sd = vsub(pOut, so);
        return spec3(Lt.I);
    }
    if (Lt.kind == B200RT_LIGHT_SPHERE) {
        // DiffuseAreaLight::Sample_Li (diffuse.cpp:68-80) over Sphere::Sample(ref, u, pdf) (sphere.cpp:232-291), reference
        // points outside the sphere (inside: uniform area sampling, not built — the host refuses scenes that need it)
        const float r = Lt.radius;
        const V3 pO = offset_ray_origin_ref(it.p, it.pError, it.n, vsub(pL, it.p));
        if (vlen2(vsub(pO, pL)) <= r * r) return zero;
        const V3 wc = vnormalize(vsub(pL, it.p));
        V3 wcX, wcY;
        coordinate_system(wc, wcX, wcY);
        const float dist2 = vlen2(vsub(it.p, pL));
        const float sinThetaMax2 = r * r / dist2;
        const float cosThetaMax = sqrtf(smax(0.f, 1.f - sinThetaMax2));
        const float cosTheta = (1.f - u0) + u0 * cosThetaMax;
        const float sinTheta = sqrtf(smax(0.f, 1.f - cosTheta * cosTheta));
        const float phi = u1 * 2.f * kPiF;
        const float dc = sqrtf(dist2);
        const float ds = dc * cosTheta - sqrtf(smax(0.f, r * r - dc * dc * sinTheta * sinTheta));
        const float cosAlpha = (dc * dc + r * r - ds * ds) / (2.f * dc * r);
        const float sinAlpha = sqrtf(smax(0.f, 1.f - cosAlpha * cosAlpha));
        // SphericalDirection(sinAlpha, cosAlpha, phi, -wcX, -wcY, -wc) (geometry.h:1809-1814)
        const float sc = sinAlpha * cos_ref(phi), ss = sinAlpha * sin_ref(phi);
        const V3 nW = vadd(vadd(vmul(vneg(wcX), sc), vmul(vneg(wcY), ss)), vmul(vneg(wc), cosAlpha));
        const V3 pW = v3(pL.x + r * nW.x, pL.y + r * nW.y, pL.z + r * nW.z);
        const float g5 = wf_gamma(5);
        const V3 peL = v3(g5 * fabsf(pW.x), g5 * fabsf(pW.y), g5 * fabsf(pW.z));
        pdf = 1.f / (2.f * kPiF * (1.f - cosThetaMax));
        const V3 dl = vsub(pW, it.p);
        if (pdf == 0 || vlen2(dl) == 0) { pdf = 0.f; return zero; }
        wi = vnormalize(dl);
        // VisibilityTester(ref, pShape) -> SpawnRayTo(const Interaction &)
        so = offset_ray_origin_ref(it.p, it.pError, it.n, dl);
        const V3 tg = offset_ray_origin_ref(pW, peL, nW, vsub(so, pW));
        sd = vsub(tg, so);
        return (vdot(nW, vneg(wi)) > 0) ? spec3(Lt.I) : zero;   // DiffuseAreaLight::L, one-sided (diffuse.h:56-58)
    }
    // InfiniteAreaLight::Sample_Li (infinite.cpp:98-123) with the constant 1x1 map: Distribution2D::SampleContinuous(u) = u, mapPdf 1
    const float theta = u1 * kPiF, phi = u0 * 2.f * kPiF;
    const float cosTheta = cos_ref(theta), sinTheta = sin_ref(theta);
    const float sinPhi = sin_ref(phi), cosPhi = cos_ref(phi);
    const V3 wl = v3(sinTheta * cosPhi, sinTheta * sinPhi, cosTheta);
    wi = v3(Lt.l2w[0] * wl.x + Lt.l2w[1] * wl.y + Lt.l2w[2] * wl.z, Lt.l2w[3] * wl.x + Lt.l2w[4] * wl.y + Lt.l2w[5] * wl.z,
            Lt.l2w[6] * wl.x + Lt.l2w[7] * wl.y + Lt.l2w[8] * wl.z);
    pdf = 1.f / (2.f * kPiF * kPiF * sinTheta);
    if (sinTheta == 0) pdf = 0.f;
    const V3 pOut = vadd(it.p, vmul(wi, 2.f * Lt.radius));
    so = offset_ray_origin_ref(it.p, it.pError, it.n, vsub(pOut, it.p));
    sd = vsub(pOut, so);
    return spec3(Lt.I);
}
// Light::Pdf_Li(ref, wi) of the lights with area
WF_HD float light_pdf(const WfLight &Lt, const WfIsect &it, V3 wi) {
    if (Lt.kind == B200RT_LIGHT_SPHERE) {   // Sphere::Pdf (sphere.cpp:293-305), reference point outside
        const V3 pL = spec3(Lt.pos);
        const float r = Lt.radius;
        const V3 pO = offset_ray_origin_ref(it.p, it.pError, it.n, vsub(pL, it.p));
        if (vlen2(vsub(pO, pL)) <= r * r) return 0.f;
        const float sinThetaMax2 = r * r / vlen2(vsub(it.p, pL));
        const float cosThetaMax = sqrtf(smax(0.f, 1.f - sinThetaMax2));
        return 1.f / (2.f * kPiF * (1.f - cosThetaMax));
    }
    // InfiniteAreaLight::Pdf_Li (infinite.cpp:125-133), constant map: distribution->Pdf = 1; WorldToLight = transpose
    const V3 wl = v3(Lt.l2w[0] * wi.x + Lt.l2w[3] * wi.y + Lt.l2w[6] * wi.z, Lt.l2w[1] * wi.x + Lt.l2w[4] * wi.y + Lt.l2w[7] * wi.z,
                     Lt.l2w[2] * wi.x + Lt.l2w[5] * wi.y + Lt.l2w[8] * wi.z);
    const float theta = acos_ref(clampf(wl.z, -1.f, 1.f));
    const float sinTheta = sin_ref(theta);
    if (sinTheta == 0) return 0.f;
    return 1.f / (2.f * kPiF * kPiF * sinTheta);
}

// ---- SpatialLightDistribution (lightdistrib.cpp:95-300) ------------------------------------------------------------------
// nVoxels (:101-110): the widest side of the scene's bound gets maxVoxels voxels, the others in proportion.
WF_HD void spatial_voxel_counts(const float *wb, int maxVoxels, uint32_t *nVox) {
    const float d[3] = {wb[3] - wb[0], wb[4] - wb[1], wb[5] - wb[2]};
    const int widest = (d[0] > d[1] && d[0] > d[2]) ? 0 : (d[1] > d[2] ? 1 : 2);   // Bounds3::MaximumExtent
    const float bmax = d[widest];
    for (int i = 0; i < 3; ++i) {
        const int v = (int)roundf(d[i] / bmax * (float)maxVoxels);
        nVox[i] = (uint32_t)(v > 1 ? v : 1);
    }
}
// Lookup's voxel of a point (:133-142): Bounds3::Offset, then Clamp(int(offset * nVoxels), 0, nVoxels - 1)
WF_HD uint32_t spatial_voxel_of(const float *wb, const uint32_t *nVox, V3 p) {
    float o[3] = {p.x - wb[0], p.y - wb[1], p.z - wb[2]};
    uint32_t pi[3];
    for (int i = 0; i < 3; ++i) {
        if (wb[3 + i] > wb[i]) o[i] /= wb[3 + i] - wb[i];
        const int v = (int)(o[i] * (float)(int)nVox[i]);
        pi[i] = (uint32_t)(v < 0 ? 0 : (v > (int)nVox[i] - 1 ? (int)nVox[i] - 1 : v));
    }
    return (pi[0] * nVox[1] + pi[1]) * nVox[2] + pi[2];
}
WF_HD float lerp_ref(float t, float v1, float v2) { return (1 - t) * v1 + t * v2; }   // pbrt::Lerp (pbrt.h:417)
// ComputeDistribution (:229-300) of voxel v: 128 points of the 3-D Halton sequence inside the voxel, every light sampled from
// each with the next two Halton dimensions, Li.y() / pdf accumulated per light, floored at a thousandth of the average, and the
// Distribution1D over that.  halton5: RadicalInverse(0..4, i) for i < 128 (5 floats per i).  contrib: nLights floats of scratch.
WF_HD void spatial_voxel_distribution(const WfLight *lights, uint32_t nLights, const float *wb, const uint32_t *nVox, uint32_t v,
                                      const float *halton5, float *contrib, float *cdfHi, float *pickPdf) {
    const uint32_t iz = v % nVox[2], iy = (v / nVox[2]) % nVox[1], ix = v / (nVox[2] * nVox[1]);
    const uint32_t pi[3] = {ix, iy, iz};
    float lo[3], hi[3];
    for (int i = 0; i < 3; ++i) {
        const float t0 = (float)(int)pi[i] / (float)(int)nVox[i], t1 = (float)(int)(pi[i] + 1) / (float)(int)nVox[i];
        const float a = lerp_ref(t0, wb[i], wb[3 + i]), b = lerp_ref(t1, wb[i], wb[3 + i]);
        lo[i] = a < b ? a : b; hi[i] = a < b ? b : a;   // Bounds3(p1, p2): per-component min / max
    }
    for (uint32_t j = 0; j < nLights; ++j) contrib[j] = 0.f;
    WfIsect it;
    it.n = v3(0.f, 0.f, 0.f); it.pError = v3(0.f, 0.f, 0.f);
    for (int i = 0; i < 128; ++i) {
        const float *h = halton5 + 5 * i;
        it.p = v3(lerp_ref(h[0], lo[0], hi[0]), lerp_ref(h[1], lo[1], hi[1]), lerp_ref(h[2], lo[2], hi[2]));
        for (uint32_t j = 0; j < nLights; ++j) {
            V3 wi, so, sd;
            float pdf = 0.f;
            const V3 Li = light_sample<true>(lights[j], it, h[3], h[4], wi, pdf, so, sd);
            if (pdf > 0) contrib[j] += (0.212671f * Li.x + 0.715160f * Li.y + 0.072169f * Li.z) / pdf;   // RGBSpectrum::y()
        }
    }
    float sum = 0.f;
    for (uint32_t j = 0; j < nLights; ++j) sum += contrib[j];
    const float avg = sum / (float)(128 * (int)nLights);
    const float minContrib = (avg > 0) ? (float)(.001 * (double)avg) : 1.f;
    for (uint32_t j = 0; j < nLights; ++j) contrib[j] = contrib[j] > minContrib ? contrib[j] : minContrib;   // std::max(a, b): a < b ? b : a
    // Distribution1D (sampling.h:57-69)
    const int n = (int)nLights;
    float c = 0.f;
    for (int j = 0; j < n; ++j) { c = c + contrib[j] / n; cdfHi[j] = c; }
    const float funcInt = c;
    for (int j = 0; j < n; ++j) {
        cdfHi[j] = (funcInt == 0) ? (float)(j + 1) / (float)n : cdfHi[j] / funcInt;
        pickPdf[j] = (funcInt > 0) ? contrib[j] / (funcInt * n) : 0.f;
    }
}

// The lobes of a textured material at this hit (out of line: it is rare and register-hungry).
// ComputeScatteringFunctions -> ComputeDifferentials(ray): only the camera ray has differentials.
WF_HD_NOINLINE void textured_material(const WfParams &P, const float *m, uint32_t depth, uint32_t x, uint32_t y, uint32_t s, V3 rayO, V3 rayD,
                                      const WfIsect &it, float *L, uint32_t &n, float &eta) {
    const uint32_t kdT = f2u(m[2]), opT = f2u(m[3]);
    WfUVDiff D = {0.f, 0.f, 0.f, 0.f};
    if (depth == 0 && P.cam_exact) {
        PathSampler cs;
        path_sampler_init(P, x, y, s, 0, cs);
        const float u0 = get1d(P, cs), u1 = get1d(P, cs);
        V3 rxO, rxD, ryO, ryD;
        camera_ray_diff_ref(P.cm, (float)cs.px + u0, (float)cs.py + u1, 1.f / sqrtf((float)P.tl.spp), rayO, rayD, rxO, rxD, ryO, ryD);
        D = compute_differentials(it.p, it.n, it.gdpdu, it.gdpdv, rxO, rxD, ryO, ryD);
    }
    V3 kdV = v3(0.f, 0.f, 0.f), opV = v3(0.f, 0.f, 0.f);
    if (kdT) kdV = tex_evaluate(P.tex[kdT - 1], it.u, it.v, D);
    if (opT) opV = (opT == kdT) ? kdV : tex_evaluate(P.tex[opT - 1], it.u, it.v, D);   // the same texture object: the same value
    textured_lobes(m, kdT != 0, kdV, opT != 0, opV, L, n, eta);
}

// One bounce of PathIntegrator::Li for a path whose ray (rayO, rayD) found `hit`.
//   bt  : beta.xyz and the (specularBounce, etaScale) code in w
//   dim : the path's next sample dimension
// The vertex in two halves, so that k_shade<true> can put a block-wide barrier between them (its warps then run the same stretch of
// this long code at the same time and share the instruction cache): shade_vertex_direct = emitted radiance, the triangle's
// interaction, the BSDF and UniformSampleOneLight; it returns false when the path ends there.  shade_vertex_bounce = the next
// direction and Russian roulette.  What the second half needs of the first lives in ShadeMid.
struct ShadeMid {
    WfIsect it;
    WfBsdf bs;
    PathSampler ps;
    float etaScale;
    float texLobes[kMaxLobes * kLobeFloats];
    // UniformSampleOneLight's choice and draws (shade_vertex_setup), for the two strategies of EstimateDirect
    const WfLight *Lp;
    float pickPdf, ul0, ul1, us0, us1;
    bool direct;
};
template <bool FULL>
WF_HD bool shade_vertex_setup(const WfParams &P, uint32_t depth, uint32_t x, uint32_t y, uint32_t s, uint64_t sidx /* path_sample_index(x, y, s) */,
                               uint32_t dim, V3 rayO, V3 rayD,
                               uint32_t prim, float t, float b0, float b1, float b2, V3 p0, V3 p1, V3 p2, bool isSphere,
                               V3 beta, float betaW, ShadeOut &o, ShadeMid &M) {
    WfIsect &it = M.it;
    WfBsdf &bs = M.bs;
    PathSampler &ps = M.ps;
    float *texLobes = M.texLobes;
    const V3 zero = v3(0.f, 0.f, 0.f);
    o.wantShadow = o.wantNext = o.wantMis = false;
    o.Le = zero;
    o.dim = dim;
    o.betaW = betaW;
    const bool multi = FULL && P.nLights > 1;
    const bool specularBounce = FULL && betaW < 0.f;
    const float etaScale = (!FULL || betaW == 0.f) ? 1.f : fabsf(betaW);
    // path.cpp:97-107: emitted radiance at camera-ray hits and after a specular bounce
    if (FULL && (depth == 0 || specularBounce)) {
        const uint32_t nL = multi ? P.nLights : 1u;
        for (uint32_t k = 0; k < nL; ++k) {
            const WfLight &Le = multi ? P.lights[k] : P.light;
            if (prim == B200RT_MISS) {
                // `for (const auto &light : scene.infiniteLights) L += beta * light->Le(ray)`: InfiniteAreaLight::Le with the constant map
                if (Le.kind == B200RT_LIGHT_INFINITE) o.Le = vadd(o.Le, spec_mul(beta, spec3(Le.I)));
            } else if (Le.kind == B200RT_LIGHT_SPHERE && prim == Le.prim) {
                // isect.Le(-ray.d) -> DiffuseAreaLight::L(intr, w): Lemit when Dot(intr.n, w) > 0; a sphere's n points outwards
                const V3 ph = v3(rayO.x + t * rayD.x, rayO.y + t * rayD.y, rayO.z + t * rayD.z);
                if (vdot(vsub(ph, spec3(Le.pos)), vneg(rayD)) > 0.f) o.Le = spec_mul(beta, spec3(Le.I));
            }
        }
    }
    if (prim == B200RT_MISS || depth >= P.max_depth) return false;   // path.cpp:110
    if (isSphere) return false;   // sphere primitives are emitters with Material "matte" Kd 0 here: no BxDF, the path ends (path.cpp:147)
    M.etaScale = etaScale;
    triangle_isect(P, prim, p0, p1, p2, b0, b1, b2, rayD, it);
    // the material's BSDF (ComputeScatteringFunctions -> BSDF(si, eta), reflection.h:170-175)
    bs.ns = it.ns; bs.ng = it.n;
    bs.ss = vnormalize(it.dpdu);
    bs.ts = vcross(bs.ns, bs.ss);
    bs.leanKd = spec3(P.kd);
    bs.lobes = nullptr; bs.nLobes = spec_black(bs.leanKd) ? 0u : 1u; bs.eta = 1.f;
    if (FULL && P.mats) {
        const float *m = P.mats + (size_t)P.primMat[prim] * kMatFloats;
        bs.nLobes = f2u(m[0]); bs.eta = m[1]; bs.lobes = m + 4;
        if (P.tex && (f2u(m[2]) | f2u(m[3]))) {
            textured_material(P, m, depth, x, y, s, rayO, rayD, it, texLobes, bs.nLobes, bs.eta);
            bs.lobes = texLobes;
        }
    }
    path_sampler_init(P, x, y, s, dim, sidx, ps);
    // ---- direct lighting: UniformSampleOneLight (integrator.cpp:86-107) when the BSDF has a non-specular lobe ----
    const uint32_t nonSpec = BSDF_ALL & ~BSDF_SPECULAR;
    // the light choice (integrator.cpp:93-97): Distribution1D::SampleDiscrete(u) = the last k with cdf[k] <= u (FindInterval,
    // pbrt.h:334-350), its pdf func[k] / (funcInt * n); one light -> 0 with pdf 1.  A pdf of 0 returns before uLight / uScattering
    // are drawn.
    const WfLight *Lp = &P.light;
    float pickPdf = 1.f;
    bool direct = bsdf_num_components(bs, nonSpec) > 0;
    if (direct) {
        const float uPick = get1d(P, ps);
        if (multi) {
            uint32_t k = 0;
            if (P.voxCdf) {   // lightDistribution->Lookup(isect.p) (path.cpp:125): this point's voxel
                const size_t v = (size_t)spatial_voxel_of(P.wb, P.nVox, it.p) * P.nLights;
                while (k + 1 < P.nLights && P.voxCdf[v + k] <= uPick) ++k;
                pickPdf = P.voxPdf[v + k];
            } else {
                while (k + 1 < P.nLights && P.lights[k].cdfHi <= uPick) ++k;
                pickPdf = P.lights[k].pickPdf;
            }
            Lp = P.lights + k;
            direct = pickPdf != 0.f;
        }
    }
    M.Lp = Lp; M.pickPdf = pickPdf; M.direct = direct;
    M.ul0 = M.ul1 = M.us0 = M.us1 = 0.f;
    if (direct) {
        M.ul0 = get1d(P, ps); M.ul1 = get1d(P, ps);   // uLight
        M.us0 = get1d(P, ps); M.us1 = get1d(P, ps);   // uScattering
    }
    o.dim = ps.dim;
    return true;
}

// EstimateDirect (integrator.cpp:109-217), the light-sampled strategy: the shadow ray and what it carries
template <bool FULL>
WF_HD void shade_vertex_light(const WfParams &P, V3 beta, const ShadeMid &M, ShadeOut &o) {
    if (!M.direct) return;
    const WfIsect &it = M.it;
    const WfBsdf &bs = M.bs;
    const uint32_t nonSpec = BSDF_ALL & ~BSDF_SPECULAR;
    const WfLight &Lt = *M.Lp;
    const float pickPdf = M.pickPdf, ul0 = M.ul0, ul1 = M.ul1;
    const bool area = FULL && Lt.kind == B200RT_LIGHT_SPHERE, infinite = FULL && Lt.kind == B200RT_LIGHT_INFINITE;
    {
        V3 wi, so, sd;
        float lightPdf = 0.f, scatPdf = 0.f;
        const V3 Li = light_sample<FULL>(Lt, it, ul0, ul1, wi, lightPdf, so, sd);
        const bool delta = !area && !infinite;
        if (lightPdf > 0 && !spec_black(Li)) {
            const V3 f = vmul(bsdf_f<FULL && B200RT_SHADE_OUTLINE>(bs, it.wo, wi, nonSpec), fabsf(vdot(wi, it.ns)));
            scatPdf = bsdf_pdf<FULL && B200RT_SHADE_OUTLINE>(bs, it.wo, wi, nonSpec);
            if (!spec_black(f)) {
                V3 Ld;
                if (delta) Ld = spec_div(spec_mul(f, Li), lightPdf);
                else {
                    const float wgt = (lightPdf * lightPdf) / (lightPdf * lightPdf + scatPdf * scatPdf);   // PowerHeuristic(1, lightPdf, 1, scatPdf)
                    Ld = spec_div(vmul(spec_mul(f, Li), wgt), lightPdf);
                }
                o.contrib = spec_mul(beta, spec_div(Ld, pickPdf));   // / lightPdf of the light choice, * beta (path.cpp:133)
                o.so = so; o.sd = sd;
                o.wantShadow = true;
            }
        }
    }
}

// ... and its BSDF-sampled strategy (integrator.cpp:166-214), for lights with area: the MIS ray and what it carries
template <bool FULL>
WF_HD void shade_vertex_mis(const WfParams &P, V3 beta, const ShadeMid &M, ShadeOut &o) {
    if (!FULL || !M.direct) return;
    const WfIsect &it = M.it;
    const WfBsdf &bs = M.bs;
    const uint32_t nonSpec = BSDF_ALL & ~BSDF_SPECULAR;
    const WfLight &Lt = *M.Lp;
    const float pickPdf = M.pickPdf, us0 = M.us0, us1 = M.us1;
    const bool area = Lt.kind == B200RT_LIGHT_SPHERE, infinite = Lt.kind == B200RT_LIGHT_INFINITE;
    const bool delta = !area && !infinite;
    float scatPdf = 0.f;
    {
        if (!delta) {
            V3 wiS;
            uint32_t st;
            V3 f = bsdf_sample<FULL && B200RT_SHADE_OUTLINE>(bs, it.wo, us0, us1, nonSpec, wiS, scatPdf, st);
            f = vmul(f, fabsf(vdot(wiS, it.ns)));
            if (!spec_black(f) && scatPdf > 0) {
                const float lp = light_pdf(Lt, it, wiS);
                if (lp != 0) {
                    const float wgt = (scatPdf * scatPdf) / (scatPdf * scatPdf + lp * lp);
                    // Ld += f * Li * Tr * weight / scatteringPdf, Li = Lemit when the ray reaches the light
                    const V3 Ld = spec_div(vmul(spec_mul(f, spec3(Lt.I)), wgt), scatPdf);
                    o.mcontrib = spec_mul(beta, spec_div(Ld, pickPdf));
                    o.misPrim = area ? Lt.prim : (uint32_t)B200RT_MISS;
                    o.md = wiS;
                    o.mo = offset_ray_origin_ref(it.p, it.pError, it.n, wiS);
                    o.wantMis = true;
                }
            }
        }
    }
}

// The first three quarters in one call (the lean kernel and the host build)
template <bool FULL>
WF_HD bool shade_vertex_direct(const WfParams &P, uint32_t depth, uint32_t x, uint32_t y, uint32_t s, uint64_t sidx, uint32_t dim, V3 rayO, V3 rayD,
                               uint32_t prim, float t, float b0, float b1, float b2, V3 p0, V3 p1, V3 p2, bool isSphere,
                               V3 beta, float betaW, ShadeOut &o, ShadeMid &M) {
    if (!shade_vertex_setup<FULL>(P, depth, x, y, s, sidx, dim, rayO, rayD, prim, t, b0, b1, b2, p0, p1, p2, isSphere, beta, betaW, o, M)) return false;
    shade_vertex_light<FULL>(P, beta, M, o);
    shade_vertex_mis<FULL>(P, beta, M, o);
    return true;
}

template <bool FULL>
WF_HD void shade_vertex_bounce(const WfParams &P, uint32_t depth, V3 rayD, V3 beta, ShadeMid &M, ShadeOut &o) {
    const WfIsect &it = M.it;
    const WfBsdf &bs = M.bs;
    PathSampler &ps = M.ps;
    const float etaScale = M.etaScale;
    // ---- the path's next direction (path.cpp:142-164) ----
    const float ub0 = get1d(P, ps), ub1 = get1d(P, ps);
    o.dim = ps.dim;
    V3 wi;
    float pdf;
    uint32_t flags;
    const V3 woPath = vneg(rayD);   // path.cpp:143 `Vector3f wo = -ray.d`: NOT isect.wo, i.e. not renormalised
    const V3 f = bsdf_sample<FULL && B200RT_SHADE_OUTLINE>(bs, woPath, ub0, ub1, BSDF_ALL, wi, pdf, flags);
    if (spec_black(f) || pdf == 0.f) return;
    V3 nb = spec_mul(beta, spec_div(vmul(f, fabsf(vdot(wi, it.ns))), pdf));
    float es = etaScale;
    const bool spec = (flags & BSDF_SPECULAR) != 0;
    if (FULL && spec && (flags & BSDF_TRANSMISSION)) {
        const float eta = bs.eta;
        es *= (vdot(woPath, it.n) > 0) ? (eta * eta) : 1.f / (eta * eta);
    }
    // Russian roulette (path.cpp:190-199)
    const V3 rr = vmul(nb, es);
    const float mx = smax(smax(rr.x, rr.y), rr.z);
    if (mx < 1.f && depth > 3) {
        const float q = smax(.05f, 1.f - mx);
        const float u = get1d(P, ps);
        o.dim = ps.dim;
        if (u < q) return;
        nb = spec_div(nb, 1.f - q);
    }
    o.beta = nb;
    o.betaW = (FULL && P.has_specular) ? (spec ? -es : es) : 0.f;
    o.nd = wi;
    o.no = offset_ray_origin_ref(it.p, it.pError, it.n, wi);
    o.wantNext = true;
}


// One bounce of PathIntegrator::Li for a path whose ray (rayO, rayD) found `hit` (both halves, as the host build and the lean
// kernel run them).
template <bool FULL>
WF_HD void shade_vertex(const WfParams &P, uint32_t depth, uint32_t x, uint32_t y, uint32_t s, uint64_t sidx, uint32_t dim, V3 rayO, V3 rayD,
                        uint32_t prim, float t, float b0, float b1, float b2, V3 p0, V3 p1, V3 p2, bool isSphere,
                        V3 beta, float betaW, ShadeOut &o) {
    ShadeMid M;
    if (shade_vertex_direct<FULL>(P, depth, x, y, s, sidx, dim, rayO, rayD, prim, t, b0, b1, b2, p0, p1, p2, isSphere, beta, betaW, o, M))
        shade_vertex_bounce<FULL>(P, depth, rayD, beta, M, o);
}

}  // namespace b200rt
