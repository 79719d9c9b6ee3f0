// `Texture "..." "spectrum" "imagemap"` on the device (host/device like the rest of the shade stage): what the reference does
// between a surface hit and the Spectrum a material reads from its texture, in the reference's float operations:
//
//   ray differentials of camera rays   PerspectiveCamera::GenerateRayDifferential (cameras/perspective.cpp:81-143, pinhole),
//                                      Transform::operator()(RayDifferential) (core/transform.h), ScaleDifferentials(1 / sqrt(spp))
//                                      (core/integrator.cpp:289-291, geometry.h:1225-1230)
//   SurfaceInteraction::ComputeDifferentials   core/interaction.cpp:102-152 (+ SolveLinearSystem2x2, transform.cpp:41-49)
//   UVMapping2D::Map                   core/texture.cpp
//   ImageTexture::Evaluate, MIPMap::Lookup (trilinear), MIPMap::triangle, MIPMap::Texel (ImageWrap::Repeat)
//                                      textures/imagemap.h:83-90, core/mipmap.h:205-275
//
// Bounce rays carry no differentials (PathIntegrator spawns plain Rays, path.cpp:151), so every hit after the camera ray's looks
// the texture up with width 0: Log2(1e-8) puts the level below 0 and the lookup is MIPMap::triangle(0, st).
//
// Log2(x) = std::log(x) * invLog2 (pbrt.h:328-331) is glibc's logf, which is not correctly rounded (0.82 ulp): log_ref restates
// its algorithm (sysdeps/ieee754/flt-32/e_logf.c = ARM optimized-routines: 16-entry table, degree-3 polynomial in double) with
// the table read from this image's libm.so.6; tests/test_shade_host.py compares it with logf on 2^26 arguments.
// Only "bool trilinear" "true" textures are supported (what scenes/living-room asks for); the EWA filter is not built.
#pragma once
#include "wf_core.cuh"

namespace b200rt {

WF_HD double u2d(uint64_t u) {
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long)u);
#else
    double d; memcpy(&d, &u, 8); return d;
#endif
}

// glibc 2.39 logf for finite x > 0 (the callers pass max(width, 1e-8))
WF_HD float log_ref(float x) {
    // __logf_data: tab[16] = {invc, logc}, ln2, poly[3]
    static const uint64_t T[32] = {
        0x3ff661ec79f8f3beull, 0xbfd57bf7808caadeull, 0x3ff571ed4aaf883dull, 0xbfd2bef0a7c06ddbull, 0x3ff49539f0f010b0ull, 0xbfd01eae7f513a67ull,
        0x3ff3c995b0b80385ull, 0xbfcb31d8a68224e9ull, 0x3ff30d190c8864a5ull, 0xbfc6574f0ac07758ull, 0x3ff25e227b0b8ea0ull, 0xbfc1aa2bc79c8100ull,
        0x3ff1bb4a4a1a343full, 0xbfba4e76ce8c0e5eull, 0x3ff12358f08ae5baull, 0xbfb1973c5a611cccull, 0x3ff0953f419900a7ull, 0xbfa252f438e10c1eull,
        0x3ff0000000000000ull, 0x0000000000000000ull, 0x3fee608cfd9a47acull, 0x3faaa5aa5df25984ull, 0x3feca4b31f026aa0ull, 0x3fbc5e53aa362eb4ull,
        0x3feb2036576afce6ull, 0x3fc526e57720db08ull, 0x3fe9c2d163a1aa2dull, 0x3fcbc2860d224770ull, 0x3fe886e6037841edull, 0x3fd1058bc8a07ee1ull,
        0x3fe767dcf5534862ull, 0x3fd4043057b6ee09ull};
    const double Ln2 = u2d(0x3fe62e42fefa39efull);
    const double A0 = u2d(0xbfd00ea348b88334ull), A1 = u2d(0x3fd5575b0be00b6aull), A2 = u2d(0xbfdffffef20a4123ull);
    uint32_t ix = f2u(x);
    if (ix == 0x3f800000u) return 0.f;
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) {   // subnormal (the other special cases cannot reach here)
        ix = f2u(x * 0x1p23f);
        ix -= 23u << 23;
    }
    const uint32_t tmp = ix - 0x3f330000u;
    const uint32_t i = (tmp >> 19) & 15u;
    const int32_t k = (int32_t)tmp >> 23;
    const uint32_t iz = ix - (tmp & 0xff800000u);
    const double invc = u2d(T[2 * i]), logc = u2d(T[2 * i + 1]);
    const double z = (double)u2f(iz);
    const double r = z * invc - 1;
    const double y0 = logc + (double)k * Ln2;
    const double r2 = r * r;
    double y = A1 * r + A2;
    y = A0 * r2 + y;
    y = y * r2 + (y0 + r);
    return (float)y;
}
WF_HD float log2_ref(float x) { return log_ref(x) * 1.442695040888963387004650940071f; }   // pbrt.h:328-331

// One image texture: the MIPMap pyramid (RGB floats, level after level, each row-major s fastest) and the UVMapping2D.
struct WfTexture {
    const float *texels;
    uint32_t nLevels, w0, h0;   // level l is max(1, w0 >> l) x max(1, h0 >> l)
    uint32_t levelOffset[16];   // in texels
    float su, sv, du, dv;
    uint32_t ewa;               // "bool trilinear" false (the ImageTexture default): the elliptically weighted average filter
    float maxAniso;             // "float maxanisotropy" (8)
    const float *lut;           // MIPMap::weightLut: WeightLUTSize = 128 Gaussian weights (mipmap.h:94-101)
};

WF_HD uint32_t smax_u(uint32_t a, uint32_t b) { return a < b ? b : a; }
WF_HD int tex_mod(int a, int b) { const int r = a - (a / b) * b; return r < 0 ? r + b : r; }   // Mod(), pbrt.h
WF_HD V3 tex_texel(const WfTexture &T, int level, int s, int t) {   // MIPMap::Texel, ImageWrap::Repeat
    const int w = (int)smax_u(1u, T.w0 >> level), h = (int)smax_u(1u, T.h0 >> level);
    s = tex_mod(s, w); t = tex_mod(t, h);
    const float *p = T.texels + 3 * ((size_t)T.levelOffset[level] + (size_t)t * w + s);
    return v3(p[0], p[1], p[2]);
}
WF_HD V3 tex_triangle(const WfTexture &T, int level, float st0, float st1) {   // MIPMap::triangle, mipmap.h:262-275
    const int nL = (int)T.nLevels;
    level = level < 0 ? 0 : (level > nL - 1 ? nL - 1 : level);
    const int w = (int)smax_u(1u, T.w0 >> level), h = (int)smax_u(1u, T.h0 >> level);
    const float s = st0 * w - 0.5f, t = st1 * h - 0.5f;
    const int s0 = (int)floorf(s), t0 = (int)floorf(t);
    const float ds = s - s0, dt = t - t0;
    const V3 a = tex_texel(T, level, s0, t0), b = tex_texel(T, level, s0, t0 + 1), c = tex_texel(T, level, s0 + 1, t0),
             d = tex_texel(T, level, s0 + 1, t0 + 1);
    const float wa = (1 - ds) * (1 - dt), wb = (1 - ds) * dt, wc = ds * (1 - dt), wd = ds * dt;
    return v3(wa * a.x + wb * b.x + wc * c.x + wd * d.x, wa * a.y + wb * b.y + wc * c.y + wd * d.y, wa * a.z + wb * b.z + wc * c.z + wd * d.z);
}
// MIPMap::Lookup(st, width) (mipmap.h:244-260); Lerp(t, a, b) = (1 - t) * a + t * b (pbrt.h)
WF_HD V3 tex_lookup(const WfTexture &T, float st0, float st1, float width) {
    const int nL = (int)T.nLevels;
    const float level = nL - 1 + log2_ref(smax(width, 1e-8f));
    if (level < 0) return tex_triangle(T, 0, st0, st1);
    if (level >= nL - 1) return tex_texel(T, nL - 1, 0, 0);
    const int iLevel = (int)floorf(level);
    const float delta = level - iLevel;
    const V3 a = tex_triangle(T, iLevel, st0, st1), b = tex_triangle(T, iLevel + 1, st0, st1);
    return v3((1 - delta) * a.x + delta * b.x, (1 - delta) * a.y + delta * b.y, (1 - delta) * a.z + delta * b.z);
}

// MIPMap::EWA (mipmap.h:305-353): the texels of `level` inside the ellipse the two differentials span, Gaussian-weighted
WF_HD V3 tex_ewa(const WfTexture &T, int level, float st0, float st1, float d00, float d01, float d10, float d11) {
    const int nL = (int)T.nLevels;
    if (level >= nL) return tex_texel(T, nL - 1, 0, 0);
    const int w = (int)smax_u(1u, T.w0 >> level), h = (int)smax_u(1u, T.h0 >> level);
    st0 = st0 * w - 0.5f; st1 = st1 * h - 0.5f;
    d00 *= w; d01 *= h; d10 *= w; d11 *= h;
    float A = d01 * d01 + d11 * d11 + 1;
    float B = -2 * (d00 * d01 + d10 * d11);
    float C = d00 * d00 + d10 * d10 + 1;
    const float invF = 1 / (A * C - B * B * 0.25f);
    A *= invF; B *= invF; C *= invF;
    const float det = -B * B + 4 * A * C;
    const float invDet = 1 / det;
    const float uSqrt = sqrtf(det * C), vSqrt = sqrtf(A * det);
    const int s0 = (int)ceilf(st0 - 2 * invDet * uSqrt), s1 = (int)floorf(st0 + 2 * invDet * uSqrt);
    const int t0 = (int)ceilf(st1 - 2 * invDet * vSqrt), t1 = (int)floorf(st1 + 2 * invDet * vSqrt);
    V3 sum = v3(0.f, 0.f, 0.f);
    float sumWts = 0;
    for (int it = t0; it <= t1; ++it) {
        const float tt = it - st1;
        for (int is = s0; is <= s1; ++is) {
            const float ss = is - st0;
            const float r2 = A * ss * ss + B * ss * tt + C * tt * tt;
            if (r2 < 1) {
                int index = (int)(r2 * 128);
                index = index < 127 ? index : 127;
                const float weight = T.lut[index];
                const V3 x = tex_texel(T, level, is, it);
                sum = v3(sum.x + x.x * weight, sum.y + x.y * weight, sum.z + x.z * weight);
                sumWts += weight;
            }
        }
    }
    return v3(sum.x / sumWts, sum.y / sumWts, sum.z / sumWts);
}
// MIPMap::Lookup(st, dst0, dst1) without "trilinear" (mipmap.h:283-303)
WF_HD V3 tex_lookup_ewa(const WfTexture &T, float st0, float st1, float d00, float d01, float d10, float d11) {
    if (d00 * d00 + d01 * d01 < d10 * d10 + d11 * d11) { float x = d00; d00 = d10; d10 = x; x = d01; d01 = d11; d11 = x; }
    const float majorLength = sqrtf(d00 * d00 + d01 * d01);
    float minorLength = sqrtf(d10 * d10 + d11 * d11);
    if (minorLength * T.maxAniso < majorLength && minorLength > 0) {   // clamp the ellipse's eccentricity
        const float scale = majorLength / (minorLength * T.maxAniso);
        d10 *= scale; d11 *= scale;
        minorLength *= scale;
    }
    if (minorLength == 0) return tex_triangle(T, 0, st0, st1);
    const float lod = smax(0.f, (float)(int)T.nLevels - 1.f + log2_ref(minorLength));
    const int ilod = (int)floorf(lod);
    const float delta = lod - ilod;
    const V3 a = tex_ewa(T, ilod, st0, st1, d00, d01, d10, d11), b = tex_ewa(T, ilod + 1, st0, st1, d00, d01, d10, d11);
    return v3((1 - delta) * a.x + delta * b.x, (1 - delta) * a.y + delta * b.y, (1 - delta) * a.z + delta * b.z);
}

// (dudx, dvdx, dudy, dvdy) of SurfaceInteraction::ComputeDifferentials; all zero without ray differentials
struct WfUVDiff { float dudx, dvdx, dudy, dvdy; };

// ImageTexture<RGBSpectrum, Spectrum>::Evaluate with "trilinear": UVMapping2D::Map, then MIPMap::Lookup(st, dst0, dst1) ->
// Lookup(st, 2 * max |dst|) (mipmap.h:277-283); convertOut (ToRGB / FromRGB) is the identity for RGBSpectrum
WF_HD V3 tex_evaluate(const WfTexture &T, float u, float v, const WfUVDiff &D) {
    const float dsdx = T.su * D.dudx, dtdx = T.sv * D.dvdx, dsdy = T.su * D.dudy, dtdy = T.sv * D.dvdy;
    const float st0 = T.su * u + T.du, st1 = T.sv * v + T.dv;
    if (T.ewa) return tex_lookup_ewa(T, st0, st1, dsdx, dtdx, dsdy, dtdy);   // dst0 = dstdx, dst1 = dstdy (imagemap.h:71-77)
    const float width = smax(smax(fabsf(dsdx), fabsf(dtdx)), smax(fabsf(dsdy), fabsf(dtdy)));
    return tex_lookup(T, st0, st1, 2 * width);
}

// The two offset rays of PerspectiveCamera::GenerateRayDifferential (pinhole; perspective.cpp:54-58,128-140) for the film point
// (fx, fy), taken to world space by Transform::operator()(RayDifferential) (the offset origins are plain point transforms, the
// main origin is the error-adjusted one camera_ray_ref returns in `o`) and scaled by ScaleDifferentials(s), s = 1 / sqrt(spp).
WF_HD V3 r2c_point(const float *m, float x, float y, float z) {   // Transform::operator()(Point3f), transform.h:222-233
    const float xp = m[0] * x + m[1] * y + m[2] * z + m[3];
    const float yp = m[4] * x + m[5] * y + m[6] * z + m[7];
    const float zp = m[8] * x + m[9] * y + m[10] * z + m[11];
    const float wp = m[12] * x + m[13] * y + m[14] * z + m[15];
    if (wp == 1.f) return v3(xp, yp, zp);
    const float inv = 1.f / wp;
    return v3(inv * xp, inv * yp, inv * zp);
}
WF_HD void camera_ray_diff_ref(const CamMat &C, float fx, float fy, float s, V3 o, V3 d, V3 &rxO, V3 &rxD, V3 &ryO, V3 &ryD) {
    const float *m = C.r2c, *w = C.c2w;
    const V3 p0 = r2c_point(m, 0.f, 0.f, 0.f);
    const V3 dxCamera = vsub(r2c_point(m, 1.f, 0.f, 0.f), p0), dyCamera = vsub(r2c_point(m, 0.f, 1.f, 0.f), p0);
    const V3 pc = r2c_point(m, fx, fy, 0.f);
    const V3 dxc = vnormalize(vadd(pc, dxCamera)), dyc = vnormalize(vadd(pc, dyCamera));
    const V3 oc = r2c_point(w, 0.f, 0.f, 0.f);   // CameraToWorld(rxOrigin = (0, 0, 0)), no error adjustment
    const V3 dxw = v3(w[0] * dxc.x + w[1] * dxc.y + w[2] * dxc.z, w[4] * dxc.x + w[5] * dxc.y + w[6] * dxc.z, w[8] * dxc.x + w[9] * dxc.y + w[10] * dxc.z);
    const V3 dyw = v3(w[0] * dyc.x + w[1] * dyc.y + w[2] * dyc.z, w[4] * dyc.x + w[5] * dyc.y + w[6] * dyc.z, w[8] * dyc.x + w[9] * dyc.y + w[10] * dyc.z);
    rxO = vadd(o, vmul(vsub(oc, o), s));
    ryO = vadd(o, vmul(vsub(oc, o), s));
    rxD = vadd(d, vmul(vsub(dxw, d), s));
    ryD = vadd(d, vmul(vsub(dyw, d), s));
}

WF_HD bool solve2x2(float a00, float a01, float a10, float a11, float b0, float b1, float &x0, float &x1) {   // transform.cpp:41-49
    const float det = a00 * a11 - a01 * a10;
    if (fabsf(det) < 1e-10f) return false;
    x0 = (a11 * b0 - a01 * b1) / det;
    x1 = (a00 * b1 - a10 * b0) / det;
    if (x0 != x0 || x1 != x1) return false;
    return true;
}
WF_HD float v3_at(V3 v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }

// SurfaceInteraction::ComputeDifferentials (interaction.cpp:102-152) for a ray WITH differentials: p, n and the geometric
// dpdu / dpdv of the hit, the two offset rays.
WF_HD WfUVDiff compute_differentials(V3 p, V3 n, V3 dpdu, V3 dpdv, V3 rxO, V3 rxD, V3 ryO, V3 ryD) {
    WfUVDiff D = {0.f, 0.f, 0.f, 0.f};
    const float d = vdot(n, p);
    const float tx = -(vdot(n, rxO) - d) / vdot(n, rxD);
    if (tx != tx || tx - tx != 0.f) return D;   // isinf || isnan
    const V3 px = v3(rxO.x + tx * rxD.x, rxO.y + tx * rxD.y, rxO.z + tx * rxD.z);
    const float ty = -(vdot(n, ryO) - d) / vdot(n, ryD);
    if (ty != ty || ty - ty != 0.f) return D;
    const V3 py = v3(ryO.x + ty * ryD.x, ryO.y + ty * ryD.y, ryO.z + ty * ryD.z);
    int d0, d1;
    if (fabsf(n.x) > fabsf(n.y) && fabsf(n.x) > fabsf(n.z)) { d0 = 1; d1 = 2; }
    else if (fabsf(n.y) > fabsf(n.z)) { d0 = 0; d1 = 2; }
    else { d0 = 0; d1 = 1; }
    const float a00 = v3_at(dpdu, d0), a01 = v3_at(dpdv, d0), a10 = v3_at(dpdu, d1), a11 = v3_at(dpdv, d1);
    const float bx0 = v3_at(px, d0) - v3_at(p, d0), bx1 = v3_at(px, d1) - v3_at(p, d1);
    const float by0 = v3_at(py, d0) - v3_at(p, d0), by1 = v3_at(py, d1) - v3_at(p, d1);
    if (!solve2x2(a00, a01, a10, a11, bx0, bx1, D.dudx, D.dvdx)) D.dudx = D.dvdx = 0.f;
    if (!solve2x2(a00, a01, a10, a11, by0, by1, D.dudy, D.dvdy)) D.dudy = D.dvdy = 0.f;
    return D;
}

}  // namespace b200rt
