// Host/device arithmetic of the wavefront shade stage, written so that it reproduces the reference's float operations in
// the reference's order (compiled with -fmad=false / -ffp-contract=off; see rt_math.cuh for the traversal's contract).
//
//   vector helpers      : src/core/geometry.h (Cross in double :1287-1293, Normalize = v * (1 / Length) :281-286,1323-1325,
//                         CoordinateSystem :1360-1367, Faceforward :1555-1557, OffsetRayOrigin :1782-1802)
//   NextFloatUp / Down  : src/core/pbrt.h
//   Transform of points / vectors / rays with error bounds : src/core/transform.h:222-300
//
// Every function is __host__ __device__: nvcc compiles the device side into libb200rt.so; tests/hostshade compiles the SAME
// text with g++ so that the restatement can be pinned against the compiled reference without a GPU (test infrastructure —
// the product library exports no host execution path of the shade stage).
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define WF_HD __host__ __device__ __forceinline__
#define WF_HD_NOINLINE __host__ __device__ __noinline__
#else
#define WF_HD inline
#define WF_HD_NOINLINE inline
#endif

namespace b200rt {

struct V3 { float x, y, z; };

constexpr float kPiF = 3.14159265358979323846f;
constexpr float kInvPiF = 0.31830988618379067154f;
constexpr float kInv2PiF = 0.15915494309189533577f;
constexpr float kPiOver2F = 1.57079632679489661923f;
constexpr float kPiOver4F = 0.78539816339744830961f;
constexpr float kOneMinusEps = 0.99999994f;       // FloatOneMinusEpsilon = 0x1.fffffep-1 (rng.h)
constexpr float kShadowEps = 0.0001f;             // pbrt.h:204
constexpr float kMachEps = 5.9604644775390625e-08f;   // std::numeric_limits<float>::epsilon() * 0.5 (pbrt.h:201)
WF_HD constexpr float wf_gamma(int n) { return (n * kMachEps) / (1 - n * kMachEps); }

WF_HD uint32_t f2u(float f) {
#ifdef __CUDA_ARCH__
    return __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
WF_HD float u2f(uint32_t u) {
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}
WF_HD float wf_inf() { return u2f(0x7f800000u); }
// std::max / std::min / Clamp with the reference's comparison forms (they differ from fmaxf/fminf for NaN operands)
WF_HD float smax(float a, float b) { return (a < b) ? b : a; }
WF_HD float smin(float a, float b) { return (b < a) ? b : a; }
WF_HD float clampf(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }
// pbrt.h NextFloatUp / NextFloatDown
WF_HD float next_float_up(float v) {
    if (isinf(v) && v > 0.f) return v;
    if (v == -0.f) v = 0.f;
    uint32_t ui = f2u(v);
    if (v >= 0) ++ui; else --ui;
    return u2f(ui);
}
WF_HD float next_float_down(float v) {
    if (isinf(v) && v < 0.f) return v;
    if (v == 0.f) v = -0.f;
    uint32_t ui = f2u(v);
    if (v > 0) --ui; else ++ui;
    return u2f(ui);
}

// sinf / cosf as the reference's libm gives them.  The reference calls std::sin / std::cos on floats (sampling.cpp:129,
// geometry.h:1805-1813, infinite.cpp:104-105), i.e. glibc's sinf / cosf — since glibc 2.28 the double-precision polynomial
// scheme of the ARM optimized routines (sysdeps/ieee754/flt-32/s_sinf.c, s_cosf.c, sincosf.h; pinned version on this image:
// glibc 2.39): reduce x by multiples of pi/2 in double, evaluate a degree-7 (sine) or degree-8 (cosine) polynomial in double,
// round once to float.  Its result is NOT the correctly rounded sine (the polynomial's error is ~2^-30: it differs from the
// correctly rounded value on ~1.5 % of the arguments), so matching the reference's sampled directions bit for bit means
// restating that algorithm — published constants below (tests/test_shade_host.py checks this text against the libm it runs
// with on 4e7 arguments).  Arguments of 120 and above (never produced by the samplers here) fall back to a double evaluation.
struct SinCosTab { double c0, c1, c2, c3, c4, s1, s2, s3; };
WF_HD float sincosf_poly(double x, double x2, bool negCos, int n) {
    // __sincosf_table[0] / [1] (sincosf_data.c): the second entry negates the cosine coefficients
    const double sg = negCos ? -1.0 : 1.0;
    const double c0 = sg * 0x1p0, c1 = sg * -0x1.ffffffd0c621cp-2, c2 = sg * 0x1.55553e1068f19p-5, c3 = sg * -0x1.6c087e89a359dp-10,
                 c4 = sg * 0x1.99343027bf8c3p-16;
    const double s1 = -0x1.555545995a603p-3, s2 = 0x1.1107605230bc4p-7, s3 = -0x1.994eb3774cf24p-13;
    if ((n & 1) == 0) {   // sinf_poly, sine branch
        const double x3 = x * x2, t1 = s2 + x2 * s3, x7 = x3 * x2, sv = x + x3 * s1;
        return (float)(sv + x7 * t1);
    }
    const double x4 = x2 * x2, t2 = c3 + x2 * c4, t1 = c0 + x2 * c1, x6 = x4 * x2, cv = t1 + x4 * c2;
    return (float)(cv + x6 * t2);
}
WF_HD uint32_t abstop12(float x) { return (f2u(x) >> 20) & 0x7ffu; }
WF_HD double sincosf_reduce(double x, int &n) {   // reduce_fast without TOINT_INTRINSICS (x86-64): hpi_inv is prescaled by 2^24
    const double r = x * 0x1.45F306DC9C883p+23;
    n = ((int32_t)r + 0x800000) >> 24;
    return x - n * 0x1.921FB54442D18p0;
}
WF_HD float sin_ref(float y) {
    double x = y;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {   // |y| < pi/4
        if (abstop12(y) < abstop12(0x1p-12f)) return y;
        return sincosf_poly(x, x * x, false, 0);
    }
    if (abstop12(y) < abstop12(120.0f)) {
        int n;
        x = sincosf_reduce(x, n);
        const double sgn = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;   // sign[n & 3] = {1, -1, -1, 1}
        return sincosf_poly(x * sgn, x * x, (n & 2) != 0, n);
    }
    return (float)sin(x);
}
WF_HD float cos_ref(float y) {
    double x = y;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {
        if (abstop12(y) < abstop12(0x1p-12f)) return 1.0f;
        return sincosf_poly(x, x * x, false, 1);
    }
    if (abstop12(y) < abstop12(120.0f)) {
        int n;
        x = sincosf_reduce(x, n);
        const double sgn = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
        return sincosf_poly(x * sgn, x * x, (n & 2) != 0, n ^ 1);
    }
    return (float)cos(x);
}
// acosf (SphericalTheta in InfiniteAreaLight::Pdf_Li only: it enters a MIS weight, never a ray): evaluated in double and
// rounded once; glibc's float routine can differ from that in the last place.
WF_HD float acos_ref(float x) { return (float)acos((double)x); }

WF_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
WF_HD V3 vadd(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
WF_HD V3 vsub(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
WF_HD V3 vmul(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
WF_HD V3 vneg(V3 a) { return v3(-a.x, -a.y, -a.z); }
WF_HD V3 vabs(V3 a) { return v3(fabsf(a.x), fabsf(a.y), fabsf(a.z)); }
WF_HD float vdot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
WF_HD float vlen2(V3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
WF_HD float vlen(V3 a) { return sqrtf(vlen2(a)); }
// Vector3::operator/(f): inv = 1 / f, then three products (geometry.h:281-286)
WF_HD V3 vdiv(V3 a, float f) { const float inv = 1.f / f; return v3(a.x * inv, a.y * inv, a.z * inv); }
WF_HD V3 vnormalize(V3 a) { return vdiv(a, vlen(a)); }
// Cross: products and differences in double, rounded once to float (geometry.h:1287-1293)
WF_HD V3 vcross(V3 a, V3 b) {
    const double ax = a.x, ay = a.y, az = a.z, bx = b.x, by = b.y, bz = b.z;
    return v3((float)((ay * bz) - (az * by)), (float)((az * bx) - (ax * bz)), (float)((ax * by) - (ay * bx)));
}
WF_HD void coordinate_system(V3 v1, V3 &v2, V3 &v3o) {
    if (fabsf(v1.x) > fabsf(v1.y)) v2 = vdiv(v3(-v1.z, 0.f, v1.x), sqrtf(v1.x * v1.x + v1.z * v1.z));
    else v2 = vdiv(v3(0.f, v1.z, -v1.y), sqrtf(v1.y * v1.y + v1.z * v1.z));
    v3o = vcross(v1, v2);
}
WF_HD V3 faceforward(V3 n, V3 v) { return (vdot(n, v) < 0.f) ? vneg(n) : n; }

// OffsetRayOrigin (geometry.h:1782-1802)
WF_HD V3 offset_ray_origin_ref(V3 p, V3 pError, V3 n, V3 w) {
    const float d = vdot(vabs(n), pError);
    V3 off = v3(d * n.x, d * n.y, d * n.z);
    if (vdot(w, n) < 0) off = vneg(off);
    V3 po = vadd(p, off);
    if (off.x > 0) po.x = next_float_up(po.x); else if (off.x < 0) po.x = next_float_down(po.x);
    if (off.y > 0) po.y = next_float_up(po.y); else if (off.y < 0) po.y = next_float_down(po.y);
    if (off.z > 0) po.z = next_float_up(po.z); else if (off.z < 0) po.z = next_float_down(po.z);
    return po;
}

// ---- camera ------------------------------------------------------------------------------------------------------
// PerspectiveCamera::GenerateRayDifferential (cameras/perspective.cpp:97-143) with lensradius 0, through the matrices the
// reference's camera holds: pCamera = RasterToCamera(pFilm) (Transform::operator()(Point3f), transform.h:222-233, with
// the projective divide), d = Normalize(pCamera), then CameraToWorld(ray) (transform.h:251-264: the origin is moved
// along d to the edge of its rounding-error box).
struct CamMat { float r2c[16], c2w[16]; };
WF_HD void camera_ray_ref(const CamMat &C, float fx, float fy, V3 &o, V3 &d) {
    const float *m = C.r2c;
    const float z = 0.f;
    const float xp = m[0] * fx + m[1] * fy + m[2] * z + m[3];
    const float yp = m[4] * fx + m[5] * fy + m[6] * z + m[7];
    const float zp = m[8] * fx + m[9] * fy + m[10] * z + m[11];
    const float wp = m[12] * fx + m[13] * fy + m[14] * z + m[15];
    V3 pc = v3(xp, yp, zp);
    if (!(wp == 1.f)) { const float inv = 1.f / wp; pc = v3(inv * xp, inv * yp, inv * zp); }
    const V3 dc = vnormalize(pc);
    // Transform::operator()(const Ray &): o = M(0,0,0) with its error bound, d = M(dc)
    const float *w = C.c2w;
    const float ox = w[0] * 0.f + w[1] * 0.f + w[2] * 0.f + w[3];
    const float oy = w[4] * 0.f + w[5] * 0.f + w[6] * 0.f + w[7];
    const float oz = w[8] * 0.f + w[9] * 0.f + w[10] * 0.f + w[11];
    const float ow = w[12] * 0.f + w[13] * 0.f + w[14] * 0.f + w[15];
    const float g3 = wf_gamma(3);
    const V3 oErr = v3(g3 * (fabsf(w[0] * 0.f) + fabsf(w[1] * 0.f) + fabsf(w[2] * 0.f) + fabsf(w[3])),
                       g3 * (fabsf(w[4] * 0.f) + fabsf(w[5] * 0.f) + fabsf(w[6] * 0.f) + fabsf(w[7])),
                       g3 * (fabsf(w[8] * 0.f) + fabsf(w[9] * 0.f) + fabsf(w[10] * 0.f) + fabsf(w[11])));
    o = v3(ox, oy, oz);
    if (!(ow == 1.f)) { const float inv = 1.f / ow; o = v3(inv * ox, inv * oy, inv * oz); }
    d = v3(w[0] * dc.x + w[1] * dc.y + w[2] * dc.z, w[4] * dc.x + w[5] * dc.y + w[6] * dc.z, w[8] * dc.x + w[9] * dc.y + w[10] * dc.z);
    const float l2 = vlen2(d);
    if (l2 > 0) {
        const float dt = vdot(vabs(d), oErr) / l2;
        o = vadd(o, vmul(d, dt));
    }
}

}  // namespace b200rt
