// Device arithmetic shared by every traversal kernel.
//
// Bit-exactness contract: this file is compiled with -fmad=false (no FMA contraction), IEEE
// division (no -use_fast_math) and keeps the operand order of the reference so that hit t,
// barycentrics and primitive ids equal the CPU reference's bit for bit:
//   slab test      : Bounds3::IntersectP(ray, invDir, dirIsNeg)   src/core/geometry.h:1754-1780
//   root clip      : Bounds3::IntersectP(ray, &t0, &t1)           src/core/geometry.h:1730-1751
//   triangle       : Triangle::Intersect / IntersectP             src/shapes/triangle.cpp:188-292, 428-530
//   constants      : MachineEpsilon, gamma(n)                     src/core/pbrt.h:201-204, 289-291
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200rt {

#define B200RT_TRI_LAST 1u        // v1.w bit 0: last record of its BVH leaf
#define B200RT_TRI_DEGENERATE 2u  // v1.w bit 1: Triangle::Intersect's tail rejects every hit (triangle.cpp:309-315)
#define B200RT_PRIM_SPHERE 4u     // v1.w bit 2: the record is a sphere; v2.w = index into the sphere parameter array

// gamma(n) = n*eps / (1 - n*eps) evaluated in float with eps = 2^-24 (pbrt.h:289-291).
__host__ __device__ constexpr float rt_gamma(int n) {
    return (n * 5.9604644775390625e-08f) / (1 - n * 5.9604644775390625e-08f);
}
// Compile-time constants (constexpr float arithmetic is IEEE binary32 round-to-nearest, i.e. the
// same values the reference computes at run time).
constexpr float kGamma2 = rt_gamma(2);
constexpr float kGamma3 = rt_gamma(3);
constexpr float kGamma5 = rt_gamma(5);
constexpr float kSlabK = 1 + 2 * rt_gamma(3);

// std::max(a, b) semantics: (a < b) ? b : a   (NOT fmaxf: differs for NaN operands)
__device__ __forceinline__ float std_max(float a, float b) { return (a < b) ? b : a; }
__device__ __forceinline__ float max3abs(float x, float y, float z) {
    return std_max(fabsf(x), std_max(fabsf(y), fabsf(z)));
}

struct RayCtx {
    float ox, oy, oz;
    float dx, dy, dz;
    float tMax;        // current (shrinking) ray.tMax
    float ix, iy, iz;  // invDir = 1/d (bvh.cpp:358, kdtreeaccel.cpp:390)
    // Triangle::Intersect per-ray invariants (triangle.cpp:208-221), hoisted: they depend on the ray only.
    int kz;
    float Sx, Sy, Sz;
};

__device__ __forceinline__ void ray_setup(RayCtx &r, const float4 a, const float4 b) {
    r.ox = a.x; r.oy = a.y; r.oz = a.z; r.tMax = a.w;
    r.dx = b.x; r.dy = b.y; r.dz = b.z;
    r.ix = 1 / r.dx; r.iy = 1 / r.dy; r.iz = 1 / r.dz;
    // MaxDimension(Abs(d)), geometry.h:1338-1340
    const float ax = fabsf(r.dx), ay = fabsf(r.dy), az = fabsf(r.dz);
    r.kz = (ax > ay) ? ((ax > az) ? 0 : 2) : ((ay > az) ? 1 : 2);
    // Permute(d, kx, ky, kz) with kx = (kz+1)%3, ky = (kx+1)%3
    const float pdx = r.kz == 0 ? r.dy : (r.kz == 1 ? r.dz : r.dx);
    const float pdy = r.kz == 0 ? r.dz : (r.kz == 1 ? r.dx : r.dy);
    const float pdz = r.kz == 0 ? r.dx : (r.kz == 1 ? r.dy : r.dz);
    r.Sx = -pdx / pdz;
    r.Sy = -pdy / pdz;
    r.Sz = 1.f / pdz;
}

// Bounds3::IntersectP(ray, invDir, dirIsNeg) split in two: the geometric part (independent of
// ray.tMax) returns the entry distance; the caller applies `tMin < ray.tMax` with the CURRENT
// tMax (needed again when a deferred far child is popped).
// lo* / hi* are already selected by dirIsNeg: lo = bounds[dirIsNeg], hi = bounds[1-dirIsNeg].
__device__ __forceinline__ bool slab_geom(const RayCtx &r, float lox, float loy, float loz, float hix, float hiy,
                                          float hiz, float &tEntry) {
    float tMin = (lox - r.ox) * r.ix;
    float tMax = (hix - r.ox) * r.ix;
    float tyMin = (loy - r.oy) * r.iy;
    float tyMax = (hiy - r.oy) * r.iy;
    tMax *= kSlabK;
    tyMax *= kSlabK;
    if (tMin > tyMax || tyMin > tMax) return false;
    if (tyMin > tMin) tMin = tyMin;
    if (tyMax < tMax) tMax = tyMax;
    float tzMin = (loz - r.oz) * r.iz;
    float tzMax = (hiz - r.oz) * r.iz;
    tzMax *= kSlabK;
    if (tMin > tzMax || tzMin > tMax) return false;
    if (tzMin > tMin) tMin = tzMin;
    if (tzMax < tMax) tMax = tzMax;
    tEntry = tMin;
    return tMax > 0;
}

// Bounds3::IntersectP(ray, &t0, &t1), geometry.h:1730-1751 (kd / RBSP root clip; own 1/d per axis).
__device__ __forceinline__ bool root_clip(const RayCtx &r, const float *b /*6*/, float &t0o, float &t1o) {
    float t0 = 0, t1 = r.tMax;
    const float o[3] = {r.ox, r.oy, r.oz};
    const float inv[3] = {r.ix, r.iy, r.iz};  // 1 / ray.d[i]: same value as the loop's own divide
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float tNear = (b[i] - o[i]) * inv[i];
        float tFar = (b[3 + i] - o[i]) * inv[i];
        if (tNear > tFar) { float s = tNear; tNear = tFar; tFar = s; }
        tFar *= kSlabK;
        t0 = tNear > t0 ? tNear : t0;
        t1 = tFar < t1 ? tFar : t1;
        if (t0 > t1) return false;
    }
    t0o = t0; t1o = t1;
    return true;
}

// Triangle::Intersect arithmetic (triangle.cpp:203-292).  v0,v1,v2 = world-space vertices.
// Returns true on a geometric hit and writes t,b0,b1,b2; the caller applies the degenerate flag.
__device__ __forceinline__ bool tri_test(const RayCtx &r, const float4 v0, const float4 v1, const float4 v2, float &tOut,
                                         float &b0o, float &b1o, float &b2o) {
    // translate (:203-205)
    const float ax = v0.x - r.ox, ay = v0.y - r.oy, az = v0.z - r.oz;
    const float bx = v1.x - r.ox, by = v1.y - r.oy, bz = v1.z - r.oz;
    const float cx = v2.x - r.ox, cy = v2.y - r.oy, cz = v2.z - r.oz;
    // permute (:208-217)
    const int kz = r.kz;
    float p0x = kz == 0 ? ay : (kz == 1 ? az : ax);
    float p0y = kz == 0 ? az : (kz == 1 ? ax : ay);
    float p0z = kz == 0 ? ax : (kz == 1 ? ay : az);
    float p1x = kz == 0 ? by : (kz == 1 ? bz : bx);
    float p1y = kz == 0 ? bz : (kz == 1 ? bx : by);
    float p1z = kz == 0 ? bx : (kz == 1 ? by : bz);
    float p2x = kz == 0 ? cy : (kz == 1 ? cz : cx);
    float p2y = kz == 0 ? cz : (kz == 1 ? cx : cy);
    float p2z = kz == 0 ? cx : (kz == 1 ? cy : cz);
    // shear (:222-227); -fmad=false keeps mul and add separate
    p0x += r.Sx * p0z; p0y += r.Sy * p0z;
    p1x += r.Sx * p1z; p1y += r.Sy * p1z;
    p2x += r.Sx * p2z; p2y += r.Sy * p2z;
    // edge functions (:230-232)
    float e0 = p1x * p2y - p1y * p2x;
    float e1 = p2x * p0y - p2y * p0x;
    float e2 = p0x * p1y - p0y * p1x;
    // double-precision fallback at edges (:235-246)
    if (e0 == 0.0f || e1 == 0.0f || e2 == 0.0f) {
        const double p2txp1ty = (double)p2x * (double)p1y;
        const double p2typ1tx = (double)p2y * (double)p1x;
        e0 = (float)(p2typ1tx - p2txp1ty);
        const double p0txp2ty = (double)p0x * (double)p2y;
        const double p0typ2tx = (double)p0y * (double)p2x;
        e1 = (float)(p0typ2tx - p0txp2ty);
        const double p1txp0ty = (double)p1x * (double)p0y;
        const double p1typ0tx = (double)p1y * (double)p0x;
        e2 = (float)(p1typ0tx - p1txp0ty);
    }
    // edge and determinant tests (:249-252)
    if ((e0 < 0 || e1 < 0 || e2 < 0) && (e0 > 0 || e1 > 0 || e2 > 0)) return false;
    const float det = e0 + e1 + e2;
    if (det == 0) return false;
    // scaled hit distance (:255-262)
    p0z *= r.Sz; p1z *= r.Sz; p2z *= r.Sz;
    const float tScaled = e0 * p0z + e1 * p1z + e2 * p2z;
    if (det < 0 && (tScaled >= 0 || tScaled < r.tMax * det)) return false;
    else if (det > 0 && (tScaled <= 0 || tScaled > r.tMax * det)) return false;
    // barycentrics and t (:265-269)
    const float invDet = 1 / det;
    const float b0 = e0 * invDet, b1 = e1 * invDet, b2 = e2 * invDet;
    const float t = tScaled * invDet;
    // conservative t > deltaT (:274-292)
    const float maxZt = max3abs(p0z, p1z, p2z);
    const float deltaZ = kGamma3 * maxZt;
    const float maxXt = max3abs(p0x, p1x, p2x);
    const float maxYt = max3abs(p0y, p1y, p2y);
    const float deltaX = kGamma5 * (maxXt + maxZt);
    const float deltaY = kGamma5 * (maxYt + maxZt);
    const float deltaE = 2 * (kGamma2 * maxXt * maxYt + deltaY * maxXt + deltaX * maxYt);
    const float maxE = max3abs(e0, e1, e2);
    const float deltaT = 3 * (kGamma3 * maxE * maxZt + deltaE * maxZt + deltaZ * maxE) * fabsf(invDet);
    if (t <= deltaT) return false;
    tOut = t; b0o = b0; b1o = b1; b2o = b2;
    return true;
}

// ---- Sphere (src/shapes/sphere.cpp:49-81, 158-190), full spheres only -----------------------------
// NextFloatUp / NextFloatDown, pbrt.h:207-253
__device__ __forceinline__ float next_up(float v) {
    if (isinf(v) && v > 0.f) return v;
    if (v == -0.f) v = 0.f;
    uint32_t ui = __float_as_uint(v);
    if (v >= 0) ++ui; else --ui;
    return __uint_as_float(ui);
}
__device__ __forceinline__ float next_down(float v) {
    if (isinf(v) && v < 0.f) return v;
    if (v == 0.f) v = -0.f;
    uint32_t ui = __float_as_uint(v);
    if (v > 0) --ui; else ++ui;
    return __uint_as_float(ui);
}
// EFloat (src/core/efloat.h:48-214, release build): value + conservative interval
struct EFloat { float v, low, high; };
__device__ __forceinline__ float std_min(float a, float b) { return (b < a) ? b : a; }
__device__ __forceinline__ EFloat ef_make(float v, float err) {
    EFloat r; r.v = v;
    if (err == 0.f) r.low = r.high = v;
    else { r.low = next_down(v - err); r.high = next_up(v + err); }
    return r;
}
__device__ __forceinline__ EFloat ef_add(EFloat a, EFloat b) {
    EFloat r; r.v = a.v + b.v; r.low = next_down(a.low + b.low); r.high = next_up(a.high + b.high); return r;
}
__device__ __forceinline__ EFloat ef_sub(EFloat a, EFloat b) {
    EFloat r; r.v = a.v - b.v; r.low = next_down(a.low - b.high); r.high = next_up(a.high - b.low); return r;
}
__device__ __forceinline__ EFloat ef_mul(EFloat a, EFloat b) {
    EFloat r; r.v = a.v * b.v;
    const float p0 = a.low * b.low, p1 = a.high * b.low, p2 = a.low * b.high, p3 = a.high * b.high;
    r.low = next_down(std_min(std_min(p0, p1), std_min(p2, p3)));
    r.high = next_up(std_max(std_max(p0, p1), std_max(p2, p3)));
    return r;
}
__device__ __forceinline__ EFloat ef_div(EFloat a, EFloat b) {
    EFloat r; r.v = a.v / b.v;
    if (b.low < 0 && b.high > 0) { r.low = -__int_as_float(0x7f800000); r.high = __int_as_float(0x7f800000); }
    else {
        const float d0 = a.low / b.low, d1 = a.high / b.low, d2 = a.low / b.high, d3 = a.high / b.high;
        r.low = next_down(std_min(std_min(d0, d1), std_min(d2, d3)));
        r.high = next_up(std_max(std_max(d0, d1), std_max(d2, d3)));
    }
    return r;
}
// sp: 4 x float4 = WorldToObject rows 0..2 (affine), (radius, -, -, -).  Transform of the ray with
// error bounds: transform.h:278-300, 337-352, 382-394; quadratic: efloat.h:267-285.  For a FULL sphere
// the clipping tests of sphere.cpp:86-103 cannot fire, so hit and tHit follow from the quadratic alone.
// noinline: cold path (one light sphere among 10^5 triangles) kept out of the traversal loops.
static __device__ __noinline__ bool sphere_test(const float4 *sp, float ox_, float oy_, float oz_, float dx_, float dy_,
                                         float dz_, float rayTMax, float &tHit) {
    const float4 m0 = __ldg(sp), m1 = __ldg(sp + 1), m2 = __ldg(sp + 2);
    const float radius = __ldg(sp + 3).x;
    float ox = m0.x * ox_ + m0.y * oy_ + m0.z * oz_ + m0.w;
    float oy = m1.x * ox_ + m1.y * oy_ + m1.z * oz_ + m1.w;
    float oz = m2.x * ox_ + m2.y * oy_ + m2.z * oz_ + m2.w;
    const float oex = kGamma3 * (fabsf(m0.x * ox_) + fabsf(m0.y * oy_) + fabsf(m0.z * oz_) + fabsf(m0.w));
    const float oey = kGamma3 * (fabsf(m1.x * ox_) + fabsf(m1.y * oy_) + fabsf(m1.z * oz_) + fabsf(m1.w));
    const float oez = kGamma3 * (fabsf(m2.x * ox_) + fabsf(m2.y * oy_) + fabsf(m2.z * oz_) + fabsf(m2.w));
    const float dex = kGamma3 * (fabsf(m0.x * dx_) + fabsf(m0.y * dy_) + fabsf(m0.z * dz_));
    const float dey = kGamma3 * (fabsf(m1.x * dx_) + fabsf(m1.y * dy_) + fabsf(m1.z * dz_));
    const float dez = kGamma3 * (fabsf(m2.x * dx_) + fabsf(m2.y * dy_) + fabsf(m2.z * dz_));
    const float dx = m0.x * dx_ + m0.y * dy_ + m0.z * dz_;
    const float dy = m1.x * dx_ + m1.y * dy_ + m1.z * dz_;
    const float dz = m2.x * dx_ + m2.y * dy_ + m2.z * dz_;
    const float lengthSquared = dx * dx + dy * dy + dz * dz;
    if (lengthSquared > 0) {
        const float dt = (fabsf(dx) * oex + fabsf(dy) * oey + fabsf(dz) * oez) / lengthSquared;
        ox += dx * dt; oy += dy * dt; oz += dz * dt;
    }
    const EFloat eox = ef_make(ox, oex), eoy = ef_make(oy, oey), eoz = ef_make(oz, oez);
    const EFloat edx = ef_make(dx, dex), edy = ef_make(dy, dey), edz = ef_make(dz, dez);
    const EFloat a = ef_add(ef_add(ef_mul(edx, edx), ef_mul(edy, edy)), ef_mul(edz, edz));
    const EFloat b = ef_mul(ef_make(2.f, 0.f), ef_add(ef_add(ef_mul(edx, eox), ef_mul(edy, eoy)), ef_mul(edz, eoz)));
    const EFloat er = ef_make(radius, 0.f);
    const EFloat c = ef_sub(ef_add(ef_add(ef_mul(eox, eox), ef_mul(eoy, eoy)), ef_mul(eoz, eoz)), ef_mul(er, er));
    const double discrim = (double)b.v * (double)b.v - 4. * (double)a.v * (double)c.v;
    if (discrim < 0.) return false;
    const double rootDiscrim = sqrt(discrim);
    const EFloat frd = ef_make((float)rootDiscrim, (float)((double)5.9604644775390625e-08f * rootDiscrim));
    EFloat q;
    if (b.v < 0) q = ef_mul(ef_make(-.5f, 0.f), ef_sub(b, frd));
    else q = ef_mul(ef_make(-.5f, 0.f), ef_add(b, frd));
    EFloat t0 = ef_div(q, a), t1 = ef_div(c, q);
    if (t0.v > t1.v) { const EFloat s = t0; t0 = t1; t1 = s; }
    if (t0.high > rayTMax || t1.low <= 0) return false;
    EFloat tShapeHit = t0;
    if (tShapeHit.low <= 0) {
        tShapeHit = t1;
        if (tShapeHit.high > rayTMax) return false;
    }
    tHit = tShapeHit.v;
    return true;
}

}  // namespace b200rt
