/*
 * oracle.c — CPU restatement of the reference's ray-intersection hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this; the product (libb200rt.so) never does.
 *
 * Parity status: PINNED.  Every function below is checked bit-for-bit against the unmodified
 * reference compiled from /root/reference (oracle/_ref/libpbrtref.so, recipe oracle/ref/Makefile)
 * by tests/test_oracle_vs_ref.py, and against the committed golden vectors that the same library
 * produced (tests/golden/, generator tests/golden/make_golden.py).
 *
 * Each function cites the reference file:line it restates (paths relative to /root/reference/src).
 * All arithmetic is IEEE binary32 with no contraction (build: gcc -O2 -ffp-contract=off, baseline
 * x86-64, like the reference's Release build) except the double-precision edge-function fallback.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { float o[3]; float tMax; float d[3]; float time; } orc_ray;
typedef struct { uint32_t prim; float t, b0, b1, b2; } orc_hit;
typedef struct {
    uint64_t rays, nodes_fetched, nodes_passed, leaf_visits, prim_tests, index_reads;
} orc_counters;
#define ORC_MISS 0xFFFFFFFFu

/* pbrt.h:201-204,289-291 */
static const float MachineEpsilon = 5.9604644775390625e-08f; /* 2^-24 = numeric_limits<float>::epsilon()*0.5 */
static inline float orc_gamma(int n) { return (n * MachineEpsilon) / (1 - n * MachineEpsilon); }

/* Default UVs of Triangle::GetUVs (shapes/triangle.h:109-121). */
static const float kDefaultUV[6] = {0, 0, 1, 0, 1, 1};

/* ------------------------------------------------------------------------------------------ */
/* Triangle::Intersect tail, shapes/triangle.cpp:294-315 (same in IntersectP only under an alpha
 * mask, :533-556, which the engine rejects at upload): the hit is "bogus" when the UV
 * parameterisation is degenerate (or dpdu x dpdv == 0) AND the geometric normal is zero.
 * Depends on the triangle only, not on the ray, so the engine precomputes it per triangle. */
int orc_tri_is_degenerate(const float *v /*9*/, const float *uv /*6 or NULL*/) {
    if (!uv) uv = kDefaultUV;
    const float *p0 = v, *p1 = v + 3, *p2 = v + 6;
    float duv02[2] = {uv[0] - uv[4], uv[1] - uv[5]};
    float duv12[2] = {uv[2] - uv[4], uv[3] - uv[5]};
    float dp02[3], dp12[3];
    for (int i = 0; i < 3; ++i) { dp02[i] = p0[i] - p2[i]; dp12[i] = p1[i] - p2[i]; }
    float determinant = duv02[0] * duv12[1] - duv02[1] * duv12[0];
    int degenerateUV = fabsf(determinant) < 1e-8; /* float < double comparison, as in the reference */
    float dpdu[3] = {0, 0, 0}, dpdv[3] = {0, 0, 0};
    if (!degenerateUV) {
        float invdet = 1 / determinant;
        for (int i = 0; i < 3; ++i) {
            /* (duv12[1] * dp02 - duv02[1] * dp12) * invdet ; (-duv12[0] * dp02 + duv02[0] * dp12) * invdet */
            dpdu[i] = (duv12[1] * dp02[i] - duv02[1] * dp12[i]) * invdet;
            dpdv[i] = (-duv12[0] * dp02[i] + duv02[0] * dp12[i]) * invdet;
        }
    }
    int need = degenerateUV;
    if (!need) {
        /* Cross() is evaluated in double (core/geometry.h:1289-1297) then narrowed. */
        double ax = dpdu[0], ay = dpdu[1], az = dpdu[2], bx = dpdv[0], by = dpdv[1], bz = dpdv[2];
        float cx = (float)((ay * bz) - (az * by));
        float cy = (float)((az * bx) - (ax * bz));
        float cz = (float)((ax * by) - (ay * bx));
        float l2 = cx * cx + cy * cy + cz * cz;
        need = (l2 == 0);
    }
    if (!need) return 0;
    float e0[3], e1[3];
    for (int i = 0; i < 3; ++i) { e0[i] = p2[i] - p0[i]; e1[i] = p1[i] - p0[i]; }
    double ax = e0[0], ay = e0[1], az = e0[2], bx = e1[0], by = e1[1], bz = e1[2];
    float nx = (float)((ay * bz) - (az * by));
    float ny = (float)((az * bx) - (ax * bz));
    float nz = (float)((ax * by) - (ay * bx));
    float l2 = nx * nx + ny * ny + nz * nz;
    return l2 == 0;
}

/* MaxComponent(Abs(Vector3f(a,b,c))) (core/geometry.h:1333-1335) = std::max(x, std::max(y, z)) with
 * std::max(a,b) = (a < b) ? b : a (differs from fmaxf for NaN operands, so the exact form is kept). */
static inline float orc_stdmax(float a, float b) { return (a < b) ? b : a; }
static inline float orc_max3abs(float x, float y, float z) {
    return orc_stdmax(fabsf(x), orc_stdmax(fabsf(y), fabsf(z)));
}

/* Triangle::Intersect, shapes/triangle.cpp:188-292 (identical arithmetic in IntersectP :428-530).
 * v = p0,p1,p2 world space.  Returns 1 and writes t,b0,b1,b2 on a hit. */
static int tri_isect(const float o[3], const float dir[3], float rayTMax, const float *v, float *tOut,
                     float *b0o, float *b1o, float *b2o) {
    float p0t[3], p1t[3], p2t[3];
    for (int i = 0; i < 3; ++i) {
        p0t[i] = v[i] - o[i];
        p1t[i] = v[3 + i] - o[i];
        p2t[i] = v[6 + i] - o[i];
    }
    float ax = fabsf(dir[0]), ay = fabsf(dir[1]), az = fabsf(dir[2]);
    int kz = (ax > ay) ? ((ax > az) ? 0 : 2) : ((ay > az) ? 1 : 2);
    int kx = kz + 1; if (kx == 3) kx = 0;
    int ky = kx + 1; if (ky == 3) ky = 0;
    float d[3] = {dir[kx], dir[ky], dir[kz]};
    float q0[3] = {p0t[kx], p0t[ky], p0t[kz]};
    float q1[3] = {p1t[kx], p1t[ky], p1t[kz]};
    float q2[3] = {p2t[kx], p2t[ky], p2t[kz]};
    float Sx = -d[0] / d[2];
    float Sy = -d[1] / d[2];
    float Sz = 1.f / d[2];
    q0[0] += Sx * q0[2]; q0[1] += Sy * q0[2];
    q1[0] += Sx * q1[2]; q1[1] += Sy * q1[2];
    q2[0] += Sx * q2[2]; q2[1] += Sy * q2[2];
    float e0 = q1[0] * q2[1] - q1[1] * q2[0];
    float e1 = q2[0] * q0[1] - q2[1] * q0[0];
    float e2 = q0[0] * q1[1] - q0[1] * q1[0];
    if (e0 == 0.0f || e1 == 0.0f || e2 == 0.0f) {
        double p2txp1ty = (double)q2[0] * (double)q1[1];
        double p2typ1tx = (double)q2[1] * (double)q1[0];
        e0 = (float)(p2typ1tx - p2txp1ty);
        double p0txp2ty = (double)q0[0] * (double)q2[1];
        double p0typ2tx = (double)q0[1] * (double)q2[0];
        e1 = (float)(p0typ2tx - p0txp2ty);
        double p1txp0ty = (double)q1[0] * (double)q0[1];
        double p1typ0tx = (double)q1[1] * (double)q0[0];
        e2 = (float)(p1typ0tx - p1txp0ty);
    }
    if ((e0 < 0 || e1 < 0 || e2 < 0) && (e0 > 0 || e1 > 0 || e2 > 0)) return 0;
    float det = e0 + e1 + e2;
    if (det == 0) return 0;
    q0[2] *= Sz; q1[2] *= Sz; q2[2] *= Sz;
    float tScaled = e0 * q0[2] + e1 * q1[2] + e2 * q2[2];
    if (det < 0 && (tScaled >= 0 || tScaled < rayTMax * det)) return 0;
    else if (det > 0 && (tScaled <= 0 || tScaled > rayTMax * det)) return 0;
    float invDet = 1 / det;
    float b0 = e0 * invDet, b1 = e1 * invDet, b2 = e2 * invDet;
    float t = tScaled * invDet;
    float maxZt = orc_max3abs(q0[2], q1[2], q2[2]);
    float deltaZ = orc_gamma(3) * maxZt;
    float maxXt = orc_max3abs(q0[0], q1[0], q2[0]);
    float maxYt = orc_max3abs(q0[1], q1[1], q2[1]);
    float deltaX = orc_gamma(5) * (maxXt + maxZt);
    float deltaY = orc_gamma(5) * (maxYt + maxZt);
    float deltaE = 2 * (orc_gamma(2) * maxXt * maxYt + deltaY * maxXt + deltaX * maxYt);
    float maxE = orc_max3abs(e0, e1, e2);
    float deltaT = 3 * (orc_gamma(3) * maxE * maxZt + deltaE * maxZt + deltaZ * maxE) * fabsf(invDet);
    if (t <= deltaT) return 0;
    *tOut = t; *b0o = b0; *b1o = b1; *b2o = b2;
    return 1;
}

/* Public single-triangle entry. degenerate = orc_tri_is_degenerate() result; pass 0 for the
 * IntersectP semantics (no tail). */
int orc_triangle(const orc_ray *r, const float *v, int degenerate, float *t, float *b0, float *b1, float *b2) {
    if (!tri_isect(r->o, r->d, r->tMax, v, t, b0, b1, b2)) return 0;
    return degenerate ? 0 : 1; /* triangle.cpp:309-315 */
}

/* ---------------------------------- Sphere ------------------------------------------------- */
/* NextFloatUp / NextFloatDown, core/pbrt.h:207-253. */
static inline float next_up(float v) {
    if (isinf(v) && v > 0.f) return v;
    if (v == -0.f) v = 0.f;
    uint32_t ui; memcpy(&ui, &v, 4);
    if (v >= 0) ++ui; else --ui;
    memcpy(&v, &ui, 4);
    return v;
}
static inline float next_down(float v) {
    if (isinf(v) && v < 0.f) return v;
    if (v == 0.f) v = -0.f;
    uint32_t ui; memcpy(&ui, &v, 4);
    if (v > 0) --ui; else ++ui;
    memcpy(&v, &ui, 4);
    return v;
}
/* EFloat, core/efloat.h:48-214 (release build: no vPrecise). */
typedef struct { float v, low, high; } efloat;
static inline float f_min4(float a, float b, float c, float d) {  /* std::min(std::min(a,b), std::min(c,d)) */
    float x = (b < a) ? b : a, y = (d < c) ? d : c;
    return (y < x) ? y : x;
}
static inline float f_max4(float a, float b, float c, float d) {
    float x = (a < b) ? b : a, y = (c < d) ? d : c;
    return (x < y) ? y : x;
}
static inline efloat ef_make(float v, float err) { /* efloat.h:52-69 */
    efloat r; r.v = v;
    if (err == 0.) r.low = r.high = v;
    else { r.low = next_down(v - err); r.high = next_up(v + err); }
    return r;
}
static inline efloat ef_add(efloat a, efloat b) { /* :76-88 */
    efloat r; r.v = a.v + b.v; r.low = next_down(a.low + b.low); r.high = next_up(a.high + b.high); return r;
}
static inline efloat ef_sub(efloat a, efloat b) { /* :100-110 */
    efloat r; r.v = a.v - b.v; r.low = next_down(a.low - b.high); r.high = next_up(a.high - b.low); return r;
}
static inline efloat ef_mul(efloat a, efloat b) { /* :111-126 */
    efloat r; r.v = a.v * b.v;
    float p0 = a.low * b.low, p1 = a.high * b.low, p2 = a.low * b.high, p3 = a.high * b.high;
    r.low = next_down(f_min4(p0, p1, p2, p3));
    r.high = next_up(f_max4(p0, p1, p2, p3));
    return r;
}
static inline efloat ef_div(efloat a, efloat b) { /* :127-149 */
    efloat r; r.v = a.v / b.v;
    if (b.low < 0 && b.high > 0) { r.low = -INFINITY; r.high = INFINITY; }
    else {
        float d0 = a.low / b.low, d1 = a.high / b.low, d2 = a.low / b.high, d3 = a.high / b.high;
        r.low = next_down(f_min4(d0, d1, d2, d3));
        r.high = next_up(f_max4(d0, d1, d2, d3));
    }
    return r;
}
/* Sphere parameters: 16 floats = WorldToObject rows 0..2 (affine 3x4, row-major), radius, 3 pad.
 * FULL spheres only (zMin = -radius, zMax = radius, phiMax = 360): then the clipping tests of
 * sphere.cpp:86-103 can never fire and hit / tHit depend on the EFloat quadratic alone.
 * Sphere::Intersect sphere.cpp:49-81 / IntersectP :158-190; returns 1 and *tHit on a hit. */
static int sphere_isect(const float *sp, const float o[3], const float d[3], float rayTMax, float *tHit) {
    const float *m0 = sp, *m1 = sp + 4, *m2 = sp + 8;
    const float radius = sp[12];
    const float g3 = orc_gamma(3);
    /* Transform::operator()(Point3f, Vector3f *pError), transform.h:278-300 (affine: wp == 1) */
    float ox = m0[0] * o[0] + m0[1] * o[1] + m0[2] * o[2] + m0[3];
    float oy = m1[0] * o[0] + m1[1] * o[1] + m1[2] * o[2] + m1[3];
    float oz = m2[0] * o[0] + m2[1] * o[1] + m2[2] * o[2] + m2[3];
    float oex = g3 * (fabsf(m0[0] * o[0]) + fabsf(m0[1] * o[1]) + fabsf(m0[2] * o[2]) + fabsf(m0[3]));
    float oey = g3 * (fabsf(m1[0] * o[0]) + fabsf(m1[1] * o[1]) + fabsf(m1[2] * o[2]) + fabsf(m1[3]));
    float oez = g3 * (fabsf(m2[0] * o[0]) + fabsf(m2[1] * o[1]) + fabsf(m2[2] * o[2]) + fabsf(m2[3]));
    /* Transform::operator()(Vector3f, Vector3f *absError), transform.h:337-352 */
    float dex = g3 * (fabsf(m0[0] * d[0]) + fabsf(m0[1] * d[1]) + fabsf(m0[2] * d[2]));
    float dey = g3 * (fabsf(m1[0] * d[0]) + fabsf(m1[1] * d[1]) + fabsf(m1[2] * d[2]));
    float dez = g3 * (fabsf(m2[0] * d[0]) + fabsf(m2[1] * d[1]) + fabsf(m2[2] * d[2]));
    float dx = m0[0] * d[0] + m0[1] * d[1] + m0[2] * d[2];
    float dy = m1[0] * d[0] + m1[1] * d[1] + m1[2] * d[2];
    float dz = m2[0] * d[0] + m2[1] * d[1] + m2[2] * d[2];
    /* Transform::operator()(Ray, oError, dError), transform.h:382-394 */
    float lengthSquared = dx * dx + dy * dy + dz * dz;
    if (lengthSquared > 0) {
        float dt = (fabsf(dx) * oex + fabsf(dy) * oey + fabsf(dz) * oez) / lengthSquared;
        ox += dx * dt; oy += dy * dt; oz += dz * dt;
    }
    /* sphere.cpp:60-65 */
    efloat eox = ef_make(ox, oex), eoy = ef_make(oy, oey), eoz = ef_make(oz, oez);
    efloat edx = ef_make(dx, dex), edy = ef_make(dy, dey), edz = ef_make(dz, dez);
    efloat a = ef_add(ef_add(ef_mul(edx, edx), ef_mul(edy, edy)), ef_mul(edz, edz));
    efloat b = ef_mul(ef_make(2.f, 0.f), ef_add(ef_add(ef_mul(edx, eox), ef_mul(edy, eoy)), ef_mul(edz, eoz)));
    efloat er = ef_make(radius, 0.f);
    efloat c = ef_sub(ef_add(ef_add(ef_mul(eox, eox), ef_mul(eoy, eoy)), ef_mul(eoz, eoz)), ef_mul(er, er));
    /* Quadratic(EFloat...), efloat.h:267-285 */
    double discrim = (double)b.v * (double)b.v - 4. * (double)a.v * (double)c.v;
    if (discrim < 0.) return 0;
    double rootDiscrim = sqrt(discrim);
    efloat frd = ef_make((float)rootDiscrim, (float)(MachineEpsilon * rootDiscrim));
    efloat q;
    if (b.v < 0) q = ef_mul(ef_make(-.5f, 0.f), ef_sub(b, frd));
    else q = ef_mul(ef_make(-.5f, 0.f), ef_add(b, frd));
    efloat t0 = ef_div(q, a), t1 = ef_div(c, q);
    if (t0.v > t1.v) { efloat s = t0; t0 = t1; t1 = s; }
    /* sphere.cpp:72-81 */
    if (t0.high > rayTMax || t1.low <= 0) return 0;
    efloat tShapeHit = t0;
    if (tShapeHit.low <= 0) {
        tShapeHit = t1;
        if (tShapeHit.high > rayTMax) return 0;
    }
    *tHit = tShapeHit.v;
    return 1;
}
int orc_sphere(const orc_ray *r, const float *params /*16*/, float *t) { return sphere_isect(params, r->o, r->d, r->tMax, t); }

/* ------------------------------------------------------------------------------------------ */
typedef struct {
    const float *verts;      /* n_prims x 9, original primitive order */
    const uint8_t *degen;    /* n_prims, orc_tri_is_degenerate flags */
    uint32_t n_prims;
    const uint8_t *kind;     /* nullable: 0 triangle, 1 sphere */
    const uint32_t *sphere_of; /* prim -> index into spheres (valid where kind == 1) */
    const float *spheres;    /* n_spheres x 16 */
} orc_prims;

/* GeometricPrimitive::Intersect, core/primitive.cpp:123-139: on a hit r.tMax = tHit. */
static inline int prim_intersect(const orc_prims *P, uint32_t prim, orc_ray *r, orc_hit *h, orc_counters *c) {
    float t, b0, b1, b2;
    if (c) c->prim_tests++;
    if (P->kind && P->kind[prim]) {
        if (!sphere_isect(P->spheres + 16 * (size_t)P->sphere_of[prim], r->o, r->d, r->tMax, &t)) return 0;
        r->tMax = t;
        h->prim = prim; h->t = t; h->b0 = h->b1 = h->b2 = 0.f; /* no barycentrics on a sphere */
        return 1;
    }
    if (!tri_isect(r->o, r->d, r->tMax, P->verts + 9 * (size_t)prim, &t, &b0, &b1, &b2)) return 0;
    if (P->degen && P->degen[prim]) return 0;
    r->tMax = t;
    h->prim = prim; h->t = t; h->b0 = b0; h->b1 = b1; h->b2 = b2;
    return 1;
}
/* GeometricPrimitive::IntersectP, core/primitive.cpp:118-121. */
static inline int prim_intersect_p(const orc_prims *P, uint32_t prim, const orc_ray *r, orc_counters *c) {
    float t, b0, b1, b2;
    if (c) c->prim_tests++;
    if (P->kind && P->kind[prim]) return sphere_isect(P->spheres + 16 * (size_t)P->sphere_of[prim], r->o, r->d, r->tMax, &t);
    if (!tri_isect(r->o, r->d, r->tMax, P->verts + 9 * (size_t)prim, &t, &b0, &b1, &b2)) return 0;
    /* NOTE: no degenerate-triangle rejection here: Triangle::IntersectP only has it inside the
     * alpha-mask branch (triangle.cpp:533-556), unlike Triangle::Intersect (:309-315). */
    return 1;
}

/* ---------------------------------- BVH ---------------------------------------------------- */
/* LinearBVHNode, accelerators/bvh.cpp:123-152: 32 bytes. */
typedef struct {
    float bmin[3], bmax[3];
    int32_t offset;    /* primitivesOffset (leaf) / secondChildOffset (interior) */
    uint32_t word;     /* axis in bits 0-1 (3 = leaf), nPrims << 2 */
} orc_bvh_node;

/* Bounds3::IntersectP(ray, invDir, dirIsNeg), core/geometry.h:1754-1780. */
static inline int slab_bvh(const orc_bvh_node *n, const orc_ray *r, const float invDir[3], const int neg[3]) {
    const float *b[2] = {n->bmin, n->bmax};
    const float k = 1 + 2 * orc_gamma(3);
    float tMin = (b[neg[0]][0] - r->o[0]) * invDir[0];
    float tMax = (b[1 - neg[0]][0] - r->o[0]) * invDir[0];
    float tyMin = (b[neg[1]][1] - r->o[1]) * invDir[1];
    float tyMax = (b[1 - neg[1]][1] - r->o[1]) * invDir[1];
    tMax *= k;
    tyMax *= k;
    if (tMin > tyMax || tyMin > tMax) return 0;
    if (tyMin > tMin) tMin = tyMin;
    if (tyMax < tMax) tMax = tyMax;
    float tzMin = (b[neg[2]][2] - r->o[2]) * invDir[2];
    float tzMax = (b[1 - neg[2]][2] - r->o[2]) * invDir[2];
    tzMax *= k;
    if (tMin > tzMax || tzMin > tMax) return 0;
    if (tzMin > tMin) tMin = tzMin;
    if (tzMax < tMax) tMax = tzMax;
    return (tMin < r->tMax) && (tMax > 0);
}

/* BVHAccel::Intersect, accelerators/bvh.cpp:354-396.  prim_order[slot] = original prim number. */
static void bvh_closest_one(const orc_bvh_node *nodes, const uint32_t *prim_order, const orc_prims *P, orc_ray *r,
                            orc_hit *h, orc_counters *c) {
    h->prim = ORC_MISS; h->t = r->tMax; h->b0 = h->b1 = h->b2 = 0;
    if (!nodes) return;
    float invDir[3] = {1 / r->d[0], 1 / r->d[1], 1 / r->d[2]};
    int neg[3] = {invDir[0] < 0, invDir[1] < 0, invDir[2] < 0};
    int toVisit = 0, cur = 0;
    int stack[64];
    for (;;) {
        const orc_bvh_node *n = &nodes[cur];
        if (c) c->nodes_fetched++;
        if (slab_bvh(n, r, invDir, neg)) {
            if (c) c->nodes_passed++;
            if ((n->word & 3) == 3) {
                if (c) c->leaf_visits++;
                uint32_t np = n->word >> 2;
                for (uint32_t i = 0; i < np; ++i) prim_intersect(P, prim_order[n->offset + i], r, h, c);
                if (toVisit == 0) break;
                cur = stack[--toVisit];
            } else {
                if (neg[n->word & 3]) { stack[toVisit++] = cur + 1; cur = n->offset; }
                else { stack[toVisit++] = n->offset; cur = cur + 1; }
            }
        } else {
            if (toVisit == 0) break;
            cur = stack[--toVisit];
        }
    }
}
/* BVHAccel::IntersectP, accelerators/bvh.cpp:398-437. */
static int bvh_any_one(const orc_bvh_node *nodes, const uint32_t *prim_order, const orc_prims *P, const orc_ray *r,
                       orc_counters *c) {
    if (!nodes) return 0;
    float invDir[3] = {1.f / r->d[0], 1.f / r->d[1], 1.f / r->d[2]};
    int neg[3] = {invDir[0] < 0, invDir[1] < 0, invDir[2] < 0};
    int toVisit = 0, cur = 0;
    int stack[64];
    for (;;) {
        const orc_bvh_node *n = &nodes[cur];
        if (c) c->nodes_fetched++;
        if (slab_bvh(n, r, invDir, neg)) {
            if (c) c->nodes_passed++;
            if ((n->word & 3) == 3) {
                if (c) c->leaf_visits++;
                uint32_t np = n->word >> 2;
                for (uint32_t i = 0; i < np; ++i)
                    if (prim_intersect_p(P, prim_order[n->offset + i], r, c)) return 1;
                if (toVisit == 0) break;
                cur = stack[--toVisit];
            } else {
                if (neg[n->word & 3]) { stack[toVisit++] = cur + 1; cur = n->offset; }
                else { stack[toVisit++] = n->offset; cur = cur + 1; }
            }
        } else {
            if (toVisit == 0) break;
            cur = stack[--toVisit];
        }
    }
    return 0;
}

/* ---------------------------------- kd / RBSP / RBSPKd -------------------------------------- */
typedef struct { uint32_t w0, w1; } orc_bsp_node; /* kdtreeaccel.cpp:146-159, rbsp.cpp:121-134 */
typedef struct {
    int kind;                 /* 1 kd, 2 rbsp, 3 rbspkd */
    const orc_bsp_node *nodes;
    const uint32_t *indices;  /* primitiveIndices */
    const float *directions;  /* M x 3 */
    uint32_t M;
    float bounds[6];
} orc_bsp_tree;

/* Bounds3::IntersectP(ray,&t0,&t1), core/geometry.h:1730-1751. */
static inline int root_clip(const float b[6], const orc_ray *r, float *hitt0, float *hitt1) {
    float t0 = 0, t1 = r->tMax;
    for (int i = 0; i < 3; ++i) {
        float invRayDir = 1 / r->d[i];
        float tNear = (b[i] - r->o[i]) * invRayDir;
        float tFar = (b[3 + i] - r->o[i]) * invRayDir;
        if (tNear > tFar) { float s = tNear; tNear = tFar; tFar = s; }
        tFar *= 1 + 2 * orc_gamma(3);
        t0 = tNear > t0 ? tNear : t0;
        t1 = tFar < t1 ? tFar : t1;
        if (t0 > t1) return 0;
    }
    *hitt0 = t0; *hitt1 = t1;
    return 1;
}
/* genericBSP.cpp:11-22: getBitOffset(M) = 32 - clz(M) ; mask = (1<<off)-1. */
static inline uint32_t bit_offset(int kind, uint32_t M) { return kind == 1 ? 2u : 32u - (uint32_t)__builtin_clz(M); }

/* intersectInterior: kdtreeaccel.cpp:95-104 + geometry.h:1833-1835 (kd);
 * rbsp.cpp:68-80 + geometry.h:1837-1843 (rbsp); rbspKd.cpp:68-92 (axis<3 fast path). */
static inline float interior(const orc_bsp_tree *T, const orc_ray *r, const float invDir[3], uint32_t axis, float split,
                             int *belowFirst) {
    if (T->kind == 1 || (T->kind == 3 && axis < 3)) {
        float tPlane = (split - r->o[axis]) * invDir[axis];
        *belowFirst = (r->o[axis] < split) || (r->o[axis] == split && r->d[axis] <= 0);
        return tPlane;
    }
    const float *dv = T->directions + 3 * axis;
    /* Dot(), core/geometry.h:1264-1267: v1.x*v2.x + v1.y*v2.y + v1.z*v2.z */
    float projectedO = dv[0] * r->o[0] + dv[1] * r->o[1] + dv[2] * r->o[2];
    float inverseProjectedD = 1 / (dv[0] * r->d[0] + dv[1] * r->d[1] + dv[2] * r->d[2]);
    *belowFirst = (projectedO < split) || (projectedO == split && inverseProjectedD <= 0);
    return (split - projectedO) * inverseProjectedD;
}

typedef struct { uint32_t node; float tMin, tMax; } orc_todo; /* genericBSP.h:195-199 */

/* KdTreeAccel::Intersect kdtreeaccel.cpp:381-452; RBSP::Intersect rbsp.cpp:405-477; RBSPKd same loop. */
static void bsp_closest_one(const orc_bsp_tree *T, const orc_prims *P, orc_ray *r, orc_hit *h, orc_counters *c) {
    h->prim = ORC_MISS; h->t = r->tMax; h->b0 = h->b1 = h->b2 = 0;
    float tMin, tMax;
    if (!root_clip(T->bounds, r, &tMin, &tMax)) return;
    float invDir[3] = {1 / r->d[0], 1 / r->d[1], 1 / r->d[2]};
    const uint32_t off = bit_offset(T->kind, T->M), mask = (1u << off) - 1u;
    const uint32_t leafTag = T->kind == 1 ? 3u : T->M;
    orc_todo todo[64];
    uint32_t todoPos = 0, cur = 0;
    for (;;) {
        if (r->tMax < tMin) break;
        if (c) c->nodes_fetched++;
        const orc_bsp_node n = T->nodes[cur];
        uint32_t axis = n.w1 & mask;
        if (axis != leafTag) {
            if (c) c->nodes_passed++;
            float split; memcpy(&split, &n.w0, 4);
            int belowFirst;
            float tPlane = interior(T, r, invDir, axis, split, &belowFirst);
            uint32_t above = n.w1 >> off, first, second;
            if (belowFirst) { first = cur + 1; second = above; } else { first = above; second = cur + 1; }
            if (tPlane > tMax || tPlane <= 0) cur = first;
            else if (tPlane < tMin) cur = second;
            else {
                todo[todoPos].node = second; todo[todoPos].tMin = tPlane; todo[todoPos].tMax = tMax; ++todoPos;
                cur = first; tMax = tPlane;
            }
        } else {
            if (c) c->leaf_visits++;
            uint32_t np = n.w1 >> off;
            if (np == 1) prim_intersect(P, n.w0, r, h, c);
            else for (uint32_t i = 0; i < np; ++i) {
                if (c) c->index_reads++;
                prim_intersect(P, T->indices[n.w0 + i], r, h, c);
            }
            if (todoPos > 0) { --todoPos; cur = todo[todoPos].node; tMin = todo[todoPos].tMin; tMax = todo[todoPos].tMax; }
            else break;
        }
    }
}
/* KdTreeAccel::IntersectP kdtreeaccel.cpp:454-521; RBSP::IntersectP rbsp.cpp:479-547. No tMax<tMin bail-out. */
static int bsp_any_one(const orc_bsp_tree *T, const orc_prims *P, const orc_ray *r, orc_counters *c) {
    float tMin, tMax;
    if (!root_clip(T->bounds, r, &tMin, &tMax)) return 0;
    float invDir[3] = {1 / r->d[0], 1 / r->d[1], 1 / r->d[2]};
    const uint32_t off = bit_offset(T->kind, T->M), mask = (1u << off) - 1u;
    const uint32_t leafTag = T->kind == 1 ? 3u : T->M;
    orc_todo todo[64];
    uint32_t todoPos = 0, cur = 0;
    for (;;) {
        if (c) c->nodes_fetched++;
        const orc_bsp_node n = T->nodes[cur];
        uint32_t axis = n.w1 & mask;
        if (axis == leafTag) {
            if (c) c->leaf_visits++;
            uint32_t np = n.w1 >> off;
            if (np == 1) { if (prim_intersect_p(P, n.w0, r, c)) return 1; }
            else for (uint32_t i = 0; i < np; ++i) {
                if (c) c->index_reads++;
                if (prim_intersect_p(P, T->indices[n.w0 + i], r, c)) return 1;
            }
            if (todoPos > 0) { --todoPos; cur = todo[todoPos].node; tMin = todo[todoPos].tMin; tMax = todo[todoPos].tMax; }
            else break;
        } else {
            if (c) c->nodes_passed++;
            float split; memcpy(&split, &n.w0, 4);
            int belowFirst;
            float tPlane = interior(T, r, invDir, axis, split, &belowFirst);
            uint32_t above = n.w1 >> off, first, second;
            if (belowFirst) { first = cur + 1; second = above; } else { first = above; second = cur + 1; }
            if (tPlane > tMax || tPlane <= 0) cur = first;
            else if (tPlane < tMin) cur = second;
            else {
                todo[todoPos].node = second; todo[todoPos].tMin = tPlane; todo[todoPos].tMax = tMax; ++todoPos;
                cur = first; tMax = tPlane;
            }
        }
    }
    return 0;
}


/* ------------------- general BSP / BSPKd (20-byte nodes with an explicit split normal) ------------------- */
/* BSPNode (BSP.h:122-184): w0 = split | onePrimitive | primitiveIndicesOffset; w1 = flags: bit 0 = leaf, the rest
 * = aboveChild << 1 or nPrims << 1 (BSP.h:13-55); splitAxis = 3 floats.
 * BSPKdNode (BSPKd.h:11-174): w1 low 3 bits: 0..2 = kd node on that axis, 3 = leaf, 4 = general node; the rest =
 * aboveChild << 3 or nPrims << 3. */
typedef struct { uint32_t w0, w1; float axis[3]; } orc_gbsp_node;
typedef struct { int kind; /* 4 BSP, 5 BSPKd */ const orc_gbsp_node *nodes; const uint32_t *indices; float bounds[6]; } orc_gbsp_tree;

static inline int gbsp_is_leaf(int kind, uint32_t w1) { return kind == 4 ? (w1 & 1u) == 1u : (w1 & 7u) == 3u; }
static inline uint32_t gbsp_shift(int kind) { return kind == 4 ? 1u : 3u; }
/* treeIntersectInterior (BSP.h:57-67) / BSPKdNode::intersectInterior (BSPKd.h:52-72) + planeDistance
 * (geometry.h:1833-1843); Dot = x*x' + y*y' + z*z' left to right. */
static inline float gbsp_interior(int kind, const orc_gbsp_node *n, const orc_ray *r, const float invDir[3], int *belowFirst) {
    float split; memcpy(&split, &n->w0, 4);
    if (kind == 5 && (n->w1 & 7u) < 4u) {
        uint32_t axis = n->w1 & 7u;
        *belowFirst = (r->o[axis] < split) || (r->o[axis] == split && r->d[axis] <= 0);
        return (split - r->o[axis]) * invDir[axis];
    }
    float projectedO = n->axis[0] * r->o[0] + n->axis[1] * r->o[1] + n->axis[2] * r->o[2];
    float inverseProjectedD = 1 / (n->axis[0] * r->d[0] + n->axis[1] * r->d[1] + n->axis[2] * r->d[2]);
    *belowFirst = (projectedO < split) || (projectedO == split && inverseProjectedD <= 0);
    return (split - projectedO) * inverseProjectedD;
}
/* BSP::Intersect BSP.cpp:27-95; BSPKd::Intersect BSPKd.cpp:25-98 */
static void gbsp_closest_one(const orc_gbsp_tree *T, const orc_prims *P, orc_ray *r, orc_hit *h, orc_counters *c) {
    h->prim = ORC_MISS; h->t = r->tMax; h->b0 = h->b1 = h->b2 = 0;
    float tMin, tMax;
    if (!root_clip(T->bounds, r, &tMin, &tMax)) return;
    float invDir[3] = {1 / r->d[0], 1 / r->d[1], 1 / r->d[2]};
    const uint32_t off = gbsp_shift(T->kind);
    orc_todo todo[64];
    uint32_t todoPos = 0, cur = 0;
    for (;;) {
        if (r->tMax < tMin) break;
        if (c) c->nodes_fetched++;
        const orc_gbsp_node *n = &T->nodes[cur];
        if (!gbsp_is_leaf(T->kind, n->w1)) {
            if (c) c->nodes_passed++;
            int belowFirst;
            float tPlane = gbsp_interior(T->kind, n, r, invDir, &belowFirst);
            uint32_t above = n->w1 >> off, first, second;
            if (belowFirst) { first = cur + 1; second = above; } else { first = above; second = cur + 1; }
            if (tPlane > tMax || tPlane <= 0) cur = first;
            else if (tPlane < tMin) cur = second;
            else {
                todo[todoPos].node = second; todo[todoPos].tMin = tPlane; todo[todoPos].tMax = tMax; ++todoPos;
                cur = first; tMax = tPlane;
            }
        } else {
            if (c) c->leaf_visits++;
            uint32_t np = n->w1 >> off;
            if (np == 1) prim_intersect(P, n->w0, r, h, c);
            else for (uint32_t i = 0; i < np; ++i) {
                if (c) c->index_reads++;
                prim_intersect(P, T->indices[n->w0 + i], r, h, c);
            }
            if (todoPos > 0) { --todoPos; cur = todo[todoPos].node; tMin = todo[todoPos].tMin; tMax = todo[todoPos].tMax; }
            else break;
        }
    }
}
/* BSP::IntersectP BSP.cpp:97-163; BSPKd::IntersectP BSPKd.cpp:100-169 (no tMax < tMin bail-out) */
static int gbsp_any_one(const orc_gbsp_tree *T, const orc_prims *P, const orc_ray *r, orc_counters *c) {
    float tMin, tMax;
    if (!root_clip(T->bounds, r, &tMin, &tMax)) return 0;
    float invDir[3] = {1 / r->d[0], 1 / r->d[1], 1 / r->d[2]};
    const uint32_t off = gbsp_shift(T->kind);
    orc_todo todo[64];
    uint32_t todoPos = 0, cur = 0;
    for (;;) {
        if (c) c->nodes_fetched++;
        const orc_gbsp_node *n = &T->nodes[cur];
        if (gbsp_is_leaf(T->kind, n->w1)) {
            if (c) c->leaf_visits++;
            uint32_t np = n->w1 >> off;
            if (np == 1) { if (prim_intersect_p(P, n->w0, r, c)) return 1; }
            else for (uint32_t i = 0; i < np; ++i) {
                if (c) c->index_reads++;
                if (prim_intersect_p(P, T->indices[n->w0 + i], r, c)) return 1;
            }
            if (todoPos > 0) { --todoPos; cur = todo[todoPos].node; tMin = todo[todoPos].tMin; tMax = todo[todoPos].tMax; }
            else break;
        } else {
            if (c) c->nodes_passed++;
            int belowFirst;
            float tPlane = gbsp_interior(T->kind, n, r, invDir, &belowFirst);
            uint32_t above = n->w1 >> off, first, second;
            if (belowFirst) { first = cur + 1; second = above; } else { first = above; second = cur + 1; }
            if (tPlane > tMax || tPlane <= 0) cur = first;
            else if (tPlane < tMin) cur = second;
            else {
                todo[todoPos].node = second; todo[todoPos].tMin = tPlane; todo[todoPos].tMax = tMax; ++todoPos;
                cur = first; tMax = tPlane;
            }
        }
    }
    return 0;
}

/* ---------------------------------- batch entry points -------------------------------------- */
#include <pthread.h>
static void add_counters(orc_counters *dst, const orc_counters *src) {
    dst->nodes_fetched += src->nodes_fetched; dst->nodes_passed += src->nodes_passed;
    dst->leaf_visits += src->leaf_visits; dst->prim_tests += src->prim_tests; dst->index_reads += src->index_reads;
}

typedef struct {
    int kind, any, want;
    const void *nodes; const uint32_t *indices;
    orc_prims P; orc_bsp_tree T; orc_gbsp_tree G;
    const orc_ray *rays; uint64_t begin, end;
    orc_hit *hits; uint8_t *occluded;
    orc_counters *per_ray; orc_counters total;
} orc_job;

static void *run_job(void *arg) {
    orc_job *j = (orc_job *)arg;
    memset(&j->total, 0, sizeof(j->total));
    for (uint64_t i = j->begin; i < j->end; ++i) {
        orc_counters c; memset(&c, 0, sizeof(c));
        orc_counters *cp = j->want ? &c : 0;
        if (j->any) {
            j->occluded[i] = (uint8_t)(j->kind == 0
                ? bvh_any_one((const orc_bvh_node *)j->nodes, j->indices, &j->P, &j->rays[i], cp)
                : j->kind >= 4 ? gbsp_any_one(&j->G, &j->P, &j->rays[i], cp)
                : bsp_any_one(&j->T, &j->P, &j->rays[i], cp));
        } else {
            orc_ray r = j->rays[i];
            if (j->kind == 0) bvh_closest_one((const orc_bvh_node *)j->nodes, j->indices, &j->P, &r, &j->hits[i], cp);
            else if (j->kind >= 4) gbsp_closest_one(&j->G, &j->P, &r, &j->hits[i], cp);
            else bsp_closest_one(&j->T, &j->P, &r, &j->hits[i], cp);
        }
        if (j->want) { c.rays = 1; add_counters(&j->total, &c); if (j->per_ray) j->per_ray[i] = c; }
    }
    return 0;
}

/* kind 0: nodes = orc_bvh_node[], indices = prim_order.  kind 1/2/3: orc_bsp_node[], primitiveIndices.
 * kind 4/5: orc_gbsp_node[] (20 B), primitiveIndices.
 * counters (nullable) is a single accumulated struct; per_ray (nullable) gets one struct per ray. */
static void run_batch(int any, int kind, const void *nodes, const uint32_t *indices, const float *directions,
                      uint32_t M, const float bounds[6], const float *verts, const uint8_t *degen, uint32_t n_prims,
                      const orc_ray *rays, uint64_t n, orc_hit *hits, uint8_t *occluded, orc_counters *counters,
                      orc_counters *per_ray, int n_threads, const uint8_t *prim_kind, const float *spheres) {
    uint32_t *sphere_of = 0;
    if (prim_kind) {
        sphere_of = (uint32_t *)calloc(n_prims ? n_prims : 1, sizeof(uint32_t));
        uint32_t k = 0;
        for (uint32_t i = 0; i < n_prims; ++i) if (prim_kind[i]) sphere_of[i] = k++;
    }
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 256) n_threads = 256;
    orc_job *jobs = (orc_job *)calloc((size_t)n_threads, sizeof(orc_job));
    pthread_t *th = (pthread_t *)calloc((size_t)n_threads, sizeof(pthread_t));
    uint64_t chunk = (n + (uint64_t)n_threads - 1) / (uint64_t)n_threads;
    for (int t = 0; t < n_threads; ++t) {
        orc_job *j = &jobs[t];
        j->kind = kind; j->any = any; j->want = counters || per_ray;
        j->nodes = nodes; j->indices = indices;
        j->P.verts = verts; j->P.degen = degen; j->P.n_prims = n_prims;
        j->P.kind = prim_kind; j->P.sphere_of = sphere_of; j->P.spheres = spheres;
        if (kind >= 4) {
            j->G.kind = kind; j->G.nodes = (const orc_gbsp_node *)nodes; j->G.indices = indices;
            memcpy(j->G.bounds, bounds, sizeof(j->G.bounds));
        } else if (kind != 0) {
            j->T.kind = kind; j->T.nodes = (const orc_bsp_node *)nodes; j->T.indices = indices;
            j->T.directions = directions; j->T.M = M;
            memcpy(j->T.bounds, bounds, sizeof(j->T.bounds));
        }
        j->rays = rays; j->hits = hits; j->occluded = occluded; j->per_ray = per_ray;
        j->begin = (uint64_t)t * chunk < n ? (uint64_t)t * chunk : n;
        j->end = j->begin + chunk < n ? j->begin + chunk : n;
        if (n_threads == 1) run_job(j); else pthread_create(&th[t], 0, run_job, j);
    }
    orc_counters tot; memset(&tot, 0, sizeof(tot));
    for (int t = 0; t < n_threads; ++t) {
        if (n_threads > 1) pthread_join(th[t], 0);
        add_counters(&tot, &jobs[t].total);
    }
    if (counters) { tot.rays = n; *counters = tot; }
    free(jobs); free(th); free(sphere_of);
}

void orc_intersect_closest(int kind, const void *nodes, const uint32_t *indices, const float *directions, uint32_t M,
                           const float bounds[6], const float *verts, const uint8_t *degen, uint32_t n_prims,
                           const orc_ray *rays, uint64_t n, orc_hit *hits, orc_counters *counters,
                           orc_counters *per_ray, int n_threads, const uint8_t *prim_kind, const float *spheres) {
    run_batch(0, kind, nodes, indices, directions, M, bounds, verts, degen, n_prims, rays, n, hits, 0, counters,
              per_ray, n_threads, prim_kind, spheres);
}
void orc_intersect_any(int kind, const void *nodes, const uint32_t *indices, const float *directions, uint32_t M,
                       const float bounds[6], const float *verts, const uint8_t *degen, uint32_t n_prims,
                       const orc_ray *rays, uint64_t n, uint8_t *occluded, orc_counters *counters,
                       orc_counters *per_ray, int n_threads, const uint8_t *prim_kind, const float *spheres) {
    run_batch(1, kind, nodes, indices, directions, M, bounds, verts, degen, n_prims, rays, n, 0, occluded, counters,
              per_ray, n_threads, prim_kind, spheres);
}

void orc_degenerate_flags(const float *verts, const float *uvs, uint32_t n_prims, uint8_t *out) {
    for (uint32_t i = 0; i < n_prims; ++i)
        out[i] = (uint8_t)orc_tri_is_degenerate(verts + 9 * (size_t)i, uvs ? uvs + 6 * (size_t)i : 0);
}
