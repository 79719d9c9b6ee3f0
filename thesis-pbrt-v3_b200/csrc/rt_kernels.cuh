// Traversal kernels (sm_100a).  One ray per thread; order-preserving so that primitive ids match the
// reference exactly: leaf primitives in array order, near/far by dirIsNeg[axis],
// strict comparisons.
//
//   BVH     : BVHAccel::Intersect / IntersectP            src/accelerators/bvh.cpp:354-437
//   kd-tree : KdTreeAccel::Intersect / IntersectP         src/accelerators/kdtreeaccel.cpp:381-521
//   RBSP    : RBSP::Intersect / IntersectP                src/accelerators/rbsp.cpp:405-547
//   RBSPKd  : same loop, intersectInterior with kd path   src/accelerators/rbspKd.cpp:68-92
//   BSP     : BSP::Intersect / IntersectP                 src/accelerators/BSP.cpp:27-163   (20-byte nodes)
//   BSPKd   : BSPKd::Intersect / IntersectP               src/accelerators/BSPKd.cpp:25-169 (20-byte nodes)
#pragma once
#include "rt_math.cuh"
#include "../../include/b200rt.h"

namespace b200rt {

// ---- device layouts -----------------------------------------------------------
// BVH: one 64-byte record per INTERIOR reference node holding BOTH children's boxes, so one fetch
// (4 x 16 B) replaces the reference's two dependent 32-byte child fetches.
//   a = (c0.min.x, c0.min.y, c0.min.z, c0.max.x)
//   b = (c0.max.y, c0.max.z, c1.min.x, c1.min.y)
//   c = (c1.min.z, c1.max.x, c1.max.y, c1.max.z)
//   m = (c0.ref, c1.ref, splitAxis, subtree prim count)
// ref: bit 31 set = leaf, low 31 bits = first triangle slot (leaf ends at the record flagged
// B200RT_TRI_LAST); bit 31 clear = index of the child's own 64-byte record.
struct __align__(16) BvhNode2 {
    float4 a, b, c;
    uint4 m;
};
#define B200RT_LEAF_BIT 0x80000000u

// Two binary levels collapsed into one 128-byte record (B200RT_BVH4): the four GRANDCHILDREN of a reference interior node,
// as two BvhNode2-shaped halves (pair 0 = the children of child 0, pair 1 = those of child 1; a child that is a leaf
// occupies slot 0 of its pair with its own box, slot 1 is then empty).  Visiting order and results stay the reference's:
// a grandchild's box lies inside its parent's, and every term of the slab test is monotone in the bounds, so a box that
// fails implies that everything below it fails — skipping the child's own test changes no outcome; pairs are ordered by
// dirIsNeg[axis of the node], slots inside a pair by dirIsNeg[axis of that child] (bvh.cpp:382-388), deferred slots are
// re-tested against the current ray.tMax when popped (bvh.cpp:367).  The chain of dependent node fetches halves.
//   half h: a, b, c as in BvhNode2; m = (slot 0 ref, slot 1 ref, meta, -), meta (half 0 only): bits 0-1 the node's split
//   axis, 2-3 child 0's, 4-5 child 1's
struct __align__(16) BvhNode4 {
    float4 q[8];
};
#define B200RT_EMPTY_SLOT 0x7FFFFFFEu

struct BvhView {
    const BvhNode4 *nodes4;   // nullptr when the tree has no interior node
    uint32_t rootRef4;
    const BvhNode2 *nodes;
    const float4 *tris;  // 3 x float4 per slot: (p0, primId) (p1, flags) (p2, sphere index)
    const float4 *spheres;  // 4 x float4 per sphere: WorldToObject rows 0..2, (radius, -, -, -)
    float root[6];       // nodes[0].bounds of the reference tree
    uint32_t rootRef;
    // The interior nodes of the top levels come first in the array, level by level: levels [0, k) are the records
    // [0, topCount[k]), so a block can stage them in shared memory with one contiguous copy (north-star item (3);
    // B200RT_TOPLEVELS, measured in profiles/r02_ab_bvh_toplevels.log).  Deeper nodes keep the reference's depth-first order.
    uint32_t topCount[12];
};
#define B200RT_TOP_MAXLEVELS 11

// kd / RBSP: the reference's own 8-byte nodes (w0 = split | onePrimitive | indicesOffset,
// w1 = axis bits + aboveChild/nPrims) + primitiveIndices + triangles in original primitive order.
// BSP / BSPKd: the reference's 20-byte nodes widened to one aligned 32-byte record
// (w0, flags, axis.x, axis.y | axis.z, 0, 0, 0) so that a node is ONE 256-bit load (a 20-byte stride would
// straddle 32-byte sectors on 3 of every 8 nodes).
struct BspView {
    const uint2 *nodes;   // kd / RBSP(/Kd): 8-byte nodes; BSP / BSPKd: 32-byte records (read through bsp_fetch)
    const uint32_t *indices;
    const float4 *tris;  // 3 x float4 per ORIGINAL primitive
    const float4 *spheres;
    float bounds[6];
    float dirs[13 * 3];  // RBSPShared.h:29-74, M <= 13
    uint32_t M, bitOffset, bitMask, leafTag;
    int kind;  // b200rt_kind
};

struct Counters {
    unsigned long long rays, nodes_fetched, nodes_passed, leaf_visits, prim_tests, index_reads;
};

struct LocalCount {
    uint32_t fetched = 0, passed = 0, leaves = 0, tests = 0, idx = 0;
};

struct HitRec {
    uint32_t prim;
    float t, b0, b1, b2;
};

// GeometricPrimitive::Intersect (primitive.cpp:123-139): on a hit ray.tMax = tHit.
// ANY = GeometricPrimitive::IntersectP (:118-121): Triangle::IntersectP has NO degenerate-triangle
// rejection outside its alpha-mask branch (triangle.cpp:533-556), unlike Intersect (:309-315).
// dptr (optional): where the ray's (d.xyz, time) float4 lives in global memory.  The BVH persistent kernel passes
// it so that the direction — needed by nothing but the rare sphere test there — is not held in three registers
// for the whole traversal.
template <bool ANY>
__device__ __forceinline__ bool prim_closest(RayCtx &r, const float4 v0, const float4 v1, const float4 v2, HitRec &h,
                                             const float4 *spheres, const float4 *dptr = nullptr) {
    float t, b0, b1, b2;
    if (__float_as_uint(v1.w) & B200RT_PRIM_SPHERE) {  // Sphere::Intersect / IntersectP (sphere.cpp:49, 158)
        float dx = r.dx, dy = r.dy, dz = r.dz;
        if (dptr) { const float4 dd = __ldg(dptr); dx = dd.x; dy = dd.y; dz = dd.z; }
        if (!sphere_test(spheres + 4 * (size_t)__float_as_uint(v2.w), r.ox, r.oy, r.oz, dx, dy, dz, r.tMax, t))
            return false;
        r.tMax = t;
        h.prim = __float_as_uint(v0.w); h.t = t; h.b0 = h.b1 = h.b2 = 0.f;  // no barycentrics on a sphere
        return true;
    }
    if (!tri_test(r, v0, v1, v2, t, b0, b1, b2)) return false;
    if (!ANY && (__float_as_uint(v1.w) & B200RT_TRI_DEGENERATE)) return false;
    r.tMax = t;
    h.prim = __float_as_uint(v0.w); h.t = t; h.b0 = b0; h.b1 = b1; h.b2 = b2;
    return true;
}

template <bool ANY>
__device__ __forceinline__ bool prim_closest(RayCtx &r, const float4 *rec, HitRec &h, const float4 *spheres) {
    return prim_closest<ANY>(r, __ldg(rec), __ldg(rec + 1), __ldg(rec + 2), h, spheres);
}

// ---- BVH --------------------------------------------------------------------------------------
template <bool ANY, bool COUNT>
__device__ __forceinline__ bool bvh_trace(const BvhView &S, RayCtx &r, HitRec &h, LocalCount &cnt) {
    const bool nx = r.ix < 0, ny = r.iy < 0, nz = r.iz < 0;  // dirIsNeg (bvh.cpp:359)
    // root box (node 0 of the reference)
    {
        float te;
        if (COUNT) cnt.fetched++;
        const bool ok = slab_geom(r, nx ? S.root[3] : S.root[0], ny ? S.root[4] : S.root[1], nz ? S.root[5] : S.root[2],
                                  nx ? S.root[0] : S.root[3], ny ? S.root[1] : S.root[4], nz ? S.root[2] : S.root[5], te);
        if (!(ok && te < r.tMax)) return false;
        if (COUNT) cnt.passed++;
    }
    bool hit = false;
    uint32_t stackRef[64];
    float stackT[64];
    int sp = 0;
    uint32_t cur = S.rootRef;
    for (;;) {
        if (cur & B200RT_LEAF_BIT) {
            // leaf: primitives in array order (bvh.cpp:372-375 / 412-418)
            if (COUNT) cnt.leaves++;
            const float4 *rec = S.tris + 3 * (size_t)(cur & ~B200RT_LEAF_BIT);
            for (;;) {
                if (COUNT) cnt.tests++;
                const float4 v0 = __ldg(rec), v1 = __ldg(rec + 1), v2 = __ldg(rec + 2);
                if (prim_closest<ANY>(r, v0, v1, v2, h, S.spheres)) {
                    hit = true;
                    if (ANY) return true;
                }
                if (__float_as_uint(v1.w) & B200RT_TRI_LAST) break;
                rec += 3;
            }
        } else {
            const BvhNode2 *n = S.nodes + cur;
            const float4 a = __ldg(&n->a), b = __ldg(&n->b), c = __ldg(&n->c);
            const uint4 m = __ldg(&n->m);
            // child boxes with lo/hi selected by dirIsNeg
            float t0, t1;
            const bool g0 = slab_geom(r, nx ? a.w : a.x, ny ? b.x : a.y, nz ? b.y : a.z, nx ? a.x : a.w, ny ? a.y : b.x,
                                      nz ? a.z : b.y, t0);
            const bool g1 = slab_geom(r, nx ? c.y : b.z, ny ? c.z : b.w, nz ? c.w : c.x, nx ? b.z : c.y, ny ? b.w : c.z,
                                      nz ? c.x : c.w, t1);
            const bool h0 = g0 && t0 < r.tMax, h1 = g1 && t1 < r.tMax;
            // near = second child when dirIsNeg[axis] (bvh.cpp:382-388)
            const bool negAxis = m.z == 0 ? nx : (m.z == 1 ? ny : nz);
            const uint32_t nearRef = negAxis ? m.y : m.x, farRef = negAxis ? m.x : m.y;
            const bool hitNear = negAxis ? h1 : h0, hitFar = negAxis ? h0 : h1;
            const float tFar = negAxis ? t0 : t1;
            if (COUNT) {
                // instrumented build: mirror the reference's fetch order exactly (far child is pushed
                // unconditionally, bvh.cpp:382-388, and its box is fetched when popped, :367)
                stackRef[sp] = farRef; stackT[sp] = hitFar ? tFar : __int_as_float(0x7f800000); ++sp;
                cnt.fetched++;
                if (hitNear) { cnt.passed++; cur = nearRef; continue; }
            } else {
                if (hitNear) {
                    if (hitFar) { stackRef[sp] = farRef; stackT[sp] = tFar; ++sp; }
                    cur = nearRef;
                    continue;
                }
                if (hitFar) { cur = farRef; continue; }
            }
        }
        // pop: the reference re-tests the far box against the CURRENT ray.tMax (bvh.cpp:367)
        bool found = false;
        while (sp > 0) {
            --sp;
            if (COUNT) cnt.fetched++;
            if ((ANY && !COUNT) || stackT[sp] < r.tMax) {
                cur = stackRef[sp];
                found = true;
                if (COUNT) cnt.passed++;
                break;
            }
        }
        if (!found) break;
    }
    return hit;
}

// ---- kd / RBSP / RBSPKd -------------------------------------------------------------------------
template <int KIND>
__device__ __forceinline__ float bsp_interior(const BspView &S, const RayCtx &r, uint32_t axis, float split,
                                              bool &belowFirst) {
    if (KIND == B200RT_KDTREE || (KIND == B200RT_RBSPKD && axis < 3)) {
        // kdtreeaccel.cpp:95-104 + geometry.h:1833-1835 ; rbspKd.cpp:70-78
        const float o = axis == 0 ? r.ox : (axis == 1 ? r.oy : r.oz);
        const float d = axis == 0 ? r.dx : (axis == 1 ? r.dy : r.dz);
        const float inv = axis == 0 ? r.ix : (axis == 1 ? r.iy : r.iz);
        belowFirst = (o < split) || (o == split && d <= 0);
        return (split - o) * inv;
    }
    // rbsp.cpp:68-80 + geometry.h:1837-1843 ; Dot = x*x' + y*y' + z*z' left to right (geometry.h:1264-1267)
    const float *dv = S.dirs + 3 * axis;
    const float projectedO = dv[0] * r.ox + dv[1] * r.oy + dv[2] * r.oz;
    const float inverseProjectedD = 1 / (dv[0] * r.dx + dv[1] * r.dy + dv[2] * r.dz);
    belowFirst = (projectedO < split) || (projectedO == split && inverseProjectedD <= 0);
    return (split - projectedO) * inverseProjectedD;
}

// One decoded tree node of any of the BSP-like kinds.
struct BspStep {
    uint32_t w0;       // leaf: onePrimitive / primitiveIndicesOffset
    uint32_t payload;  // leaf: nPrims; interior: aboveChild
    float tPlane;      // interior
    bool leaf, belowFirst;
};
template <int KIND>
struct BspCodec {  // bit layout of the flags word
    static constexpr bool wide = KIND == B200RT_BSP || KIND == B200RT_BSPKD;
};

// Decode the two words of an 8-byte kd / RBSP / RBSPKd node and, for an interior node, intersect its split plane.
template <int KIND>
__device__ __forceinline__ BspStep bsp_decode(const BspView &S, const RayCtx &r, uint2 n, uint32_t off, uint32_t mask, uint32_t leafTag) {
    BspStep st;
    const uint32_t axis = n.y & mask;
    st.w0 = n.x;
    st.payload = n.y >> off;
    st.leaf = axis == leafTag;
    st.tPlane = 0.f; st.belowFirst = false;
    if (!st.leaf) st.tPlane = bsp_interior<KIND>(S, r, axis, __uint_as_float(n.x), st.belowFirst);
    return st;
}

// Fetch + decode node `cur` and, for an interior node, intersect its split plane.
//   kd / RBSP / RBSPKd : kdtreeaccel.cpp:43-104, rbsp.cpp:15-80, rbspKd.cpp:68-92 (axis field `off` bits wide)
//   BSP                : BSP.h:13-67 (bit 0 = leaf, payload << 1), treeIntersectInterior
//   BSPKd              : BSPKd.h:13-72 (low 3 bits: 0..2 kd axis, 3 leaf, 4 general; payload << 3)
template <int KIND>
__device__ __forceinline__ BspStep bsp_fetch(const BspView &S, const RayCtx &r, uint32_t cur, uint32_t off, uint32_t mask,
                                             uint32_t leafTag) {
    BspStep st;
    if (!BspCodec<KIND>::wide) return bsp_decode<KIND>(S, r, __ldg(S.nodes + cur), off, mask, leafTag);
    float4 lo, hi;
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(lo.x), "=f"(lo.y), "=f"(lo.z), "=f"(lo.w), "=f"(hi.x), "=f"(hi.y), "=f"(hi.z), "=f"(hi.w)
                 : "l"(reinterpret_cast<const float4 *>(S.nodes) + 2 * (size_t)cur));
    const uint32_t flags = __float_as_uint(lo.y);
    const float split = lo.x;
    st.w0 = __float_as_uint(lo.x);
    st.tPlane = 0.f; st.belowFirst = false;
    if (KIND == B200RT_BSP) { st.leaf = (flags & 1u) == 1u; st.payload = flags >> 1; }
    else { st.leaf = (flags & 7u) == 3u; st.payload = flags >> 3; }
    if (st.leaf) return st;
    if (KIND == B200RT_BSPKD && (flags & 7u) < 3u) {  // isKdNode with a valid axis (BSPKd.h:52-60)
        const uint32_t axis = flags & 7u;
        const float o = axis == 0 ? r.ox : (axis == 1 ? r.oy : r.oz);
        const float d = axis == 0 ? r.dx : (axis == 1 ? r.dy : r.dz);
        const float inv = axis == 0 ? r.ix : (axis == 1 ? r.iy : r.iz);
        st.belowFirst = (o < split) || (o == split && d <= 0);
        st.tPlane = (split - o) * inv;
        return st;
    }
    // planeDistance(axis, split, ray, projectedO, inverseProjectedD) (geometry.h:1837-1843)
    const float projectedO = lo.z * r.ox + lo.w * r.oy + hi.x * r.oz;
    const float inverseProjectedD = 1 / (lo.z * r.dx + lo.w * r.dy + hi.x * r.dz);
    st.belowFirst = (projectedO < split) || (projectedO == split && inverseProjectedD <= 0);
    st.tPlane = (split - projectedO) * inverseProjectedD;
    return st;
}

template <int KIND, bool ANY, bool COUNT>
__device__ __forceinline__ bool bsp_trace(const BspView &S, RayCtx &r, HitRec &h, LocalCount &cnt) {
    float tMin, tMax;
    if (!root_clip(r, S.bounds, tMin, tMax)) return false;
    const uint32_t off = KIND == B200RT_KDTREE ? 2u : S.bitOffset;
    const uint32_t mask = KIND == B200RT_KDTREE ? 3u : S.bitMask;
    const uint32_t leafTag = KIND == B200RT_KDTREE ? 3u : S.leafTag;
    uint32_t todoNode[64];
    float todoMin[64], todoMax[64];
    int todoPos = 0;
    uint32_t cur = 0;
    bool hit = false;
    for (;;) {
        if (!ANY && r.tMax < tMin) break;  // kdtreeaccel.cpp:400 (closest-hit only)
        if (COUNT) cnt.fetched++;
        const BspStep n = bsp_fetch<KIND>(S, r, cur, off, mask, leafTag);
        if (!n.leaf) {
            if (COUNT) cnt.passed++;
            const float tPlane = n.tPlane;
            const uint32_t above = n.payload;
            const uint32_t first = n.belowFirst ? cur + 1 : above, second = n.belowFirst ? above : cur + 1;
            if (tPlane > tMax || tPlane <= 0) cur = first;
            else if (tPlane < tMin) cur = second;
            else {
                todoNode[todoPos] = second; todoMin[todoPos] = tPlane; todoMax[todoPos] = tMax; ++todoPos;
                cur = first;
                tMax = tPlane;
            }
        } else {
            if (COUNT) cnt.leaves++;
            const uint32_t np = n.payload;
            if (np == 1) {
                if (COUNT) cnt.tests++;
                if (prim_closest<ANY>(r, S.tris + 3 * (size_t)n.w0, h, S.spheres)) { hit = true; if (ANY) return true; }
            } else {
                for (uint32_t i = 0; i < np; ++i) {
                    if (COUNT) { cnt.tests++; cnt.idx++; }
                    const uint32_t p = __ldg(S.indices + n.w0 + i);
                    if (prim_closest<ANY>(r, S.tris + 3 * (size_t)p, h, S.spheres)) { hit = true; if (ANY) return true; }
                }
            }
            if (todoPos > 0) {
                --todoPos;
                cur = todoNode[todoPos]; tMin = todoMin[todoPos]; tMax = todoMax[todoPos];
            } else break;
        }
    }
    return hit;
}

// ---- kernels ------------------------------------------------------------------------------------
#define B200RT_BLOCK 128

template <int KIND, bool ANY, bool COUNT>
__global__ void __launch_bounds__(B200RT_BLOCK)
trace_kernel(const __grid_constant__ BvhView bvh, const __grid_constant__ BspView bsp, const float4 *__restrict__ rays,
             uint64_t n, const uint32_t *__restrict__ n_dev, b200rt_hit *__restrict__ hits,
             uint8_t *__restrict__ occluded, Counters *counters, uint4 *__restrict__ per_ray = nullptr) {
    LocalCount cnt;
    if (n_dev) { const uint64_t m = *n_dev; n = m < n ? m : n; }
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t base = (uint64_t)blockIdx.x * blockDim.x; base < n; base += stride) {
        const uint64_t i = base + threadIdx.x;
        if (i < n) {
            RayCtx r;
            ray_setup(r, __ldg(rays + 2 * i), __ldg(rays + 2 * i + 1));
            HitRec h;
            h.prim = B200RT_MISS; h.t = r.tMax; h.b0 = h.b1 = h.b2 = 0.f;
            bool hit;
            const LocalCount before = cnt;
            if (KIND == B200RT_BVH) hit = bvh_trace<ANY, COUNT>(bvh, r, h, cnt);
            else hit = bsp_trace<KIND, ANY, COUNT>(bsp, r, h, cnt);
            if (COUNT && per_ray) {
                // the reference's per-ray GeneralStats (geometry.h:1078-1173): primitiveIntersections(P),
                // {bvh,kd,bsp}TreeNodeTraversals(P) (interior nodes only: bvh.cpp:379, kdtreeaccel.cpp:404), leafNodeTraversals(P)
                const uint32_t leaves = cnt.leaves - before.leaves, passed = cnt.passed - before.passed;
                per_ray[i] = make_uint4(cnt.tests - before.tests, KIND == B200RT_BVH ? passed - leaves : passed, leaves,
                                        cnt.fetched - before.fetched);
            }
            if (ANY) occluded[i] = hit ? 1 : 0;
            else {
                b200rt_hit o;
                o.prim = h.prim; o.t = h.t; o.b0 = h.b0; o.b1 = h.b1; o.b2 = h.b2;
                hits[i] = o;
            }
        }
    }
    if (COUNT) {
        const uint64_t first = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
        // every launched thread takes part in the warp reductions
        uint64_t mine = 0;
        for (uint64_t i = first; i < n; i += stride) ++mine;
        LocalCount c2 = cnt;
        uint32_t v[5] = {c2.fetched, c2.passed, c2.leaves, c2.tests, c2.idx};
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            const uint32_t s = __reduce_add_sync(0xFFFFFFFFu, v[k]);
            if ((threadIdx.x & 31) == 0 && s) atomicAdd(&(&counters->nodes_fetched)[k], (unsigned long long)s);
        }
        const uint32_t s = __reduce_add_sync(0xFFFFFFFFu, (uint32_t)mine);
        if ((threadIdx.x & 31) == 0 && s) atomicAdd(&counters->rays, (unsigned long long)s);
    }
}

}  // namespace b200rt
