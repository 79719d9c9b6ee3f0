// The k-DOP arithmetic of the RBSP builder, host/device (every function compiles for the B200 in build_rbsp.cu and, as test
// infrastructure, for the host in tests/hostbuild — the same text, so the restatement is pinned on the CPU against the
// reference's own trees before it runs on the GPU).
//
// What is restated, with the reference's float operations in the reference's order (-fmad=false / -ffp-contract=off):
//   KDOPEdge::getBounds, KDOPCutAddEdge, KDOPCut, KDOPSurfaceArea   src/accelerators/kDOPMesh.h:14-27, 59-97, 99-176, 178-227
//   Bounds3::toKDOPMesh (the root box as 12 edges with face ids)    src/core/geometry.h:1001-1027
//   Triangle::getBounds (projection of a triangle on a direction)   src/shapes/triangle.cpp:661-676
//   getDirections (the 3 / 7 / 9 / 13 split normals)                src/accelerators/RBSPShared.h:30-76
//   the SAH of one candidate plane                                  src/accelerators/rbsp.cpp:306-318
//
// Storage is fixed-capacity (no allocation per candidate): a convex polytope bounded by 2M <= 26 planes has at most 72 edges;
// kMaxEdges leaves room for the duplicate / degenerate edges the reference's cut can leave behind.  Overflow is reported,
// never truncated silently.
#pragma once
#include "wf_core.cuh"

namespace b200rt {
namespace rbsp {

constexpr int kMaxDirs = 16;
constexpr int kMaxFaces = 2 * kMaxDirs;
constexpr int kMaxEdges = 96;
constexpr int kMaxFacePoints = 8;

struct P3 { float x, y, z; };
WF_HD bool p3_eq(const P3 &a, const P3 &b) { return a.x == b.x && a.y == b.y && a.z == b.z; }   // Point3::operator==
// Dot(Vector3f, Point3f), geometry.h:1275-1278
WF_HD float dot_dp(const float *d, const P3 &p) { return d[0] * p.x + d[1] * p.y + d[2] * p.z; }

struct KdopEdge {
    P3 v1, v2;
    uint16_t f1, f2;
};
struct KdopMesh {
    uint32_t n;
    KdopEdge e[kMaxEdges];
};

WF_HD bool mesh_add(KdopMesh &m, const P3 &a, const P3 &b, uint32_t f1, uint32_t f2) {
    if (m.n >= (uint32_t)kMaxEdges) return false;
    KdopEdge &e = m.e[m.n++];
    e.v1 = a; e.v2 = b; e.f1 = (uint16_t)f1; e.f2 = (uint16_t)f2;
    return true;
}
// KDOPMeshBase::addEdgeIfNeeded, kDOPMesh.h:42-49
WF_HD bool mesh_add_if_needed(KdopMesh &m, const P3 &a, const P3 &b, uint32_t f1, uint32_t f2) {
    for (uint32_t i = 0; i < m.n; ++i) {
        const KdopEdge &e = m.e[i];
        if ((p3_eq(e.v1, b) && p3_eq(e.v2, a)) || (p3_eq(e.v1, a) && p3_eq(e.v2, b))) return true;
    }
    return mesh_add(m, a, b, f1, f2);
}

// The per-face cut points (faceVertices of KDOPCut): a point is appended once (KDOPCutHelper, kDOPMesh.h:53-57)
struct FacePoints {
    uint8_t n[kMaxFaces];
    P3 p[kMaxFaces][kMaxFacePoints];
};
WF_HD bool face_point(FacePoints &fp, uint32_t face, const P3 &pt) {
    const uint32_t n = fp.n[face];
    for (uint32_t i = 0; i < n; ++i)
        if (p3_eq(fp.p[face][i], pt)) return true;
    if (n >= (uint32_t)kMaxFacePoints) return false;
    fp.p[face][n] = pt;
    fp.n[face] = (uint8_t)(n + 1);
    return true;
}

// KDOPCut<KDOPMesh> (kDOPMesh.h:99-176): the mesh `in` cut by the plane Dot(direction, p) = t into the part below (left, new
// face 2 * directionId) and the part above (right, new face 2 * directionId + 1).  Returns false on capacity overflow.
WF_HD_NOINLINE bool kdop_cut(const KdopMesh &in, uint32_t M, float t, const float *direction, uint32_t directionId,
                             KdopMesh &left, KdopMesh &right, FacePoints &fp) {
    left.n = 0;
    right.n = 0;
    for (uint32_t i = 0; i < 2 * M; ++i) fp.n[i] = 0;
    // coincident edges are kept as indices into `in` (with the orientation they were seen in)
    uint8_t coincident[kMaxEdges];
    uint8_t coincidentSwapped[kMaxEdges];
    uint32_t nCoincident = 0;
    bool ok = true;
    for (uint32_t k = 0; k < in.n; ++k) {
        const KdopEdge &src = in.e[k];
        float t1 = dot_dp(direction, src.v1);
        float t2 = dot_dp(direction, src.v2);
        P3 v1 = src.v1, v2 = src.v2;
        bool swapped = false;
        if (t1 > t2) {   // KDOPEdge(edge.v2, edge.v1, ...) with (t2, t1)
            const P3 tp = v1; v1 = v2; v2 = tp;
            const float tt = t1; t1 = t2; t2 = tt;
            swapped = true;
        }
        const uint32_t f1 = src.f1, f2 = src.f2;
        // KDOPCutAddEdge, kDOPMesh.h:59-97
        if (t1 < t && t2 < t) {
            ok &= mesh_add(left, v1, v2, f1, f2);
        } else if (t1 > t && t2 > t) {
            ok &= mesh_add(right, v1, v2, f1, f2);
        } else if (t1 < t && t == t2) {
            ok &= mesh_add(left, v1, v2, f1, f2);
            ok &= face_point(fp, f1, v2);
            ok &= face_point(fp, f2, v2);
        } else if (t1 == t && t < t2) {
            ok &= mesh_add(right, v1, v2, f1, f2);
            ok &= face_point(fp, f1, v1);
            ok &= face_point(fp, f2, v1);
        } else if (t1 < t && t < t2) {
            const float dx = v2.x - v1.x, dy = v2.y - v1.y, dz = v2.z - v1.z;   // Vector3f d = edge.v2 - edge.v1
            const float tAlongEdge = (-(t1 - t)) / (t2 - t1);
            const P3 vs = {v1.x + tAlongEdge * dx, v1.y + tAlongEdge * dy, v1.z + tAlongEdge * dz};
            ok &= mesh_add(left, v1, vs, f1, f2);
            ok &= mesh_add(right, vs, v2, f1, f2);
            ok &= face_point(fp, f1, vs);
            ok &= face_point(fp, f2, vs);
        } else if (t1 == t && t == t2) {
            coincident[nCoincident] = (uint8_t)k;
            coincidentSwapped[nCoincident++] = swapped ? 1 : 0;
        }
    }
    // edges inside the plane go to the side whose faces they continue (kDOPMesh.h:126-143)
    for (uint32_t c = 0; c < nCoincident; ++c) {
        const KdopEdge &src = in.e[coincident[c]];
        const P3 v1 = coincidentSwapped[c] ? src.v2 : src.v1, v2 = coincidentSwapped[c] ? src.v1 : src.v2;
        const uint32_t f1 = src.f1, f2 = src.f2;
        const uint32_t nLeft = left.n;   // the scan stops at the first match, so edges appended by it are never visited
        for (uint32_t i = 0; i < nLeft; ++i) {
            const uint32_t l1 = left.e[i].f1, l2 = left.e[i].f2;
            if (l1 == f1 || l2 == f1) {
                ok &= mesh_add(left, v1, v2, f1, 2 * directionId);
                ok &= mesh_add(right, v1, v2, f2, 2 * directionId + 1);
                break;
            } else if (l1 == f2 || l2 == f2) {
                ok &= mesh_add(left, v1, v2, f2, 2 * directionId);
                ok &= mesh_add(right, v1, v2, f1, 2 * directionId + 1);
                break;
            }
        }
    }
    // the cut polygon: one edge per face that was crossed exactly twice (kDOPMesh.h:164-173)
    for (uint32_t i = 0; i < 2 * M; ++i) {
        if (fp.n[i] == 2) {
            ok &= mesh_add_if_needed(left, fp.p[i][0], fp.p[i][1], i, 2 * directionId);
            ok &= mesh_add_if_needed(right, fp.p[i][0], fp.p[i][1], i, 2 * directionId + 1);
        }
    }
    return ok;
}

// KDOPSurfaceArea (kDOPMesh.h:178-227).  Walks each face's edges as a chain, re-orienting edges in place as it goes — the
// mesh is MODIFIED (v1 / v2 swaps), and the modified mesh is what the builder keeps for the children, so callers keep it too.
WF_HD_NOINLINE float kdop_area(KdopMesh &m, const float *directions, uint32_t M) {
    float SA = 0;
    for (uint32_t i = 0; i < 2 * M; ++i) {
        float fx = 0, fy = 0, fz = 0;   // Vector3f FSA
        uint8_t face[2 * kMaxEdges];
        uint32_t nf = 0;
        for (uint32_t k = 0; k < m.n; ++k) {   // an edge is listed once per face id it carries (twice when f1 == f2)
            if (m.e[k].f1 == i) face[nf++] = (uint8_t)k;
            if (m.e[k].f2 == i) face[nf++] = (uint8_t)k;
        }
        if (nf) {
            uint64_t used[(2 * kMaxEdges + 63) / 64] = {0, 0, 0};
            uint32_t edgeId = 0;
            do {
                if ((used[edgeId >> 6] >> (edgeId & 63)) & 1) break;
                used[edgeId >> 6] |= 1ull << (edgeId & 63);
                const uint32_t cur = face[edgeId];
                {   // FSA += Cross(v1, v2): products and differences in double (geometry.h:1314-1320), the sum in float
                    const double ax = m.e[cur].v1.x, ay = m.e[cur].v1.y, az = m.e[cur].v1.z;
                    const double bx = m.e[cur].v2.x, by = m.e[cur].v2.y, bz = m.e[cur].v2.z;
                    fx += (float)((ay * bz) - (az * by));
                    fy += (float)((az * bx) - (ax * bz));
                    fz += (float)((ax * by) - (ay * bx));
                }
                for (uint32_t j = 0; j < nf; ++j) {
                    if (j == edgeId) continue;
                    KdopEdge &ej = m.e[face[j]];
                    // currentEdge->v2 is re-read every time: face[j] may be the same edge as `cur` (listed twice)
                    if (p3_eq(ej.v2, m.e[cur].v2)) { const P3 tp = ej.v1; ej.v1 = ej.v2; ej.v2 = tp; }
                    if (p3_eq(ej.v1, m.e[cur].v2) && !((used[j >> 6] >> (j & 63)) & 1)) {
                        edgeId = j;
                        break;
                    }
                }
            } while (edgeId != 0);
        }
        const float *d = directions + 3 * (i / 2);
        SA += fabsf(d[0] * fx + d[1] * fy + d[2] * fz);   // std::abs(Dot(directions[i / 2], FSA))
    }
    return SA / 2.0f;
}

// Bounds3::toKDOPMesh, geometry.h:1001-1023
WF_HD void kdop_from_box(const float *b /* pMin.xyz pMax.xyz */, KdopMesh &m) {
    const P3 v1 = {b[0], b[1], b[2]}, v2 = {b[0], b[1], b[5]}, v3 = {b[0], b[4], b[2]}, v4 = {b[3], b[1], b[2]};
    const P3 v5 = {b[0], b[4], b[5]}, v6 = {b[3], b[1], b[5]}, v7 = {b[3], b[4], b[2]}, v8 = {b[3], b[4], b[5]};
    m.n = 0;
    mesh_add(m, v1, v2, 1, 3); mesh_add(m, v1, v3, 1, 5); mesh_add(m, v1, v4, 3, 5); mesh_add(m, v2, v5, 1, 4);
    mesh_add(m, v2, v6, 3, 4); mesh_add(m, v3, v5, 1, 2); mesh_add(m, v3, v7, 2, 5); mesh_add(m, v4, v6, 0, 3);
    mesh_add(m, v4, v7, 0, 5); mesh_add(m, v5, v8, 2, 4); mesh_add(m, v6, v8, 0, 4); mesh_add(m, v7, v8, 0, 2);
}

// Union over the mesh's edges of KDOPEdge::getBounds(direction) (rbsp.cpp:365-377): the child's extent along one direction
WF_HD void kdop_extent(const KdopMesh &m, const float *direction, float &mn, float &mx) {
    mn = 3.402823466e+38f;    // Bounds(): max / lowest
    mx = -3.402823466e+38f;
    for (uint32_t k = 0; k < m.n; ++k) {
        const float t1 = dot_dp(direction, m.e[k].v1), t2 = dot_dp(direction, m.e[k].v2);
        const float emx = t1 < t2 ? t2 : t1, emn = t2 < t1 ? t2 : t1;   // std::max(t1, t2), std::min(t1, t2)
        mn = emn < mn ? emn : mn;   // std::min(b1.min, b2.min) with b1 = the running bound
        mx = mx < emx ? emx : mx;
    }
}

// Triangle::getBounds, triangle.cpp:661-676 (v: 9 floats)
WF_HD void tri_extent(const float *v, const float *d, float &mn, float &mx) {
    float t = d[0] * v[0] + d[1] * v[1] + d[2] * v[2];
    mn = t; mx = t;
    for (int i = 1; i < 3; ++i) {
        t = d[0] * v[3 * i] + d[1] * v[3 * i + 1] + d[2] * v[3 * i + 2];
        if (t > mx) mx = t;
        else if (t < mn) mn = t;
    }
}

// rbsp.cpp:312-318
WF_HD float sah_cost(uint32_t traversalCost, uint32_t isectCost, float emptyBonus, float areaBelow, float areaAbove,
                     float invTotalSA, uint32_t nBelow, uint32_t nAbove) {
    const float pBelow = areaBelow * invTotalSA;
    const float pAbove = areaAbove * invTotalSA;
    const float eb = (nAbove == 0 || nBelow == 0) ? emptyBonus : 0;
    return traversalCost + isectCost * (1 - eb) * (pBelow * nBelow + pAbove * nAbove);
}

// getBitOffset(M) = log2_fast(M + 1), kDOPMesh.h:229-236
WF_HD uint32_t bit_offset(uint32_t M) {
    uint32_t b = 0;
    while ((1u << b) <= M) ++b;   // == 32 - clz(M)
    return b;
}

// order-preserving map float -> uint32 (used for sort keys and arg-min; -0 is folded into +0 by the callers)
WF_HD uint32_t float_order(float f) {
    const uint32_t u = f2u(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
WF_HD float float_unorder(uint32_t o) {
    const uint32_t u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
    return u2f(u);
}

}  // namespace rbsp
}  // namespace b200rt
