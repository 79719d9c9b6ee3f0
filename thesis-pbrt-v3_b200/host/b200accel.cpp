// b200accel.cpp — `Accelerator "b200"`: a pbrt Aggregate whose Intersect/IntersectP run on the B200 through
// the C ABI of libb200rt.so.  See b200accel.h for the interface it mirrors.
//
// Geometry / tree capture: GeometricPrimitive::shape (primitive.h:135-140), Triangle::mesh / v
// (triangle.h:119-123), Sphere::radius/zMin/zMax/phiMax (sphere.h:71-75), BVHAccel::nodes / primNumMapping
// (bvh.h:76-83) and GenericBSP<>::nodes / nextFreeNode / primitiveIndices / directions / bounds
// (genericBSP.h:180-193) are private or protected in the reference and have no accessors.  Inside the reference
// tree the one-line fix is `friend class B200Accel;` in those five classes; because this repo never edits the
// reference, this TU opens member access for exactly those headers instead (all standard headers are included
// first so that the macro cannot touch the standard library).
#include <algorithm>
#include <array>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <functional>
#include <iostream>
#include <limits>
#include <list>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <sstream>
#include <string>
#include <thread>
#include <unordered_map>
#include <utility>
#include <vector>
#include <glog/logging.h>

#define private public
#define protected public
#include "core/primitive.h"
#include "shapes/triangle.h"
#include "shapes/sphere.h"
#include "accelerators/bvh.h"
#include "accelerators/kdtreeaccel.h"
#include "accelerators/rbsp.h"
#include "accelerators/rbspKd.h"
#include "accelerators/bspPaper.h"
#include "accelerators/bspPaperKd.h"
#include "accelerators/bspCluster.h"
#include "accelerators/bspClusterFastKd.h"
#undef private
#undef protected
#include "core/interaction.h"
#include "core/paramset.h"
#include "core/transform.h"

#include "b200accel.h"

namespace pbrt {

static std::atomic<const B200Accel *> g_current{nullptr};
const B200Accel *B200Accel::current() { return g_current.load(); }

namespace {

struct Captured {
    std::vector<float> verts, uvs, spheres, normals;
    std::vector<uint8_t> kind, flags;   // flags: b200rt_scene_desc::prim_flags
    bool anyUV = false, anyN = false, anyFlag = false;
    uint32_t nSpheres = 0;
};

// One GeometricPrimitive per triangle / sphere, as pbrtShape creates them (api.cpp:1576-1597).
bool capture(const std::vector<std::shared_ptr<Primitive>> &prims, Captured *c) {
    const size_t n = prims.size();
    c->verts.assign(9 * n, 0.f);
    c->uvs.assign(6 * n, 0.f);
    c->kind.assign(n, 0);
    c->flags.assign(n, 0);
    for (size_t i = 0; i < n; ++i) {
        auto *gp = dynamic_cast<const GeometricPrimitive *>(prims[i].get());
        if (!gp) {
            Error("b200: primitive %d is not a GeometricPrimitive (instancing / motion is not on the b200 path)", (int)i);
            return false;
        }
        const Shape *sh = gp->shape.get();
        if (auto *tri = dynamic_cast<const Triangle *>(sh)) {
            const TriangleMesh *m = tri->mesh.get();
            if (m->alphaMask || m->shadowAlphaMask) {
                Error("b200: alpha-masked triangle meshes are unsupported (triangle.cpp:334,533)");
                return false;
            }
            for (int k = 0; k < 3; ++k) {
                const Point3f &q = m->p[tri->v[k]];
                c->verts[9 * i + 3 * k] = q.x; c->verts[9 * i + 3 * k + 1] = q.y; c->verts[9 * i + 3 * k + 2] = q.z;
            }
            // Triangle::GetUVs (triangle.h:109-118): mesh uv or the (0,0),(1,0),(1,1) default
            float *uv = &c->uvs[6 * i];
            if (m->uv) {
                c->anyUV = true;
                for (int k = 0; k < 3; ++k) { uv[2 * k] = m->uv[tri->v[k]].x; uv[2 * k + 1] = m->uv[tri->v[k]].y; }
            } else {
                uv[0] = 0; uv[1] = 0; uv[2] = 1; uv[3] = 0; uv[4] = 1; uv[5] = 1;
            }
            // what the shading-geometry tail of Triangle::Intersect reads (triangle.cpp:346-424): per-vertex normals
            // (interpolated shading normal), the orientation flag; per-vertex tangents (mesh->s) are not on the device
            if (m->s) {
                Error("b200: triangle meshes with per-vertex tangents (\"S\") are unsupported");
                return false;
            }
            if (m->n) {
                if (!c->anyN) { c->normals.assign(9 * n, 0.f); c->anyN = true; }
                for (int k = 0; k < 3; ++k) {
                    const Normal3f &q = m->n[tri->v[k]];
                    c->normals[9 * i + 3 * k] = q.x; c->normals[9 * i + 3 * k + 1] = q.y; c->normals[9 * i + 3 * k + 2] = q.z;
                }
                c->flags[i] |= 2;
            }
            if (tri->reverseOrientation ^ tri->transformSwapsHandedness) c->flags[i] |= 1;
            if (c->flags[i]) c->anyFlag = true;
        } else if (auto *sp = dynamic_cast<const Sphere *>(sh)) {
            const Float twoPi = 2 * Pi;
            if (sp->zMin > -sp->radius || sp->zMax < sp->radius || sp->phiMax < twoPi) {
                Error("b200: partial spheres are unsupported");
                return false;
            }
            c->kind[i] = 1;
            const Matrix4x4 &w2o = sp->WorldToObject->GetMatrix();
            for (int r = 0; r < 3; ++r)
                for (int col = 0; col < 4; ++col) c->spheres.push_back(w2o.m[r][col]);
            c->spheres.push_back(sp->radius);
            c->spheres.push_back(0); c->spheres.push_back(0); c->spheres.push_back(0);
            c->nSpheres++;
        } else {
            Error("b200: primitive %d is neither a Triangle nor a Sphere", (int)i);
            return false;
        }
    }
    return true;
}

template <class Tree>
void describeBSP(const Tree *t, b200rt_scene_desc *d, std::vector<float> *dirs) {
    static_assert(sizeof(Float) == 4, "b200rt computes in binary32");
    d->nodes = (const void *)t->nodes;          // 8-byte nodes (kdtreeaccel.cpp:43-160, rbsp.cpp:15-135) or 20-byte BSP nodes
    d->node_count = t->nextFreeNode;
    d->prim_indices = t->primitiveIndices.data();
    d->n_indices = (uint32_t)t->primitiveIndices.size();
    for (const Vector3f &v : t->directions) { dirs->push_back(v.x); dirs->push_back(v.y); dirs->push_back(v.z); }
    d->directions = dirs->data();
    d->n_directions = (uint32_t)t->directions.size();
    const Bounds3f &b = t->bounds;
    d->bounds[0] = b.pMin.x; d->bounds[1] = b.pMin.y; d->bounds[2] = b.pMin.z;
    d->bounds[3] = b.pMax.x; d->bounds[4] = b.pMax.y; d->bounds[5] = b.pMax.z;
}

inline b200rt_ray toDeviceRay(const Ray &r) {
    b200rt_ray q;
    q.o[0] = r.o.x; q.o[1] = r.o.y; q.o[2] = r.o.z; q.tMax = r.tMax;
    q.d[0] = r.d.x; q.d[1] = r.d.y; q.d[2] = r.d.z; q.time = r.time;
    return q;
}

}  // namespace

B200Accel::B200Accel(std::vector<std::shared_ptr<Primitive>> p, const ParamSet &ps) : prims(std::move(p)) {
    if (prims.empty()) { Error("b200: no primitives"); return; }
    Captured cap;
    if (!capture(prims, &cap)) return;

    const std::string tree = ps.FindOneString("tree", "bvh");
    const std::string builder = ps.FindOneString("builder", "reference");
    b200rt_scene_desc d;
    std::memset(&d, 0, sizeof(d));
    std::shared_ptr<Primitive> host;   // keeps the reference-built tree alive until the upload has copied it
    std::vector<float> dirs;
    void *ownNodes = nullptr;          // builder "b200": arrays malloc'ed by the library
    uint32_t *ownIdx = nullptr;

    if (builder == "b200") {
        if (cap.nSpheres) { Error("b200: builder \"b200\" handles triangle-only aggregates"); return; }
        uint32_t nNodes = 0, nIdx = 0;
        int rc;
        if (tree == "bvh") {            // parameter names and defaults of CreateBVHAccelerator (bvh.cpp:529-535)
            rc = b200rt_build_bvh_device(cap.verts.data(), (uint32_t)prims.size(), ps.FindOneInt("maxnodeprims", 4),
                                         ps.FindOneInt("intersectcost", 8), ps.FindOneInt("traversalcost", 1), &ownNodes,
                                         &nNodes, &ownIdx, nullptr);
            d.kind = B200RT_BVH;
            nIdx = (uint32_t)prims.size();
        } else if (tree == "kdtree") {  // CreateKdTreeAccelerator (kdtreeaccel.cpp:523-545)
            rc = b200rt_build_kdtree_device(cap.verts.data(), (uint32_t)prims.size(), ps.FindOneInt("intersectcost", 80),
                                            ps.FindOneInt("traversalcost", 1), ps.FindOneFloat("emptybonus", 0.f),
                                            ps.FindOneInt("maxprims", 1), ps.FindOneInt("maxdepth", -1), &ownNodes, &nNodes,
                                            &ownIdx, &nIdx, d.bounds, nullptr);
            d.kind = B200RT_KDTREE;
            d.n_directions = 3;
        } else if (tree == "rbsp" || tree == "rbspkd") {   // CreateRBSP(Kd)TreeAccelerator (rbsp.cpp:549-570, rbspKd.cpp:640-665)
            const int M = ps.FindOneInt("nbDirections", 3);
            dirs.assign(3 * 16, 0.f);
            if (tree == "rbsp")
                rc = b200rt_build_rbsp(cap.verts.data(), (uint32_t)prims.size(), (uint32_t)M, ps.FindOneInt("intersectcost", 80),
                                       ps.FindOneInt("traversalcost", 5), ps.FindOneFloat("emptybonus", 0.f),
                                       ps.FindOneInt("maxprims", 1), ps.FindOneInt("maxdepth", -1), &ownNodes, &nNodes, &ownIdx,
                                       &nIdx, d.bounds, dirs.data(), nullptr);
            else
                rc = b200rt_build_rbspkd(cap.verts.data(), (uint32_t)prims.size(), (uint32_t)M, ps.FindOneInt("intersectcost", 80),
                                         ps.FindOneInt("traversalcost", 5), ps.FindOneInt("kdtraversalcost", 1),
                                         ps.FindOneFloat("emptybonus", 0.f), ps.FindOneInt("maxprims", 1),
                                         ps.FindOneInt("maxdepth", -1), &ownNodes, &nNodes, &ownIdx, &nIdx, d.bounds, dirs.data(), nullptr);
            d.kind = tree == "rbsp" ? B200RT_RBSP : B200RT_RBSPKD;
            d.directions = dirs.data();
            d.n_directions = (uint32_t)M;
        } else {
            Error("b200: builder \"b200\" knows \"bvh\", \"kdtree\", \"rbsp\" and \"rbspkd\"; \"%s\" needs the reference's builder", tree.c_str());
            return;
        }
        if (rc) { Error("b200: %s", b200rt_last_error()); return; }
        d.nodes = ownNodes; d.node_count = nNodes; d.prim_indices = ownIdx; d.n_indices = nIdx;
    } else if (tree == "bvh") {
        std::shared_ptr<BVHAccel> b = CreateBVHAccelerator(prims, ps);
        host = b;
        d.kind = B200RT_BVH;
        d.nodes = (const void *)b->nodes;
        // LinearBVHNode is private to bvh.cpp; its 32-byte layout is the ABI contract (bvh.cpp:123-152): word 6 =
        // secondChildOffset, word 7 bits 0-1 = axis (3 = leaf).  Depth-first order => the node count is the index
        // of the right-most leaf + 1.
        const uint32_t *w = (const uint32_t *)d.nodes;
        uint32_t i = 0;
        while (w && (w[8 * i + 7] & 3u) != 3u) i = w[8 * i + 6];
        d.node_count = w ? i + 1 : 0;
        d.prim_indices = b->primNumMapping.data();
        d.n_indices = (uint32_t)b->primNumMapping.size();
    } else if (tree == "kdtree") {
        std::shared_ptr<KdTreeAccel> t = CreateKdTreeAccelerator(prims, ps);
        host = t; d.kind = B200RT_KDTREE; describeBSP(t.get(), &d, &dirs);
    } else if (tree == "rbsp") {
        std::shared_ptr<RBSP> t = CreateRBSPTreeAccelerator(prims, ps);
        host = t; d.kind = B200RT_RBSP; describeBSP(t.get(), &d, &dirs);
    } else if (tree == "rbspkd") {
        std::shared_ptr<RBSPKd> t = CreateRBSPKdTreeAccelerator(prims, ps);
        host = t; d.kind = B200RT_RBSPKD; describeBSP(t.get(), &d, &dirs);
    } else if (tree == "bsppaper" || tree == "bspcluster") {
        // general BSP families over 20-byte BSPNode (BSP.h:122-184); the traversal never reads `directions`
        std::shared_ptr<BSP> t;
        if (tree == "bsppaper") t = CreateBSPPaperTreeAccelerator(prims, ps);
        else t = CreateBSPClusterTreeAccelerator(prims, ps);
        host = t; d.kind = B200RT_BSP; describeBSP(t.get(), &d, &dirs);
        d.directions = nullptr; d.n_directions = 0;
    } else if (tree == "bsppaperkd" || tree == "bspclusterfastkd") {
        std::shared_ptr<BSPKd> t;   // 20-byte BSPKdNode (BSPKd.h:11-174)
        if (tree == "bsppaperkd") t = CreateBSPPaperKdTreeAccelerator(prims, ps);
        else t = CreateBSPClusterFastKdTreeAccelerator(prims, ps);
        host = t; d.kind = B200RT_BSPKD; describeBSP(t.get(), &d, &dirs);
        d.directions = nullptr; d.n_directions = 0;
    } else {
        Error("b200: unknown \"tree\" \"%s\" (bvh | kdtree | rbsp | rbspkd | bsppaper | bsppaperkd | bspcluster | "
              "bspclusterfastkd)", tree.c_str());
        return;
    }
    d.tri_verts = cap.verts.data();
    d.tri_uvs = cap.anyUV ? cap.uvs.data() : nullptr;
    d.tri_normals = cap.anyN ? cap.normals.data() : nullptr;
    d.prim_flags = cap.anyFlag ? cap.flags.data() : nullptr;
    d.n_prims = (uint32_t)prims.size();
    if (cap.nSpheres) { d.prim_kind = cap.kind.data(); d.sphere_params = cap.spheres.data(); d.n_spheres = cap.nSpheres; }

    int rc = b200rt_init(ps.FindOneInt("device", 0));
    if (!rc) rc = b200rt_scene_create(&d, &scene);
    if (ownNodes) b200rt_free(ownNodes);
    if (ownIdx) b200rt_free(ownIdx);
    if (rc) { Error("b200: %s", b200rt_last_error()); scene = nullptr; return; }
    kind = (int)d.kind;
    for (const auto &q : prims) bounds = Union(bounds, q->WorldBound());
    g_current.store(this);
}

B200Accel::~B200Accel() {
    const B200Accel *self = this;
    g_current.compare_exchange_strong(self, nullptr);
    if (scene) b200rt_scene_destroy(scene);
}

Bounds3f B200Accel::WorldBound() const { return bounds; }

bool B200Accel::FillInteraction(const Ray &ray, const b200rt_hit &h, SurfaceInteraction *isect) const {
    if (h.prim == B200RT_MISS) return false;
    // The winning primitive's own Intersect fills dpdu/dpdv, pError, uv, the shading frame and
    // isect->primitive, and sets ray.tMax (primitive.cpp:123-139); the device t must be that very float.
    const Float tIn = ray.tMax;
    bool ok = prims[h.prim]->Intersect(ray, isect);
    if (!ok || ray.tMax != h.t) {
        Error("b200: host re-intersection of primitive %u disagrees with the device (t %.9g vs %.9g)", h.prim,
              ok ? ray.tMax : tIn, h.t);
        return false;
    }
    return true;
}

bool B200Accel::Intersect(const Ray &ray, SurfaceInteraction *isect) const {
    b200rt_ray r = toDeviceRay(ray);
    b200rt_hit h;
    if (b200rt_intersect_closest(scene, &r, 1, &h, nullptr)) { Error("b200: %s", b200rt_last_error()); return false; }
    return FillInteraction(ray, h, isect);
}

bool B200Accel::IntersectP(const Ray &ray) const {
    b200rt_ray r = toDeviceRay(ray);
    uint8_t occ = 0;
    if (b200rt_intersect_any(scene, &r, 1, &occ, nullptr)) { Error("b200: %s", b200rt_last_error()); return false; }
    return occ != 0;
}

bool B200Accel::IntersectBatch(const Ray *rays, size_t n, b200rt_hit *hits) const {
    std::vector<b200rt_ray> q(n);
    for (size_t i = 0; i < n; ++i) q[i] = toDeviceRay(rays[i]);
    if (b200rt_intersect_closest(scene, q.data(), n, hits, nullptr)) { Error("b200: %s", b200rt_last_error()); return false; }
    for (size_t i = 0; i < n; ++i) rays[i].tMax = hits[i].t;   // Ray::tMax is mutable (geometry.h:1200)
    return true;
}

bool B200Accel::IntersectPBatch(const Ray *rays, size_t n, uint8_t *occluded) const {
    std::vector<b200rt_ray> q(n);
    for (size_t i = 0; i < n; ++i) q[i] = toDeviceRay(rays[i]);
    if (b200rt_intersect_any(scene, q.data(), n, occluded, nullptr)) { Error("b200: %s", b200rt_last_error()); return false; }
    return true;
}

std::shared_ptr<Primitive> CreateB200Accelerator(std::vector<std::shared_ptr<Primitive>> prims, const ParamSet &ps) {
    auto a = std::make_shared<B200Accel>(std::move(prims), ps);
    if (!a->ok()) return nullptr;   // MakeScene falls back to the reference's BVHAccel (api.cpp:1886)
    return a;
}

}  // namespace pbrt
