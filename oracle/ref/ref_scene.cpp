// Shim TU: builds pbrt Primitives from flat arrays and drives the reference's own
// Aggregate::Intersect / IntersectP over ray arrays.  TEST INFRASTRUCTURE ONLY (see ref_api.h).
#include "std_first.h"
#include "ref_internal.h"
#define private public
#define protected public
#include "shapes/triangle.h"
#include "core/primitive.h"
#undef private
#undef protected
#include "shapes/sphere.h"
#include "shapes/loopsubdiv.h"
#include "core/paramset.h"
#include "core/interaction.h"
#include "core/parser.h"
#include "core/api.h"

namespace pbrt {
// Defined in api.cpp / parser.cpp, which are outside the hot path and not compiled here.
Options PbrtOptions;
Loc *parserLoc = nullptr;
}  // namespace pbrt
using namespace pbrt;

static Transform g_identity;

extern "C" ref_scene *ref_scene_create(const float *P, int n_verts, const int *indices, int n_tris,
                                       const float *uv, const float *spheres, int n_spheres) {
    ref_scene *s = new ref_scene;
    PbrtOptions.quiet = true;  // no progress bars from the builders
    MediumInterface mi;
    if (n_tris > 0) {
        std::vector<Point3f> p(n_verts);
        for (int i = 0; i < n_verts; ++i) p[i] = Point3f(P[3 * i], P[3 * i + 1], P[3 * i + 2]);
        std::vector<Point2f> uvs;
        if (uv) {
            uvs.resize(n_verts);
            for (int i = 0; i < n_verts; ++i) uvs[i] = Point2f(uv[2 * i], uv[2 * i + 1]);
        }
        std::vector<std::shared_ptr<Shape>> shapes =
            CreateTriangleMesh(&g_identity, &g_identity, false, n_tris, indices, n_verts, p.data(), nullptr,
                               nullptr, uv ? uvs.data() : nullptr, nullptr, nullptr);
        for (auto &sh : shapes)
            s->prims.push_back(std::make_shared<GeometricPrimitive>(sh, nullptr, nullptr, mi));
    }
    for (int i = 0; i < n_spheres; ++i) {
        const float *c = spheres + 4 * i;
        std::unique_ptr<Transform> o2w(new Transform(Translate(Vector3f(c[0], c[1], c[2]))));
        std::unique_ptr<Transform> w2o(new Transform(Inverse(*o2w)));
        auto sh = std::make_shared<Sphere>(o2w.get(), w2o.get(), false, c[3], -c[3], c[3], 360.f);
        s->transforms.push_back(std::move(o2w));
        s->transforms.push_back(std::move(w2o));
        s->prims.push_back(std::make_shared<GeometricPrimitive>(sh, nullptr, nullptr, mi));
    }
    for (size_t i = 0; i < s->prims.size(); ++i) s->index[s->prims[i].get()] = (int)i;
    return s;
}
// ref_scene_create + per-vertex normals N (n_verts x 3, nullable; kept by the mesh as TriangleMesh::n) and the shape's
// reverseOrientation flag (`ReverseOrientation` in a scene file): what Triangle::Intersect's shading-geometry tail reads.
extern "C" ref_scene *ref_scene_create2(const float *P, int n_verts, const int *indices, int n_tris, const float *uv,
                                        const float *N, int reverse_orientation, const float *spheres, int n_spheres) {
    ref_scene *s = ref_scene_create(nullptr, 0, nullptr, 0, nullptr, spheres, n_spheres);
    std::vector<std::shared_ptr<Primitive>> sph;
    sph.swap(s->prims);
    MediumInterface mi;
    if (n_tris > 0) {
        std::vector<Point3f> p(n_verts);
        for (int i = 0; i < n_verts; ++i) p[i] = Point3f(P[3 * i], P[3 * i + 1], P[3 * i + 2]);
        std::vector<Point2f> uvs;
        if (uv) { uvs.resize(n_verts); for (int i = 0; i < n_verts; ++i) uvs[i] = Point2f(uv[2 * i], uv[2 * i + 1]); }
        std::vector<Normal3f> ns;
        if (N) { ns.resize(n_verts); for (int i = 0; i < n_verts; ++i) ns[i] = Normal3f(N[3 * i], N[3 * i + 1], N[3 * i + 2]); }
        std::vector<std::shared_ptr<Shape>> shapes =
            CreateTriangleMesh(&g_identity, &g_identity, reverse_orientation != 0, n_tris, indices, n_verts, p.data(), nullptr,
                               N ? ns.data() : nullptr, uv ? uvs.data() : nullptr, nullptr, nullptr);
        for (auto &sh : shapes) s->prims.push_back(std::make_shared<GeometricPrimitive>(sh, nullptr, nullptr, mi));
    }
    for (auto &q : sph) s->prims.push_back(q);
    s->index.clear();
    for (size_t i = 0; i < s->prims.size(); ++i) s->index[s->prims[i].get()] = (int)i;
    return s;
}
extern "C" void ref_scene_destroy(ref_scene *s) { delete s; }
extern "C" int ref_scene_num_prims(const ref_scene *s) { return (int)s->prims.size(); }
extern "C" void ref_free(void *p) { std::free(p); }

static std::vector<float> g_lastSubdivN;
extern "C" int ref_loopsubdiv(const float *P, int n_verts, const int *indices, int n_indices, int n_levels,
                              float **outP, int *out_n_verts, int **out_indices, int *out_n_tris) {
    ParamSet ps;
    std::unique_ptr<int[]> lv(new int[1]);
    lv[0] = n_levels;
    ps.AddInt("nlevels", std::move(lv), 1);
    std::unique_ptr<int[]> idx(new int[n_indices]);
    std::memcpy(idx.get(), indices, sizeof(int) * n_indices);
    ps.AddInt("indices", std::move(idx), n_indices);
    std::unique_ptr<Point3f[]> pts(new Point3f[n_verts]);
    for (int i = 0; i < n_verts; ++i) pts[i] = Point3f(P[3 * i], P[3 * i + 1], P[3 * i + 2]);
    ps.AddPoint3f("P", std::move(pts), n_verts);
    std::vector<std::shared_ptr<Shape>> shapes = CreateLoopSubdiv(&g_identity, &g_identity, false, ps);
    if (shapes.empty()) return -1;
    const Triangle *t0 = static_cast<const Triangle *>(shapes[0].get());
    const TriangleMesh *m = t0->mesh.get();
    *out_n_verts = m->nVertices;
    *out_n_tris = m->nTriangles;
    *outP = (float *)std::malloc(sizeof(float) * 3 * m->nVertices);
    *out_indices = (int *)std::malloc(sizeof(int) * 3 * m->nTriangles);
    for (int i = 0; i < m->nVertices; ++i) {
        (*outP)[3 * i] = m->p[i].x; (*outP)[3 * i + 1] = m->p[i].y; (*outP)[3 * i + 2] = m->p[i].z;
    }
    std::memcpy(*out_indices, m->vertexIndices.data(), sizeof(int) * 3 * m->nTriangles);
    // the limit-surface normals LoopSubdivide hands to CreateTriangleMesh (loopsubdiv.cpp:395-398), as the mesh stores them
    g_lastSubdivN.clear();
    if (m->n)
        for (int i = 0; i < m->nVertices; ++i) { g_lastSubdivN.push_back(m->n[i].x); g_lastSubdivN.push_back(m->n[i].y); g_lastSubdivN.push_back(m->n[i].z); }
    return 0;
}
// Per-vertex normals of the LAST ref_loopsubdiv result (n_verts x 3; returns 0 when the mesh has none).
extern "C" int ref_loopsubdiv_normals(float *out) {
    if (out) std::memcpy(out, g_lastSubdivN.data(), sizeof(float) * g_lastSubdivN.size());
    return (int)(g_lastSubdivN.size() / 3);
}

// Transform::operator()(Normal3f) of the Transform pbrtTransform builds from a row-major 4x4 (its inverse computed by the
// reference's own Inverse()), i.e. what TriangleMesh's constructor applies to a plymesh's normals (triangle.cpp:61-64); and
// Transform::SwapsHandedness (transform.cpp:238-243).  Used by tests/golden/make_living_room.py.
extern "C" int ref_transform_normals(const float *m16, const float *N, int n, float *out) {
    Matrix4x4 m(m16[0], m16[1], m16[2], m16[3], m16[4], m16[5], m16[6], m16[7], m16[8], m16[9], m16[10], m16[11], m16[12],
                m16[13], m16[14], m16[15]);
    Transform t(m);
    for (int i = 0; i < n; ++i) {
        const Normal3f r = t(Normal3f(N[3 * i], N[3 * i + 1], N[3 * i + 2]));
        out[3 * i] = r.x; out[3 * i + 1] = r.y; out[3 * i + 2] = r.z;
    }
    return t.SwapsHandedness() ? 1 : 0;
}

extern "C" ref_accel *ref_accel_create(ref_scene *s, const char *kind, const char *const *int_names,
                                       const int *int_vals, const char *const *float_names,
                                       const float *float_vals) {
    ParamSet ps;
    for (int i = 0; int_names && int_names[i]; ++i) {
        std::unique_ptr<int[]> v(new int[1]);
        v[0] = int_vals[i];
        ps.AddInt(int_names[i], std::move(v), 1);
    }
    for (int i = 0; float_names && float_names[i]; ++i) {
        std::unique_ptr<Float[]> v(new Float[1]);
        v[0] = float_vals[i];
        ps.AddFloat(float_names[i], std::move(v), 1);
    }
    std::string k(kind);
    if (k == "bvh") return ref_make_bvh(s, ps);
    if (k == "kdtree") return ref_make_kd(s, ps);
    if (k == "rbsp") return ref_make_rbsp(s, ps);
    if (k == "rbspkd") return ref_make_rbspkd(s, ps);
    return ref_make_bsp(s, kind, ps);
}
extern "C" void ref_accel_destroy(ref_accel *a) { delete a; }
extern "C" int ref_accel_info(const ref_accel *a, ref_tree_info *ti) { return a->info(a, ti); }
extern "C" int ref_accel_export(const ref_accel *a, void *nodes, uint32_t *idx, float *dirs) {
    return a->exp(a, nodes, idx, dirs);
}

static inline Ray make_ray(const float *r) {
    return Ray(Point3f(r[0], r[1], r[2]), Vector3f(r[4], r[5], r[6]), r[3], r[7]);
}

template <typename F>
static void run_chunks(uint64_t n, int n_threads, F f) {
    if (n_threads <= 1) { f(0, n); return; }
    std::vector<std::thread> th;
    uint64_t chunk = (n + n_threads - 1) / n_threads;
    for (int t = 0; t < n_threads; ++t) {
        uint64_t b = std::min<uint64_t>(n, t * chunk), e = std::min<uint64_t>(n, b + chunk);
        if (b < e) th.emplace_back([=] { f(b, e); });
    }
    for (auto &t : th) t.join();
}

static void do_intersect(const Primitive *agg, const ref_scene *sc, const float *rays, uint64_t n, ref_hit *hits,
                         ref_counters *cnt, int n_threads) {
    run_chunks(n, n_threads, [=](uint64_t b, uint64_t e) {
        for (uint64_t i = b; i < e; ++i) {
            Ray ray = make_ray(rays + 8 * i);
            SurfaceInteraction isect;
            bool hit = agg->Intersect(ray, &isect);
            ref_hit h;
            std::memset(&h, 0, sizeof(h));
            h.prim = -1;
            h.t = ray.tMax;
            if (hit) {
                auto it = sc->index.find(isect.primitive);
                h.prim = it == sc->index.end() ? -2 : it->second;
                h.uv[0] = isect.uv.x; h.uv[1] = isect.uv.y;
                h.p[0] = isect.p.x; h.p[1] = isect.p.y; h.p[2] = isect.p.z;
            }
            hits[i] = h;
            if (cnt) {
                cnt[i].prim_tests = (uint32_t)ray.stats.primitiveIntersections;
                cnt[i].interior_visits = (uint32_t)(ray.stats.bvhTreeNodeTraversals + ray.stats.kdTreeNodeTraversals +
                                                    ray.stats.bspTreeNodeTraversals);
                cnt[i].leaf_visits = (uint32_t)ray.stats.leafNodeTraversals;
            }
        }
    });
}
static void do_intersect_p(const Primitive *agg, const float *rays, uint64_t n, uint8_t *occ, ref_counters *cnt,
                           int n_threads) {
    run_chunks(n, n_threads, [=](uint64_t b, uint64_t e) {
        for (uint64_t i = b; i < e; ++i) {
            Ray ray = make_ray(rays + 8 * i);
            occ[i] = agg->IntersectP(ray) ? 1 : 0;
            if (cnt) {
                cnt[i].prim_tests = (uint32_t)ray.stats.primitiveIntersectionsP;
                cnt[i].interior_visits = (uint32_t)(ray.stats.bvhTreeNodeTraversalsP + ray.stats.kdTreeNodeTraversalsP +
                                                    ray.stats.bspTreeNodeTraversalsP);
                cnt[i].leaf_visits = (uint32_t)ray.stats.leafNodeTraversalsP;
            }
        }
    });
}
extern "C" void ref_intersect(const ref_accel *a, const float *rays, uint64_t n, ref_hit *hits, ref_counters *cnt,
                              int n_threads) {
    do_intersect(a->agg.get(), a->scene, rays, n, hits, cnt, n_threads);
}
extern "C" void ref_intersect_p(const ref_accel *a, const float *rays, uint64_t n, uint8_t *occ, ref_counters *cnt,
                                int n_threads) {
    do_intersect_p(a->agg.get(), rays, n, occ, cnt, n_threads);
}
extern "C" void ref_prim_intersect(const ref_scene *s, int prim, const float *rays, uint64_t n, ref_hit *hits) {
    do_intersect(s->prims[prim].get(), s, rays, n, hits, nullptr, 1);
}
extern "C" void ref_prim_intersect_p(const ref_scene *s, int prim, const float *rays, uint64_t n, uint8_t *occ) {
    do_intersect_p(s->prims[prim].get(), rays, n, occ, nullptr, 1);
}
