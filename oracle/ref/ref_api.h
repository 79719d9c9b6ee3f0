// C API of oracle/_ref/libpbrtref.so — TEST INFRASTRUCTURE ONLY.
//
// libpbrtref.so is the UNMODIFIED reference (jhoobergs/Thesis-pbrt-v3) hot path compiled with plain
// g++ from the sources where they lie under /root/reference/src (recipe: oracle/ref/Makefile) plus
// the thin shim TUs in this directory, which only (a) build pbrt Primitives from flat arrays,
// (b) call the reference's own Create*Accelerator / Aggregate::Intersect / IntersectP, and
// (c) export the reference's flattened node arrays.  Nothing here is used by the product path.
#ifndef B200RT_REF_API_H
#define B200RT_REF_API_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct ref_scene ref_scene;   // vector<shared_ptr<Primitive>> + ptr->index map
typedef struct ref_accel ref_accel;   // shared_ptr<Primitive> aggregate built by the reference

// Triangle soup -> one GeometricPrimitive per triangle (api.cpp:1576-1597), identity transform.
// uv may be NULL.  Spheres: n_spheres x (cx,cy,cz,radius), appended AFTER the triangles
// (pure Translate object-to-world, as in scenes/killeroo-simple:20-24).
ref_scene *ref_scene_create(const float *P, int n_verts, const int *indices, int n_tris,
                            const float *uv, const float *spheres, int n_spheres);
ref_scene *ref_scene_create2(const float *P, int n_verts, const int *indices, int n_tris, const float *uv,
                             const float *N, int reverse_orientation, const float *spheres, int n_spheres);
void ref_scene_destroy(ref_scene *);
int ref_scene_num_prims(const ref_scene *);

// Reference loop subdivision (shapes/loopsubdiv.cpp:402 CreateLoopSubdiv). Caller frees with ref_free.
int ref_loopsubdiv(const float *P, int n_verts, const int *indices, int n_indices, int n_levels,
                   float **outP, int *out_n_verts, int **out_indices, int *out_n_tris);
void ref_free(void *);

// kind: "bvh" | "kdtree" | "rbsp" | "rbspkd".  int_params / float_params are NULL-terminated
// name lists with parallel value arrays, forwarded through a pbrt ParamSet to the reference's
// own Create*Accelerator (same names as the .pbrt Accelerator line).
ref_accel *ref_accel_create(ref_scene *, const char *kind,
                            const char *const *int_names, const int *int_vals,
                            const char *const *float_names, const float *float_vals);
void ref_accel_destroy(ref_accel *);

// Export of the reference's flattened tree.  Query sizes first (pass NULL buffers).
//  BVH : nodes = node_count x 32 B LinearBVHNode (bvh.cpp:123-152); prim_order[slot] = original
//        primitive number (primNumMapping, bvh.cpp:227,322); indices unused.
//  kd/RBSP(/Kd): nodes = node_count x 8 B (kdtreeaccel.cpp:43-160, rbsp.cpp:15-135);
//        indices = primitiveIndices; directions = M x 3 floats; bounds = 6 floats.
typedef struct {
    uint32_t node_count, node_stride, n_indices, n_directions, n_prims;
    float bounds[6];
} ref_tree_info;
int ref_accel_info(const ref_accel *, ref_tree_info *);
int ref_accel_export(const ref_accel *, void *nodes, uint32_t *indices_or_prim_order, float *directions);

// Per-ray results of the reference's own Intersect / IntersectP.
// rays: n x 8 floats (ox,oy,oz,tMax,dx,dy,dz,time).
typedef struct {
    int32_t prim;      // original primitive number, -1 = miss
    float t;           // ray.tMax after the call
    float uv[2];       // SurfaceInteraction::uv
    float p[3];        // SurfaceInteraction::p
} ref_hit;
typedef struct {       // from Ray::stats (geometry.h:1078-1173)
    uint32_t prim_tests, interior_visits, leaf_visits;
} ref_counters;
void ref_intersect(const ref_accel *, const float *rays, uint64_t n, ref_hit *hits,
                   ref_counters *counters /*nullable*/, int n_threads);
void ref_intersect_p(const ref_accel *, const float *rays, uint64_t n, uint8_t *occluded,
                     ref_counters *counters /*nullable*/, int n_threads);

// Direct single-shape tests (Triangle::Intersect / IntersectP through GeometricPrimitive).
void ref_prim_intersect(const ref_scene *, int prim, const float *rays, uint64_t n, ref_hit *hits);
void ref_prim_intersect_p(const ref_scene *, int prim, const float *rays, uint64_t n, uint8_t *occluded);

#ifdef __cplusplus
}
#endif
#endif
