/*
 * b200rt.h — C ABI of the B200-native ray-intersection engine (libb200rt.so).
 *
 * This is the drop-in boundary for ONE hot path of jhoobergs/Thesis-pbrt-v3:
 *   Aggregate::Intersect / IntersectP over the flattened BVH / kd-tree / RBSP(/Kd) node arrays
 *   plus the watertight Triangle::Intersect, batched over ray arrays.
 *
 * Each entry point cites the reference interface it replaces (paths relative to the reference
 * tree).  Conventions: plain pointers and sizes only; every call returns 0 on success or a
 * B200RT_E* code (never throws across the ABI; b200rt_last_error() has the text); the caller keeps
 * ownership of every host buffer it passes in; a scene is immutable after creation; intersect
 * calls are thread-safe on distinct streams.  There is NO CPU fallback: without a CUDA device every
 * compute entry point fails with B200RT_ENODEVICE.
 */
#ifndef B200RT_H
#define B200RT_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ABI history (b200rt_version()): 8 the reference's sample streams, camera matrices and the triangle tail's inputs; 9 image textures
 * (trilinear); 10 several lights (b200rt_light, light_strategy: uniform / power / spatial), the EWA texture filter
 * (b200rt_texture::filter, max_anisotropy), b200rt_intersect_closest_async / _any_async / _wait, b200rt_ewa_weights.
 * Structures only grow at their end; zero in a new field means the previous version's behaviour. */
#define B200RT_VERSION 10

enum {
    B200RT_OK = 0,
    B200RT_EINVAL = 1,     /* bad argument / malformed tree */
    B200RT_ENODEVICE = 2,  /* no CUDA device (there is no CPU path) */
    B200RT_ECUDA = 3,      /* CUDA runtime error, see b200rt_last_error() */
    B200RT_EUNSUPPORTED = 4, /* e.g. tree deeper than the reference's 64-entry stack */
    B200RT_ENOMEM = 5
};

/* Accelerator kinds = the `Accelerator "<name>"` strings of MakeAccelerator (src/core/api.cpp:790-1015). */
typedef enum {
    B200RT_BVH = 0,    /* "bvh"    src/accelerators/bvh.cpp          (32-byte LinearBVHNode) */
    B200RT_KDTREE = 1, /* "kdtree" src/accelerators/kdtreeaccel.cpp  (8-byte KdAccelNode)    */
    B200RT_RBSP = 2,   /* "rbsp"   src/accelerators/rbsp.cpp         (8-byte RBSPNode)       */
    B200RT_RBSPKD = 3, /* "rbspkd" src/accelerators/rbspKd.cpp       (RBSP + axis<3 kd path) */
    /* General BSP families: every node carries its own split normal (20-byte nodes).  One traversal serves all
     * of the thesis's builders of that node format: "bsppaper", "bspcluster", "bsprandom", "bsparbitrary",
     * "bsp*withkd" (api.cpp:847-1010) build BSPNode trees; "bsppaperkd" and the "*fastkd" ones BSPKdNode trees. */
    B200RT_BSP = 4,    /* src/accelerators/BSP.cpp:27-163   over BSPNode   (BSP.h:122-184)   */
    B200RT_BSPKD = 5   /* src/accelerators/BSPKd.cpp:25-169 over BSPKdNode (BSPKd.h:11-174)  */
} b200rt_kind;

/* Ray = the fields of pbrt::Ray the hot path reads (src/core/geometry.h:1176-1203): 32 bytes. */
typedef struct {
    float o[3];
    float tMax; /* Ray::tMax (mutable in the reference; the shrunk value is returned in b200rt_hit::t) */
    float d[3];
    float time; /* carried, unused by triangles */
} b200rt_ray;

/* Closest-hit record, 20 bytes.  Replaces the (bool, ray.tMax, SurfaceInteraction) triple of
 * Aggregate::Intersect (src/core/primitive.h:49-56); b0,b1,b2 are Triangle::Intersect's barycentrics
 * (src/shapes/triangle.cpp:266-270) from which the shading stage rebuilds the SurfaceInteraction. */
#define B200RT_MISS 0xFFFFFFFFu
typedef struct {
    uint32_t prim; /* ORIGINAL primitive number (index into the prims vector handed to the
                      accelerator, i.e. primNumMapping[slot] for the BVH), B200RT_MISS if none */
    float t;       /* ray.tMax after the call (unchanged input tMax on a miss) */
    float b0, b1, b2;
} b200rt_hit;

/* Scene description: the host-built flattened tree in the REFERENCE'S OWN node format plus
 * world-space triangles in original primitive order.  The library re-lays both out for HBM. */
typedef struct {
    uint32_t kind;         /* b200rt_kind */
    const void *nodes;     /* BVH: node_count x 32 B (bvh.cpp:123-152); kd/RBSP(/Kd): node_count x 8 B
                              (kdtreeaccel.cpp:43-160, rbsp.cpp:15-135); BSP/BSPKd: node_count x 20 B
                              (split|prim|offset, flags, splitAxis.xyz — BSP.h:170-183, BSPKd.h:161-173) */
    uint32_t node_count;
    const uint32_t *prim_indices; /* BVH: primNumMapping[slot] -> original prim (bvh.cpp:227,322),
                                     n_indices == n_prims.  kd/RBSP: primitiveIndices
                                     (genericBSP.h:187). */
    uint32_t n_indices;
    const float *directions; /* RBSP/RBSPKd: M x 3 split normals (RBSPShared.h:29-74); else NULL */
    uint32_t n_directions;   /* M (3 for kd) */
    float bounds[6];         /* kd/RBSP root bounds pMin.xyz pMax.xyz (genericBSP.h:190); BVH: ignored */
    const float *tri_verts;  /* n_prims x 9 floats: world-space p0,p1,p2 (TriangleMesh::p, triangle.h:50-69) */
    const float *tri_uvs;    /* n_prims x 6 floats or NULL (default (0,0),(1,0),(1,1), triangle.h:109-121): the
                                degenerate-triangle rejection of triangle.cpp:302-315 and the shade stage's dpdu frame */
    uint32_t n_prims;
    /* Non-triangle primitives in the aggregate (Sphere::Intersect / IntersectP, src/shapes/sphere.cpp:49-156,
     * 158-215; killeroo-simple's area light and scenes/simple put spheres next to the triangles).
     * prim_kind: NULL (all triangles) or n_prims bytes, 0 = triangle, 1 = FULL sphere (zmin/zmax/phimax at
     * their defaults; partial spheres are rejected with B200RT_EUNSUPPORTED by the caller's shim).  The k-th
     * sphere primitive (in primitive order) uses sphere_params[16*k .. 16*k+15] = WorldToObject rows 0..2 of
     * the shape's transform (affine 3x4, row-major), radius, 3 unused floats.  tri_verts rows of sphere
     * primitives are ignored.  A sphere hit reports b0 = b1 = b2 = 0. */
    const uint8_t *prim_kind;
    const float *sphere_params;
    uint32_t n_spheres;
    /* ---- appended in ABI v8: what Triangle::Intersect's tail (triangle.cpp:294-425) reads besides the vertices; used by the
     * wavefront shade stage only (the traversal never looks at them) ---- */
    const float *tri_normals;   /* n_prims x 9 floats: TriangleMesh::n at the triangle's three vertices (world space, as the mesh
                                   stores them), or NULL.  Rows of primitives without normals are ignored (prim_flags bit 1). */
    const uint8_t *prim_flags;  /* n_prims bytes or NULL (= all 0): bit 0 = Shape::reverseOrientation ^ transformSwapsHandedness
                                   (triangle.cpp:421-424, interaction.cpp:62-66), bit 1 = the primitive's mesh has vertex normals */
} b200rt_scene_desc;

typedef struct b200rt_scene b200rt_scene;

/* Per-launch traversal counters (optional instrumented kernels; = the reference's per-ray
 * GeneralStats, src/core/geometry.h:1078-1173, summed over the batch). */
typedef struct {
    uint64_t rays;
    uint64_t nodes_fetched;  /* BVH: nodes whose box was tested; kd/RBSP: loop iterations (kdtreeaccel.cpp:401) */
    uint64_t nodes_passed;   /* BVH: boxes passed (bvh.cpp:368); kd/RBSP: interior nodes */
    uint64_t leaf_visits;
    uint64_t prim_tests;     /* GeometricPrimitive::Intersect(P) calls (primitive.cpp:118-126) */
    uint64_t index_reads;    /* primitiveIndices reads in multi-prim kd/RBSP leaves */
} b200rt_counters;

/* ---- device / lifetime -------------------------------------------------------------------- */
int b200rt_init(int device);              /* cudaSetDevice + context warm-up; replaces nothing (pbrtInit has no device) */
const char *b200rt_last_error(void);      /* thread-local text of the last failure */
int b200rt_version(void);

/* MakeAccelerator(name, prims, paramSet) (src/core/api.cpp:790) AFTER the host builder ran:
 * copies the tree + triangles to HBM in the device layout. */
int b200rt_scene_create(const b200rt_scene_desc *desc, b200rt_scene **out);
void b200rt_scene_destroy(b200rt_scene *scene);
/* Aggregate::WorldBound() (src/core/primitive.h:53; bvh.cpp:187, genericBSP.h:97). */
int b200rt_scene_bounds(const b200rt_scene *scene, float out_bounds[6]);

/* Aggregate::Intersect over a batch (bvh.cpp:354, kdtreeaccel.cpp:381, rbsp.cpp:405, rbspKd.cpp).
 * rays/hits may each be host or device pointers (detected per pointer); with host pointers the
 * copies are part of the call.  stream: a cudaStream_t or NULL for the default stream.
 * The call is synchronous w.r.t. host buffers, asynchronous for device buffers. */
int b200rt_intersect_closest(b200rt_scene *scene, const b200rt_ray *rays, uint64_t n, b200rt_hit *hits,
                             void *stream);
/* Aggregate::IntersectP over a batch (bvh.cpp:398, kdtreeaccel.cpp:454, rbsp.cpp:479): occluded[i] = 0/1. */
int b200rt_intersect_any(b200rt_scene *scene, const b200rt_ray *rays, uint64_t n, uint8_t *occluded,
                         void *stream);

/* The two calls above split into enqueue + wait, for HOST rays and HOST results only.  The reference has no counterpart
 * (Aggregate::Intersect returns its answer); a host-side wavefront integrator has, after shading one bounce, two independent
 * queues in hand — that bounce's shadow rays (IntersectP) and the next bounce's rays (Intersect) — and these calls let it keep
 * both in flight: every call appends its chunks to the scene's copy / kernel / copy pipeline (three streams, a ring of staging
 * slots) and returns; the pipeline does not drain between calls.  `rays` is read from the moment of the call on (work pending on
 * `stream` is NOT waited for); `rays` and the result buffer must stay alive and untouched
 * until b200rt_intersect_wait(scene) returns (or until `stream`, which is made to wait for the results, has been synchronised).
 * Results are those of the synchronous calls, bit for bit (same kernels, same chunks). */
int b200rt_intersect_closest_async(b200rt_scene *scene, const b200rt_ray *rays, uint64_t n, b200rt_hit *hits, void *stream);
int b200rt_intersect_any_async(b200rt_scene *scene, const b200rt_ray *rays, uint64_t n, uint8_t *occluded, void *stream);
int b200rt_intersect_wait(b200rt_scene *scene);
/* Instrumented variants (slower; counters accumulate into *counters, a HOST struct). */
int b200rt_intersect_closest_counted(b200rt_scene *scene, const b200rt_ray *rays, uint64_t n,
                                     b200rt_hit *hits, b200rt_counters *counters, void *stream);
int b200rt_intersect_any_counted(b200rt_scene *scene, const b200rt_ray *rays, uint64_t n,
                                 uint8_t *occluded, b200rt_counters *counters, void *stream);
/* Per-RAY traversal statistics = the reference's Ray::stats after the call (GeneralStats, src/core/geometry.h:1078-1173;
 * incremented at bvh.cpp:370,379 / kdtreeaccel.cpp:404,414 / rbsp.cpp / BSP.cpp:49,79 / primitive.cpp:118-126).  The
 * thesis sums them per pixel into its traversal heat maps (integrator.cpp:328, film.cpp:170-264). */
typedef struct {
    uint32_t prim_tests;      /* primitiveIntersections (closest) / primitiveIntersectionsP (any) */
    uint32_t interior_visits; /* bvh-/kd-/bspTreeNodeTraversals(P): interior nodes processed */
    uint32_t leaf_visits;     /* leafNodeTraversals(P) */
    uint32_t nodes_fetched;   /* node records read (not counted by the reference; feeds the roofline's N_fetch) */
} b200rt_ray_stats;
/* Instrumented Intersect / IntersectP that also return per_ray[i] (HOST or DEVICE buffer of n records). */
int b200rt_intersect_closest_stats(b200rt_scene *scene, const b200rt_ray *rays, uint64_t n, b200rt_hit *hits,
                                   b200rt_ray_stats *per_ray, void *stream);
int b200rt_intersect_any_stats(b200rt_scene *scene, const b200rt_ray *rays, uint64_t n, uint8_t *occluded,
                               b200rt_ray_stats *per_ray, void *stream);
/* b200rt_intersect_closest for callers that cannot allocate.  NO COUNTERPART IN THE REFERENCE: added
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  rays / hits are
 * HOST buffers of n records; the device-side staging the ordinary entry point cudaMalloc's on first use lives here in a
 * file-scope static __device__ array (csrc/no_alloc.cu; 3.25 MiB, held by every GPU that loads the library until exit), so the
 * call makes no cudaMalloc / cudaFree / allocating library call.  n > B200RT_NOALLOC_MAX_RAYS is rejected with B200RT_EINVAL.
 * Same kernels, same results (bit for bit) as b200rt_intersect_closest, every tree kind.  The scene's device must be current.
 * The array is shared by all calls and keeps the last call's contents (this call's hit records are cleared before the launch);
 * calls that overlap on the GPU would race on it: the entry point serialises its callers on a mutex and returns only after
 * `stream` has drained. */
#define B200RT_NOALLOC_MAX_RAYS 65536u
int b200rt_intersect_closest_noalloc(b200rt_scene *scene, const b200rt_ray *rays, uint64_t n, b200rt_hit *hits, void *stream);
/* Number of kernel launches issued by this library since b200rt_init (bench.py's gpu_launches). */
uint64_t b200rt_launch_count(void);

/* ---- host builders (C++; node-for-node restatement of the reference's builders) ------------ */
/* BVHAccel::BVHAccel / iterativeBuild / flattenBVHTree (bvh.cpp:155-352) on triangle bounds.
 * Params = CreateBVHAccelerator's (bvh.cpp:529-535): maxnodeprims 4, intersectcost 8, traversalcost 1.
 * Outputs malloc'ed arrays the caller frees with b200rt_free. */
int b200rt_build_bvh(const float *tri_verts, uint32_t n_prims, int max_node_prims, int isect_cost,
                     int traversal_cost, void **out_nodes, uint32_t *out_node_count,
                     uint32_t **out_prim_order);
/* KdTreeAccel::buildTree (kdtreeaccel.cpp:208-379); params of CreateKdTreeAccelerator (:523-545):
 * intersectcost 80, traversalcost 1, emptybonus 0, maxprims 1, maxdepth -1. */
int b200rt_build_kdtree(const float *tri_verts, uint32_t n_prims, int isect_cost, int traversal_cost,
                        float empty_bonus, int max_prims, int max_depth, void **out_nodes,
                        uint32_t *out_node_count, uint32_t **out_prim_indices, uint32_t *out_n_indices,
                        float out_bounds[6]);
/* ---- device builder -------------------------------------------------------------------------- */
/* RBSP::buildTree (rbsp.cpp:181-403) with kDOPMesh.h's cut / surface-area arithmetic, level-synchronous on the B200: every
 * candidate plane of every node of a tree level is cut and measured in parallel (the reference: 38-319 s per tree on the CPU),
 * then the nodes are numbered in the reference's depth-first order and emitted as 8-byte RBSPNodes (rbsp.cpp:15-135).
 * Params = CreateRBSPTreeAccelerator's (rbsp.cpp:549-570): intersectcost 80, traversalcost 5, emptybonus 0, maxprims 1,
 * maxdepth -1 (= round(2 + 1.6 log2 n), geometry.h:1845), nbDirections 3 | 7 | 9 | 13 (RBSPShared.h:50-71).
 * The node array equals the reference's byte for byte; the order of the primitive indices inside a leaf (and, in the rare
 * leaf where the reference's choice hangs on std::sort's order among equal keys, its members) may differ: csrc/rbsp_build.cuh.
 * Triangles only.  out_directions: nb_directions x 3 floats.  Outputs are malloc'ed; free with b200rt_free.  stats may be NULL.
 * No host path: B200RT_ENODEVICE without a CUDA device. */
typedef struct b200rt_build_stats {
    uint32_t levels;            /* tree levels processed (<= maxdepth + 1) */
    uint64_t candidates;        /* bound edges examined: 2 n M summed over the levels (statNbSplitTests' upper bound) */
    uint64_t kernel_launches;
    float build_ms;             /* CUDA events around upload, build and download */
} b200rt_build_stats;
int b200rt_build_rbsp(const float *tri_verts, uint32_t n_prims, uint32_t nb_directions, int isect_cost, int traversal_cost,
                      float empty_bonus, int max_prims, int max_depth, void **out_nodes, uint32_t *out_node_count,
                      uint32_t **out_prim_indices, uint32_t *out_n_indices, float out_bounds[6], float *out_directions,
                      b200rt_build_stats *stats);
/* RBSPKd::buildTree (rbspKd.cpp:187-470; CreateRBSPKdTreeAccelerator :640-665 adds "kdtraversalcost", default 1) through the
 * same builder: axis-aligned planes are charged kd_traversal_cost, the others 0.1 * isect_cost * (n - 1) + kd_traversal_cost. */
int b200rt_build_rbspkd(const float *tri_verts, uint32_t n_prims, uint32_t nb_directions, int isect_cost, int traversal_cost,
                        int kd_traversal_cost, float empty_bonus, int max_prims, int max_depth, void **out_nodes,
                        uint32_t *out_node_count, uint32_t **out_prim_indices, uint32_t *out_n_indices, float out_bounds[6],
                        float *out_directions, b200rt_build_stats *stats);
/* KdTreeAccel::buildTree (kdtreeaccel.cpp:208-379) through the same level-synchronous device builder (box areas in closed
 * form instead of k-DOP cuts).  Same outputs and parameters as b200rt_build_kdtree; the node array is identical to it (and to
 * the reference's), the order of the indices inside a leaf follows the stable radix sort instead of std::sort's. */
int b200rt_build_kdtree_device(const float *tri_verts, uint32_t n_prims, int isect_cost, int traversal_cost,
                               float empty_bonus, int max_prims, int max_depth, void **out_nodes, uint32_t *out_node_count,
                               uint32_t **out_prim_indices, uint32_t *out_n_indices, float out_bounds[6],
                               b200rt_build_stats *stats);
/* BVHAccel's sweep-SAH build (bvh.cpp:155-352) on the device: every axis sorted once for the whole scene, then per tree level
 * two segmented box scans per axis, one thread per candidate position, stable partitions that keep the runs sorted, and
 * primitiveInfo permuted exactly as std::partition leaves it (csrc/bvh_build.cuh).  Outputs and parameters as
 * b200rt_build_bvh; nodes AND primNumMapping equal its output (and the reference's) byte for byte. */
int b200rt_build_bvh_device(const float *tri_verts, uint32_t n_prims, int max_node_prims, int isect_cost, int traversal_cost,
                            void **out_nodes, uint32_t *out_node_count, uint32_t **out_prim_order, b200rt_build_stats *stats);
void b200rt_free(void *p);

/* ---- wavefront driver (SamplerIntegrator::Render restructured, src/core/integrator.cpp:230-360
 *      + PathIntegrator::Li, src/integrators/path.cpp:64-204) ---------------------------------
 * See b200rt_wavefront.h. */

#ifdef __cplusplus
}
#endif
#endif /* B200RT_H */
