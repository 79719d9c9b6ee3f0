// b200accel.h — the pbrt-side half of the drop-in: `Accelerator "b200"` for Thesis-pbrt-v3.
//
// This file is written against the REFERENCE'S OWN headers (it is compiled with -I<reference>/src and
// linked into pbrt next to the other accelerators; nothing under /root/reference is edited).  It mirrors
// the reference's accelerator interface for the hot path:
//
//   std::shared_ptr<Primitive> MakeAccelerator(name, prims, paramSet)     src/core/api.cpp:790-1015
//   class Aggregate : public Primitive                                     src/core/primitive.h:188-198
//     Bounds3f WorldBound() const                                          src/core/primitive.h:53
//     bool Intersect(const Ray &, SurfaceInteraction *) const              src/core/primitive.h:54
//     bool IntersectP(const Ray &) const                                   src/core/primitive.h:55
//
// and forwards the work to libb200rt.so through the C ABI of include/b200rt.h.  The tree is built by the
// reference's own builder selected with "string tree" (CreateBVHAccelerator bvh.cpp:529, CreateKdTreeAccelerator
// kdtreeaccel.cpp:523, CreateRBSPTreeAccelerator rbsp.cpp:549, CreateRBSPKdTreeAccelerator rbspKd.cpp) with the
// SAME ParamSet names, so `Accelerator "b200" "string tree" "rbsp" "integer nbDirections" [7]` reads like the
// reference's own scene lines; "string builder" "b200" switches bvh/kdtree to this repo's node-for-node
// restatements of those builders (b200rt_build_bvh / b200rt_build_kdtree).
//
// There is no CPU fallback inside B200Accel: if the device or the library is unavailable,
// CreateB200Accelerator reports through pbrt's Error() and returns nullptr, which is exactly what
// MakeAccelerator does for an unknown name (api.cpp:1011-1014) — RenderOptions::MakeScene then falls back to
// the reference's BVHAccel (api.cpp:1886).
#ifndef B200RT_HOST_B200ACCEL_H
#define B200RT_HOST_B200ACCEL_H

#include <memory>
#include <vector>

#include "pbrt.h"
#include "primitive.h"
#include "b200rt.h"

namespace pbrt {

class B200Accel : public Aggregate {
  public:
    // Construction = geometry capture + reference builder + upload (b200rt_scene_create).  ok() is false when
    // any step failed; the message went through Error().
    B200Accel(std::vector<std::shared_ptr<Primitive>> prims, const ParamSet &ps);
    ~B200Accel();
    bool ok() const { return scene != nullptr; }

    // ---- the reference's per-ray interface (batch of one; kept for API completeness and stray callers) ----
    Bounds3f WorldBound() const;
    // Sets ray.tMax, fills the full SurfaceInteraction (by re-running the ONE winning primitive's own
    // Intersect on the host, triangle.cpp:294-425 / sphere.cpp:100-156) and isect->primitive.
    bool Intersect(const Ray &ray, SurfaceInteraction *isect) const;
    bool IntersectP(const Ray &ray) const;

    // ---- the batched interface the wavefront integrator uses ----
    // hits[i].prim indexes the prims vector handed to the constructor (B200RT_MISS = none); rays[i].tMax is
    // updated like the per-ray call does.  occluded[i] = IntersectP(rays[i]).
    bool IntersectBatch(const Ray *rays, size_t n, b200rt_hit *hits) const;
    bool IntersectPBatch(const Ray *rays, size_t n, uint8_t *occluded) const;
    // Finishes a batched hit into a SurfaceInteraction (same host re-intersection as Intersect()).
    bool FillInteraction(const Ray &ray, const b200rt_hit &hit, SurfaceInteraction *isect) const;

    const std::shared_ptr<Primitive> &primitive(uint32_t i) const { return prims[i]; }
    size_t primitiveCount() const { return prims.size(); }
    b200rt_scene *deviceScene() const { return scene; }
    int treeKind() const { return kind; }   // B200RT_BVH ... B200RT_BSPKD of the uploaded tree
    // The process-global the `b200wavefront` integrator looks up (Scene::aggregate is private, scene.h:78).
    static const B200Accel *current();

  private:
    std::vector<std::shared_ptr<Primitive>> prims;
    b200rt_scene *scene = nullptr;
    int kind = B200RT_BVH;
    Bounds3f bounds;
};

// The branch a maintainer adds to MakeAccelerator (api.cpp, before the `else Warning(...)` at :1011):
//     else if (name == "b200") accel = CreateB200Accelerator(std::move(prims), paramSet);
std::shared_ptr<Primitive> CreateB200Accelerator(std::vector<std::shared_ptr<Primitive>> prims,
                                                 const ParamSet &ps);

}  // namespace pbrt
#endif
