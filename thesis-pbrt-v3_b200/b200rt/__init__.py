"""b200rt — host-side mirror of the reference's Accelerator interface for the ray-intersection hot
path, over the C ABI of include/b200rt.h (libb200rt.so, hand-written CUDA for sm_100a).

The reference's interface for this path (paths relative to the reference tree):
  MakeAccelerator(name, prims, paramSet)              src/core/api.cpp:790-1015
  Aggregate::WorldBound / Intersect / IntersectP      src/core/primitive.h:49-56
  CreateBVHAccelerator / CreateKdTreeAccelerator ...  src/accelerators/*.cpp (ParamSet names + defaults)

There is no CPU fallback: importing works anywhere (the library loads without a GPU so symbol
checks can run), but every compute call raises B200RTError without a CUDA device."""
from .api import (B200RTError, B200Accel, MakeAccelerator, RAY_DTYPE, HIT_DTYPE, MISS, lib, lib_path,
                  build_bvh, build_kdtree, build_rbsp, build_kdtree_device, build_bvh_device, make_rays, init, launch_count, KINDS)

from .wavefront import WavefrontPathIntegrator, film_to_rgb  # noqa: E402

__all__ = ["WavefrontPathIntegrator", "film_to_rgb", "B200RTError", "B200Accel", "MakeAccelerator", "RAY_DTYPE", "HIT_DTYPE", "MISS", "lib", "lib_path",
           "build_bvh", "build_kdtree", "build_rbsp", "build_kdtree_device", "build_bvh_device", "make_rays", "init", "launch_count", "KINDS"]
