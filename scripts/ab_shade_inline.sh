#!/bin/bash
# Same-box A/B of the shade stage with the BSDF evaluated inline (the shipped build) or out of line
# (-DB200RT_BSDF_NOINLINE -> gpurun_out/libb200rt_noinline.so, built here on the CPU box):
#   build:  scripts/ab_shade_inline.sh build      measure (on a GPU box):  scripts/ab_shade_inline.sh run
set -e
cd "$(dirname "$0")/.."
P=thesis-pbrt-v3_b200
if [ "$1" = build ]; then
  mkdir -p gpurun_out
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC,-ffp-contract=off,-fno-fast-math \
       -cudart static -DB200RT_BSDF_NOINLINE -c -o gpurun_out/wavefront_noinline.o $P/csrc/wavefront.cu
  nvcc -gencode arch=compute_100a,code=sm_100a --shared -cudart static -o $P/b200rt/libb200rt_noinline.so \
       $P/_obj/b200rt.cu.o gpurun_out/wavefront_noinline.o $P/_obj/build_bvh.cpp.o $P/_obj/build_kd.cpp.o
else
  for i in 1 2 3; do
    for v in inline noinline; do
      L=""; [ $v = noinline ] && L=$PWD/$P/b200rt/libb200rt_noinline.so
      B200RT_LIB=$L python scripts/living_room.py --xres 1280 --yres 720 --spp 64 --frames 2 2>/dev/null |
        python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$v', round(d['ms_per_frame'],1), 'ms/frame, non-traversal', round(d['ms_per_frame']*(1-d['rank0_trace_share_of_frame']),1), 'ms')"
    done
  done
fi
