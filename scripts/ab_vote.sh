#!/bin/bash
# A/B of the persistent kernels' vote weights (rt_persist.cuh pack_steps): interior round when nI * wI >= nL * wL.
# kd-tree on the museum / conference stand-ins (config 3), BVH on the headline frame.  Results are bit-identical by
# construction (schedules only); every line carries the film mean so that this can be seen.
#   bash scripts/ab_vote.sh > gpurun_out/ab_vote.log 2>&1
cd "$(dirname "$0")/.."
for v in 1,1 2,1 3,1 4,1 6,1 8,1 1,2; do
  echo "== B200RT_BSP_VOTE=$v museum (kd)"
  B200RT_BSP_VOTE=$v python scripts/config3_museum.py --frames 3 | cut -c1-700
done
for v in 1,1 3,1 6,1; do
  echo "== B200RT_BSP_VOTE=$v conference (kd)"
  B200RT_BSP_VOTE=$v python scripts/config3_museum.py --frames 2 --scene conference | cut -c1-700
done
for v in 1,1 2,1 3,1 4,1 1,2; do
  echo "== B200RT_BVH_VOTE=$v headline (bvh)"
  B200RT_BVH_VOTE=$v python bench.py --steps 5 --warmup 3 --no-cpu-baseline | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
r = d['roofline']
print(json.dumps({'ms_per_step': d['ms_per_step'], 'value': d['value'], 'closest_kernel_mrays_s': r.get('closest_kernel_mrays_s'), 'any_kernel_mrays_s': r.get('any_kernel_mrays_s')}))"
done
