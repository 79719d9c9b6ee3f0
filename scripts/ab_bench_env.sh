#!/bin/bash
# Same-box A/B of whole-frame bench numbers under different B200RT_* environment settings (alternating runs).
#   scripts/ab_bench_env.sh "B200RT_TOPLEVELS=0" "B200RT_TOPLEVELS=5" ...
for rep in 1 2; do
  for cfg in "$@"; do
    env $cfg python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
l = json.loads(sys.stdin.readline()); r = l['roofline']
print('%-40s frame %.2f ms  value %.0f  closest %.0f  any %.0f Mrays/s  kernel share %.3f' % ('$cfg', l['ms_per_step'], l['value'], r['closest_kernel_mrays_s'], r['any_kernel_mrays_s'], r['kernel_share_of_step']))"
  done
done
