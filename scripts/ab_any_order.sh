#!/bin/bash
# Child order of the BVH any-hit kernel (IntersectP's answer does not depend on it): headline frame.
# NOTE: the B200RT_BVH_ANY_ORDER switch this script drives was an experiment and is NOT in the tree any more (it cost the default
# order 6 % just by being compiled in, and no order won: profiles/r02_ab_any_order.log); kept as the record of how
# the log was made.
cd "$(dirname "$0")/.."
for m in 0 1 2 3; do
  B200RT_BVH_ANY_ORDER=$m python bench.py --steps 5 --warmup 3 --no-cpu-baseline | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); r = d['roofline']
print(json.dumps({'B200RT_BVH_ANY_ORDER': $m, 'ms_per_step': d['ms_per_step'], 'value': d['value'], 'closest_kernel_mrays_s': r.get('closest_kernel_mrays_s'), 'any_kernel_mrays_s': r.get('any_kernel_mrays_s')}))"
done
python -m pytest tests/test_gpu_parity.py tests/test_gpu_wavefront.py -x -q 2>&1 | tail -2
for m in 1 2; do B200RT_BVH_ANY_ORDER=$m python -m pytest tests/test_gpu_parity.py tests/test_gpu_halton.py -x -q 2>&1 | tail -1; done
