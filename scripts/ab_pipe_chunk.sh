#!/bin/bash
# e2e (host rays in, host hits out through the C ABI) against the pipeline's stage size and its number of kernel streams.
#   bash scripts/ab_pipe_chunk.sh > gpurun_out/ab_pipe_chunk.log 2>&1
cd "$(dirname "$0")/.."
for k in 1 2 3; do
for c in 131072 262144 524288; do
  B200RT_PIPE_KSTREAMS=$k B200RT_PIPE_CHUNK=$c python bench.py --steps 3 --warmup 3 --no-cpu-baseline | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); e = d['e2e']
print(json.dumps({'B200RT_PIPE_KSTREAMS': $k, 'B200RT_PIPE_CHUNK': $c, 'rays': e['rays_per_step'], 'e2e_mrays_s': e['value'], 'one_blocking_call_per_queue_mrays_s': e['one_blocking_call_per_queue_mrays_s'], 'copies_only_mrays_s': e['copies_only_mrays_s'], 'h2d_gbs': e['rank0_link']['h2d_gbs']}))"
done
done
