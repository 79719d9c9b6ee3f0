#!/bin/bash
# Where the e2e pipeline's time goes: the same chunked copies and event chain WITHOUT the kernels (B200RT_PIPE_NOKERNEL=1).
cd "$(dirname "$0")/.."
for nk in 0 1; do
for c in 262144 1048576; do
  B200RT_PIPE_NOKERNEL=$nk B200RT_PIPE_CHUNK=$c python bench.py --steps 3 --warmup 3 --no-cpu-baseline | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); e = d['e2e']
print(json.dumps({'B200RT_PIPE_NOKERNEL': $nk, 'B200RT_PIPE_CHUNK': $c, 'e2e_mrays_s': e['value'], 'one_blocking_call_per_queue_mrays_s': e['one_blocking_call_per_queue_mrays_s'], 'copies_only_mrays_s': e['copies_only_mrays_s'], 'h2d_gbs': e['rank0_link']['h2d_gbs']}))"
done
done
