#!/bin/bash
# Experiment: the traversal kernels read pinned host rays over PCIe themselves (B200RT_PIPE_ZEROCOPY=1) instead of an H2D stage.
cd "$(dirname "$0")/.."
for z in 0 1; do
for c in 262144 1048576; do
  B200RT_PIPE_ZEROCOPY=$z B200RT_PIPE_CHUNK=$c python bench.py --steps 3 --warmup 3 --no-cpu-baseline | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); e = d['e2e']
print(json.dumps({'B200RT_PIPE_ZEROCOPY': $z, 'B200RT_PIPE_CHUNK': $c, 'e2e_mrays_s': e['value'], 'one_blocking_call_per_queue_mrays_s': e['one_blocking_call_per_queue_mrays_s'], 'copies_only_mrays_s': e['copies_only_mrays_s']}))"
done
done
