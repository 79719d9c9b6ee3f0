#!/bin/bash
# Wavefront batch size (paths in flight per batch, B200RT_WF_BATCH; default 2^22): headline frame and living-room as shipped.
cd "$(dirname "$0")/.."
for b in 4194304 8388608 16777216 33554432; do
  echo "== B200RT_WF_BATCH=$b"
  B200RT_WF_BATCH=$b python bench.py --steps 5 --warmup 3 --no-cpu-baseline | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(json.dumps({'headline_ms_per_step': d['ms_per_step'], 'value': d['value']}))"
  B200RT_WF_BATCH=$b python scripts/living_room.py --xres 1280 --yres 720 --spp 64 | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(json.dumps({'living_room_ms_per_frame': d['ms_per_frame'], 'mrays_per_s': d['mrays_per_s'], 'film_mean_rgb': d['film_mean_rgb']}))"
done
