#!/bin/bash
# k_shade<true> (scenes/living-room as shipped): block size of its launches (B200RT_WF_TB_FULL) and, second loop, resident blocks
# per SM through __launch_bounds__ (B200RT_WF_SHADE_MINB — that switch was an experiment and is not in the tree any more, see the
# comment above k_shade in csrc/wavefront.cu; the loop is kept as the record of how profiles/r02_ab_shade_full.log was made).
cd "$(dirname "$0")/.."
for tb in 256 128 64; do
  B200RT_WF_TB_FULL=$tb python scripts/living_room.py --xres 1280 --yres 720 --spp 64 | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(json.dumps({'B200RT_WF_TB_FULL': $tb, 'ms_per_frame': d['ms_per_frame'], 'mrays_per_s': d['mrays_per_s'], 'trace_share': d['rank0_trace_share_of_frame'], 'film_mean_rgb': d['film_mean_rgb']}))"
done
for mb in 2 3 4; do
  B200RT_WF_SHADE_MINB=$mb python scripts/living_room.py --xres 1280 --yres 720 --spp 64 | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(json.dumps({'B200RT_WF_SHADE_MINB': $mb, 'ms_per_frame': d['ms_per_frame'], 'mrays_per_s': d['mrays_per_s'], 'trace_share': d['rank0_trace_share_of_frame'], 'film_mean_rgb': d['film_mean_rgb']}))"
done
