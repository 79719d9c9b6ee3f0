#!/bin/bash
# Same-box A/B of the kd persistent kernel's below-child prefetch (B200RT_KD_PREFETCH): parity first, then config 3 alternating.
B200RT_KD_PREFETCH=1 python -m pytest tests/test_gpu_parity.py tests/test_living_room.py -m gpu -q -k "kdtree or small or edge or degenerate" 2>&1 | tail -2
for rep in 1 2; do
  for pf in 0 1; do
    for scene in museum conference; do
      B200RT_KD_PREFETCH=$pf python scripts/config3_museum.py --scene $scene --frames 3 2>/dev/null | head -1 | python -c "
import sys, json
l = json.loads(sys.stdin.readline())
print('prefetch=$pf %-10s frame %.2f ms  closest %.0f  any %.0f Mrays/s  film %.7f' % ('$scene', l['ms_per_frame'], l['closest_kernel_mrays_per_s'], l['any_kernel_mrays_per_s'], l['film_mean']))"
    done
  done
done
