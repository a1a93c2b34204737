#!/bin/bash
# The n = 32768 step with the last Strassen level (a) fused and grouped into one launch (default),
# (b) fused, seven launches, (c) unfused Winograd schedule.
for mode in grouped seven unfused; do
  unset M4RI_B200_NO_FUSED_STRASSEN M4RI_B200_NO_GROUPED_STRASSEN
  [ $mode = seven ] && export M4RI_B200_NO_GROUPED_STRASSEN=1
  [ $mode = unfused ] && export M4RI_B200_NO_FUSED_STRASSEN=1
  python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('$mode ms/step', round(d['ms_per_step'],3), 'kernel_ms', round(r['kernel_ms'],4), 'x', r['kernel_launches_per_step'], 'share', round(r['kernel_share_of_step'],4), 'frac', round(r['frac'],4), 'launches', d['gpu_launches'], d['step_ms_rank0'])"
done
