#!/bin/bash
# Burst vs sustained step time under forced Strassen levels (M4RI_B200_STRASSEN): the board's power
# management takes a few percent after ~50 ms of tensor load, and fewer MMAs per product draw less.
#   usage: bash tools/sustained_levels.sh
run() {  # n steps levels
  M4RI_B200_STRASSEN=$3 python bench.py --n $1 --steps $2 --warmup 3 --no-cpu-baseline --no-e2e --no-extras 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); s=d['step_ms_rank0']; k=max(1,len(s)//2)
print(json.dumps({'n': $1, 'levels': $3, 'steps': len(s), 'mean_ms': round(d['ms_per_step'],4), 'first5_ms': round(sum(s[:5])/5,4), 'second_half_ms': round(sum(s[-k:])/k,4)}))"
}
for lv in 1 2 3; do run 32768 30 $lv; done
for lv in 0 1 2; do run 16384 200 $lv; done
for lv in 1 2 3; do run 65536 6 $lv; done
