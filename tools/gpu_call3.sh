#!/bin/bash
mkdir -p gpurun_out
C="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --algo mma --n 16384"
$C > gpurun_out/plain_v2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:mma2_kernel -s 1 -c 1 -o gpurun_out/prof_mma2 $C > gpurun_out/ncu_v2.log 2>&1
python bench.py --steps 5 --warmup 3 --algo mma --no-cpu-baseline > gpurun_out/bench_mma2.json 2> gpurun_out/bench_mma2.err
tail -3 gpurun_out/ncu_v2.log; cat gpurun_out/bench_mma2.json
