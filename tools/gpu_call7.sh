#!/bin/bash
mkdir -p gpurun_out
SWEEP_ALGOS=mma timeout 600 python tools/size_sweep.py 4096 8192 16384 32768 > gpurun_out/size_sweep_fp4.jsonl 2> gpurun_out/size_sweep.err
C1="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e"
export M4RI_B200_STRASSEN=0
timeout 300 $C1 > gpurun_out/plain_fp4.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:mma4_kernel -s 1 -c 1 -o gpurun_out/prof_mma4_n32768 $C1 > gpurun_out/ncu_fp4.log 2>&1
timeout 300 $C1 --n 8192 > gpurun_out/plain_fp4_8192.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:mma4_kernel -s 1 -c 1 -o gpurun_out/prof_mma4_n8192 $C1 --n 8192 > gpurun_out/ncu_fp4_8192.log 2>&1
cat gpurun_out/size_sweep_fp4.jsonl; cut -c1-400 gpurun_out/plain_fp4.log
