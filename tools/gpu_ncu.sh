#!/bin/bash
# ncu full capture of the product kernel at n (default 32768): bash tools/gpu_ncu.sh <kernel-regex> <tag> [n]
mkdir -p gpurun_out
K=${1:-mma5_kernel}; TAG=${2:-mma5}; N=${3:-32768}
export M4RI_B200_STRASSEN=0
C1="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --n $N"
timeout 300 $C1 > gpurun_out/plain_$TAG.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 1 -c 1 -o gpurun_out/prof_${TAG}_n$N $C1 > gpurun_out/ncu_$TAG.log 2>&1
cut -c1-300 gpurun_out/plain_$TAG.log; tail -2 gpurun_out/ncu_$TAG.log
