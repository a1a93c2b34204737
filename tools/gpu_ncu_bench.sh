#!/bin/bash
# ncu full capture of the dominant kernel inside the default bench step (the grouped launch of the
# seven Strassen leaves): bash tools/gpu_ncu_bench.sh <tag> [bench args]
TAG=${1:-bench}; shift
mkdir -p gpurun_out
C1="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline $@"
timeout 300 $C1 > gpurun_out/plain_$TAG.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mma6_kernel -s 1 -c 1 -f -o gpurun_out/prof_$TAG $C1 > gpurun_out/ncu_$TAG.log 2>&1
tail -2 gpurun_out/ncu_$TAG.log
python tools/ncu_summary.py gpurun_out/prof_$TAG.ncu-rep > gpurun_out/ncu_$TAG.txt 2>&1; head -30 gpurun_out/ncu_$TAG.txt
