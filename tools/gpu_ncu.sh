#!/bin/bash
# ncu full capture of one launch of the default tensor kernel at a cube size, Strassen off:
#   bash tools/gpu_ncu.sh <kernel-regex> <tag> <n>      -> gpurun_out/prof_<tag>_n<n>.ncu-rep + .txt summary
# (n may also be m:k:n for a rectangular launch, e.g. the N-GPU row-block leaves 16384:32768:32768)
mkdir -p gpurun_out
K=${1:-mma6_kernel}; TAG=${2:-mma6}; N=${3:-16384}
C1="python tools/mma_variant_check.py ${VAR:-fp4s256} --time-only"
case $N in *:*) export SHAPES=$N;; *) export SIZES=$N;; esac
N=${N//:/x}
timeout 300 $C1 > gpurun_out/plain_${TAG}_n$N.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 -f -o gpurun_out/prof_${TAG}_n$N $C1 > gpurun_out/ncu_${TAG}_n$N.log 2>&1
tail -1 gpurun_out/plain_${TAG}_n$N.log; tail -2 gpurun_out/ncu_${TAG}_n$N.log
python tools/ncu_summary.py gpurun_out/prof_${TAG}_n$N.ncu-rep > gpurun_out/ncu_${TAG}_n$N.txt 2>&1; head -30 gpurun_out/ncu_${TAG}_n$N.txt
