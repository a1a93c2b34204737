#!/bin/bash
# quick parity + timing of one MMA variant: bash tools/gpu_quick.sh fp4
mkdir -p gpurun_out
v=${1:-fp4}
timeout 300 python tools/mma_variant_check.py $v --time > gpurun_out/quick_$v.log 2>&1; echo "exit $?" >> gpurun_out/quick_$v.log
tail -7 gpurun_out/quick_$v.log
