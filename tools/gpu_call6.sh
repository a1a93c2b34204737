#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/fp4_check.py > gpurun_out/fp4_check.log 2>&1; echo "exit $?" >> gpurun_out/fp4_check.log
timeout 300 python tools/mma_variant_check.py fp4cg1 --time > gpurun_out/check_fp4cg1.log 2>&1; echo "exit $?" >> gpurun_out/check_fp4cg1.log
timeout 300 python tools/mma_variant_check.py fp4 --time > gpurun_out/check_fp4.log 2>&1; echo "exit $?" >> gpurun_out/check_fp4.log
timeout 300 python tools/mma_variant_check.py v2 --time > gpurun_out/check_v2.log 2>&1; echo "exit $?" >> gpurun_out/check_v2.log
tail -12 gpurun_out/fp4_check.log; tail -8 gpurun_out/check_fp4cg1.log; tail -8 gpurun_out/check_fp4.log; tail -5 gpurun_out/check_v2.log
