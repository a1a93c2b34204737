#!/bin/bash
mkdir -p gpurun_out
for v in v2cg1 v2 v1; do
  timeout 300 python tools/mma_variant_check.py $v --time > gpurun_out/check_$v.log 2>&1; echo "exit $?" >> gpurun_out/check_$v.log
done
tail -n 8 gpurun_out/check_*.log
