#!/bin/bash
# Isolation experiment on a tensor kernel variant, perf-hook build only
# (python m4ri-rust_b200/build.py --perf): DBG bit 0 = no operand expansion, bit 2 = no accumulator
# drain; 5 = MMA issue + barriers only = the measured FP4 pipe rate under the same clocks.
# usage: VAR=fp4s256 DBGS="0 1 4 5" SIZES=16384,32768 bash tools/gpu_dbg.sh
mkdir -p gpurun_out
export M4RI_B200_LIB=$PWD/m4ri-rust_b200/lib/libm4ri_b200_perf.so
for d in ${DBGS:-0 1 4 5}; do
  echo "DBG=$d"; M4RI_B200_DBG=$d timeout 180 python tools/mma_variant_check.py ${VAR:-fp4s256} --time-only 2>&1 | tail -${TAILN:-2}
done
