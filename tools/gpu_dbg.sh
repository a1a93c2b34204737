#!/bin/bash
mkdir -p gpurun_out
for d in ${DBGS:-0 1 2}; do
echo "DBG=$d"; M4RI_B200_DBG=$d timeout 120 python tools/mma_variant_check.py ${VAR:-fp4ldg} --time-only 2>&1 | tail -1
done
