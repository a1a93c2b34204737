#!/bin/bash
# times every A/B build of the tensor kernel under m4ri-rust_b200/lib/exp/ (tools/mma6_explore.py);
# boxes differ by a few per cent, so round 1's kernel (fp4ldg) is timed first in the same call
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit,temperature.gpu --format=csv,noheader
echo "== baseline mma5 (fp4ldg)"
SIZES=${SIZES:-16384,32768} timeout 120 python tools/mma_variant_check.py fp4ldg --time-only 2>&1 | tail -${TAILN:-2}
for so in m4ri-rust_b200/lib/exp/*.so; do
  for d in ${DBGS:-0}; do
    echo "== $(basename $so .so) DBG=$d"
    M4RI_B200_DBG=$d M4RI_B200_LIB=$PWD/$so SIZES=${SIZES:-16384,32768} timeout 120 python tools/mma_variant_check.py fp4s256 --time-only 2>&1 | tail -${TAILN:-2}
  done
done
echo "== baseline mma5 again"
SIZES=${SIZES:-16384,32768} timeout 120 python tools/mma_variant_check.py fp4ldg --time-only 2>&1 | tail -${TAILN:-2}
