#!/bin/bash
# ncu --set full capture of ONE grouped leaf launch (the seven leaf products of a fused Strassen
# level) of the row-block products of the N-GPU bench: bash tools/gpu_ncu_group.sh <tag>
#   N = 8: 8192:65536:65536 (one level: 1 grouped launch per product)
#   N = 4: 16384:65536:65536, N = 2: 32768:65536:65536 (two levels: 7 grouped launches per product)
TAG=${1:-r2}
mkdir -p gpurun_out
for spec in "8192:65536:65536 2" "16384:65536:65536 14" "32768:65536:65536 14"; do
  set -- $spec; S=$1; SKIP=$2; NAME=${S//:/x}
  timeout 300 python tools/ncu_group_launch.py $S > gpurun_out/plain_${TAG}_group_$NAME.log 2>&1 || { echo "plain run failed for $S"; continue; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:mma6_kernel -s $SKIP -c 1 -f \
    -o gpurun_out/prof_${TAG}_group_$NAME python tools/ncu_group_launch.py $S > gpurun_out/ncu_${TAG}_group_$NAME.log 2>&1
  python tools/ncu_summary.py gpurun_out/prof_${TAG}_group_$NAME.ncu-rep > gpurun_out/${TAG}_ncu_mma6_group7_block_$NAME.txt 2>&1
  head -12 gpurun_out/${TAG}_ncu_mma6_group7_block_$NAME.txt
done
