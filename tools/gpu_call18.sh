#!/bin/bash
# default bench line at N ranks (what the driver's scaling run launches)
mkdir -p gpurun_out
N=${1:-8}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
echo "default N=$N rc=$?"; tail -1 gpurun_out/scale_n$N.json | cut -c1-2500; grep -v "^\*\|OMP_NUM" gpurun_out/scale_n$N.err | tail -5
