#!/bin/bash
# N=2: fused-exchange vs NCCL-broadcast sharded product, then the driver-shaped default bench line
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
N=${1:-2}
bash tools/gpu_call11.sh $N
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
echo "default N=$N rc=$?"; tail -1 gpurun_out/scale_n$N.json | cut -c1-2500; tail -5 gpurun_out/scale_n$N.err
