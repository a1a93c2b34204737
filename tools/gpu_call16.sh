#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 tools/exchange_probe.py > gpurun_out/exchange_probe_n$N.json 2> gpurun_out/exchange_probe_n$N.err
echo "rc=$?"; cat gpurun_out/exchange_probe_n$N.json; tail -5 gpurun_out/exchange_probe_n$N.err
