#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
for ex in fused nccl; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29800+RANDOM%100)) bench.py --gpus $N --steps 5 --warmup 3 --no-e2e --exchange $ex > gpurun_out/ex_$ex.json 2> gpurun_out/ex_$ex.err
echo "$ex rc=$?"; tail -1 gpurun_out/ex_$ex.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$ex', round(d['value'],1), round(d['ms_per_step'],2), round(d['roofline']['kernel_ms'],3), d['roofline']['kernel_launches_per_step'], d['parity'])" || tail -15 gpurun_out/ex_$ex.err
done
