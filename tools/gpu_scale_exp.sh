#!/bin/bash
# N=8 experiments: SMs reserved for NCCL x chunk plan
mkdir -p gpurun_out
run() { # reserve chunks
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29700+RANDOM%200)) bench.py --gpus 8 --steps 5 --warmup 3 --no-e2e --reserve $1 --chunks $2 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('reserve=$1 chunks=$2', round(d['value'],1), round(d['ms_per_step'],2), round(d['roofline']['kernel_ms'],3), d['roofline']['kernel_launches_per_step'])"
}
run 8 ramp
run 16 ramp
run 32 ramp
run 16 ramp4
run 0 1
run 8 1
