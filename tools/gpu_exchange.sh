#!/bin/bash
# Multi-GPU exchange experiments at N ranks:  gpurun --gpus N -- bash tools/gpu_exchange.sh N [probe|compare|both]
#   probe    tools/exchange_probe.py: where the fused exchange's time goes
#   compare  the device-resident bench line with the fused exchange and with the NCCL broadcast
mkdir -p gpurun_out
N=${1:-2}; WHAT=${2:-both}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500+RANDOM%400))"
if [ "$WHAT" != "compare" ]; then
  timeout 300 $TR tools/exchange_probe.py > gpurun_out/exchange_probe_n$N.json 2> gpurun_out/exchange_probe_n$N.err
  echo "probe rc=$?"; tail -1 gpurun_out/exchange_probe_n$N.json | cut -c1-2500
fi
if [ "$WHAT" != "probe" ]; then
  for ex in fused nccl; do
    timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 --no-e2e --exchange $ex > gpurun_out/ex_${ex}_n$N.json 2> gpurun_out/ex_${ex}_n$N.err
    echo "$ex rc=$?"; tail -1 gpurun_out/ex_${ex}_n$N.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$ex', round(d['value'],1), round(d['ms_per_step'],2), round(d['roofline']['kernel_ms'],3), d['roofline']['kernel_launches_per_step'], d['parity'])" || tail -15 gpurun_out/ex_${ex}_n$N.err
  done
fi
