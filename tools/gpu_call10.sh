#!/bin/bash
# scaling run as the driver does it: N = 1, 2, 4, 8 back to back
mkdir -p gpurun_out
NG=${1:-8}
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err
for N in 2 4 8; do
  [ $N -le $NG ] || continue
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
  echo "N=$N rc=$?"
done
for N in 1 2 4 8; do [ -f gpurun_out/scale_n$N.json ] && python -c "
import json,sys
d=json.loads(open('gpurun_out/scale_n$N.json').read().strip().splitlines()[-1])
print($N, round(d['value'],1), round(d['ms_per_step'],2), d['roofline']['kernel_ms'], d.get('e2e',{}).get('value'))"; done
