#!/bin/bash
# Repo arm of the scaling bench at N GPUs (the driver runs both arms; the reference arm is CPU
# work and is not worth N x the GPU charge here) + the in-library multi-device check and tests.
#   gpurun --gpus N -- bash tools/gpu_scale.sh N [tag]
N=${1:-2}; TAG=${2:-scale}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29700+RANDOM%200)) \
  bench.py --gpus $N --steps ${STEPS:-8} --warmup 3 > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err
echo "bench rc=$?"; tail -1 gpurun_out/${TAG}_bench_n$N.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('N=%d value=%.0f ms/step=%.3f e2e_ms=%.2f kernel_ms=%.3f share=%.3f frac=%.3f launches/step=%s traffic=%s' % (d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], r['kernel_ms'], r['kernel_share_of_step'], r['frac'], r['kernel_launches_per_step'], r.get('traffic')))"
timeout 600 python tools/multi_device_check.py $N > gpurun_out/${TAG}_multi_device_n$N.json 2>&1; tail -1 gpurun_out/${TAG}_multi_device_n$N.json | cut -c1-900
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "multi_device or set_devices" > gpurun_out/${TAG}_pytest_multi_n$N.log 2>&1; tail -3 gpurun_out/${TAG}_pytest_multi_n$N.log
