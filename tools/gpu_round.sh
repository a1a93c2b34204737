#!/bin/bash
# Round evidence on one B200: GPU tests, the default bench line (both arms), the LPN line,
# the launch list of the bench command. usage: bash tools/gpu_round.sh <tag> [steps]
TAG=${1:-r2}; STEPS=${2:-10}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_gpu.log
tail -3 gpurun_out/${TAG}_pytest_gpu.log
timeout 600 python bench.py --steps $STEPS --warmup 3 > gpurun_out/${TAG}_bench_n1.json 2> gpurun_out/${TAG}_bench_n1.err; echo "bench exit $?"
cut -c1-1500 gpurun_out/${TAG}_bench_n1.json
timeout 300 python bench.py --workload lpn --steps 20 --warmup 3 > gpurun_out/${TAG}_bench_lpn.json 2> gpurun_out/${TAG}_bench_lpn.err; echo "lpn exit $?"
cut -c1-3000 gpurun_out/${TAG}_bench_lpn.json
if [ "${NCU:-1}" = "1" ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_bench_n1.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/${TAG}_ncu_launches.log 2>&1; echo "ncu exit $?"
fi
