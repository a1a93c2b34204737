#!/bin/bash
# round evidence (1 GPU): parity, bench (both arms), LPN line, launch list
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_auto.json 2> gpurun_out/bench_auto.err
timeout 300 python bench.py --workload lpn --steps 20 --warmup 3 > gpurun_out/bench_lpn.json 2> gpurun_out/bench_lpn.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
C1="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_auto.csv $C1 > gpurun_out/ncu1.log 2>&1
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log; cat gpurun_out/bench_auto.json | cut -c1-3000; tail -3 gpurun_out/bench_auto.err
