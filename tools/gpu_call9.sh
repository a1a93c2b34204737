#!/bin/bash
# round evidence: parity, bench (both arms), launch list, ncu full capture of the dominant kernel as the bench runs it
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_auto.json 2> gpurun_out/bench_auto.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
timeout 300 python bench.py --workload lpn --steps 10 --warmup 3 > gpurun_out/bench_lpn.json 2> gpurun_out/bench_lpn.err
C1="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e"
timeout 300 $C1 > gpurun_out/plain1.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_auto.csv $C1 > gpurun_out/ncu1.log 2>&1
timeout 300 $C1 > gpurun_out/plain2.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:mma5_kernel -s 8 -c 1 -o gpurun_out/prof_mma5_leaf16384 $C1 > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_auto.json gpurun_out/bench_reference.json | cut -c1-3000
