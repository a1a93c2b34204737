#!/bin/bash
# GPU call: parity tests, bench both paths, ncu launch list + full capture of each product kernel
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_m4rm.json 2> gpurun_out/bench_m4rm.err
python bench.py --steps 5 --warmup 3 --algo mma --no-cpu-baseline > gpurun_out/bench_mma.json 2> gpurun_out/bench_mma.err
python bench.py --steps 5 --warmup 3 --algo mma --n 8192 --no-cpu-baseline --no-e2e > gpurun_out/bench_mma_8192.json 2>> gpurun_out/bench_mma.err
python bench.py --steps 5 --warmup 3 --algo m4rm --n 8192 --no-cpu-baseline --no-e2e > gpurun_out/bench_m4rm_8192.json 2>> gpurun_out/bench_m4rm.err
C1="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --algo m4rm"
$C1 > gpurun_out/plain1.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_m4rm.csv $C1 > gpurun_out/ncu1.log 2>&1
$C1 > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:m4rm_kernel -s 1 -c 1 -o gpurun_out/prof_m4rm $C1 > gpurun_out/ncu2.log 2>&1
C2="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --algo mma --n 8192"
$C2 > gpurun_out/plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:mma_i8 -s 1 -c 1 -o gpurun_out/prof_mma $C2 > gpurun_out/ncu3.log 2>&1
echo done
