#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
for n in 32768 16384 12288; do
timeout 300 python tools/e2e_sweep.py $n > gpurun_out/e2e_sweep_$n.jsonl 2> gpurun_out/e2e_sweep_$n.err; cat gpurun_out/e2e_sweep_$n.jsonl; tail -3 gpurun_out/e2e_sweep_$n.err
done
