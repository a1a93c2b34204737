#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 200 python tools/transpose_time.py > gpurun_out/transpose_time.jsonl 2> gpurun_out/transpose_time.err; cat gpurun_out/transpose_time.jsonl; tail -3 gpurun_out/transpose_time.err
