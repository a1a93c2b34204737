#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/mma_variant_check.py v2 --time > gpurun_out/check2_v2.log 2>&1; echo "exit $?" >> gpurun_out/check2_v2.log
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; echo "exit $?" >> gpurun_out/pytest_gpu2.log
timeout 600 python tools/size_sweep.py > gpurun_out/size_sweep.jsonl 2> gpurun_out/size_sweep.err
tail -7 gpurun_out/check2_v2.log; tail -3 gpurun_out/pytest_gpu2.log; cat gpurun_out/size_sweep.jsonl
