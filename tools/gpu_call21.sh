#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
M4RI_B200_HOST_CACHE_MB=0 timeout 120 python tools/null_result_time.py > gpurun_out/null_result_nocache.json 2>&1; cat gpurun_out/null_result_nocache.json
timeout 120 python tools/null_result_time.py > gpurun_out/null_result_cache.json 2>&1; cat gpurun_out/null_result_cache.json
