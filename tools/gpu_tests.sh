#!/bin/bash
# GPU test suite + smoke on one B200: bash tools/gpu_tests.sh [tag] [pytest args...]
mkdir -p gpurun_out
TAG=${1:-r2}; shift
timeout 1500 python -m pytest tests -m gpu -q "$@" > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_gpu.log
tail -15 gpurun_out/${TAG}_pytest_gpu.log
