#!/bin/bash
# Memory-safety evidence -> gpurun_out/sanitize_*.log
# compute-sanitizer is closed on this pool (its wrapper prints a refusal); the guarded sweep is the substitute.
mkdir -p gpurun_out
for tool in ${TOOLS:-memcheck}; do
  timeout 300 compute-sanitizer --tool $tool --error-exitcode 3 python tools/sanitize_sweep.py m4rm > gpurun_out/sanitize_$tool.log 2>&1
  echo "$tool rc=$?: $(head -c 200 gpurun_out/sanitize_$tool.log)"
done
timeout ${SAN_TIMEOUT:-1500} python tools/sanitize_sweep.py --guard --repeat ${REPEAT:-5} ${FAMS:-} > gpurun_out/sanitize_guard.log 2>&1
echo "guard rc=$?"; tail -3 gpurun_out/sanitize_guard.log
