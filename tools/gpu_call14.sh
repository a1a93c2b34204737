#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "matvec or pipelined" > gpurun_out/pytest_k.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_k.log
tail -15 gpurun_out/pytest_k.log
timeout 300 python tools/e2e_sweep.py > gpurun_out/e2e_sweep.jsonl 2> gpurun_out/e2e_sweep.err; cat gpurun_out/e2e_sweep.jsonl; tail -3 gpurun_out/e2e_sweep.err
timeout 300 python bench.py --workload lpn --steps 20 --warmup 3 > gpurun_out/bench_lpn.json 2> gpurun_out/bench_lpn.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_lpn.json").read().strip().splitlines()[-1])
print({k: (round(v["ms"], 4), round(v["roofline"]["frac"], 3)) for k, v in d.items() if k.startswith("matvec")})
PY
