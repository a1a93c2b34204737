#!/bin/bash
# bench line with the two-strategy CPU arm and the full-size CPU parity check; reference arm
mkdir -p gpurun_out
timeout 400 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_auto.json 2> gpurun_out/bench_auto.err
echo "rc=$?"; cut -c1-400 gpurun_out/bench_auto.json; tail -3 gpurun_out/bench_auto.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_auto.json").read().strip().splitlines()[-1])
print(d["parity"], d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"], {k: round(v["value"], 1) for k, v in d["cpu_baseline"]["strategies"].items()}, d["e2e"]["ms_per_step"])
PY
