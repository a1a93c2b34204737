#!/bin/bash
# mat-vec kernels: parity (new coalesced kernel vs thread-per-row vs oracle), LPN bench both ways, ncu of mat-vec + transpose
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "matvec or lpn or golden or known" > gpurun_out/pytest_matvec.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_matvec.log
tail -3 gpurun_out/pytest_matvec.log
timeout 300 python bench.py --workload lpn --steps 20 --warmup 3 > gpurun_out/bench_lpn.json 2> gpurun_out/bench_lpn.err
M4RI_B200_MULV=rows timeout 300 python bench.py --workload lpn --steps 20 --warmup 3 > gpurun_out/bench_lpn_rows.json 2> gpurun_out/bench_lpn_rows.err
python - <<'PY'
import json
for f in ("bench_lpn", "bench_lpn_rows"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, {k: (round(v["ms"], 4), round(v["roofline"]["frac"], 3)) for k, v in d.items() if k.startswith("matvec")})
    except Exception as e:
        print(f, "failed", e)
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"mul_v|transpose" -c 12 -o gpurun_out/prof_matvec python bench.py --workload lpn --steps 1 --warmup 1 > gpurun_out/ncu_matvec.log 2>&1
echo "ncu rc=$?"
