#!/bin/bash
# Elimination: tests, timings, and a warm-cache per-kernel launch list (n = 8192) -> gpurun_out/
mkdir -p gpurun_out
timeout 1000 python -m pytest tests/test_gpu_elim.py -m gpu -q 2>&1 | tail -3
python tools/elim_time.py ${SIZES:-4096 8192 16384 32768} | tee gpurun_out/elim_time.log
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file gpurun_out/elim_launches.csv python tools/elim_time.py 8192 > /dev/null 2>&1
python - <<PY | tee gpurun_out/elim_launches_summary.txt
import csv,collections
rows=list(csv.reader(l for l in open("gpurun_out/elim_launches.csv") if l.startswith('"')))
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value"); ui=h.index("Metric Unit")
d=collections.defaultdict(lambda:[0,0.0])
for r in rows[1:]:
    v=float(r[vi].replace(",","")); v = v/1000 if r[ui]=="ns" else v
    d[r[ki][:48]][0]+=1; d[r[ki][:48]][1]+=v
for k,(c,t) in sorted(d.items(), key=lambda x:-x[1][1])[:12]: print(f"{t:10.1f} us {c:6d} launches {t/c:8.2f} us avg  {k}")
PY
