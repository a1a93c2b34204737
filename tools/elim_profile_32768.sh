python tools/elim_time.py 32768 | cut -c1-200
M4RI_B200_ELIM_BLOCK=2048 python tools/elim_time.py 32768 | cut -c1-200
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file gpurun_out/elim_launches_32768.csv python tools/elim_time.py 32768 > /dev/null 2>&1
python - <<PY
import csv,collections
rows=list(csv.reader(l for l in open("gpurun_out/elim_launches_32768.csv") if l.startswith('"')))
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value"); ui=h.index("Metric Unit")
d=collections.defaultdict(lambda:[0,0.0])
for r in rows[1:]:
    v=float(r[vi].replace(",","")); v = v/1000 if r[ui]=="ns" else v
    d[r[ki][:48]][0]+=1; d[r[ki][:48]][1]+=v
for k,(c,t) in sorted(d.items(), key=lambda x:-x[1][1])[:12]: print(f"{t:10.1f} us {c:6d} launches {t/c:8.2f} us avg  {k}")
PY
