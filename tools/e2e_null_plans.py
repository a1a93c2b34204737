"""The call m4ri-rust makes — c = mzd_mul(NULL, a, b, 0); mzd_free(c) — under different block plans
of the host pipeline (the result block comes from the pinned-block cache and goes back to it).
usage: python tools/e2e_null_plans.py [n]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import m4ri_rust_b200 as m4
L = m4.lib()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
hA, hB = L.mzd_init(n, n), L.mzd_init(n, n)
dA, dB = m4.DMat.init(n, n).fill_random(1), m4.DMat.init(n, n).fill_random(2)
L.m4ri_b200_download(hA, dA.h)
L.m4ri_b200_download(hB, dB.h)
L.m4ri_b200_synchronize()
del dA, dB
for rows, cols in [(0, 0), (4, 4), (16, 2), (8, 2), (4, 2), (6, 3), (0, 0)]:
    if rows == 0:
        os.environ.pop("M4RI_B200_PIPE_ROWS", None), os.environ.pop("M4RI_B200_PIPE_COLS", None)
    else:
        os.environ["M4RI_B200_PIPE_ROWS"], os.environ["M4RI_B200_PIPE_COLS"] = str(rows), str(cols)
    for _ in range(2):
        L.mzd_free(L.mzd_mul(None, hA, hB, 0))
    ts = []
    for _ in range(6):
        t0 = time.perf_counter()
        L.mzd_free(L.mzd_mul(None, hA, hB, 0))
        ts.append((time.perf_counter() - t0) * 1e3)
    hC = L.mzd_init(n, n)
    tp = []
    for _ in range(4):
        t0 = time.perf_counter()
        L.mzd_mul(hC, hA, hB, 0)
        tp.append((time.perf_counter() - t0) * 1e3)
    L.mzd_free(hC)
    print(json.dumps({"n": n, "plan": [rows, cols], "null_result_ms": [round(t, 3) for t in ts],
                      "preallocated_ms": [round(t, 3) for t in tp]}), flush=True)
