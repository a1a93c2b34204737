"""Host-ABI mzd_mul at n = 32768 under different block plans of the copy/compute pipeline
(M4RI_B200_PIPE_ROWS x M4RI_B200_PIPE_COLS). Usage: python tools/e2e_sweep.py [n]"""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import m4ri_rust_b200 as m4

L = m4.lib()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
dA = m4.DMat.init(n, n).fill_random(1)
dB = m4.DMat.init(n, n).fill_random(2)
hA, hB, hC = L.mzd_init(n, n), L.mzd_init(n, n), L.mzd_init(n, n)
L.m4ri_b200_download(hA, dA.h)
L.m4ri_b200_download(hB, dB.h)
L.m4ri_b200_synchronize()
dC = m4.DMat.init(n, n)
m4.dmul(dC, dA, dB)
ref = dC.checksum()
import ctypes
pa, pb = ctypes.c_int(), ctypes.c_int()
L.m4ri_b200_pipeline_plan(n, n, n, 1, ctypes.byref(pa), ctypes.byref(pb))
plans = [(0, 0)] + ([(8, 1), (8, 2), (8, 4), (8, 8), (4, 4), (16, 4), (4, 8), (16, 8), (12, 6), (4, 2), (6, 3), (16, 2), (12, 2), (16, 3), (16, 1), (12, 3)] if n == 32768 else
                    [(1, 1), (2, 2), (4, 2), (4, 4), (8, 4), (2, 4), (1, 6), (1, 4), (1, 8), (2, 8), (6, 6), (8, 8)])
for rows, cols in plans:
    if rows == 0:
        os.environ.pop("M4RI_B200_PIPE_ROWS", None), os.environ.pop("M4RI_B200_PIPE_COLS", None)
        rows, cols = "auto:%d" % pa.value, "auto:%d" % pb.value
    else:
        os.environ["M4RI_B200_PIPE_ROWS"], os.environ["M4RI_B200_PIPE_COLS"] = str(rows), str(cols)
    for _ in range(2):
        L.mzd_mul(hC, hA, hB, 0)
    ts = []
    for _ in range(5):
        t0 = time.perf_counter()
        L.mzd_mul(hC, hA, hB, 0)
        ts.append((time.perf_counter() - t0) * 1e3)
    d = m4.DMat.from_host(hC)
    print(json.dumps({"n": n, "rows": rows, "cols": cols, "ms_min": round(min(ts), 3), "ms_mean": round(sum(ts) / len(ts), 3),
                      "tbitops": round(2.0 * n ** 3 / (sum(ts) / len(ts)) / 1e9, 1), "ok": d.checksum() == ref}), flush=True)
