"""One host mzd_mul over G GPUs (m4ri_b200_set_devices) at n = 65536 under different row-chunk
counts per GPU (M4RI_B200_PIPE_ROWS; auto = plan_row_chunks). usage: python tools/multi_device_chunks.py G [n]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import m4ri_rust_b200 as m4
L = m4.lib()
G = int(sys.argv[1]) if len(sys.argv) > 1 else 2
n = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
dA, dB = m4.DMat.init(n, n).fill_random(1), m4.DMat.init(n, n).fill_random(2)
hA, hB = L.mzd_init(n, n), L.mzd_init(n, n)
L.m4ri_b200_download(hA, dA.h); L.m4ri_b200_download(hB, dB.h); L.m4ri_b200_synchronize()
del dA, dB
L.m4ri_b200_set_devices(G)
for rows in (0, 1, 2, 4, 8, 16, 0):
    if rows:
        os.environ["M4RI_B200_PIPE_ROWS"] = str(rows)
    else:
        os.environ.pop("M4RI_B200_PIPE_ROWS", None)
    for _ in range(2):
        L.mzd_free(L.mzd_mul(None, hA, hB, 0))
    ts = []
    for _ in range(4):
        t0 = time.perf_counter()
        L.mzd_free(L.mzd_mul(None, hA, hB, 0))
        ts.append(round((time.perf_counter() - t0) * 1e3, 2))
    print(json.dumps({"G": G, "n": n, "row_chunks": rows or "auto", "ms": ts}), flush=True)
