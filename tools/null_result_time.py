"""mzd_mul(NULL, a, b, 0) + mzd_free(c): the call m4ri-rust's `&a * &b` makes (binary_matrix.rs:465),
result allocated by the library every time. Run with M4RI_B200_HOST_CACHE_MB=0 for the uncached cost."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import m4ri_rust_b200 as m4
L = m4.lib()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
dA, dB = m4.DMat.init(n, n).fill_random(1), m4.DMat.init(n, n).fill_random(2)
hA, hB = L.mzd_init(n, n), L.mzd_init(n, n)
L.m4ri_b200_download(hA, dA.h); L.m4ri_b200_download(hB, dB.h); L.m4ri_b200_synchronize()
dC = m4.DMat.init(n, n); m4.dmul(dC, dA, dB); ref = dC.checksum()
ts = []
for i in range(7):
    t0 = time.perf_counter()
    c = L.mzd_mul(None, hA, hB, 0)
    t1 = time.perf_counter()
    if i == 6:
        ok = m4.DMat.from_host(c).checksum() == ref
    t2 = time.perf_counter()
    L.mzd_free(c)
    ts.append((t1 - t0 + time.perf_counter() - t2) * 1e3)
print(json.dumps({"n": n, "cache_mb": os.environ.get("M4RI_B200_HOST_CACHE_MB", "default"),
                  "ms_first": round(ts[0], 2), "ms_steady": round(sum(ts[2:]) / len(ts[2:]), 2), "ok": ok}))
