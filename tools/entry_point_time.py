"""The reference's three multiplication entry points through the host ABI (result allocated and
freed per call, as m4ri-rust does) over small and medium n: where the AND+popcount kernel, the
Four-Russians kernel and the tensor path cross. usage: python tools/entry_point_time.py [n ...]"""
import sys, time, json, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import m4ri_rust_b200 as m4
L = m4.lib()
for n in [int(x) for x in sys.argv[1:]] or [256, 512, 768, 1024, 1536, 2048, 4096, 8192]:
    dA, dB = m4.DMat.init(n, n).fill_random(1), m4.DMat.init(n, n).fill_random(2)
    hA, hB = L.mzd_init(n, n), L.mzd_init(n, n)
    L.m4ri_b200_download(hA, dA.h); L.m4ri_b200_download(hB, dB.h); L.m4ri_b200_synchronize()
    rec = {"n": n}
    for name, f in (("mzd_mul", lambda: L.mzd_mul(None, hA, hB, 0)), ("mzd_mul_naive", lambda: L.mzd_mul_naive(None, hA, hB)),
                    ("mzd_mul_m4rm", lambda: L.mzd_mul_m4rm(None, hA, hB, 0))):
        for _ in range(5):
            L.mzd_free(f())
        reps = 30 if n <= 2048 else 5
        ts = []
        for _ in range(reps):
            t0 = time.perf_counter()
            L.mzd_free(f())
            ts.append((time.perf_counter() - t0) * 1e3)
        ts.sort()
        rec[name + "_ms"] = round(ts[len(ts) // 2], 4)
    print(json.dumps(rec), flush=True)
    L.mzd_free(hA); L.mzd_free(hB)
