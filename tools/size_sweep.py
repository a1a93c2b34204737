"""Device-resident timing of C = A*B (m4ri_b200_dmul, includes the transpose of B on
the MMA path) over sizes, algorithms and Strassen levels; every result is checked
with Freivalds (2^-64). usage: python tools/size_sweep.py [sizes...]"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import m4ri_rust_b200 as m4
L = m4.lib()
sizes = [int(x) for x in sys.argv[1:]] or [1024, 2048, 4096, 8192, 16384, 32768]
import ctypes, time
ev = ctypes.CDLL("libcudart.so.12") if False else None
out = []
for n in sizes:
    dA, dB, dC = (m4.DMat.init(n, n) for _ in range(3))
    dA.fill_random(1); dB.fill_random(2)
    for algo in (os.environ.get("SWEEP_ALGOS", "m4rm,mma").split(",")):
        for lv in (0, 1, 2, 3):
            if n // (1 << lv) < 2048:
                continue
            L.m4ri_b200_set_strassen_levels(lv)
            for _ in range(2):
                m4.dmul(dC, dA, dB, clear=True, algo=m4.ALGO_NAMES[algo])
            L.m4ri_b200_synchronize()
            reps = 5 if n >= 16384 else 20
            t0 = time.perf_counter()
            for _ in range(reps):
                m4.dmul(dC, dA, dB, clear=True, algo=m4.ALGO_NAMES[algo])
            L.m4ri_b200_synchronize()
            ms = (time.perf_counter() - t0) * 1e3 / reps
            ok = L.m4ri_b200_dfreivalds(dC.h, dA.h, dB.h, 99)
            rec = {"n": n, "algo": algo, "strassen_levels": lv, "ms": round(ms, 4),
                   "tbitops": round(2.0 * n ** 3 / ms / 1e9, 1), "freivalds": ok}
            out.append(rec)
            print(json.dumps(rec), flush=True)
L.m4ri_b200_set_strassen_levels(-1)
