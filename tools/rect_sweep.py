"""Device-resident C = A*Bt^T (m4ri_b200_dmul_abt, no transpose) on rectangular shapes with 0 / 1 / 2
Strassen levels forced: where the fused level pays (the pieces of the host pipeline and the
row-blocks of the N-GPU product are rectangular). usage: python tools/rect_sweep.py [m:k:n ...]"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import m4ri_rust_b200 as m4
L = m4.lib()
shapes = [tuple(int(v) for v in a.split(":")) for a in sys.argv[1:]] or [
    (4096, 32768, 4096), (8192, 32768, 8192), (8192, 32768, 16384), (16384, 32768, 16384), (8192, 32768, 32768),
    (8192, 16384, 8192), (16384, 16384, 16384), (4096, 16384, 32768), (8192, 8192, 8192), (32768, 8192, 32768)]
for (m, k, n) in shapes:
    dA, dBt, dC = m4.DMat.init(m, k), m4.DMat.init(n, k), m4.DMat.init(m, n)
    dA.fill_random(1); dBt.fill_random(2)
    dB = m4.DMat.init(k, n)
    L.m4ri_b200_dtranspose(dB.h, dBt.h)
    for lv in (0, 1, 2):
        if min(m, k, n) >> lv < 2048:
            continue
        L.m4ri_b200_set_strassen_levels(lv)
        for _ in range(2):
            m4.dmul_abt(dC, dA, dBt, clear=True, algo=m4.ALGO_MMA)
        L.m4ri_b200_synchronize()
        reps = 10
        t0 = time.perf_counter()
        for _ in range(reps):
            m4.dmul_abt(dC, dA, dBt, clear=True, algo=m4.ALGO_MMA)
        L.m4ri_b200_synchronize()
        ms = (time.perf_counter() - t0) * 1e3 / reps
        ok = L.m4ri_b200_dfreivalds(dC.h, dA.h, dB.h, 99)
        print(json.dumps({"m": m, "k": k, "n": n, "strassen_levels": lv, "ms": round(ms, 4),
                          "tbitops": round(2.0 * m * k * n / ms / 1e9, 1), "freivalds": ok}), flush=True)
    del dA, dBt, dC, dB
L.m4ri_b200_set_strassen_levels(-1)
