"""Device-resident C ^= A*B with a SHORT contraction (the trailing updates of the elimination:
k = block width; the LPN product: k = 512), default kernel vs a variant build
(M4RI_B200_LIB=m4ri-rust_b200/lib/exp/<name>.so). usage: python tools/shortk_sweep.py"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import m4ri_rust_b200 as m4
L = m4.lib()
for (m, k, n) in [(32768, 1024, 16384), (32768, 1024, 32768), (32768, 2048, 16384), (32768, 512, 16384),
                  (1 << 20, 512, 512), (16384, 16384, 16384)]:
    dA, dB, dC = m4.DMat.init(m, k), m4.DMat.init(k, n), m4.DMat.init(m, n)
    dA.fill_random(1); dB.fill_random(2)
    for clear in (False, True):
        for _ in range(2):
            m4.dmul(dC, dA, dB, clear=clear, algo=m4.ALGO_MMA)
        L.m4ri_b200_synchronize()
        t0 = time.perf_counter()
        for _ in range(10):
            m4.dmul(dC, dA, dB, clear=clear, algo=m4.ALGO_MMA)
        L.m4ri_b200_synchronize()
        ms = (time.perf_counter() - t0) * 100
        ok = L.m4ri_b200_dfreivalds(dC.h, dA.h, dB.h, 99) if clear else -1
        print(json.dumps({"lib": os.path.basename(m4.LIB_PATH), "m": m, "k": k, "n": n, "clear": clear, "ms": round(ms, 4),
                          "tbitops": round(2.0 * m * k * n / ms / 1e9, 1), "freivalds": ok}), flush=True)
    del dA, dB, dC
