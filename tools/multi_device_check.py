"""In-library multi-GPU host products (m4ri_b200_set_devices): same words as one GPU, timing.
Usage: python tools/multi_device_check.py G"""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import m4ri_rust_b200 as m4
import oracle
L = m4.lib()
G = int(sys.argv[1]) if len(sys.argv) > 1 else 2
out = {"devices_requested": G}

def timed(fn, reps=4):
    for _ in range(2):
        fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    return (time.perf_counter() - t0) * 1e3 / reps

# 1. ragged shapes against the oracle, every form, accumulate
for (m, k, n) in ([(5000, 20000, 9000), (2049, 30001, 4099)] if len(sys.argv) < 3 else []):
    A, B, C0 = oracle.random(m, k, 21), oracle.random(k, n, 22), oracle.random(m, n, 23)
    want = oracle.mul_m4rm(A, B, k, n, threads=0)[:, : oracle.width(n)]
    a, b = m4.mzd_from_words(A, k), m4.mzd_from_words(B, n)
    L.m4ri_b200_set_devices(G)
    c = L.mzd_mul(None, a, b, 0)
    ok1 = bool((m4.mzd_words(c)[:, : oracle.width(n)] == want).all())
    L.mzd_free(c)
    c = m4.mzd_from_words(C0, n)
    L.mzd_addmul(c, a, b, 0)
    ok2 = bool((m4.mzd_words(c)[:, : oracle.width(n)] == (C0[:, : oracle.width(n)] ^ want)).all())
    L.mzd_free(c)
    bt = L.mzd_transpose(None, b)
    L.m4ri_b200_set_algo(m4.ALGO_MMA)
    c = L._mzd_mul_naive(None, a, bt, 1)
    L.m4ri_b200_set_algo(m4.ALGO_AUTO)
    ok3 = bool((m4.mzd_words(c)[:, : oracle.width(n)] == want).all())
    for p in (c, bt, a, b):
        L.mzd_free(p)
    out[f"parity_{m}x{k}x{n}"] = [ok1, ok2, ok3]
out["devices_effective"] = L.m4ri_b200_get_devices()
print(json.dumps(out), flush=True)

# 2. square products: one GPU vs G GPUs, same words
for n in ((32768, 65536) if len(sys.argv) < 3 else (int(sys.argv[2]),)):
    dA, dB = m4.DMat.init(n, n).fill_random(1), m4.DMat.init(n, n).fill_random(2)
    hA, hB = L.mzd_init(n, n), L.mzd_init(n, n)
    L.m4ri_b200_download(hA, dA.h); L.m4ri_b200_download(hB, dB.h); L.m4ri_b200_synchronize()
    del dA, dB
    L.m4ri_b200_set_devices(1)
    c1 = L.mzd_mul(None, hA, hB, 0)
    t1 = timed(lambda: L.mzd_free(L.mzd_mul(None, hA, hB, 0)), reps=2)
    L.m4ri_b200_set_devices(G)
    cg = L.mzd_mul(None, hA, hB, 0)
    tg = timed(lambda: L.mzd_free(L.mzd_mul(None, hA, hB, 0)), reps=2)
    same = bool(np.array_equal(m4.mzd_words(c1), m4.mzd_words(cg)))
    print(json.dumps({"n": n, "ms_1gpu": round(t1, 2), f"ms_{G}gpu": round(tg, 2), "same_words": same,
                      "pbitops_multi": round(2.0 * n ** 3 / tg / 1e12, 2)}), flush=True)
    for p in (c1, cg, hA, hB):
        L.mzd_free(p)
