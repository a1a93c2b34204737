"""BASELINE configs[0] and [1] through the host ABI the way m4ri-rust calls it:
c = mzd_mul(NULL, a, b, 0); mzd_free(c), host operands, small to medium n — next to the CPU oracle
(Four Russians / Strassen, 1 thread and all threads) on the same seeded operands."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import m4ri_rust_b200 as m4
import oracle
L = m4.lib()
for n in (64, 256, 1024, 2048, 4096, 8192):
    A, B = oracle.random(n, n, 1), oracle.random(n, n, 2)
    hA, hB = m4.mzd_from_words(A, n), m4.mzd_from_words(B, n)
    want = oracle.mul_m4rm(A, B, n, n, threads=0)
    for _ in range(3):
        L.mzd_free(L.mzd_mul(None, hA, hB, 0))
    reps = 50 if n <= 2048 else 10
    t0 = time.perf_counter()
    for _ in range(reps - 1):
        L.mzd_free(L.mzd_mul(None, hA, hB, 0))
    c = L.mzd_mul(None, hA, hB, 0)
    gpu_ms = (time.perf_counter() - t0) * 1e3 / reps
    ok = bool((m4.mzd_words(c)[:, : oracle.width(n)] == want[:, : oracle.width(n)]).all())
    L.mzd_free(c)
    def cpu(fn, **kw):
        fn(A, B, n, n, **kw)
        r = 3 if n >= 4096 else 10
        t0 = time.perf_counter()
        for _ in range(r):
            fn(A, B, n, n, **kw)
        return (time.perf_counter() - t0) * 1e3 / r
    row = {"n": n, "gpu_e2e_ms": round(gpu_ms, 4), "bit_exact": ok,
           "cpu_m4rm_1t_ms": round(cpu(oracle.mul_m4rm, threads=1), 4),
           "cpu_m4rm_all_ms": round(cpu(oracle.mul_m4rm, threads=0), 4),
           "cpu_strassen_all_ms": round(cpu(oracle.mul_strassen, threads=0), 4),
           "cpu_threads": oracle.max_threads()}
    row["gpu_tbitops"] = round(2.0 * n ** 3 / gpu_ms / 1e9, 2)
    print(json.dumps(row), flush=True)
    L.mzd_free(hA); L.mzd_free(hB)
