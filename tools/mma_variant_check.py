"""Parity sweep + timing of one path-(b) kernel variant against the CPU oracle.
usage: python tools/mma_variant_check.py v1|v2cg1|v2 [--time]
(test/diagnostic tool: the oracle is the checker, never the thing measured)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import m4ri_rust_b200 as m4
import oracle

variant = sys.argv[1]
L = m4.lib()
L.m4ri_b200_set_mma_variant(m4.MMA_VARIANTS[variant])
SWEEP = [(128, 256, 128), (256, 256, 256), (128, 1024, 128), (256, 2048, 256), (1, 1, 1), (1, 64, 1),
         (2, 63, 65), (7, 129, 3), (65, 64, 64), (64, 65, 64), (64, 64, 65), (33, 31, 200),
         (191, 193, 129), (200, 257, 321), (5, 700, 1100), (1030, 40, 1030), (1024, 1024, 1024),
         (257, 2050, 1500), (256, 4096, 128), (1500, 96, 70), (3000, 5000, 2500), (700, 9000, 300)]
bad = 0
for (m, k, n) in ([] if '--time-only' in sys.argv else SWEEP):
    A, B = oracle.random(m, k, 77 + m), oracle.random(k, n, 99 + n)
    want = oracle.mul_m4rm(A, B, k, n, threads=0)[:, : oracle.width(n)]
    a, b = m4.mzd_from_words(A, k), m4.mzd_from_words(B, n)
    dA, dB, dC = m4.DMat.from_host(a), m4.DMat.from_host(b), m4.DMat.init(m, n)
    m4.dmul(dC, dA, dB, clear=True, algo=m4.ALGO_MMA)
    got = dC.to_words()[:, : oracle.width(n)]
    ok = bool((got == want).all())
    m4.dmul(dC, dA, dB, clear=False, algo=m4.ALGO_MMA)
    ok2 = not dC.to_words().any()
    if not (ok and ok2):
        bad += 1
        nbad = int((got != want).sum())
        rows = np.flatnonzero((got != want).any(axis=1))[:8]
        print(f"{variant} {m}x{k}x{n}: MISMATCH words={nbad} addmul_ok={ok2} first bad rows {rows.tolist()}", flush=True)
    else:
        print(f"{variant} {m}x{k}x{n}: ok", flush=True)
print(f"{variant}: {len(SWEEP) - bad}/{len(SWEEP)} shapes bit-exact", flush=True)
if ("--time" in sys.argv or "--time-only" in sys.argv) and bad == 0:
    import ctypes
    L.m4ri_b200_set_strassen_levels(int(os.environ.get("LEVELS", "0")))  # time the kernel itself
    sizes = [int(x) for x in os.environ["SIZES"].split(",")] if os.environ.get("SIZES") else None
    # SHAPES=m:k:n,... times C (m x n) = A (m x k) * Bt^T (Bt n x k) instead of cubes
    shapes = [tuple(int(v) for v in x.split(":")) for x in os.environ["SHAPES"].split(",")] if os.environ.get("SHAPES") else None
    for n in (shapes or sizes or ((32768,) if '--time-only' in sys.argv else (4096, 8192, 16384, 32768))):
        if isinstance(n, tuple):
            (m_, k_, n_) = n
            dA, dBt, dC = m4.DMat.init(m_, k_), m4.DMat.init(n_, k_), m4.DMat.init(m_, n_)
            dA.fill_random(1); dBt.fill_random(2)
            for _ in range(2):
                m4.dmul_abt(dC, dA, dBt, clear=True, algo=m4.ALGO_MMA)
            L.m4ri_b200_synchronize()
            L.m4ri_b200_timing_reset()
            for _ in range(5):
                m4.dmul_abt(dC, dA, dBt, clear=True, algo=m4.ALGO_MMA)
            ms = L.m4ri_b200_timing_ms() / 5
            print(f"{variant} {m_}x{k_}x{n_}: {ms:.3f} ms  {2.0 * m_ * k_ * n_ / ms / 1e9:.1f} Tbit-op/s", flush=True)
            continue
        dA, dB, dBt, dC = (m4.DMat.init(n, n) for _ in range(4))
        dA.fill_random(1); dB.fill_random(2)
        L.m4ri_b200_dtranspose(dBt.h, dB.h)
        for _ in range(2):
            m4.dmul_abt(dC, dA, dBt, clear=True, algo=m4.ALGO_MMA)
        L.m4ri_b200_synchronize()
        L.m4ri_b200_timing_reset()
        reps = 5
        for _ in range(reps):
            m4.dmul_abt(dC, dA, dBt, clear=True, algo=m4.ALGO_MMA)
        ms = L.m4ri_b200_timing_ms() / reps
        ok = L.m4ri_b200_dfreivalds(dC.h, dA.h, dB.h, 123)
        print(f"{variant} n={n}: {ms:.3f} ms  {2.0 * n ** 3 / ms / 1e9:.1f} Tbit-op/s  freivalds={ok}", flush=True)
