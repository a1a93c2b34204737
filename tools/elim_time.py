"""Timing of the GPU elimination (csrc/elim.cu) next to the CPU oracle: RREF of a random n x n
matrix and inversion, device-resident and through the host ABI. usage: python tools/elim_time.py [n...]"""
import ctypes, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import m4ri_rust_b200 as m4
import oracle
L = m4.lib()
for n in [int(x) for x in sys.argv[1:]] or [1024, 4096, 8192, 16384, 32768]:
    rec = {"n": n, "block": os.environ.get("M4RI_B200_ELIM_BLOCK", "default")}
    W = m4.DMat.init(n, n)
    for rep in range(2):
        W.fill_random(7)
        L.m4ri_b200_synchronize()
        t0 = time.perf_counter()
        l0 = m4.launch_count()
        r = L.m4ri_b200_drref(W.h, 0, None, 0)
        L.m4ri_b200_synchronize()
        rec["rref_device_ms"] = (time.perf_counter() - t0) * 1e3
        rec["rref_launches"] = m4.launch_count() - l0
    rec["rank"] = r
    # PLE decomposition, device-resident (P, Q come back as host arrays), checked at the end by
    # the GPU's own product: rows of the input permuted by P must equal L * E
    Pv, Qv = (ctypes.c_int * n)(), (ctypes.c_int * n)()
    for rep in range(2):
        W.fill_random(7)
        L.m4ri_b200_synchronize()
        t0 = time.perf_counter()
        rp = L.m4ri_b200_dple(W.h, Pv, Qv, 0)
        L.m4ri_b200_synchronize()
        rec["ple_device_ms"] = (time.perf_counter() - t0) * 1e3
    rec["ple_rank"] = rp
    if n <= 16384:
        packed = oracle.to_bits(np.ascontiguousarray(m4.mzd_words(W.to_host())), n)
        lower = np.zeros_like(packed)
        upper = np.zeros_like(packed)
        q = list(Qv)
        for i in range(n):
            k = min(i, rp)
            lower[i, :k] = packed[i, :k]
            if i < rp:
                lower[i, i] = 1
                upper[i, q[i]:] = packed[i, q[i]:]
        l, u = m4.mzd_from_words(oracle.from_bits(lower), n), m4.mzd_from_words(oracle.from_bits(upper), n)
        prod = L.mzd_mul(None, l, u, 0)
        a0 = W.fill_random(7).to_host()
        P = L.mzp_init(n)
        for i in range(n):
            P.contents.values[i] = Pv[i]
        L.mzd_apply_p_left(a0, P)
        rec["ple_identity_holds"] = bool(L.mzd_equal(prod, a0) == 1)
        for x in (l, u, prod, a0):
            L.mzd_free(x)
        L.mzp_free(P)
    # inversion through the host ABI (H2D + [A|I] elimination + D2H)
    a = W.fill_random(7).to_host()
    for rep in range(2):
        t0 = time.perf_counter()
        p = L.mzd_inv_m4ri(None, a, 0)
        rec["inv_host_ms"] = (time.perf_counter() - t0) * 1e3
        rec["invertible"] = bool(p)
        if p:
            L.mzd_free(p)
    if n <= 4096:
        A = m4.mzd_words(a).copy()
        t0 = time.perf_counter()
        oracle.rref(A, n)
        rec["rref_cpu_oracle_ms"] = (time.perf_counter() - t0) * 1e3
    L.mzd_free(a)
    print(json.dumps(rec), flush=True)
