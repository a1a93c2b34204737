"""Concurrent entry into the host ABI on the GPU (SURVEY 8b "Threading").

BinMatrix is Send + Sync (m4ri-rust/src/friendly/binary_matrix.rs:38-39) and the reference builds
M4RI with --enable-thread-safe (m4ri-sys/build.rs:44): a Rust program may multiply from several
threads at once, on distinct operands or on one shared read-only matrix. Here eight host threads
call the library at the same time — ctypes drops the GIL for the duration of a foreign call, so
the calls really overlap inside libm4ri_b200.so — mixing every size class of mzd_mul (the
256 KiB single-copy path, the plain upload/compute/download path, the pipelined path from 24 MiB
of operands on), mzd_addmul, the mul_slice shape of mzd_mul_naive, mzd_echelonize, and mzd_free
of results the library allocated. Every result is compared word for word with the CPU oracle."""
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

THREADS, ITERATIONS = 8, 20


def test_products_from_eight_threads(gpu, m4, orc):
    L = m4.lib()
    W = orc.width

    # one B per size class, shared read-only by all threads; every thread owns its A's
    shapes = {"tiny": (100, 120, 90), "mid": (1000, 1500, 1100), "big": (4096, 16384, 8192)}
    Bw = {c: orc.random(k, n, 1000 + i) for i, (c, (m, k, n)) in enumerate(shapes.items())}
    Bh = {c: m4.mzd_from_words(Bw[c], shapes[c][2]) for c in shapes}
    big_A = orc.random(shapes["big"][0], shapes["big"][1], 77)  # the pipelined size: one shared A as well
    big_Ah = m4.mzd_from_words(big_A, shapes["big"][1])
    big_want = orc.mul_m4rm(big_A, Bw["big"], shapes["big"][1], shapes["big"][2], threads=0)[:, : W(shapes["big"][2])]
    assert (shapes["big"][0] * W(shapes["big"][1]) + shapes["big"][1] * W(shapes["big"][2])) * 8 >= 24 << 20

    plans = []
    for t in range(THREADS):
        p = {}
        for c in ("tiny", "mid"):
            m, k, n = shapes[c]
            A = orc.random(m, k, 10 * t + len(c))
            p[c] = (A, orc.mul_m4rm(A, Bw[c], k, n)[:, : W(n)])
        # mul_slice: A * x with x an n x 1 column (binary_matrix.rs:426-427)
        ms, ns = 700 + 13 * t, 500 + 7 * t
        As, x = orc.random(ms, ns, 300 + t), orc.random(1, ns, 400 + t)
        y = orc.mul_v(As, x[0], ns)
        p["slice"] = (As, ns, x, np.array([(int(y[i // 64]) >> (i % 64)) & 1 for i in range(ms)], dtype=np.uint64))
        # echelon form of a rank-deficient matrix
        me, ne, re_ = 300 + t, 400, 150
        E = orc.mul_m4rm(orc.random(me, re_, 500 + t), orc.random(re_, ne, 600 + t), re_, ne)
        p["rref"] = (E, ne) + tuple(orc.rref(E, ne)[:2])
        p["acc"] = orc.random(shapes["mid"][0], shapes["mid"][2], 700 + t)
        plans.append(p)

    errors = []
    start = threading.Barrier(THREADS)

    def worker(t):
        try:
            p = plans[t]
            hA = {c: m4.mzd_from_words(p[c][0], shapes[c][1]) for c in ("tiny", "mid")}
            hs = m4.mzd_from_words(p["slice"][0], p["slice"][1])
            xcol = L.mzd_transpose(None, m4.mzd_from_words(p["slice"][2], p["slice"][1]))
            start.wait()
            for it in range(ITERATIONS):
                kind = (it + t) % 6
                if kind in (0, 1):
                    c = "tiny" if kind == 0 else "mid"
                    n = shapes[c][2]
                    r = L.mzd_mul(None, hA[c], Bh[c], 0)
                    ok = (m4.mzd_words(r)[:, : W(n)] == p[c][1]).all()
                    L.mzd_free(r)
                elif kind == 2:  # C ^= A*B on a caller-owned C
                    n = shapes["mid"][2]
                    hc = m4.mzd_from_words(p["acc"], n)
                    assert L.mzd_addmul(hc, hA["mid"], Bh["mid"], 0)
                    ok = (m4.mzd_words(hc)[:, : W(n)] == (p["acc"][:, : W(n)] ^ p["mid"][1])).all()
                    L.mzd_free(hc)
                elif kind == 3:
                    r = L.mzd_mul_naive(None, hs, xcol)
                    ok = (m4.mzd_words(r)[:, 0] == p["slice"][3]).all()
                    L.mzd_free(r)
                elif kind == 4:
                    E, ne, want, rank = p["rref"]
                    he = m4.mzd_from_words(E, ne)
                    ok = L.mzd_echelonize(he, 1) == rank and (m4.mzd_words(he)[:, : W(ne)] == want[:, : W(ne)]).all()
                    L.mzd_free(he)
                else:  # both operands shared with the other threads
                    r = L.mzd_mul(None, big_Ah, Bh["big"], 0)
                    ok = (m4.mzd_words(r)[:, : W(shapes["big"][2])] == big_want).all()
                    L.mzd_free(r)
                if not ok:
                    errors.append((t, it, kind))
            for h in list(hA.values()) + [hs, xcol]:
                L.mzd_free(h)
        except Exception as ex:  # a thread that dies must fail the test, not vanish
            errors.append((t, "exception", repr(ex)))

    threads = [threading.Thread(target=worker, args=(t,)) for t in range(THREADS)]
    for th in threads:
        th.start()
    for th in threads:
        th.join(timeout=600)
    assert not any(th.is_alive() for th in threads), "a thread is stuck inside the library"
    assert not errors, errors[:10]
    for h in list(Bh.values()) + [big_Ah]:
        L.mzd_free(h)
