"""Randomised stress of the host ABI against the CPU oracle (a tool, not a test: run it on a GPU box
for as long as you like). Every iteration draws a shape with mass on the word / tile / small-path
boundaries and checks one product entry point (fresh and accumulating) and one consumer
(echelon form, PLE or PLUQ identity, inverse, solve). usage: python tools/fuzz_abi.py [seconds] [seed] [largest dimension]"""
import os, random, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import m4ri_rust_b200 as m4
import oracle as orc
L = m4.lib()
budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
rng = random.Random(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
HI = int(sys.argv[3]) if len(sys.argv) > 3 else 2600
EDGES = [1, 2, 3, 63, 64, 65, 127, 128, 129, 255, 256, 257, 511, 512, 513, 1023, 1024, 1025, 1447, 1448, 1449, 2047, 2048, 2049]

def dim(hi=None):
    hi = hi or HI
    r = rng.random()
    if r < 0.35:
        return rng.choice(EDGES)
    if r < 0.6:
        return rng.randint(1, 130)
    return rng.randint(1, hi)

def words(p, n):
    return m4.mzd_words(p)[:, : orc.width(n)]

t_end, it, fails = time.time() + budget, 0, 0
while time.time() < t_end:
    it += 1
    m, k, n = dim(), dim(), dim()
    seed = rng.randint(0, 2 ** 31)
    A, B, C0 = orc.random(m, k, seed), orc.random(k, n, seed + 1), orc.random(m, n, seed + 2)
    want = orc.mul_m4rm(A, B, k, n, threads=4)
    a, b = m4.mzd_from_words(A, k), m4.mzd_from_words(B, n)
    name = rng.choice(["mzd_mul", "mzd_mul_m4rm", "mzd_mul_naive", "_mzd_mul_naive", "mzd_mul_mp"])
    try:
        if name == "_mzd_mul_naive":
            bt = L.mzd_transpose(None, b)
            c = L._mzd_mul_naive(None, a, bt, 1)
            c2 = m4.mzd_from_words(C0, n)
            L._mzd_mul_naive(c2, a, bt, 0)
            L.mzd_free(bt)
        elif name in ("mzd_mul", "mzd_mul_m4rm", "mzd_mul_mp"):
            c = getattr(L, name)(None, a, b, 0)
            c2 = m4.mzd_from_words(C0, n)
            getattr(L, name.replace("mul", "addmul", 1))(c2, a, b, 0)
        else:
            c = L.mzd_mul_naive(None, a, b)
            c2 = m4.mzd_from_words(C0, n)
            L.mzd_addmul_naive(c2, a, b)
        w = orc.width(n)
        assert (words(c, n) == want[:, :w]).all(), "product"
        assert (words(c2, n) == (C0 ^ want)[:, :w]).all(), "accumulate"
        assert not m4.mzd_words(c)[:, w:].any(), "padding"
        for q in (c, c2):
            L.mzd_free(q)
        # one consumer on a matrix of random rank
        em, en = dim(1800), dim(1800)
        rk = rng.choice([0, 1, min(em, en) // 3, min(em, en)])
        E = orc.random(em, en, seed + 3) if rk == min(em, en) else (
            orc.zeros(em, en) if rk == 0 else orc.mul_m4rm(orc.random(em, rk, seed + 4), orc.random(rk, en, seed + 5), rk, en))
        want_e, rank, piv = orc.rref(E, en)
        what = rng.choice(["echelonize", "ple", "pluq", "solve"])
        e = m4.mzd_from_words(E, en)
        ew = orc.width(en)
        if what == "echelonize":
            assert L.mzd_echelonize(e, 1) == rank and (words(e, en) == want_e[:, :ew]).all(), "rref"
        elif what in ("ple", "pluq"):
            P, Q = L.mzp_init(em), L.mzp_init(en)
            r = (L.mzd_pluq if what == "pluq" else L.mzd_ple)(e, P, Q, 0)
            Pv = [int(P.contents.values[i]) for i in range(em)]
            Qv = [int(Q.contents.values[i]) for i in range(en)]
            assert r == rank and Qv[:r] == [int(x) for x in piv], "rank / pivot columns"
            back = orc.ple_reconstruct(np.ascontiguousarray(m4.mzd_words(e)), en, r, Pv, Qv, what == "pluq")
            assert (back[:, :ew] == E[:, :ew]).all(), "P L E != A"
            L.mzp_free(P); L.mzp_free(Q)
        else:
            kk = rng.randint(1, 130)
            X0 = orc.random(en, kk, seed + 6)
            Bc = orc.mul_m4rm(E, X0, en, kk)
            rows = max(em, en)
            Bf = orc.zeros(rows, kk)
            Bf[:em] = Bc
            bb = m4.mzd_from_words(Bf, kk)
            assert L.mzd_solve_left(e, bb, 0, 1) == 0, "consistent system reported inconsistent"
            X = np.ascontiguousarray(m4.mzd_words(bb)[:en])
            assert (orc.mul_m4rm(E, X, en, kk)[:, : orc.width(kk)] == Bc[:, : orc.width(kk)]).all(), "A X != B"
            L.mzd_free(bb)
        L.mzd_free(e)
    except AssertionError as ex:
        fails += 1
        print(f"FAIL it={it} {name} {m}x{k}x{n} seed={seed}: {ex}", flush=True)
    L.mzd_free(a); L.mzd_free(b)
print(f"fuzz_abi: {it} iterations, {fails} failures")
sys.exit(1 if fails else 0)
