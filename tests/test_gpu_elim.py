"""GPU parity of the consumers of multiplication behind the reference's names —
mzd_echelonize / _pluq / _m4ri (m4ri-sys/src/echelonform.rs:16,27,36), mzd_inv_m4ri
(brilliantrussian.rs:208), mzd_solve_left (solve.rs:23) and their BinMatrix wrappers
(binary_matrix.rs:246-268, 575-586) — against the CPU oracle's Gauss-Jordan. The reduced row
echelon form is unique, so the comparison is bit for bit."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def host(m4, words, ncols):
    return m4.mzd_from_words(words, ncols)


def low_rank(orc, m, n, r, seed):
    """m x n matrix of rank <= r: product of random m x r and r x n factors."""
    return orc.mul_m4rm(orc.random(m, r, seed), orc.random(r, n, seed + 1), r, n)


SHAPES = [(1, 1), (5, 7), (64, 64), (65, 63), (100, 300), (300, 100), (1000, 1000), (1030, 2050), (2050, 1030),
          (1500, 1500), (3000, 2500), (70, 5000), (5000, 70)]


@pytest.mark.parametrize("entry", ["mzd_echelonize", "mzd_echelonize_pluq", "mzd_echelonize_m4ri"])
def test_echelonize_matches_the_oracle(gpu, m4, orc, entry):
    L = m4.lib()
    for (m, n) in SHAPES:
        for kind in ("random", "lowrank", "zero_cols"):
            if kind == "random":
                A = orc.random(m, n, 11 + m)
            elif kind == "lowrank":
                A = low_rank(orc, m, n, max(1, min(m, n) // 3), 5 + n)
            else:  # pivots far apart: zero column blocks and duplicated rows
                A = orc.random(m, n, 7)
                A[:, : orc.width(n) // 2] = 0
                A[m // 2:] = A[: m - m // 2]
            want, r_want, piv = orc.rref(A, n)
            a = host(m4, A, n)
            r = L.mzd_echelonize_m4ri(a, 1, 0) if entry == "mzd_echelonize_m4ri" else getattr(L, entry)(a, 1)
            assert r == r_want, (entry, m, n, kind, r, r_want)
            got = m4.mzd_words(a)
            assert not got[:, orc.width(n):].any()
            assert (got[:, : orc.width(n)] == want[:, : orc.width(n)]).all(), (entry, m, n, kind)
            L.mzd_free(a)


def test_rank_and_echelonize_wrappers(gpu, m4, orc):
    # BinMatrix::rank / echelonize (binary_matrix.rs:246-261)
    I = m4.BinMatrix.identity(10)
    assert I.rank() == 10 and I.echelonize() == 10 and I == m4.BinMatrix.identity(10)
    A = low_rank(orc, 500, 700, 123, 3)
    M = m4.BinMatrix(host(m4, A, 700))
    before = M.clone()
    assert M.rank() == orc.rref(A, 700)[1]
    assert M == before  # rank() works on a clone


def test_block_widths_agree(gpu, m4, orc, monkeypatch):
    """The block width is a performance knob: every width gives the same (unique) form — run in
    child processes because the width is read once."""
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = ("import sys; sys.path.insert(0, %r); import numpy as np, m4ri_rust_b200 as m4, oracle as orc; L = m4.lib();"
            "ok = True\n"
            "for (m, n, r) in [(1500, 2100, 0), (2000, 2000, 700), (900, 3000, 0)]:\n"
            "    A = orc.random(m, n, 3) if r == 0 else orc.mul_m4rm(orc.random(m, r, 1), orc.random(r, n, 2), r, n)\n"
            "    want, rk, piv = orc.rref(A, n); a = m4.mzd_from_words(A, n)\n"
            "    ok &= L.mzd_echelonize(a, 1) == rk and bool((m4.mzd_words(a)[:, : orc.width(n)] == want[:, : orc.width(n)]).all())\n"
            "print('ok' if ok else 'MISMATCH')") % root
    for width in ("128", "256", "512", "2048", "4096"):
        env = dict(os.environ, M4RI_B200_ELIM_BLOCK=width)
        p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=600)
        assert p.returncode == 0 and p.stdout.strip().endswith("ok"), (width, p.stdout[-500:], p.stderr[-2000:])


def test_inverse(gpu, m4, orc):
    L = m4.lib()
    for n in (1, 2, 10, 64, 65, 200, 1000, 1025, 2048, 3000):
        # random matrices over GF(2) are invertible with probability ~0.29: try seeds until one is
        for seed in range(40):
            A = orc.random(n, n, 100 * n + seed)
            want = orc.inverse(A, n)
            a = host(m4, A, n)
            p = L.mzd_inv_m4ri(None, a, 0)
            if want is None:
                assert not p, (n, seed)
                L.mzd_free(a)
                continue
            assert p, (n, seed)
            got = m4.mzd_words(p)
            assert not got[:, orc.width(n):].any()
            assert (got[:, : orc.width(n)] == want[:, : orc.width(n)]).all(), (n, seed)
            # A * A^-1 == I through the GPU product, and inverted() of the mirror
            prod = L.mzd_mul(None, a, p, 0)
            eye = L.mzd_init(n, n)
            L.mzd_set_ui(eye, 1)
            assert L.mzd_equal(prod, eye) == 1
            # preallocated destination
            d = L.mzd_init(n, n)
            L.mzd_randomize(d)
            assert L.mzd_inv_m4ri(d, a, 0) and L.mzd_equal(d, p) == 1
            for q in (a, p, prod, eye, d):
                L.mzd_free(q)
            break
        else:
            pytest.fail(f"no invertible matrix found at n={n}")
    with pytest.raises(ValueError):
        m4.BinMatrix.zero(5, 5).inverted()
    assert m4.BinMatrix.identity(70).inverted() == m4.BinMatrix.identity(70)


def test_solve_left(gpu, m4, orc):
    L = m4.lib()
    for (m, n, k) in [(10, 10, 1), (64, 64, 64), (200, 200, 70), (1000, 1000, 300), (1500, 1000, 10), (700, 1000, 129),
                      (2100, 2100, 1)]:
        A = orc.random(m, n, 3 + m)
        X0 = orc.random(n, k, 4 + k)
        Bc = orc.mul_m4rm(A, X0, n, k)  # consistent by construction
        want = orc.solve_left(A, n, Bc, k)
        assert want is not None
        rows = max(m, n)
        Bfull = np.zeros((rows, Bc.shape[1]), dtype=np.uint64)
        Bfull[:m] = Bc
        a, b = host(m4, A, n), host(m4, Bfull, k)
        assert L.mzd_solve_left(a, b, 0, 1) == 0
        got = m4.mzd_words(b)
        assert (got[:n, : orc.width(k)] == want[:, : orc.width(k)]).all(), (m, n, k)
        assert not got[n:].any() and not got[:, orc.width(k):].any()
        assert (m4.mzd_words(a)[:, : orc.width(n)] == A[:, : orc.width(n)]).all()  # A untouched
        # A * X == B
        x = host(m4, got[:n].copy(), k)
        ax = L.mzd_mul(None, a, x, 0)
        assert (m4.mzd_words(ax)[:, : orc.width(k)] == Bc[:, : orc.width(k)]).all()
        for q in (a, b, x, ax):
            L.mzd_free(q)
    # inconsistent system: rank-deficient A, right-hand side outside its column space
    m = n = 300
    A = low_rank(orc, m, n, 100, 9)
    Bbad = orc.random(m, 5, 77)
    assert orc.solve_left(A, n, Bbad, 5) is None
    a, b = host(m4, A, n), host(m4, Bbad, 5)
    assert L.mzd_solve_left(a, b, 0, 1) == -1
    assert (m4.mzd_words(b)[:, :1] == Bbad[:, :1]).all()  # untouched
    # the mirror of friendly::solve_left (binary_matrix.rs:575-586)
    Am, Bm = m4.BinMatrix.identity(20), m4.BinMatrix.random(20, 3)
    want = Bm.clone()
    assert m4.solve_left(Am, Bm) and Bm == want
    for q in (a, b):
        L.mzd_free(q)


def test_device_rref_large(gpu, m4):
    """n = 8192 device-resident: A^-1 from [A | I], checked by A * A^-1 == I with the product
    kernels (size-independent property; the oracle's cubic loop stops at a few thousand)."""
    import ctypes
    L = m4.lib()
    n = 8192
    for seed in range(1, 30):
        W = m4.DMat.init(n, 2 * n)
        left = m4.DMat.wrap(W.ptr, n, n, W.rowstride, keepalive=W)
        right = m4.DMat.wrap(W.ptr + (n // 64) * 8, n, n, W.rowstride, keepalive=W)
        left.fill_random(seed)
        A = m4.DMat.init(n, n)
        L.m4ri_b200_dcopy(A.h, left.h)
        heye = m4.BinMatrix.identity(n)  # (kept alive: the upload reads its rows[])
        eye = m4.DMat.from_host(heye.mzd)
        L.m4ri_b200_synchronize()
        L.m4ri_b200_dcopy(right.h, eye.h)
        piv = (ctypes.c_int * n)()
        r = L.m4ri_b200_drref(W.h, n, piv, n)
        if r < n:
            continue
        assert list(piv) == list(range(n))
        L.m4ri_b200_dadd(left.h, left.h, eye.h)
        assert L.m4ri_b200_dis_zero(left.h) == 1  # left half became the identity
        P = m4.DMat.init(n, n)
        m4.dmul(P, A, right)
        L.m4ri_b200_dadd(P.h, P.h, eye.h)
        assert L.m4ri_b200_dis_zero(P.h) == 1
        return
    pytest.fail("no invertible 8192 x 8192 matrix among the seeds")


# ---- PLE / PLUQ (m4ri-sys/src/ple.rs:35,65) and the permutation calls (mzp.rs) -----------------
def _perm_values(P):
    return [int(P.contents.values[i]) for i in range(P.contents.length)]


def _ple_cases(orc):
    for (m, n) in SHAPES + [(4100, 700), (700, 4100), (2500, 2500)]:
        for kind in ("random", "lowrank", "zero_cols"):
            if kind == "random":
                A = orc.random(m, n, 21 + m)
            elif kind == "lowrank":
                A = low_rank(orc, m, n, max(1, min(m, n) // 3), 9 + n)
            else:
                A = orc.random(m, n, 13)
                A[:, : orc.width(n) // 2] = 0
                A[m // 2:] = A[: m - m // 2]
            yield m, n, kind, A


@pytest.mark.parametrize("entry", ["mzd_ple", "mzd_pluq"])
def test_ple_pluq_defining_identity(gpu, m4, orc, entry):
    """The decomposition is not unique (the pivot ROW is a free choice), so the check is its
    defining identity through the storage contract of include/m4ri_b200.h: unpack L and E (or U),
    verify their shapes, and P * L * E must be the input bit for bit; the rank and the pivot
    columns Q (the column rank profile, which IS unique) must equal the oracle's."""
    L = m4.lib()
    pluq = entry == "mzd_pluq"
    for m, n, kind, A in _ple_cases(orc):
        w = orc.width(n)
        _, r_want, Pw, Qw = orc.ple(A, n, pluq)
        a = host(m4, A, n)
        P, Q = L.mzp_init(m), L.mzp_init(n)
        r = getattr(L, entry)(a, P, Q, 0)
        assert r == r_want, (entry, m, n, kind, r, r_want)
        Pv, Qv = _perm_values(P), _perm_values(Q)
        assert Qv == [int(q) for q in Qw], (entry, m, n, kind)
        got = m4.mzd_words(a)
        assert not got[:, w:].any()
        back = orc.ple_reconstruct(np.ascontiguousarray(got), n, r, Pv, Qv, pluq)
        assert (back[:, :w] == A[:, :w]).all(), (entry, m, n, kind)
        # same input, same answer (pivot rows are chosen deterministically)
        a2 = host(m4, A, n)
        P2, Q2 = L.mzp_init(m), L.mzp_init(n)
        assert getattr(L, entry)(a2, P2, Q2, 0) == r and _perm_values(P2) == Pv
        assert L.mzd_equal(a, a2) == 1
        for q in (P, Q, P2, Q2):
            L.mzp_free(q)
        L.mzd_free(a)
        L.mzd_free(a2)


def test_ple_through_the_products(gpu, m4, orc):
    """mzd_apply_p_left(A, P) == L * E with the product done by mzd_mul on the GPU; full rank
    square input, where PLE is the LU decomposition with row pivoting."""
    L = m4.lib()
    n = 1536
    for seed in range(40):
        A = orc.random(n, n, 500 + seed)
        if orc.rref(A, n)[1] == n:
            break
    a, lu = host(m4, A, n), host(m4, A, n)
    P, Q = L.mzp_init(n), L.mzp_init(n)
    assert L.mzd_ple(lu, P, Q, 0) == n and _perm_values(Q) == list(range(n))
    packed = orc.to_bits(np.ascontiguousarray(m4.mzd_words(lu)), n)
    lower = np.tril(packed, -1) + np.eye(n, dtype=np.uint8)
    upper = np.triu(packed)
    l, u = host(m4, orc.from_bits(lower), n), host(m4, orc.from_bits(upper), n)
    prod = L.mzd_mul(None, l, u, 0)
    L.mzd_apply_p_left(a, P)
    assert L.mzd_equal(prod, a) == 1
    L.mzd_apply_p_left_trans(a, P)  # and back
    assert (m4.mzd_words(a)[:, : orc.width(n)] == A[:, : orc.width(n)]).all()
    for q in (a, lu, l, u, prod):
        L.mzd_free(q)
    L.mzp_free(P)
    L.mzp_free(Q)


def test_apply_permutations(gpu, m4, orc):
    L = m4.lib()
    rng = np.random.default_rng(5)
    for (m, n) in [(1, 1), (7, 5), (64, 130), (300, 257), (1000, 70)]:
        A = orc.random(m, n, 31 + n)
        bits = orc.to_bits(A, n)
        for which, length in (("left", m), ("left_trans", m), ("right", n), ("right_trans", n)):
            P = L.mzp_init(length)
            vals = [int(rng.integers(i, length)) for i in range(length)]
            for i, v in enumerate(vals):
                P.contents.values[i] = v
            want = bits.copy()
            order = range(length) if which in ("left", "right_trans") else reversed(range(length))
            for i in order:
                if which.startswith("left"):
                    want[[i, vals[i]]] = want[[vals[i], i]]
                else:
                    want[:, [i, vals[i]]] = want[:, [vals[i], i]]
            a = host(m4, A, n)
            getattr(L, "mzd_apply_p_" + which)(a, P)
            got = orc.to_bits(np.ascontiguousarray(m4.mzd_words(a)), n)
            assert (got == want).all(), (which, m, n)
            L.mzd_free(a)
            L.mzp_free(P)
    # windows and copies of permutations (host bookkeeping)
    P = L.mzp_init(10)
    assert _perm_values(P) == list(range(10))
    W = L.mzp_init_window(P, 3, 8)
    assert W.contents.length == 5
    W.contents.values[0] = 7
    assert P.contents.values[3] == 7
    C = L.mzp_copy(None, P)
    assert _perm_values(C) == _perm_values(P)
    L.mzp_set_ui(P, 1)
    assert _perm_values(P) == list(range(10))
    L.Mzp_free_window(W)
    L.mzp_free(C)
    L.mzp_free(P)


def test_m4ri_sys_aliases(gpu, m4, orc):
    """mzd_mul_mp / mzd_addmul_mp (mp.rs:6,8), mzd_invert_naive (mzd.rs:218) and
    mzd_top_echelonize_m4ri (brilliantrussian.rs:199) are the GPU calls behind other names."""
    L = m4.lib()
    m, k, n = 700, 1300, 900
    A, B, C0 = orc.random(m, k, 1), orc.random(k, n, 2), orc.random(m, n, 3)
    want = orc.mul_m4rm(A, B, k, n)
    a, b = host(m4, A, k), host(m4, B, n)
    c = L.mzd_mul_mp(None, a, b, 0)
    assert (m4.mzd_words(c)[:, : orc.width(n)] == want[:, : orc.width(n)]).all()
    c0 = host(m4, C0, n)
    L.mzd_addmul_mp(c0, a, b, 0)
    assert (m4.mzd_words(c0)[:, : orc.width(n)] == (C0 ^ want)[:, : orc.width(n)]).all()
    for q in (a, b, c, c0):
        L.mzd_free(q)
    nn = 300
    for seed in range(40):
        S = orc.random(nn, nn, 900 + seed)
        inv = orc.inverse(S, nn)
        if inv is not None:
            break
    s_ = host(m4, S, nn)
    p = L.mzd_invert_naive(None, s_, None)
    assert p and (m4.mzd_words(p)[:, : orc.width(nn)] == inv[:, : orc.width(nn)]).all()
    E = low_rank(orc, 500, 800, 200, 11)
    want_e, rank, _ = orc.rref(E, 800)
    e = host(m4, E, 800)
    assert L.mzd_top_echelonize_m4ri(e, 0) == rank
    assert (m4.mzd_words(e)[:, : orc.width(800)] == want_e[:, : orc.width(800)]).all()
    for q in (s_, p, e):
        L.mzd_free(q)


def test_pluq_solve_left(gpu, m4, orc):
    """mzd_pluq_solve_left (solve.rs:53): solve from the packed factors mzd_pluq leaves; every
    solution is verified by substituting it back with the oracle's product."""
    L = m4.lib()
    cases = [(10, 10, 3, None), (64, 64, 64, None), (300, 300, 70, None), (700, 500, 33, None), (500, 700, 129, None),
             (900, 900, 10, 400), (1300, 1100, 65, 500), (1100, 1300, 200, 300), (2100, 2100, 1, None)]
    for (m, n, k, rk) in cases:
        A = orc.random(m, n, 7 + m) if rk is None else low_rank(orc, m, n, rk, 3 + n)
        X0 = orc.random(n, k, 5 + k)
        Bc = orc.mul_m4rm(A, X0, n, k)  # consistent by construction
        rows = max(m, n)
        Bfull = orc.zeros(rows, k)
        Bfull[:m] = Bc
        lu = host(m4, A, n)
        P, Q = L.mzp_init(m), L.mzp_init(n)
        r = L.mzd_pluq(lu, P, Q, 0)
        assert r == orc.rref(A, n)[1]
        b = host(m4, Bfull, k)
        assert L.mzd_pluq_solve_left(lu, r, P, Q, b, 0, 1) == 0, (m, n, k, rk)
        X = np.ascontiguousarray(m4.mzd_words(b))
        assert not X[n:].any(), "rows beyond ncols(A) are zeroed"
        back = orc.mul_m4rm(A, np.ascontiguousarray(X[:n]), n, k)
        assert (back[:, : orc.width(k)] == Bc[:, : orc.width(k)]).all(), (m, n, k, rk)
        # the same answer as mzd_solve_left gives when the solution is unique
        if r == n:
            assert (X[:n, : orc.width(k)] == X0[:, : orc.width(k)]).all()
        L.mzd_free(b)
        # an inconsistent right-hand side (rank-deficient rows): -1 and B untouched
        if r < m:
            Bbad = orc.random(rows, k, 99)
            Bbad[m:] = 0
            if orc.solve_left(A, n, np.ascontiguousarray(Bbad[:m]), k) is None:
                bb = host(m4, Bbad, k)
                assert L.mzd_pluq_solve_left(lu, r, P, Q, bb, 0, 1) == -1
                assert (m4.mzd_words(bb)[:, : orc.width(k)] == Bbad[:, : orc.width(k)]).all()
                L.mzd_free(bb)
        L.mzd_free(lu)
        L.mzp_free(P)
        L.mzp_free(Q)
