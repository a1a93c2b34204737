"""Hypothesis-generated (m, k, n) sweep of every product entry point of the C-ABI against the
CPU oracle (SURVEY 8c "hypothesis-generated (m,k,n) incl. non-multiples of 64, width odd >= 3
(stride padding), m = 1, n = 1"). Bit-exact, padding and excess bits included."""
import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

pytestmark = pytest.mark.gpu

# dimensions biased towards the edges: 1, word boundaries, odd widths >= 3 (129..192, 257..320)
dim = st.one_of(st.integers(1, 8), st.sampled_from([63, 64, 65, 127, 128, 129, 191, 192, 193, 255, 256, 257, 320, 321]),
                st.integers(1, 700), st.integers(129, 192), st.integers(257, 320), st.integers(700, 2600))
COMMON = dict(deadline=None, suppress_health_check=list(HealthCheck), derandomize=True)


def words_ok(m4, p, want, n, orc):
    w = m4.mzd_words(p)
    width = orc.width(n)
    assert not w[:, width:].any(), "padding words must be zero"
    assert (w[:, :width] == want[:, :width]).all()


@settings(max_examples=60, **COMMON)
@given(m=dim, k=dim, n=dim, seed=st.integers(0, 2 ** 31))
def test_every_entry_point_on_generated_shapes(gpu, m4, orc, m, k, n, seed):
    L = m4.lib()
    A, B, C0 = orc.random(m, k, seed), orc.random(k, n, seed + 1), orc.random(m, n, seed + 2)
    want = orc.mul_m4rm(A, B, k, n, threads=4)
    acc = C0 ^ want
    a, b = m4.mzd_from_words(A, k), m4.mzd_from_words(B, n)
    bt = L.mzd_transpose(None, b)
    assert (m4.mzd_words(bt)[:, : orc.width(k)] == orc.transpose(B, n)[:, : orc.width(k)]).all()
    fresh = {
        "mzd_mul": lambda: L.mzd_mul(None, a, b, 0),
        "mzd_mul_m4rm": lambda: L.mzd_mul_m4rm(None, a, b, 0),
        "_mzd_mul_m4rm": lambda: L._mzd_mul_m4rm(None, a, b, 0, 1),
        "mzd_mul_naive": lambda: L.mzd_mul_naive(None, a, b),
        "_mzd_mul_naive": lambda: L._mzd_mul_naive(None, a, bt, 1),
    }
    for name, f in fresh.items():
        p = f()
        try:
            words_ok(m4, p, want, n, orc)
        except AssertionError as e:
            raise AssertionError(f"{name} {m}x{k}x{n}: {e}")
        L.mzd_free(p)
    accum = {
        "mzd_addmul": lambda c: L.mzd_addmul(c, a, b, 0),
        "mzd_addmul_m4rm": lambda c: L.mzd_addmul_m4rm(c, a, b, 0),
        "_mzd_mul_m4rm(0)": lambda c: L._mzd_mul_m4rm(c, a, b, 0, 0),
        "mzd_addmul_naive": lambda c: L.mzd_addmul_naive(c, a, b),
        "_mzd_mul_naive(0)": lambda c: L._mzd_mul_naive(c, a, bt, 0),
    }
    for name, f in accum.items():
        c = m4.mzd_from_words(C0, n)
        f(c)
        try:
            words_ok(m4, c, acc, n, orc)
        except AssertionError as e:
            raise AssertionError(f"{name} {m}x{k}x{n}: {e}")
        L.mzd_free(c)
    if m <= 4:
        p = L._mzd_mul_va(None, a, b, 1)
        words_ok(m4, p, want, n, orc)
        L.mzd_free(p)
    if n == 1:
        p = L.mzd_mul_v(None, a, b)
        words_ok(m4, p, want, n, orc)
        L.mzd_free(p)
    for p in (a, b, bt):
        L.mzd_free(p)


@settings(max_examples=40, **COMMON)
@given(m=dim, k=dim, n=dim, algo=st.sampled_from(["auto", "m4rm", "mma", "simple"]), seed=st.integers(0, 2 ** 31))
def test_device_products_on_generated_shapes(gpu, m4, orc, m, k, n, algo, seed):
    L = m4.lib()
    A, B = orc.random(m, k, seed), orc.random(k, n, seed + 1)
    want = orc.mul_m4rm(A, B, k, n, threads=4)[:, : orc.width(n)]
    a, b = m4.mzd_from_words(A, k), m4.mzd_from_words(B, n)
    dA, dB, dC = m4.DMat.from_host(a), m4.DMat.from_host(b), m4.DMat.init(m, n).fill_random(5)
    m4.dmul(dC, dA, dB, clear=True, algo=m4.ALGO_NAMES[algo])
    got = dC.to_words()
    assert not got[:, orc.width(n):].any()
    assert (got[:, : orc.width(n)] == want).all(), (algo, m, k, n)
    dBt = m4.DMat.init(n, k)
    L.m4ri_b200_dtranspose(dBt.h, dB.h)
    m4.dmul_abt(dC, dA, dBt, clear=False, algo=m4.ALGO_NAMES[algo])
    assert not dC.to_words().any(), (algo, m, k, n)
    L.mzd_free(a)
    L.mzd_free(b)


@settings(max_examples=40, **COMMON)
@given(m=st.one_of(dim, st.integers(2600, 70000)), n=dim, seed=st.integers(0, 2 ** 31))
def test_matvec_on_generated_shapes(gpu, m4, orc, m, n, seed):
    L = m4.lib()
    A, x = orc.random(m, n, seed), orc.random(1, n, seed + 1)
    want = orc.mul_v(A, x[0], n)
    a, hx = m4.mzd_from_words(A, n), m4.mzd_from_words(x, n)
    dA, dx, dy = m4.DMat.from_host(a), m4.DMat.from_host(hx), m4.DMat.init(1, m).fill_random(3)
    L.m4ri_b200_dmul_v(dy.h, dA.h, dx.h, 1)
    got = dy.to_words()
    assert (got[0, : orc.width(m)] == want).all() and not got[0, orc.width(m):].any(), (m, n)
    col = m4.BinMatrix(a).mul_slice(x[0][: orc.width(n)])
    assert (m4.mzd_words(col.mzd)[:, 0] == [(int(want[i // 64]) >> (i % 64)) & 1 for i in range(m)]).all()
    L.mzd_free(hx)


# ---- the consumers: generated shapes and rank structures for PLE / PLUQ / echelon form ----------
edim = st.one_of(st.integers(1, 8), st.sampled_from([63, 64, 65, 127, 128, 129, 1023, 1024, 1025]),
                 st.integers(1, 400), st.integers(400, 1500))


@settings(max_examples=40, **COMMON)
@given(m=edim, n=edim, rank_frac=st.sampled_from([0.0, 0.1, 0.5, 1.0]), seed=st.integers(0, 2 ** 31),
       pluq=st.booleans())
def test_ple_and_echelon_form_on_generated_shapes(gpu, m4, orc, m, n, rank_frac, seed, pluq):
    """mzd_ple / mzd_pluq (ple.rs:65,35) satisfy their defining identity and agree with the oracle
    on rank and pivot columns; mzd_echelonize gives the oracle's reduced form of the same matrix."""
    L = m4.lib()
    r_target = int(round(rank_frac * min(m, n)))
    if rank_frac >= 1.0:
        A = orc.random(m, n, seed)
    elif r_target == 0:
        A = orc.zeros(m, n)
    else:
        A = orc.mul_m4rm(orc.random(m, r_target, seed), orc.random(r_target, n, seed + 1), r_target, n)
    w = orc.width(n)
    want, rank, piv = orc.rref(A, n)
    a = m4.mzd_from_words(A, n)
    P, Q = L.mzp_init(m), L.mzp_init(n)
    r = (L.mzd_pluq if pluq else L.mzd_ple)(a, P, Q, 0)
    assert r == rank, (m, n, r, rank)
    Pv = [int(P.contents.values[i]) for i in range(m)]
    Qv = [int(Q.contents.values[i]) for i in range(n)]
    assert Qv[:r] == [int(c) for c in piv]
    got = np.ascontiguousarray(m4.mzd_words(a))
    assert not got[:, w:].any()
    back = orc.ple_reconstruct(got, n, r, Pv, Qv, pluq)
    assert (back[:, :w] == A[:, :w]).all(), (m, n, rank_frac, pluq)
    L.mzp_free(P)
    L.mzp_free(Q)
    L.mzd_free(a)
    e = m4.mzd_from_words(A, n)
    assert L.mzd_echelonize(e, 1) == rank
    assert (m4.mzd_words(e)[:, :w] == want[:, :w]).all()
    L.mzd_free(e)
