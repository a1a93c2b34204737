"""GPU parity of the mat-vec surface (BinMatrix::mul_slice, Mul<&BinVector> for
&BinMatrix, Mul<&BinMatrix> for &BinVector; BASELINE configs[4] "batched mzd_mul_v") against the
CPU oracle, and proof that the reference's real calls reach the HBM-bound kernels: mul_slice is
mzd_mul_naive(NULL, A, x_column) (binary_matrix.rs:426-427) and must launch mul_v_*, v * A must
launch mul_va — not a product tile."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def host(m4, words, ncols):
    return m4.mzd_from_words(words, ncols)


def kernel_used(L, call):
    """Name of the (last) tagged kernel `call` launched and the launches it made."""
    L.m4ri_b200_timing_reset()
    before = L.m4ri_b200_launch_count()
    out = call()
    L.m4ri_b200_synchronize()
    n = L.m4ri_b200_timing_count()
    return out, (L.m4ri_b200_timing_name().decode() if n else ""), L.m4ri_b200_launch_count() - before


def test_mul_slice_reaches_the_matvec_kernels(gpu, m4, orc):
    L = m4.lib()
    for (m, n) in [(1 << 16, 512), (4096, 4096), (34, 128), (5000, 8192), (100, 129)]:
        A, x = orc.random(m, n, 3), orc.random(1, n, 4)
        want = orc.mul_v(A, x[0], n)
        M = m4.BinMatrix(host(m4, A, n))
        for strategy in ("strassen_mul", "m4rm_mul", "naive_mul"):  # mul_slice ignores the feature flag
            m4.BinMatrix.mul_strategy = strategy
            col, name, launches = kernel_used(L, lambda: M.mul_slice(x[0][: orc.width(n)]))
            assert name.startswith("mul_v"), (m, n, name)
            assert launches <= 5, (m, n, launches)  # transpose of the x row (mask + kernel), mask of A, mul_v
            bits = m4.mzd_words(col.mzd)[:, 0]
            assert (bits == [(int(want[i // 64]) >> (i % 64)) & 1 for i in range(m)]).all()
        m4.BinMatrix.mul_strategy = "strassen_mul"
        # every product entry point given the n x 1 column takes the same route
        xcol = M.from_slices([x[0][: orc.width(n)]], n).transposed()
        for f in (lambda: L.mzd_mul(None, M.mzd, xcol.mzd, 0), lambda: L.mzd_mul_m4rm(None, M.mzd, xcol.mzd, 0),
                  lambda: L.mzd_mul_naive(None, M.mzd, xcol.mzd), lambda: L.mzd_mul_v(None, M.mzd, xcol.mzd)):
            p, name, _ = kernel_used(L, f)
            assert name.startswith("mul_v"), name
            assert (m4.mzd_words(p)[:, 0] == [(int(want[i // 64]) >> (i % 64)) & 1 for i in range(m)]).all()
            L.mzd_free(p)
        # accumulate form: y ^= A*x twice restores y
        y0 = orc.random(m, 1, 9)
        y = host(m4, y0, 1)
        L.mzd_addmul_naive(y, M.mzd, xcol.mzd)
        once = m4.mzd_words(y)[:, 0].copy()
        assert (once == (y0[:, 0] ^ np.array([(int(want[i // 64]) >> (i % 64)) & 1 for i in range(m)], dtype=np.uint64))).all()
        L.mzd_addmul_naive(y, M.mzd, xcol.mzd)
        assert (m4.mzd_words(y)[:, 0] == y0[:, 0]).all()
        # pre-transposed form with a one-row Bt (= the packed vector itself)
        xr = host(m4, x, n)
        p, name, _ = kernel_used(L, lambda: L._mzd_mul_naive(None, M.mzd, xr, 1))
        assert name.startswith("mul_v"), name
        assert (m4.mzd_words(p)[:, 0] == [(int(want[i // 64]) >> (i % 64)) & 1 for i in range(m)]).all()
        for q in (p, xr, y):
            L.mzd_free(q)


def test_vector_times_matrix_reaches_mul_va(gpu, m4, orc):
    L = m4.lib()
    for (k, n) in [(10, 3), (5000, 8192), (8192, 5000), (700, 20000)]:
        v, A = orc.random(1, k, 5), orc.random(k, n, 6)
        want = orc.mul_va(v, A, k, n)[:, : orc.width(n)]
        a, hv = host(m4, A, n), host(m4, v, k)
        for f in (lambda: L.mzd_mul(None, hv, a, 0), lambda: L.mzd_mul_m4rm(None, hv, a, 0),
                  lambda: L.mzd_mul_naive(None, hv, a)):
            p, name, _ = kernel_used(L, f)
            assert name == "mul_va_kernel", (k, n, name)
            assert (m4.mzd_words(p)[:, : orc.width(n)] == want).all()
            L.mzd_free(p)
        # three rows of v, device API, AUTO
        v3 = orc.random(3, k, 7)
        h3 = host(m4, v3, k)
        dV, dA, dC = m4.DMat.from_host(h3), m4.DMat.from_host(a), m4.DMat.init(3, n)
        _, name, _ = kernel_used(L, lambda: m4.dmul(dC, dV, dA))
        assert name == "mul_va_kernel"
        assert (dC.to_words()[:, : orc.width(n)] == orc.mul_va(v3, A, k, n)[:, : orc.width(n)]).all()
        for q in (a, hv, h3):
            L.mzd_free(q)


BATCH_SHAPES = [(1 << 16, 512), (70000, 512), (4097, 128), (100003, 64), (33, 256), (5000, 1024),
                (4100, 4096), (2049, 5000), (300, 8192), (65, 200), (1000, 960), (64, 64)]


@pytest.mark.parametrize("b", [1, 2, 3, 4, 5, 8, 9, 16, 31, 32, 40])
def test_batched_matvec(gpu, m4, orc, b):
    """Y = X * A^T: row q of Y = A * x_q, for every row-length class of the kernels (coalesced
    short rows, generic rows), ragged m, clear and accumulate; b > 32 takes the tensor path."""
    L = m4.lib()
    for (m, n) in BATCH_SHAPES:
        A, X = orc.random(m, n, 11), orc.random(b, n, 12 + b)
        want = np.stack([orc.mul_v(A, X[q], n) for q in range(b)])
        a, hx = host(m4, A, n), host(m4, X, n)
        dA, dX, dY = m4.DMat.from_host(a), m4.DMat.from_host(hx), m4.DMat.init(b, m).fill_random(99)
        L.m4ri_b200_dmul_v_batched(dY.h, dA.h, dX.h, 1)
        got = dY.to_words()
        assert (got[:, : orc.width(m)] == want).all(), (b, m, n)
        assert not got[:, orc.width(m):].any()
        L.m4ri_b200_dmul_v_batched(dY.h, dA.h, dX.h, 0)
        assert not dY.to_words().any(), (b, m, n)
        L.mzd_free(a)
        L.mzd_free(hx)


def test_thin_device_products_take_the_hbm_bound_route(gpu, m4, orc):
    """AUTO on device operands: a few rows of A -> mul_va; a few columns of B, or a few rows of A
    against a pre-transposed Bt -> the batched mat-vec (one pass over the big operand)."""
    L = m4.lib()
    m, k = 50000, 2048
    A = orc.random(m, k, 21)
    a = host(m4, A, k)
    dA = m4.DMat.from_host(a)
    for n in (2, 7, 32):
        B = orc.random(k, n, 22 + n)
        want = orc.mul_m4rm(A, B, k, n, threads=0)[:, : orc.width(n)]
        hb = host(m4, B, n)
        dB, dC = m4.DMat.from_host(hb), m4.DMat.init(m, n)
        _, name, _ = kernel_used(L, lambda: m4.dmul(dC, dA, dB))
        assert name in ("mul_vb_kernel", "mul_v_coal_kernel", "mul_v_long_kernel", "mul_v_kernel"), name
        got = dC.to_words()
        assert (got[:, : orc.width(n)] == want).all() and not got[:, orc.width(n):].any()
        m4.dmul(dC, dA, dB, clear=False)
        assert not dC.to_words().any()
        # host ABI, same shape
        p = L.mzd_mul(None, a, hb, 0)
        assert (m4.mzd_words(p)[:, : orc.width(n)] == want).all()
        L.mzd_free(p)
        L.mzd_free(hb)
    # few rows against Bt: C (b x m) = X * A^T through dmul_abt
    X = orc.random(6, k, 31)
    hx = host(m4, X, k)
    dX, dC = m4.DMat.from_host(hx), m4.DMat.init(6, m)
    _, name, _ = kernel_used(L, lambda: m4.dmul_abt(dC, dX, dA))
    assert name == "mul_vb_kernel"
    want = np.stack([orc.mul_v(A, X[q], k) for q in range(6)])
    assert (dC.to_words()[:, : orc.width(m)] == want).all()
    L.mzd_free(a)
    L.mzd_free(hx)
