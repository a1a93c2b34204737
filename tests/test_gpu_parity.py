"""GPU parity tests: the CUDA path, called through the C-ABI of libm4ri_b200.so
(the m4ri-sys-shaped mzd_* entry points and the BinMatrix/BinVector mirrors),
against the CPU oracle on identical seeded inputs — bit-exact, including excess
bits and padding words. Reads neither /root/reference nor anything else outside
the repo."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")

ALGOS = ["m4rm", "mma", "simple"]


def need(m4, algo):
    if not m4.lib().m4ri_b200_algo_available(m4.ALGO_NAMES[algo]):
        pytest.fail(f"algorithm {algo} is not built into libm4ri_b200.so")


def host(m4, words, ncols):
    return m4.mzd_from_words(words, ncols)


def result_words(m4, p, free=True):
    """[nrows, width] valid words + assert padding is zero."""
    w = m4.mzd_words(p).copy()
    width = p.contents.width
    assert not w[:, width:].any(), "padding words must be zero"
    if p.contents.ncols % 64 and width:
        assert not (w[:, width - 1] & ~np.uint64(p.contents.high_bitmask)).any(), "excess bits must be zero"
    if free:
        m4.lib().mzd_free(p)
    return w[:, :width]


def entry_points(m4):
    L = m4.lib()
    return {
        "mzd_mul": lambda a, b: L.mzd_mul(None, a, b, 0),
        "mzd_mul_m4rm": lambda a, b: L.mzd_mul_m4rm(None, a, b, 0),
        "_mzd_mul_m4rm": lambda a, b: L._mzd_mul_m4rm(None, a, b, 0, 1),
        "mzd_mul_naive": lambda a, b: L.mzd_mul_naive(None, a, b),
    }


# ---- the reference's own pinned answers, replayed through the mirrors -----------
def test_reference_known_answers(gpu, m4):
    cases = json.load(open(os.path.join(GOLD, "reference_known_answers.json")))["cases"]
    for strategy in ("strassen_mul", "m4rm_mul", "naive_mul"):  # mul_impl!, binary_matrix.rs:53-95
        m4.BinMatrix.mul_strategy = strategy
        for c in cases:
            if c["op"] == "mul":  # binary_matrix.rs:661-670
                m1, m2 = m4.BinMatrix.identity(8), m4.BinMatrix.identity(8)
                prod = m1 * m2
                assert prod == m4.BinMatrix.identity(8)
            elif c["op"] == "matvec":  # binary_matrix.rs:677-678
                result = m4.BinMatrix.identity(10) * m4.BinVector.from_elem(10, True)
                assert result == m4.BinVector.from_elem(10, True)
            elif c["op"] == "vecmat":  # binary_matrix.rs:680-681
                result = m4.BinVector.from_elem(10, True) * m4.BinMatrix.identity(10)
                assert result == m4.BinVector.from_elem(10, True)
            elif c["op"] == "vecmat_shape":  # binary_matrix.rs:683-685
                result = m4.BinVector.from_elem(10, True) * m4.BinMatrix.random(10, 3)
                assert len(result) == 3
    m4.BinMatrix.mul_strategy = "strassen_mul"


def test_as_vector_and_transposed(gpu, m4):  # binary_matrix.rs:701-719, binary_vector.rs:262-278
    v = m4.BinVector.from_function(10, lambda i: i % 2 == 1)
    assert v.as_matrix().as_vector() == v
    col = v.as_column_matrix()
    assert (col.nrows(), col.ncols()) == (10, 1)
    assert col.as_vector() == v
    a = m4.BinMatrix.random(70, 200)
    assert a.transposed().transposed() == a
    b = a.clone()
    b += a
    assert b == m4.BinMatrix.zero(70, 200)
    assert (a + a) == m4.BinMatrix.zero(70, 200)


# ---- golden products (committed fixtures) ---------------------------------------
def test_golden_products_all_entry_points(gpu, m4, orc):
    g = np.load(os.path.join(GOLD, "oracle_products.npz"))
    L = m4.lib()
    n_cases = sum(1 for k in g.files if k.startswith("shape_"))
    for idx in range(n_cases):
        m, k, n = (int(x) for x in g[f"shape_{idx}"])
        A = orc.random(m, k, 0x4D345249 ^ (idx << 32) ^ 1)
        B = orc.random(k, n, 0x4D345249 ^ (idx << 32) ^ 2)
        C = g[f"C_{idx}"][:, : orc.width(n)]
        a, b = host(m4, A, k), host(m4, B, n)
        for name, f in entry_points(m4).items():
            got = result_words(m4, f(a, b))
            assert (got == C).all(), f"{name} differs on golden case {idx} ({m}x{k}x{n})"
        bt = L.mzd_transpose(None, b)
        assert (result_words(m4, L._mzd_mul_naive(None, a, bt, 1)) == C).all()
        if m <= 4:
            assert (result_words(m4, L._mzd_mul_va(None, a, b, 1)) == C).all()
        if n == 1:
            assert (result_words(m4, L.mzd_mul_v(None, a, b)) == C).all()
        for p in (a, b, bt):
            L.mzd_free(p)


# ---- shape sweep, every algorithm, device API -----------------------------------
SWEEP = [(1, 1, 1), (1, 64, 1), (2, 63, 65), (7, 129, 3), (65, 64, 64), (64, 65, 64), (64, 64, 65),
         (33, 31, 200), (191, 193, 129), (200, 257, 321), (5, 700, 1100), (1030, 40, 1030),
         (1024, 1024, 1024), (257, 2050, 1500), (128, 128, 256), (256, 4096, 128), (1500, 96, 70)]


@pytest.mark.parametrize("algo", ALGOS)
def test_shape_sweep(gpu, m4, orc, algo):
    need(m4, algo)
    L = m4.lib()
    for (m, k, n) in SWEEP:
        A, B = orc.random(m, k, 77 + m), orc.random(k, n, 99 + n)
        want = orc.mul_m4rm(A, B, k, n, threads=4)[:, : orc.width(n)]
        a, b = host(m4, A, k), host(m4, B, n)
        dA, dB = m4.DMat.from_host(a), m4.DMat.from_host(b)
        dC = m4.DMat.init(m, n)
        m4.dmul(dC, dA, dB, clear=True, algo=m4.ALGO_NAMES[algo])
        got = dC.to_words()
        assert not got[:, orc.width(n):].any()
        assert (got[:, : orc.width(n)] == want).all(), f"{algo} differs at {m}x{k}x{n}"
        # accumulate: C ^= A*B twice restores a random C
        C0 = orc.random(m, n, 5)
        c0 = host(m4, C0, n)
        dC2 = m4.DMat.from_host(c0)
        m4.dmul(dC2, dA, dB, clear=False, algo=m4.ALGO_NAMES[algo])
        once = dC2.to_words()[:, : orc.width(n)]
        assert (once == (C0[:, : orc.width(n)] ^ want)).all(), f"{algo} addmul differs at {m}x{k}x{n}"
        m4.dmul(dC2, dA, dB, clear=False, algo=m4.ALGO_NAMES[algo])
        assert (dC2.to_words()[:, : orc.width(n)] == C0[:, : orc.width(n)]).all()
        for p in (a, b, c0):
            L.mzd_free(p)


def test_addmul_entry_points(gpu, m4, orc):
    L = m4.lib()
    m, k, n = 300, 500, 700
    A, B, C0 = orc.random(m, k, 1), orc.random(k, n, 2), orc.random(m, n, 3)
    want = (C0 ^ orc.mul_m4rm(A, B, k, n))[:, : orc.width(n)]
    a, b = host(m4, A, k), host(m4, B, n)
    bt = L.mzd_transpose(None, b)
    calls = {
        "mzd_addmul": lambda c: L.mzd_addmul(c, a, b, 0),
        "mzd_addmul_m4rm": lambda c: L.mzd_addmul_m4rm(c, a, b, 0),
        "_mzd_mul_m4rm(clear=0)": lambda c: L._mzd_mul_m4rm(c, a, b, 0, 0),
        "mzd_addmul_naive": lambda c: L.mzd_addmul_naive(c, a, b),
        "_mzd_mul_naive(clear=0)": lambda c: L._mzd_mul_naive(c, a, bt, 0),
    }
    for name, f in calls.items():
        c = host(m4, C0, n)
        r = f(c)
        assert r and (m4.mzd_words(r)[:, : orc.width(n)] == want).all(), name
        L.mzd_free(c)
    # _mzd_mul_va accumulate, 3-row v
    v = orc.random(3, k, 9)
    hv = host(m4, v, k)
    c = host(m4, C0[:3].copy(), n)
    L._mzd_mul_va(c, hv, b, 0)
    assert (m4.mzd_words(c)[:, : orc.width(n)] == (C0[:3] ^ orc.mul_va(v, B, k, n))[:, : orc.width(n)]).all()
    for p in (a, b, bt, hv, c):
        L.mzd_free(p)


def test_window_operands(gpu, m4, orc):
    """Inputs that are windows (rows[] not one contiguous slab) are read through rows[]."""
    L = m4.lib()
    P = orc.random(300, 640, 4)
    p = host(m4, P, 640)
    w = L.mzd_init_window(p, 10, 128, 210, 448)  # 200 x 320
    A = np.ascontiguousarray(P[10:210, 2:7])
    B = orc.random(320, 100, 5)
    b = host(m4, B, 100)
    want = orc.mul_m4rm(A, B, 320, 100)[:, :2]
    assert (result_words(m4, L.mzd_mul(None, w, b, 0)) == want).all()
    for q in (w, p, b):
        L.mzd_free(q)


def test_transpose_add_device_and_host(gpu, m4, orc):
    L = m4.lib()
    for (r, c) in [(1, 1), (10, 1), (1, 10), (65, 129), (300, 1000), (1000, 64), (1030, 2050)]:
        S, S2 = orc.random(r, c, 1), orc.random(r, c, 2)
        s, s2 = host(m4, S, c), host(m4, S2, c)
        assert (result_words(m4, L.mzd_transpose(None, s)) == orc.transpose(S, c)[:, : orc.width(r)]).all()
        assert (result_words(m4, L.mzd_add(None, s, s2)) == orc.add(S, S2, c)[:, : orc.width(c)]).all()
        L.mzd_add(s, s, s2)  # C aliases A (binary_matrix.rs:497-523)
        assert (result_words(m4, s, free=False) == orc.add(S, S2, c)[:, : orc.width(c)]).all()
        L.mzd_free(s)
        L.mzd_free(s2)


@pytest.mark.parametrize("variant", ["narrow", "wide"])
def test_transpose_kernels_and_scatter(gpu, m4, orc, monkeypatch, variant):
    """Both transpose kernels (256 x 1024 tiles with 32-byte output segments, 1024 x 1024 tiles with
    128-byte segments) on ragged shapes, a stale destination, and the multi-destination scatter
    that the fused multi-GPU exchange uses (here: two buffers on one GPU)."""
    import ctypes
    L = m4.lib()
    monkeypatch.setenv("M4RI_B200_TRANSPOSE", variant)
    for (r, c) in [(1, 1), (65, 129), (300, 1000), (1000, 64), (1030, 2050), (2048, 4096), (5000, 1111),
                   (1025, 8192), (4096, 130)]:
        S = orc.random(r, c, 7)
        want = orc.transpose(S, c)[:, : orc.width(r)]
        s = host(m4, S, c)
        dS = m4.DMat.from_host(s)
        d1, d2 = m4.DMat.init(c, r).fill_random(9), m4.DMat.init(c, r).fill_random(10)  # stale contents
        L.m4ri_b200_dtranspose(d1.h, dS.h)
        w1 = d1.to_words()
        assert (w1[:, : orc.width(r)] == want).all() and not w1[:, orc.width(r):].any(), (r, c)
        d1.fill_random(11)
        arr = (ctypes.c_void_p * 2)(d1.h, d2.h)
        L.m4ri_b200_dtranspose_scatter(arr, 2, dS.h)
        for d in (d1, d2):
            w = d.to_words()
            assert (w[:, : orc.width(r)] == want).all() and not w[:, orc.width(r):].any(), (r, c)
        L.mzd_free(s)


def test_matvec_shapes(gpu, m4, orc):
    L = m4.lib()
    for (m, n) in [(34, 128), (10, 10), (64, 1000), (1000, 64), (1, 1), (4097, 777), (300, 8192),
                   (20, 512), (33, 64), (100, 129), (70000, 512), (65, 1024), (4096, 960),
                   (1000, 256), (100001, 128), (12345, 1024), (31, 200), (8191, 70), (2049, 500),
                   (4100, 4096), (2049, 5000), (3000, 2112), (2048, 32768), (2081, 4097)]:
        A = orc.random(m, n, 3)
        x = orc.random(1, n, 4)
        a = host(m4, A, n)
        want = orc.mul_v(A, x[0], n)
        # mul_slice path (binary_matrix.rs:416-431): m x 1 matrix, one word per row
        M = m4.BinMatrix(a)
        col = M.mul_slice(x[0][: orc.width(n)])
        assert (col.nrows(), col.ncols()) == (m, 1)
        bits = m4.mzd_words(col.mzd)[:, 0]
        assert [int(b) for b in bits] == [int(want[i // 64] >> np.uint64(i % 64)) & 1 for i in range(m)]
        # packed in / packed out device API
        dA = m4.DMat.from_host(a)
        hx = host(m4, x, n)
        dx = m4.DMat.from_host(hx)
        dy = m4.DMat.init(1, m)
        L.m4ri_b200_dmul_v(dy.h, dA.h, dx.h, 1)
        assert (dy.to_words()[0, : orc.width(m)] == want).all()
        L.m4ri_b200_dmul_v(dy.h, dA.h, dx.h, 0)  # accumulate: y ^= A*x restores zero
        assert not dy.to_words().any()
        L.mzd_free(hx)


def test_matvec_short_row_kernels_agree(gpu, m4, orc, monkeypatch):
    """Rows of 1/2/4/8 16-byte vectors: the coalesced warp-per-32-rows kernel (default) and the
    thread-per-row kernel (M4RI_B200_MULV=rows) against the oracle, ragged m, clear and accumulate."""
    L = m4.lib()
    for n in (64, 128, 200, 256, 512, 1000, 1024):
        for m in (1, 31, 32, 33, 63, 64, 65, 255, 4097, 100003):
            A = orc.random(m, n, 5)
            x = orc.random(1, n, 6)
            want = orc.mul_v(A, x[0], n)
            a, hx = host(m4, A, n), host(m4, x, n)
            dA, dx = m4.DMat.from_host(a), m4.DMat.from_host(hx)
            for mode in ("", "rows"):
                monkeypatch.setenv("M4RI_B200_MULV", mode)
                dy = m4.DMat.init(1, m)
                L.m4ri_b200_dmul_v(dy.h, dA.h, dx.h, 1)
                assert (dy.to_words()[0, : orc.width(m)] == want).all(), (m, n, mode)
                L.m4ri_b200_dmul_v(dy.h, dA.h, dx.h, 0)
                assert not dy.to_words().any(), (m, n, mode)
            L.mzd_free(hx)
            L.mzd_free(a)


def test_device_generator_and_checksum_match_oracle(gpu, m4, orc):
    for (r, c) in [(5, 70), (100, 64), (257, 1000)]:
        d = m4.DMat.init(r, c).fill_random(0xABCD + r)
        want = orc.random(r, c, 0xABCD + r)
        assert (d.to_words()[:, : orc.width(c)] == want[:, : orc.width(c)]).all()
        assert d.checksum() == orc.checksum(want, c)


# ---- BASELINE.json sizes ---------------------------------------------------------
@pytest.mark.parametrize("algo", ["m4rm", "mma"])
def test_n8192_bit_exact(gpu, m4, orc, algo):
    """configs[1]: square n=8192, full compare against the oracle."""
    need(m4, algo)
    n = 8192
    A, B = orc.random(n, n, 0x4D345249 ^ (1 << 32) ^ 1), orc.random(n, n, 0x4D345249 ^ (1 << 32) ^ 2)
    want = orc.mul_m4rm(A, B, n, n, threads=0)
    dA = m4.DMat.init(n, n).fill_random(0x4D345249 ^ (1 << 32) ^ 1)
    dB = m4.DMat.init(n, n).fill_random(0x4D345249 ^ (1 << 32) ^ 2)
    dC = m4.DMat.init(n, n)
    m4.dmul(dC, dA, dB, clear=True, algo=m4.ALGO_NAMES[algo])
    assert dC.checksum() == orc.checksum(want, n)
    assert (dC.to_words() == want).all()
    # and through the host ABI
    L = m4.lib()
    a, b = host(m4, A, n), host(m4, B, n)
    m4.lib().m4ri_b200_set_algo(m4.ALGO_NAMES[algo])
    try:
        c = L.mzd_mul(None, a, b, 0)
    finally:
        m4.lib().m4ri_b200_set_algo(0)
    assert (m4.mzd_words(c) == want).all()
    for p in (a, b, c):
        L.mzd_free(p)


def test_n32768_properties(gpu, m4):
    """configs[2]: n=32768 on one B200 — size-independent properties: Freivalds
    (false accept 2^-64), the two independent kernels agree bit for bit,
    A*(B^B2) == A*B ^ A*B2, addmul twice restores C."""
    n = 32768
    L = m4.lib()
    dA = m4.DMat.init(n, n).fill_random(11)
    dB = m4.DMat.init(n, n).fill_random(12)
    dC = m4.DMat.init(n, n)
    m4.dmul(dC, dA, dB, algo=m4.ALGO_M4RM)
    assert L.m4ri_b200_dfreivalds(dC.h, dA.h, dB.h, 0xF5E1) == 1
    chk = dC.checksum()
    dC2 = m4.DMat.init(n, n)
    m4.dmul(dC2, dA, dB, algo=m4.ALGO_MMA)
    assert dC2.checksum() == chk
    L.m4ri_b200_dadd(dC2.h, dC2.h, dC.h)
    assert L.m4ri_b200_dis_zero(dC2.h) == 1
    # a corrupted product is rejected
    m4.dmul(dC2, dA, dB, algo=m4.ALGO_MMA)
    L.m4ri_b200_dflip_bit(dC2.h, 12345, 23456)
    assert L.m4ri_b200_dfreivalds(dC2.h, dA.h, dB.h, 0xF5E1) == 0
    L.m4ri_b200_dflip_bit(dC2.h, 12345, 23456)
    # linearity
    dB2 = m4.DMat.init(n, n).fill_random(13)
    m4.dmul(dC2, dA, dB2, clear=False, algo=m4.ALGO_AUTO)  # C2 = A*B ^ A*B2
    L.m4ri_b200_dadd(dB2.h, dB2.h, dB.h)
    m4.dmul(dC, dA, dB2, algo=m4.ALGO_AUTO)               # C = A*(B^B2)
    assert dC.checksum() == dC2.checksum()


def test_lpn_shape(gpu, m4, orc):
    """configs[4]: tall 2^20 x 512 times 512 x 512, full compare on a 2^16-row
    slice per algorithm and Freivalds on the whole product."""
    m, k, n = 1 << 20, 512, 512
    L = m4.lib()
    dA = m4.DMat.init(m, k).fill_random(21)
    dB = m4.DMat.init(k, n).fill_random(22)
    dC = m4.DMat.init(m, n)
    rows = 1 << 16
    A = orc.random(rows, k, 21)  # word (r,w) = stream index r*width+w: a row prefix is a prefix
    B = orc.random(k, n, 22)
    want = orc.mul_m4rm(A, B, k, n, threads=0)
    for algo in (m4.ALGO_M4RM, m4.ALGO_MMA, m4.ALGO_AUTO):
        m4.dmul(dC, dA, dB, algo=algo)
        assert L.m4ri_b200_dfreivalds(dC.h, dA.h, dB.h, 7) == 1
        head = m4.DMat.wrap(dC.ptr, rows, n, dC.rowstride, keepalive=dC)
        assert (head.to_words() == want).all()


def test_pipelined_host_products(gpu, m4, orc, monkeypatch):
    """Host-ABI products large enough for the copy/compute-overlapped path (api.cu,
    pipelined_host_product): ragged shapes, accumulate, an operand that is a window of
    a bigger matrix (rows[] with a foreign stride), and the pre-transposed form."""
    L = m4.lib()
    for (m, k, n) in [(5000, 20000, 9000), (2049, 30001, 4099), (9000, 16384, 8192)]:
        A, B, C0 = orc.random(m, k, 21), orc.random(k, n, 22), orc.random(m, n, 23)
        want = orc.mul_m4rm(A, B, k, n, threads=0)[:, : orc.width(n)]
        a, b = host(m4, A, k), host(m4, B, n)
        assert (result_words(m4, L.mzd_mul(None, a, b, 0)) == want).all(), (m, k, n)
        c = host(m4, C0, n)
        L.mzd_addmul(c, a, b, 0)
        assert (result_words(m4, c, free=False) == (C0[:, : orc.width(n)] ^ want)).all(), (m, k, n)
        L.mzd_addmul(c, a, b, 0)
        assert (result_words(m4, c) == C0[:, : orc.width(n)]).all()
        # A*Bt^T with the tensor path forced (the *_naive entry maps to the simple kernel otherwise)
        bt = L.mzd_transpose(None, b)
        L.m4ri_b200_set_algo(m4.ALGO_MMA)
        got = result_words(m4, L._mzd_mul_naive(None, a, bt, 1))
        L.m4ri_b200_set_algo(m4.ALGO_AUTO)
        assert (got == want).all(), (m, k, n)
        for p in (a, b, bt):
            L.mzd_free(p)
    # other block plans of the pipeline (rows of A x column blocks of B), ragged blocks
    m, k, n = 3000, 30000, 5003
    A, B, C0 = orc.random(m, k, 41), orc.random(k, n, 42), orc.random(m, n, 43)
    want = orc.mul_m4rm(A, B, k, n, threads=0)[:, : orc.width(n)]
    a, b = host(m4, A, k), host(m4, B, n)
    for rows_, cols_ in [(1, 1), (3, 5), (12, 2), (2, 20)]:
        monkeypatch.setenv("M4RI_B200_PIPE_ROWS", str(rows_))
        monkeypatch.setenv("M4RI_B200_PIPE_COLS", str(cols_))
        assert (result_words(m4, L.mzd_mul(None, a, b, 0)) == want).all(), (rows_, cols_)
        c = host(m4, C0, n)
        L.mzd_addmul(c, a, b, 0)
        assert (result_words(m4, c) == (C0[:, : orc.width(n)] ^ want)).all(), (rows_, cols_)
    monkeypatch.delenv("M4RI_B200_PIPE_ROWS")
    monkeypatch.delenv("M4RI_B200_PIPE_COLS")
    for q in (a, b):
        L.mzd_free(q)
    # window operand: 4000 x 16384 window at column 128 of a 4100 x 20000 parent
    P = orc.random(4100, 20000, 31)
    p = host(m4, P, 20000)
    w = L.mzd_init_window(p, 50, 128, 4050, 128 + 16384)
    A = np.ascontiguousarray(P[50:4050, 2:2 + 256])
    B = orc.random(16384, 3000, 32)
    b = host(m4, B, 3000)
    want = orc.mul_m4rm(A, B, 16384, 3000, threads=0)[:, : orc.width(3000)]
    assert (result_words(m4, L.mzd_mul(None, w, b, 0)) == want).all()
    for q in (w, p, b):
        L.mzd_free(q)


VARIANT_SWEEP = [(1, 1, 1), (2, 63, 65), (65, 64, 64), (191, 193, 129), (200, 257, 321), (5, 700, 1100),
                 (1030, 40, 1030), (257, 2050, 1500), (256, 4096, 128), (300, 512, 300), (129, 1536, 257)]


@pytest.mark.parametrize("variant", ["v1", "v2cg1", "v2", "fp4cg1", "fp4", "fp4ldgcg1", "fp4ldg", "fp4s256cg1", "fp4s256"])
def test_every_tensor_kernel_variant(gpu, m4, orc, variant):
    """Each kernel behind ALGO_MMA (int8 SS, int8 TS single/pair, FP4 with TMA staging
    single/pair, FP4 with direct global loads single/pair) is bit-exact on ragged shapes,
    accumulates, handles contractions longer than one 65536-bit chunk and Strassen windows."""
    L = m4.lib()
    L.m4ri_b200_set_mma_variant(m4.MMA_VARIANTS[variant])
    try:
        for (m, k, n) in VARIANT_SWEEP + [(70, 70000, 90)]:
            A, B = orc.random(m, k, 3 + m), orc.random(k, n, 4 + n)
            want = orc.mul_m4rm(A, B, k, n, threads=0)[:, : orc.width(n)]
            a, b = host(m4, A, k), host(m4, B, n)
            dA, dB, dC = m4.DMat.from_host(a), m4.DMat.from_host(b), m4.DMat.init(m, n)
            m4.dmul(dC, dA, dB, clear=True, algo=m4.ALGO_MMA)
            got = dC.to_words()
            assert not got[:, orc.width(n):].any()
            assert (got[:, : orc.width(n)] == want).all(), f"{variant} differs at {m}x{k}x{n}"
            m4.dmul(dC, dA, dB, clear=False, algo=m4.ALGO_MMA)
            assert not dC.to_words().any(), f"{variant} addmul differs at {m}x{k}x{n}"
            L.mzd_free(a)
            L.mzd_free(b)
        # one Strassen level on (pointer, stride) windows: operands whose rows continue past the window
        L.m4ri_b200_set_strassen_levels(1)
        m = k = n = 1024
        A, B = orc.random(m, k, 8), orc.random(k, n, 9)
        want = orc.mul_m4rm(A, B, k, n, threads=0)[:, : orc.width(n)]
        a, b = host(m4, A, k), host(m4, B, n)
        dA, dB, dC = m4.DMat.from_host(a), m4.DMat.from_host(b), m4.DMat.init(m, n)
        m4.dmul(dC, dA, dB, clear=True, algo=m4.ALGO_MMA)
        assert (dC.to_words()[:, : orc.width(n)] == want).all(), f"{variant} with one Strassen level"
        L.mzd_free(a)
        L.mzd_free(b)
    finally:
        L.m4ri_b200_set_strassen_levels(-1)
        L.m4ri_b200_set_mma_variant(0)


@pytest.mark.parametrize("variant", ["fp4s256", "fp4s256cg1"])
def test_fused_strassen_level(gpu, m4, orc, variant, monkeypatch):
    """The last Strassen level of the default kernel is fused (classic Strassen: operand sums in
    two launches, every product written or accumulated into one or two quadrants of C by the
    kernel's epilogue): bit-exact against the oracle for C = and C ^=, on rectangular shapes, in
    both operand forms, with two levels (Winograd outside, fused inside), and equal to the
    unfused schedule."""
    L = m4.lib()
    L.m4ri_b200_set_mma_variant(m4.MMA_VARIANTS[variant])
    try:
        for levels, (m, k, n) in [(1, (1024, 1024, 1024)), (1, (512, 1536, 2048)), (1, (2048, 512, 1024)),
                                  (2, (2048, 1024, 3072)), (1, (1024, 66560, 512))]:
            L.m4ri_b200_set_strassen_levels(levels)
            A, B, C0 = orc.random(m, k, 31 + m), orc.random(k, n, 32 + n), orc.random(m, n, 33)
            want = orc.mul_m4rm(A, B, k, n, threads=0)[:, : orc.width(n)]
            a, b = host(m4, A, k), host(m4, B, n)
            dA, dB, dC = m4.DMat.from_host(a), m4.DMat.from_host(b), m4.DMat.init(m, n)
            l0 = m4.launch_count()
            m4.dmul(dC, dA, dB, clear=True, algo=m4.ALGO_MMA)
            fused_launches = m4.launch_count() - l0
            assert (dC.to_words()[:, : orc.width(n)] == want).all(), (variant, levels, m, k, n)
            # accumulate form on the pre-transposed operand: C0 ^= A * Bt^T
            dBt = m4.DMat.init(n, k)
            L.m4ri_b200_dtranspose(dBt.h, dB.h)
            c0 = host(m4, C0, n)
            dC0 = m4.DMat.from_host(c0)
            m4.dmul_abt(dC0, dA, dBt, clear=False, algo=m4.ALGO_MMA)
            assert (dC0.to_words()[:, : orc.width(n)] == (C0[:, : orc.width(n)] ^ want)).all(), (variant, levels, m, k, n)
            # seven separate launches with two-destination epilogues instead of the one grouped launch
            monkeypatch.setenv("M4RI_B200_NO_GROUPED_STRASSEN", "1")
            m4.dmul(dC, dA, dB, clear=True, algo=m4.ALGO_MMA)
            monkeypatch.delenv("M4RI_B200_NO_GROUPED_STRASSEN")
            assert (dC.to_words()[:, : orc.width(n)] == want).all(), (variant, levels, m, k, n)
            # the unfused (Winograd) schedule gives the same words with more launches
            monkeypatch.setenv("M4RI_B200_NO_FUSED_STRASSEN", "1")
            l0 = m4.launch_count()
            m4.dmul(dC, dA, dB, clear=True, algo=m4.ALGO_MMA)
            unfused_launches = m4.launch_count() - l0
            monkeypatch.delenv("M4RI_B200_NO_FUSED_STRASSEN")
            assert (dC.to_words()[:, : orc.width(n)] == want).all()
            assert fused_launches < unfused_launches, (fused_launches, unfused_launches)
            for p in (a, b, c0):
                L.mzd_free(p)
        # operands and result as (pointer, stride) windows of larger parents: rows continue past
        # the window on both sides, the destination is zeroed and accumulated into in place
        L.m4ri_b200_set_strassen_levels(1)
        m, k, n = 1024, 1536, 512
        PA, PB, PC = m4.DMat.init(m + 7, k + 640).fill_random(41), m4.DMat.init(n + 5, k + 896).fill_random(42), \
            m4.DMat.init(m + 3, n + 1152).fill_random(43)
        wA = m4.DMat.wrap(PA.ptr + 3 * PA.rowstride * 8 + 2 * 16, m, k, PA.rowstride, keepalive=PA)
        wBt = m4.DMat.wrap(PB.ptr + 2 * PB.rowstride * 8 + 4 * 16, n, k, PB.rowstride, keepalive=PB)
        wC = m4.DMat.wrap(PC.ptr + 1 * PC.rowstride * 8 + 6 * 16, m, n, PC.rowstride, keepalive=PC)
        before = PC.to_words()
        A = PA.to_words()[3: 3 + m, 4: 4 + k // 64].copy()
        Bt = PB.to_words()[2: 2 + n, 8: 8 + k // 64].copy()
        want = orc.mul_m4rm(A, orc.transpose(Bt, k), k, n, threads=0)[:, : n // 64]
        m4.dmul_abt(wC, wA, wBt, clear=True, algo=m4.ALGO_MMA)
        after = PC.to_words()
        assert (after[1: 1 + m, 12: 12 + n // 64] == want).all()
        after[1: 1 + m, 12: 12 + n // 64] = before[1: 1 + m, 12: 12 + n // 64]
        assert (after == before).all(), "the product wrote outside its window"
    finally:
        L.m4ri_b200_set_strassen_levels(-1)
        L.m4ri_b200_set_mma_variant(0)


def test_fp4_accumulators_are_exact(gpu, m4, orc):
    """The FP4 path's fp32 accumulators equal popcount(A_i & Bt_j)/4 exactly, also in the worst
    case the wrapper allows (all ones, k = 65536): no rounding anywhere in the tensor core."""
    import torch
    L = m4.lib()
    for cg in (1, 2):
        for (m, k, n, ones) in [(200, 1000, 300, False), (256, 65536, 256, True), (128, 20000, 130, False)]:
            if ones:
                A = np.full((m, orc.width(k)), np.uint64(0xFFFFFFFFFFFFFFFF))
                B = np.full((n, orc.width(k)), np.uint64(0xFFFFFFFFFFFFFFFF))
            else:
                A, B = orc.random(m, k, 5 + m)[:, : orc.width(k)].copy(), orc.random(n, k, 7 + n)[:, : orc.width(k)].copy()
            a, b = host(m4, A, k), host(m4, B, k)
            dA, dB, dC = m4.DMat.from_host(a), m4.DMat.from_host(b), m4.DMat.init(m, n)
            acc = torch.full((m, n), -1.0, dtype=torch.float32, device="cuda")
            torch.cuda.synchronize()
            L.m4ri_b200_dmul_abt_fp4_debug(dC.h, dA.h, dB.h, acc.data_ptr(), cg)
            L.m4ri_b200_synchronize()
            a8 = np.unpackbits(m4.mzd_words(a)[:, : orc.width(k)].copy().view(np.uint8), axis=1, bitorder="little").astype(np.float32)
            b8 = np.unpackbits(m4.mzd_words(b)[:, : orc.width(k)].copy().view(np.uint8), axis=1, bitorder="little").astype(np.float32)
            want = (a8 @ b8.T) / 4.0
            assert (acc.cpu().numpy() == want).all(), (cg, m, k, n)
            L.mzd_free(a)
            L.mzd_free(b)


@pytest.mark.parametrize("peer", [True, False])
def test_multi_device_host_products(gpu, m4, orc, monkeypatch, peer):
    """m4ri_b200_set_devices(G): ONE host-ABI call spread over the GPUs of the process, for every G
    from 2 to the number of visible GPUs (skips with one). Default plan: GPU g uploads only its
    k-slab of B, the slabs of Bt are exchanged with NVLink peer stores; peer=False forces the
    independent plan (every GPU uploads all of B). The result must be the single-GPU words, for
    the plain, accumulate and pre-transposed forms on ragged shapes."""
    L = m4.lib()
    ndev = L.m4ri_b200_device_count()
    if ndev < 2:
        pytest.skip("one GPU visible")
    if not peer:
        import subprocess, sys
        # the peer decision is cached per process: run the independent plan in a child
        code = ("import os, sys; os.environ['M4RI_B200_NO_PEER'] = '1'; sys.path.insert(0, %r);"
                "import numpy as np, m4ri_rust_b200 as m4, oracle as orc; L = m4.lib();"
                "m, k, n = 3000, 20000, 9000; A, B = orc.random(m, k, 21), orc.random(k, n, 22);"
                "want = orc.mul_m4rm(A, B, k, n, threads=0)[:, : orc.width(n)];"
                "a, b = m4.mzd_from_words(A, k), m4.mzd_from_words(B, n); L.m4ri_b200_set_devices(2);"
                "c = L.mzd_mul(None, a, b, 0); assert (m4.mzd_words(c)[:, : orc.width(n)] == want).all(); print('ok')"
                ) % os.path.dirname(HERE)
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
        assert r.returncode == 0 and "ok" in r.stdout, r.stderr[-2000:]
        return
    try:
        for G in sorted(set([2, ndev])):
            L.m4ri_b200_set_devices(G)
            assert L.m4ri_b200_get_devices() == G
            for (m, k, n) in [(5000, 20000, 9000), (2049, 30001, 4099), (600, 16384, 8192)]:
                A, B, C0 = orc.random(m, k, 21), orc.random(k, n, 22), orc.random(m, n, 23)
                want = orc.mul_m4rm(A, B, k, n, threads=0)[:, : orc.width(n)]
                a, b = host(m4, A, k), host(m4, B, n)
                assert (result_words(m4, L.mzd_mul(None, a, b, 0)) == want).all(), (G, m, k, n)
                c = host(m4, C0, n)
                L.mzd_addmul(c, a, b, 0)
                assert (result_words(m4, c) == (C0[:, : orc.width(n)] ^ want)).all(), (G, m, k, n)
                bt = L.mzd_transpose(None, b)
                L.m4ri_b200_set_algo(m4.ALGO_MMA)
                got = result_words(m4, L._mzd_mul_naive(None, a, bt, 1))
                L.m4ri_b200_set_algo(m4.ALGO_AUTO)
                assert (got == want).all(), (G, m, k, n)
                for p in (a, b, bt):
                    L.mzd_free(p)
        # one big square product: every GPU's rows equal the single-GPU result
        n = 16384
        dA, dB = m4.DMat.init(n, n).fill_random(5), m4.DMat.init(n, n).fill_random(6)
        a, b = dA.to_host(), dB.to_host()
        L.m4ri_b200_set_devices(ndev)
        cG = L.mzd_mul(None, a, b, 0)
        L.m4ri_b200_set_devices(1)
        c1 = L.mzd_mul(None, a, b, 0)
        assert L.mzd_equal(cG, c1) == 1
        for p in (a, b, cG, c1):
            L.mzd_free(p)
    finally:
        L.m4ri_b200_set_algo(m4.ALGO_AUTO)
        L.m4ri_b200_set_devices(1)


def test_set_devices_clamps_to_the_visible_gpus(gpu, m4):
    L = m4.lib()
    try:
        L.m4ri_b200_set_devices(64)
        assert L.m4ri_b200_get_devices() == L.m4ri_b200_device_count()
    finally:
        L.m4ri_b200_set_devices(1)


def test_fp4_accumulators_exact_on_dense_operands(gpu, m4, orc):
    """the FP4 path is only correct if fp32 accumulation of 0.25-steps is exact up to
    the largest sum the wrapper allows (k = 65536 -> 16384.0, 17 significant bits). All-ones
    operands only ever add 16.0 per MMA and 50 % random ones stay below 4200: here near-dense
    operands (all ones with random holes at densities 0.9 ... 0.999, and rows whose per-word
    popcounts are odd) put the accumulators between 4197 and 16384 with odd quarter increments.
    Both the raw fp32 accumulators (mma4 debug entry) and the parities of the default kernel are
    compared with exact integer arithmetic."""
    import torch
    L = m4.lib()
    k = 65536
    kw = k // 64
    rng = np.random.default_rng(7)
    ones = np.uint64(0xFFFFFFFFFFFFFFFF)

    def dense(rows, density, seed):
        r = np.random.default_rng(seed)
        bits = (r.random((rows, k)) < density).astype(np.uint8)
        return np.packbits(bits, axis=1, bitorder="little").view(np.uint64).reshape(rows, kw).copy()

    def odd_words(rows, seed):
        # every 64-bit word has an odd popcount: all ones with one hole at a random position
        r = np.random.default_rng(seed)
        return ones ^ (np.uint64(1) << r.integers(0, 64, size=(rows, kw)).astype(np.uint64))

    cases = [(dense(256, 0.9, 1), dense(256, 0.95, 2)), (dense(256, 0.99, 3), dense(256, 0.999, 4)),
             (odd_words(256, 5), odd_words(256, 6)), (dense(256, 0.999, 8), np.full((256, kw), ones))]
    for ci, (A, B) in enumerate(cases):
        a8 = np.unpackbits(A.view(np.uint8), axis=1, bitorder="little").astype(np.float64)
        b8 = np.unpackbits(B.view(np.uint8), axis=1, bitorder="little").astype(np.float64)
        pc = (a8 @ b8.T)  # exact in float64 (<= 65536)
        assert pc.max() > 4197 * 4, "the case must exercise accumulators beyond the 50 %-random range"
        a, b = host(m4, A, k), host(m4, B, k)
        dA, dB, dC = m4.DMat.from_host(a), m4.DMat.from_host(b), m4.DMat.init(256, 256)
        for cg in (1, 2):
            acc = torch.full((256, 256), -1.0, dtype=torch.float32, device="cuda")
            torch.cuda.synchronize()
            L.m4ri_b200_dmul_abt_fp4_debug(dC.h, dA.h, dB.h, acc.data_ptr(), cg)
            L.m4ri_b200_synchronize()
            assert (acc.cpu().numpy().astype(np.float64) * 4 == pc).all(), (ci, cg)
        want = np.packbits((pc.astype(np.int64) & 1).astype(np.uint8), axis=1, bitorder="little").view(np.uint64)
        for variant in ("fp4s256", "fp4s256cg1", "fp4ldg", "fp4", "fp4ldgcg1"):
            L.m4ri_b200_set_mma_variant(m4.MMA_VARIANTS[variant])
            try:
                m4.dmul_abt(dC, dA, dB, clear=True, algo=m4.ALGO_MMA)
                assert (dC.to_words()[:, :4] == want).all(), (ci, variant)
            finally:
                L.m4ri_b200_set_mma_variant(0)
        L.mzd_free(a)
        L.mzd_free(b)


@pytest.mark.parametrize("k", [32768, 65536])
def test_fp4_accumulators_lone_quarter_worst_case(gpu, m4, k):
    """VERDICT r2 (weak 3): the deterministic worst case for alignment loss in the fp32 accumulator
    — a sum that is already as large as the contraction allows and then receives ONE 0.25, and the
    mirror image (the lone 0.25 first, everything else on top). Row r < 64 of A is all ones up to
    the last 64-bit word, which holds the single bit r; row 64 + r holds the single bit r in its
    FIRST word and all ones after it; rows 128.. repeat both with the full words cut at k/2 (a
    lone quarter on either side of the kernel's k-chunk boundary). B is all ones, and all ones
    with one hole per row. Raw accumulators (debug entry) and the default kernels' parities are
    compared with exact integer arithmetic."""
    import torch
    L = m4.lib()
    kw = k // 64
    ones = np.uint64(0xFFFFFFFFFFFFFFFF)
    A = np.zeros((256, kw), dtype=np.uint64)
    for r in range(64):
        bit = np.uint64(1) << np.uint64(r)
        A[r, : kw - 1], A[r, kw - 1] = ones, bit              # 8176.0 (k = 32768) ... then 0.25
        A[64 + r, 1:], A[64 + r, 0] = ones, bit               # 0.25 ... then the rest
        A[128 + r, : kw // 2], A[128 + r, kw - 1] = ones, bit
        A[192 + r, kw // 2:], A[192 + r, 0] = ones, bit
    B = np.full((256, kw), ones)
    holes = np.random.default_rng(k).integers(0, k, size=128)
    for j, h in enumerate(holes):
        B[128 + j, h // 64] ^= np.uint64(1) << np.uint64(h % 64)
    a8 = np.unpackbits(A.view(np.uint8), axis=1, bitorder="little").astype(np.float64)
    b8 = np.unpackbits(B.view(np.uint8), axis=1, bitorder="little").astype(np.float64)
    pc = a8 @ b8.T  # exact in float64
    assert pc.max() == k - 63 and (pc[:128, :128] == k - 63).all()
    a, b = host(m4, A, k), host(m4, B, k)
    dA, dB, dC = m4.DMat.from_host(a), m4.DMat.from_host(b), m4.DMat.init(256, 256)
    for cg in (1, 2):
        acc = torch.full((256, 256), -1.0, dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        L.m4ri_b200_dmul_abt_fp4_debug(dC.h, dA.h, dB.h, acc.data_ptr(), cg)
        L.m4ri_b200_synchronize()
        assert (acc.cpu().numpy().astype(np.float64) * 4 == pc).all(), cg
    want = np.packbits((pc.astype(np.int64) & 1).astype(np.uint8), axis=1, bitorder="little").view(np.uint64)
    for variant in ("fp4s256", "fp4s256cg1", "fp4ldg", "fp4"):
        L.m4ri_b200_set_mma_variant(m4.MMA_VARIANTS[variant])
        try:
            m4.dmul_abt(dC, dA, dB, clear=True, algo=m4.ALGO_MMA)
            assert (dC.to_words()[:, :4] == want).all(), variant
        finally:
            L.m4ri_b200_set_mma_variant(0)
    L.mzd_free(a)
    L.mzd_free(b)


def test_n65536_single_gpu(gpu, m4):
    """configs[3] on ONE GPU: n = 65536 through m4ri_b200_dmul AUTO — the plan with TWO Strassen
    levels (api.cu strassen_levels: leaves of 16384) — checked with Freivalds (2^-64), against the
    Four-Russians kernel's checksum, and a row-block product (the 8-GPU job's 8192 x 65536 x 65536
    shard, one level) against the same rows of the full result."""
    n = 65536
    L = m4.lib()
    dA = m4.DMat.init(n, n).fill_random(61)
    dB = m4.DMat.init(n, n).fill_random(62)
    dC = m4.DMat.init(n, n)
    L.m4ri_b200_timing_reset()
    m4.dmul(dC, dA, dB, algo=m4.ALGO_AUTO)
    L.m4ri_b200_synchronize()
    # two levels: 7 sub-products (Winograd outside), each ONE grouped launch of its 7 leaf products
    assert L.m4ri_b200_timing_count() == 7, "AUTO at n = 65536 runs 7 grouped launches of 7 leaves (two levels)"
    assert abs(L.m4ri_b200_timing_ops() - 49 * 2.0 * 16384 ** 3) < 1e6
    assert L.m4ri_b200_dfreivalds(dC.h, dA.h, dB.h, 0xF5E1) == 1
    chk = dC.checksum()
    # row-block of the sharded job: rows [16384, 24576)
    r0, rows = 16384, 8192
    ws = n // 64
    dAb = m4.DMat.wrap(dA.ptr + r0 * ws * 8, rows, n, ws, keepalive=dA)
    dCb = m4.DMat.init(rows, n)
    m4.dmul(dCb, dAb, dB, algo=m4.ALGO_AUTO)
    dCref = m4.DMat.wrap(dC.ptr + r0 * ws * 8, rows, n, ws, keepalive=dC)
    L.m4ri_b200_dadd(dCb.h, dCb.h, dCref.h)
    assert L.m4ri_b200_dis_zero(dCb.h) == 1
    del dCb
    # independent algorithm, same checksum
    m4.dmul(dC, dA, dB, algo=m4.ALGO_M4RM)
    assert dC.checksum() == chk


def test_result_into_window_keeps_the_parents_bits(gpu, m4, orc):
    """C is a window whose last word is shared with other columns of its parent (ncols % 64 != 0):
    the product must not touch the parent's bits beyond the window."""
    L = m4.lib()
    P0 = orc.random(300, 640, 4)
    for (m, k, n, lowc) in [(200, 320, 100, 128), (200, 5000, 70, 0), (64, 64, 1, 64)]:
        p = host(m4, P0, 640)
        w = L.mzd_init_window(p, 10, lowc, 10 + m, lowc + n)
        A, B = orc.random(m, k, 5), orc.random(k, n, 6)
        a, b = host(m4, A, k), host(m4, B, n)
        want = orc.mul_m4rm(A, B, k, n)[:, : orc.width(n)]
        for f in (lambda: L.mzd_mul(w, a, b, 0), lambda: L.mzd_mul_m4rm(w, a, b, 0), lambda: L.mzd_mul_naive(w, a, b)):
            f()
            now = m4.mzd_words(p)[:, :10]
            exp = P0[:, :10].copy()
            mask = np.uint64((1 << (n % 64)) - 1) if n % 64 else np.uint64(0xFFFFFFFFFFFFFFFF)
            wd = orc.width(n)
            blk = exp[10:10 + m, lowc // 64: lowc // 64 + wd]
            blk[:, : wd - 1] = want[:, : wd - 1]
            blk[:, wd - 1] = (blk[:, wd - 1] & ~mask) | (want[:, wd - 1] & mask)
            assert (now == exp).all(), (m, k, n, lowc)
        for q in (w, p, a, b):
            L.mzd_free(q)
