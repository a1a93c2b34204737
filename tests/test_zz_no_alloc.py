"""GPU test of m4ri_b200_mul_noalloc, the product entry point for callers that cannot allocate.

Who those callers are: programs in which the GPU's memory already belongs to somebody else — a
torch process that hands this library views of its own tensors, or a service that filled HBM
with its matrices — so that a cudaMallocAsync inside m4ri_b200_dmul (for B^T) may fail or eat
memory the owner had counted on; and tight loops of small products that must not run into
cudaMalloc's device-wide synchronisation. For them the B^T temporary is a fixed 8 MiB array in
the library's device image (B up to 8192 x 8192 bits), shared by every call on that GPU.

What is checked: the new entry point, called through the C-ABI with raw device pointers, gives
word-for-word the result of the existing m4ri_b200_dmul on the same device operands — on two
calls in a row, the second smaller and ragged so that it runs over the first call's leftovers in
the shared array, and accumulating (clear = 0). Then the largest B the array can hold, and the
three kinds of rejected arguments. The operands are allocated by the TEST (m4ri_b200_dmat_init);
the entry point itself only ever sees pointers.

This file is meant to be collected last (pytest -x would otherwise hide later tests behind a
failure here); it passes alone in a fresh process and after the rest of the suite.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def noalloc(L, C, A, B, m, k, n, clear):
    return L.m4ri_b200_mul_noalloc(C.ptr, C.rowstride, A.ptr, A.rowstride, B.ptr, B.rowstride, m, k, n, int(clear), None)


def pair(m4, m, k, n, seed, prefill):
    """A, B and two equal destinations (random when the product will be accumulated)."""
    A = m4.DMat.init(m, k).fill_random(seed)
    B = m4.DMat.init(k, n).fill_random(seed + 1)
    C_new, C_old = m4.DMat.init(m, n), m4.DMat.init(m, n)
    if prefill:
        C_new.fill_random(seed + 2)
        C_old.fill_random(seed + 2)
    return A, B, C_new, C_old


def test_two_calls_in_a_row_match_dmul(gpu, m4, orc):
    L = m4.lib()
    # first: a fresh product on the tensor path, k words per row odd (stride padding in B^T)
    # second: smaller in every dimension, ragged, accumulating into a random C
    for (m, k, n, clear) in [(1000, 3000, 2000, True), (333, 1037, 700, False)]:
        A, B, C_new, C_old = pair(m4, m, k, n, 100 + m, prefill=not clear)
        assert noalloc(L, C_new, A, B, m, k, n, clear) == 0
        m4.dmul(C_old, A, B, clear=clear)
        got, want = C_new.to_words(), C_old.to_words()
        assert got.shape == want.shape
        assert (got == want).all(), (m, k, n, clear)
        assert got.any()
    # and once against the CPU oracle, small enough for it: after the two calls above the
    # shared array is full of their bits
    m, k, n = 130, 200, 70
    a, b = orc.random(m, k, 7), orc.random(k, n, 8)
    A, B, C = m4.DMat.from_host(m4.mzd_from_words(a, k)), m4.DMat.from_host(m4.mzd_from_words(b, n)), m4.DMat.init(m, n)
    assert noalloc(L, C, A, B, m, k, n, True) == 0
    assert (C.to_words()[:, : orc.width(n)] == orc.mul_m4rm(a, b, k, n)[:, : orc.width(n)]).all()


def test_largest_b_and_repeat(gpu, m4):
    L = m4.lib()
    m, k, n = 2048, 8192, 8192  # B^T fills the array exactly
    A, B, C_new, C_old = pair(m4, m, k, n, 5, prefill=False)
    m4.dmul(C_old, A, B, clear=True)
    want = C_old.to_words()
    for _ in range(2):  # the same call twice: the second starts from the first one's B^T
        assert noalloc(L, C_new, A, B, m, k, n, True) == 0
        assert (C_new.to_words() == want).all()
    # C ^= A*B twice restores C
    assert noalloc(L, C_new, A, B, m, k, n, False) == 0
    assert not C_new.to_words().any()


def test_rejected_arguments(gpu, m4, capfd):
    L = m4.lib()
    A, B, C = m4.DMat.init(64, 8256), m4.DMat.init(8256, 8192), m4.DMat.init(64, 8192)
    C.fill_random(3)
    before = C.to_words()
    launches = L.m4ri_b200_launch_count()
    assert noalloc(L, C, A, B, 64, 8256, 8192, True) == -1  # one 64-bit word per row too many
    assert "scratch" in capfd.readouterr().err
    assert L.m4ri_b200_mul_noalloc(C.ptr + 8, C.rowstride, A.ptr, A.rowstride, B.ptr, B.rowstride, 64, 128, 128, 1, None) == -2
    assert L.m4ri_b200_mul_noalloc(C.ptr, C.rowstride, A.ptr, A.rowstride, B.ptr, 1, 64, 128, 128, 1, None) == -3
    assert L.m4ri_b200_mul_noalloc(C.ptr, C.rowstride, A.ptr, A.rowstride, B.ptr, B.rowstride, 64, -1, 128, 1, None) == -3
    assert L.m4ri_b200_launch_count() == launches  # a rejected call launches nothing
    assert (C.to_words() == before).all()
    # empty shapes are fine and leave C alone / zero it for k = 0
    assert noalloc(L, C, A, B, 0, 8256, 8192, True) == 0
    assert noalloc(L, C, A, B, 64, 0, 8192, False) == 0
    assert (C.to_words() == before).all()
    assert noalloc(L, C, A, B, 64, 0, 8192, True) == 0
    assert not C.to_words().any()
