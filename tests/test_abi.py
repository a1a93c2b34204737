"""CPU tests of the C-ABI boundary: the library loads without a GPU, exports
every symbol include/m4ri_b200.h declares, and its host-side mzd_t allocator set
honours the layout contract the reference's own tests pin
(m4ri-sys/src/mzd.rs:381-456). No product is computed here (no GPU)."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "m4ri_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"\b(_?mzd_[a-z0-9_]+|[mM]zp_[a-z0-9_]+|m4ri_b200_[a-z0-9_]+)\s*\(", src)
    return sorted(set(n for n in names if n not in ("m4ri_b200_dmat",)))


def test_every_declared_symbol_is_exported(m4):
    L = m4.lib()
    names = declared_symbols()
    assert len(names) >= 40, names
    for must in ["mzd_mul", "mzd_addmul", "mzd_mul_m4rm", "mzd_addmul_m4rm", "_mzd_mul_m4rm",
                 "mzd_mul_naive", "mzd_addmul_naive", "_mzd_mul_naive", "_mzd_mul_va", "mzd_mul_v",
                 "mzd_init", "mzd_free", "mzd_transpose", "mzd_copy", "mzd_equal", "mzd_add",
                 "mzd_randomize", "mzd_set_ui", "mzd_ple", "mzd_pluq", "mzp_init", "mzp_free",
                 "mzd_apply_p_left", "mzd_apply_p_right_trans", "Mzp_free_window"]:
        assert must in names
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/m4ri_b200.h but not exported"
    # and through the dynamic symbol table, not only ctypes' lazy lookup
    out = subprocess.run(["nm", "-D", "--defined-only", m4.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(line.split()[-1] for line in out.splitlines() if line.strip())
    assert set(names) <= exported


def test_no_torch_or_oracle_dependency(m4):
    out = subprocess.run(["ldd", m4.LIB_PATH], capture_output=True, text=True).stdout
    assert "torch" not in out and "oracle" not in out and "m4ri.so" not in out


def test_sizeof_mzd_is_64(m4):  # m4ri-sys/src/mzd.rs:385
    assert ctypes.sizeof(m4.Mzd) == 64
    assert m4.Mzd.high_bitmask.offset == 40 and m4.Mzd.blocks.offset == 48 and m4.Mzd.rows.offset == 56


def test_init_layout(m4):  # mzd.rs:381-422
    L = m4.lib()
    for (r, c) in [(10, 10), (1000, 1000), (3, 129), (7, 193), (5, 64), (64, 321)]:
        p = L.mzd_init(r, c)
        m = p.contents
        w = (c + 63) // 64
        assert (m.nrows, m.ncols, m.width) == (r, c, w)
        assert m.rowstride == (w if (w < 3 or w % 2 == 0) else w + 1)  # mzd.rs:35-37
        assert bool(m.blocks) and bool(m.rows)  # mzd.rs:387-388
        assert m.high_bitmask == ((1 << (c % 64)) - 1 if c % 64 else (1 << 64) - 1)
        base = ctypes.addressof(m.blocks[0].begin.contents)
        for i in range(r):  # mzd_row / mzd_first_row address rule, mzd.rs:277-313
            assert ctypes.addressof(m.rows[i].contents) == base + 8 * (m.offset_vector + i * m.rowstride)
        assert not m4.mzd_words(p).any()
        assert L.mzd_equal(p, p) == 1
        L.mzd_free(p)


def test_set_ui_write_read_bit(m4):  # mzd.rs:425-456
    L = m4.lib()
    p = L.mzd_init(1000, 1000)
    L.mzd_set_ui(p, 1)
    M = m4.BinMatrix(p)
    for i in range(0, 1000, 7):
        for j in (0, i, 999 - i, 63, 64):
            assert M.bit(i, j) == (i == j)
    # identity(3) words: rows [1],[2],[4] (binary_matrix.rs:693-699)
    I3 = m4.BinMatrix.identity(3)
    assert [int(I3.words()[r, 0]) for r in range(3)] == [1, 2, 4]


def test_randomize_copy_equal(m4):  # mzd.rs:389-394
    L = m4.lib()
    a = L.mzd_init(100, 131)
    L.mzd_randomize(a)
    w = m4.mzd_words(a)
    assert w.any()
    assert not (w[:, 2] >> np.uint64(131 - 128)).any()  # excess bits masked
    assert not w[:, 3:].any()                          # padding word zero
    b = L.mzd_copy(None, a)
    assert L.mzd_equal(a, b) == 1
    L.mzd_randomize(b)
    assert L.mzd_equal(a, b) == 0
    L.mzd_free(a)
    L.mzd_free(b)


def test_window_shares_words(m4):  # mzd.rs:121-127
    L = m4.lib()
    a = L.mzd_init(20, 300)
    L.mzd_randomize(a)
    w = L.mzd_init_window(a, 5, 64, 15, 200)
    assert (w.contents.nrows, w.contents.ncols) == (10, 136)
    for i in range(10):
        assert ctypes.addressof(w.contents.rows[i].contents) == ctypes.addressof(a.contents.rows[5 + i].contents) + 8
    L.mzd_free(w)
    L.mzd_free(a)


def test_binvector_mirror(m4):  # binary_vector.rs:217-288
    v = m4.BinVector.from_bytes(bytes([0b10000000]))
    assert [int(b) for b in v.bits] == [1, 0, 0, 0, 0, 0, 0, 0]
    a = m4.BinVector.from_function(70, lambda i: i % 3 == 0)
    assert a.count_ones() == 24 and len(a) == 70
    st = a.get_storage()
    assert m4.BinVector.from_storage(st, 70) == a
    assert (a + a).count_ones() == 0
    assert (a * a) == (24 % 2 == 1)


def test_product_without_gpu_dies_loudly(m4):
    """No CPU fallback: on a box without a GPU the product aborts the process."""
    if m4.device_count() > 0:
        return
    code = ("import sys; sys.path.insert(0, %r); import m4ri_rust_b200 as m; "
            "a = m.BinMatrix.identity(8); a * a" % ROOT)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert r.returncode != 0
    assert "no CPU fallback" in r.stderr


def test_pipeline_plan_is_host_arithmetic():
    """m4ri_b200_pipeline_plan needs no device: the block plan of the host-ABI pipeline for the
    benchmark shapes is the one measured fastest (profiles/r2p_e2e_sweep*.jsonl; at n = 65536 within
    0.5 % of it), and every plan
    has at least one block each way."""
    import ctypes
    import m4ri_rust_b200 as m4
    L = m4.lib()
    def plan(m, k, n, clear=1):
        a, b = ctypes.c_int(), ctypes.c_int()
        L.m4ri_b200_pipeline_plan(m, k, n, clear, ctypes.byref(a), ctypes.byref(b))
        return a.value, b.value
    assert plan(32768, 32768, 32768) == (16, 2)
    assert plan(16384, 16384, 16384) == (2, 4)
    assert plan(12288, 12288, 12288) == (2, 4)
    assert plan(65536, 65536, 65536) == (8, 4)
    for shape in [(1, 1, 1), (300, 70000, 129), (1 << 20, 512, 512), (8192, 65536, 65536), (5000, 20000, 9000)]:
        a, b = plan(*shape)
        assert 1 <= a <= 16 and 1 <= b <= 16
        a, b = plan(*shape, clear=0)
        assert 1 <= a <= 16 and 1 <= b <= 16


def test_host_block_cache_reuses_and_zeroes():
    """mzd_free keeps big blocks for the next mzd_init of about the same size (m4ri-rust allocates
    and drops a result per product): the block comes back zeroed, a slightly smaller request
    takes it too, a much smaller one does not, and trim empties the cache."""
    import ctypes
    import numpy as np
    import m4ri_rust_b200 as m4
    L = m4.lib()
    def stats():
        h, m_, b = ctypes.c_ulonglong(), ctypes.c_ulonglong(), ctypes.c_ulonglong()
        L.m4ri_b200_host_cache_stats(ctypes.byref(h), ctypes.byref(m_), ctypes.byref(b))
        return h.value, m_.value, b.value
    L.m4ri_b200_host_cache_trim()
    h0, m0, b0 = stats()
    assert b0 == 0
    a = L.mzd_init(2048, 4096)  # 1 MiB
    first = a.contents.blocks[0].begin
    addr = ctypes.cast(first, ctypes.c_void_p).value
    L.mzd_randomize(a)
    assert m4.mzd_words(a).any()
    L.mzd_free(a)
    assert stats()[2] == 2048 * 64 * 8
    b = L.mzd_init(2048, 4096)
    assert ctypes.cast(b.contents.blocks[0].begin, ctypes.c_void_p).value == addr
    assert not m4.mzd_words(b).any(), "a recycled block must come back zeroed"
    assert stats()[0] == h0 + 1 and stats()[2] == 0
    L.mzd_free(b)
    c = L.mzd_init(2000, 4096)  # 2.3 % smaller: same block
    assert ctypes.cast(c.contents.blocks[0].begin, ctypes.c_void_p).value == addr
    assert c.contents.blocks[0].size == 2000 * 64 * 8 and not m4.mzd_words(c).any()
    L.mzd_free(c)
    d = L.mzd_init(1024, 4096)  # half the size: a block of its own
    assert ctypes.cast(d.contents.blocks[0].begin, ctypes.c_void_p).value != addr
    L.mzd_free(d)
    assert stats()[2] == (2048 + 1024) * 64 * 8
    L.m4ri_b200_host_cache_trim()
    assert stats()[2] == 0
    # small matrices never enter the cache
    e = L.mzd_init(10, 10)
    L.mzd_free(e)
    assert stats()[2] == 0


def test_allocator_is_thread_safe():
    """BinMatrix is Send + Sync (binary_matrix.rs:38-39): mzd_init / mzd_free from several host
    threads at once, through the block cache, never hand one block to two owners."""
    import ctypes
    import threading
    import m4ri_rust_b200 as m4
    L = m4.lib()
    errors = []

    def worker(tag):
        try:
            for it in range(40):
                a = L.mzd_init(1024 + 64 * (it % 3), 4096)  # 512-576 KiB: cached sizes
                w = m4.mzd_words(a)
                if w.any():
                    errors.append("block not zeroed")
                w[:] = tag  # every word of the block carries the owner's tag
                if not (m4.mzd_words(a) == tag).all():
                    errors.append("block shared between owners")
                L.mzd_free(a)
        except Exception as e:  # pragma: no cover
            errors.append(repr(e))

    threads = [threading.Thread(target=worker, args=(t + 1,)) for t in range(6)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors[:3]
    L.m4ri_b200_host_cache_trim()


def _bits(m4, p):
    m = p.contents
    return np.array([[(int(m.rows[r][c // 64]) >> (c % 64)) & 1 for c in range(m.ncols)] for r in range(m.nrows)],
                    dtype=np.uint8).reshape(m.nrows, m.ncols)


def test_concat_and_stack(m4):
    """mzd_concat / mzd_stack (mzd.rs:195,201; BinMatrix::augmented / stacked,
    binary_matrix.rs:221-244): host copies at arbitrary bit offsets, NULL and preallocated C."""
    L = m4.lib()
    for (r, ca, cb) in [(5, 3, 4), (7, 64, 64), (9, 70, 130), (4, 1, 200), (3, 127, 1), (6, 128, 65), (2, 200, 300)]:
        a, b = L.mzd_init(r, ca), L.mzd_init(r, cb)
        L.mzd_randomize(a)
        L.mzd_randomize(b)
        c = L.mzd_concat(None, a, b)
        assert (c.contents.nrows, c.contents.ncols) == (r, ca + cb)
        want = np.concatenate([_bits(m4, a), _bits(m4, b)], axis=1)
        assert (_bits(m4, c) == want).all(), (r, ca, cb)
        w = m4.mzd_words(c)
        assert not (w[:, c.contents.width - 1] & ~np.uint64(c.contents.high_bitmask)).any()
        assert not w[:, c.contents.width:].any()
        c2 = L.mzd_init(r, ca + cb)
        L.mzd_randomize(c2)  # stale contents are overwritten
        assert ctypes.addressof(L.mzd_concat(c2, a, b).contents) == ctypes.addressof(c2.contents)
        assert L.mzd_equal(c, c2) == 1
        for p in (a, b, c, c2):
            L.mzd_free(p)
    for (ra, rb, c_) in [(3, 4, 5), (1, 1, 64), (10, 20, 130), (7, 2, 200)]:
        a, b = L.mzd_init(ra, c_), L.mzd_init(rb, c_)
        L.mzd_randomize(a)
        L.mzd_randomize(b)
        s = L.mzd_stack(None, a, b)
        assert (s.contents.nrows, s.contents.ncols) == (ra + rb, c_)
        assert (_bits(m4, s) == np.concatenate([_bits(m4, a), _bits(m4, b)], axis=0)).all()
        for p in (a, b, s):
            L.mzd_free(p)


def test_concat_and_stack_into_a_window_keep_the_parents_bits(m4):
    """C may be a window (mzd.rs:121-127) that ends inside a word of its parent: the parent's
    columns in that word, and everything outside the window, survive."""
    L = m4.lib()
    for (r, ca, cb) in [(6, 30, 40), (5, 64, 1), (4, 100, 27)]:
        n = ca + cb
        for op in ("concat", "stack"):
            parent = L.mzd_init(r + 4, 64 + n + 50)
            L.mzd_randomize(parent)
            before = _bits(m4, parent)
            win = L.mzd_init_window(parent, 2, 64, 2 + r, 64 + n)
            if op == "concat":
                a, b = L.mzd_init(r, ca), L.mzd_init(r, cb)
            else:
                a, b = L.mzd_init(2, n), L.mzd_init(r - 2, n)
            L.mzd_randomize(a)
            L.mzd_randomize(b)
            (L.mzd_concat if op == "concat" else L.mzd_stack)(win, a, b)
            want = before.copy()
            want[2:2 + r, 64:64 + n] = np.concatenate([_bits(m4, a), _bits(m4, b)], axis=1 if op == "concat" else 0)
            assert (_bits(m4, parent) == want).all(), (op, r, ca, cb)
            L.mzd_free(win)
            for p in (a, b, parent):
                L.mzd_free(p)


def test_free_rejects_foreign_descriptors(m4):
    """A descriptor this library did not create (e.g. from another M4RI in the process) must not
    reach free()/cudaFreeHost: mzd_free dies loudly instead."""
    code = ("import sys; sys.path.insert(0, %r); import ctypes, m4ri_rust_b200 as m4; L = m4.lib();"
            "p = L.mzd_init(4, 4); q = m4.Mzd(); ctypes.memmove(ctypes.byref(q), p, 64);"
            "q.padding[9] = 0; L.mzd_free(ctypes.pointer(q))") % ROOT
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert r.returncode != 0 and "not created by libm4ri_b200" in r.stderr


def test_row_and_column_utilities(m4, orc):
    """mzd_row_swap / col_swap / copy_row / is_zero / row_clear_offset / submatrix (m4ri-sys
    mzd.rs:130-240) against numpy on unpacked bits; pure host work, no device needed."""
    L = m4.lib()
    for (m, n) in [(5, 7), (64, 130), (33, 257), (70, 64)]:
        A = orc.random(m, n, 3 + n)
        bits = orc.to_bits(A, n)
        a = m4.mzd_from_words(A, n)
        L.mzd_row_swap(a, 1, m - 1)
        bits[[1, m - 1]] = bits[[m - 1, 1]]
        L.mzd_col_swap(a, 0, n - 1)
        bits[:, [0, n - 1]] = bits[:, [n - 1, 0]]
        L.mzd_col_swap(a, 2, min(n - 1, 66))
        bits[:, [2, min(n - 1, 66)]] = bits[:, [min(n - 1, 66), 2]]
        assert (orc.to_bits(np.ascontiguousarray(m4.mzd_words(a)), n) == bits).all(), (m, n)
        L.mzd_row_clear_offset(a, 2, n // 2)
        bits[2, n // 2:] = 0
        L.mzd_copy_row(a, 0, a, 3)
        bits[0] = bits[3]
        got = m4.mzd_words(a)
        assert (orc.to_bits(np.ascontiguousarray(got), n) == bits).all(), (m, n)
        assert not got[:, orc.width(n):].any()
        assert L.mzd_is_zero(a) == 0
        for (r0, c0, r1, c1) in [(0, 0, m, n), (1, 3, m - 1, n - 1), (2, min(n - 2, 65), m, n), (0, 1, 1, 2)]:
            sub = L.mzd_submatrix(None, a, r0, c0, r1, c1)
            sb = orc.to_bits(np.ascontiguousarray(m4.mzd_words(sub)), c1 - c0)
            assert (sb == bits[r0:r1, c0:c1]).all(), (m, n, r0, c0, r1, c1)
            assert not m4.mzd_words(sub)[:, orc.width(c1 - c0):].any()
            L.mzd_free(sub)
        z = L.mzd_init(m, n)
        assert L.mzd_is_zero(z) == 1
        L.mzd_free(z)
        L.mzd_free(a)


def test_permutation_bookkeeping(m4):
    """mzp_init / mzp_init_window / mzp_copy / mzp_set_ui / mzp_free (m4ri-sys/src/mzp.rs:18-36):
    pure host bookkeeping on M4RI's published {values, length} layout — no device needed."""
    L = m4.lib()
    P = L.mzp_init(10)
    assert P.contents.length == 10 and [P.contents.values[i] for i in range(10)] == list(range(10))
    W = L.mzp_init_window(P, 3, 8)
    assert W.contents.length == 5 and W.contents.values[0] == 3
    W.contents.values[0] = 7  # a window shares its parent's values
    assert P.contents.values[3] == 7
    C = L.mzp_copy(None, P)
    assert [C.contents.values[i] for i in range(10)] == [0, 1, 2, 7, 4, 5, 6, 7, 8, 9]
    C.contents.values[0] = 9  # a copy does not
    assert P.contents.values[0] == 0
    L.mzp_set_ui(P, 1)
    assert [P.contents.values[i] for i in range(10)] == list(range(10))
    Z = L.mzp_init(0)
    assert Z.contents.length == 0
    for q in (C, Z):
        L.mzp_free(q)
    L.mzp_free_window(W)
    L.mzp_free(P)
