"""m4ri-rust_b200 — Python host side of libm4ri_b200.so (ctypes over the C-ABI).

The product is the shared library (C++/CUDA, include/m4ri_b200.h). This module
is the thin binding the tests and bench.py use, plus `BinMatrix` / `BinVector`
mirrors of m4ri-rust's friendly API (m4ri-rust/src/friendly/binary_matrix.rs,
binary_vector.rs) with the same names, argument meaning and error behaviour, so
the parity tests read like the reference's own tests. The C++ mirror of the
same API for compiled hosts is include/m4ri_b200.hpp.

The directory name contains a hyphen (it is the name the build contract asks
for); import it as `m4ri_rust_b200` through the shim at the repository root.

There is no CPU implementation of any product here: if the library is missing
the import fails, and without a GPU every multiplication entry point aborts.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# M4RI_B200_LIB selects another build of the same library (the perf-hook build of build.py --perf)
LIB_PATH = os.environ.get("M4RI_B200_LIB") or os.path.join(_HERE, "lib", "libm4ri_b200.so")
HEADER_PATH = os.path.join(_HERE, "..", "include", "m4ri_b200.h")

ALGO_AUTO, ALGO_M4RM, ALGO_MMA, ALGO_SIMPLE = 0, 1, 2, 3
ALGO_NAMES = {"auto": 0, "m4rm": 1, "mma": 2, "simple": 3}
MMA_VARIANTS = {"default": 0, "v1": 1, "v2cg1": 2, "v2": 3, "fp4cg1": 4, "fp4": 5, "fp4ldgcg1": 6, "fp4ldg": 7, "fp4s256cg1": 8, "fp4s256": 9}


class MzdBlock(ctypes.Structure):
    _fields_ = [("size", ctypes.c_size_t), ("begin", ctypes.POINTER(ctypes.c_uint64)),
                ("end", ctypes.POINTER(ctypes.c_uint64))]


class Mzd(ctypes.Structure):
    """m4ri-sys/src/mzd.rs:24-79, 64 bytes."""
    _fields_ = [("nrows", ctypes.c_int), ("ncols", ctypes.c_int), ("width", ctypes.c_int),
                ("rowstride", ctypes.c_int), ("offset_vector", ctypes.c_int),
                ("row_offset", ctypes.c_int), ("flags", ctypes.c_uint8),
                ("blockrows_log", ctypes.c_uint8), ("padding", ctypes.c_uint8 * 14),
                ("high_bitmask", ctypes.c_uint64), ("blocks", ctypes.POINTER(MzdBlock)),
                ("rows", ctypes.POINTER(ctypes.POINTER(ctypes.c_uint64)))]


MzdP = ctypes.POINTER(Mzd)


class Mzp(ctypes.Structure):
    """M4RI's mzp_t (opaque in m4ri-sys/src/mzp.rs:9-12): LAPACK-style transposition list."""
    _fields_ = [("values", ctypes.POINTER(ctypes.c_int)), ("length", ctypes.c_int)]


MzpP = ctypes.POINTER(Mzp)
_lib = None


def lib():
    """Load libm4ri_b200.so and declare the C-ABI. Fails loudly when it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is not built: run `python m4ri-rust_b200/build.py` "
            "(there is no Python or CPU fallback for the GF(2) product)")
    L = ctypes.CDLL(LIB_PATH)
    i, u, vp, u64 = ctypes.c_int, ctypes.c_uint, ctypes.c_void_p, ctypes.c_uint64
    sig = {
        "mzd_init": (MzdP, [i, i]), "mzd_free": (None, [MzdP]),
        "mzd_init_window": (MzdP, [MzdP, i, i, i, i]),
        "mzd_copy": (MzdP, [MzdP, MzdP]), "mzd_equal": (i, [MzdP, MzdP]),
        "mzd_set_ui": (None, [MzdP, u]), "mzd_randomize": (None, [MzdP]),
        "mzd_transpose": (MzdP, [MzdP, MzdP]),
        "mzd_add": (MzdP, [MzdP, MzdP, MzdP]), "mzd_sub": (MzdP, [MzdP, MzdP, MzdP]),
        "mzd_echelonize": (i, [MzdP, i]), "mzd_echelonize_pluq": (i, [MzdP, i]),
        "mzd_echelonize_m4ri": (i, [MzdP, i, i]), "mzd_inv_m4ri": (MzdP, [MzdP, MzdP, i]),
        "mzd_solve_left": (i, [MzdP, MzdP, i, i]),
        "m4ri_b200_drref": (i, [vp, i, ctypes.POINTER(i), i]),
        "mzp_init": (MzpP, [i]), "mzp_free": (None, [MzpP]), "mzp_init_window": (MzpP, [MzpP, i, i]),
        "mzp_free_window": (None, [MzpP]), "Mzp_free_window": (None, [MzpP]), "mzp_copy": (MzpP, [MzpP, MzpP]),
        "mzp_set_ui": (None, [MzpP, u]), "mzp_print": (None, [MzpP]),
        "mzd_apply_p_left": (None, [MzdP, MzpP]), "mzd_apply_p_left_trans": (None, [MzdP, MzpP]),
        "mzd_apply_p_right": (None, [MzdP, MzpP]), "mzd_apply_p_right_trans": (None, [MzdP, MzpP]),
        "mzd_ple": (i, [MzdP, MzpP, MzpP, i]), "mzd_pluq": (i, [MzdP, MzpP, MzpP, i]),
        "mzd_pluq_solve_left": (i, [MzdP, i, MzpP, MzpP, MzdP, i, i]),
        "m4ri_b200_dple": (i, [vp, ctypes.POINTER(i), ctypes.POINTER(i), i]),
        "mzd_concat": (MzdP, [MzdP, MzdP, MzdP]), "mzd_stack": (MzdP, [MzdP, MzdP, MzdP]),
        "mzd_row_swap": (None, [MzdP, i, i]), "mzd_col_swap": (None, [MzdP, i, i]),
        "mzd_copy_row": (None, [MzdP, i, MzdP, i]), "mzd_is_zero": (i, [MzdP]),
        "mzd_row_clear_offset": (None, [MzdP, i, i]), "mzd_submatrix": (MzdP, [MzdP, MzdP, i, i, i, i]),
        "mzd_mul_mp": (MzdP, [MzdP, MzdP, MzdP, i]), "mzd_addmul_mp": (MzdP, [MzdP, MzdP, MzdP, i]),
        "mzd_invert_naive": (MzdP, [MzdP, MzdP, MzdP]), "mzd_top_echelonize_m4ri": (i, [MzdP, i]),
        "mzd_mul": (MzdP, [MzdP, MzdP, MzdP, i]), "mzd_addmul": (MzdP, [MzdP, MzdP, MzdP, i]),
        "mzd_mul_m4rm": (MzdP, [MzdP, MzdP, MzdP, i]),
        "mzd_addmul_m4rm": (MzdP, [MzdP, MzdP, MzdP, i]),
        "_mzd_mul_m4rm": (MzdP, [MzdP, MzdP, MzdP, i, i]),
        "mzd_mul_naive": (MzdP, [MzdP, MzdP, MzdP]),
        "mzd_addmul_naive": (MzdP, [MzdP, MzdP, MzdP]),
        "_mzd_mul_naive": (MzdP, [MzdP, MzdP, MzdP, i]),
        "_mzd_mul_va": (MzdP, [MzdP, MzdP, MzdP, i]),
        "mzd_mul_v": (MzdP, [MzdP, MzdP, MzdP]),
        "m4ri_b200_set_algo": (None, [i]), "m4ri_b200_get_algo": (i, []),
        "m4ri_b200_set_strassen_levels": (None, [i]),
        "m4ri_b200_set_mma_variant": (None, [i]),
        "m4ri_b200_device_count": (i, []),
        "m4ri_b200_device_info": (i, [ctypes.c_char_p, i, ctypes.POINTER(i), ctypes.POINTER(i), ctypes.POINTER(i)]),
        "m4ri_b200_set_stream": (None, [vp]), "m4ri_b200_use_own_stream": (None, []), "m4ri_b200_synchronize": (None, []),
        "m4ri_b200_launch_count": (ctypes.c_ulonglong, []),
        "m4ri_b200_version": (ctypes.c_char_p, []),
        "m4ri_b200_dmat_init": (vp, [i, i]), "m4ri_b200_dmat_wrap": (vp, [vp, i, i, i]),
        "m4ri_b200_dmat_free": (None, [vp]),
        "m4ri_b200_dmat_nrows": (i, [vp]), "m4ri_b200_dmat_ncols": (i, [vp]),
        "m4ri_b200_dmat_rowstride": (i, [vp]), "m4ri_b200_dmat_ptr": (vp, [vp]),
        "m4ri_b200_upload": (None, [vp, MzdP]), "m4ri_b200_download": (None, [MzdP, vp]),
        "m4ri_b200_dmul": (None, [vp, vp, vp, i, i]),
        "m4ri_b200_dmul_abt": (None, [vp, vp, vp, i, i]),
        "m4ri_b200_mul_noalloc": (i, [vp, i, vp, i, vp, i, i, i, i, i, vp]),
        "m4ri_b200_dmul_abt_fp4_debug": (None, [vp, vp, vp, vp, i]),
        "m4ri_b200_dmul_v": (None, [vp, vp, vp, i]),
        "m4ri_b200_dmul_va": (None, [vp, vp, vp, i]),
        "m4ri_b200_dmul_v_batched": (None, [vp, vp, vp, i]),
        "m4ri_b200_dtranspose": (None, [vp, vp]), "m4ri_b200_dadd": (None, [vp, vp, vp]),
        "m4ri_b200_dcopy": (None, [vp, vp]),
        "m4ri_b200_dfill_random": (None, [vp, u64]),
        "m4ri_b200_dchecksum": (u64, [vp]),
        "m4ri_b200_dfreivalds": (i, [vp, vp, vp, u64]),
        "m4ri_b200_dis_zero": (i, [vp]), "m4ri_b200_dflip_bit": (None, [vp, i, i]),
        "m4ri_b200_algo_available": (i, [i]),
        "m4ri_b200_dfill_random_at": (None, [vp, u64, u64]),
        "m4ri_b200_set_sm_limit": (None, [i]), "m4ri_b200_set_devices": (None, [i]), "m4ri_b200_get_devices": (i, []), "m4ri_b200_host_cache_trim": (None, []), "m4ri_b200_host_cache_stats": (None, [ctypes.POINTER(ctypes.c_ulonglong)] * 3), "m4ri_b200_pipeline_plan": (None, [i, i, i, i, ctypes.POINTER(i), ctypes.POINTER(i)]), "m4ri_b200_dtranspose_scatter": (None, [ctypes.POINTER(vp), i, vp]), "m4ri_b200_timing_reset": (None, []), "m4ri_b200_timing_ms": (ctypes.c_double, []),
        "m4ri_b200_timing_count": (i, []), "m4ri_b200_timing_ops": (ctypes.c_double, []), "m4ri_b200_timing_name": (ctypes.c_char_p, []),
    }
    for name, (res, args) in sig.items():
        f = getattr(L, name)  # AttributeError = symbol missing: loud
        f.restype, f.argtypes = res, args
    _lib = L
    return L


def device_count():
    return int(lib().m4ri_b200_device_count())


def device_info():
    name = ctypes.create_string_buffer(256)
    sms, maj, mnr = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    ok = lib().m4ri_b200_device_info(name, 256, ctypes.byref(sms), ctypes.byref(maj), ctypes.byref(mnr))
    return {"name": name.value.decode(), "sms": sms.value, "cc": (maj.value, mnr.value)} if ok else None


def launch_count():
    return int(lib().m4ri_b200_launch_count())


def mzd_words(p):
    """numpy view [nrows, rowstride] of the words of a (non-window) mzd_t."""
    m = p.contents
    n = m.nrows * m.rowstride
    if n == 0:
        return np.zeros((m.nrows, m.rowstride), dtype=np.uint64)
    addr = ctypes.addressof(m.blocks[0].begin.contents) + 8 * m.offset_vector
    buf = (ctypes.c_uint64 * n).from_address(addr)
    return np.frombuffer(buf, dtype=np.uint64).reshape(m.nrows, m.rowstride)


def mzd_from_words(words, ncols):
    """New mzd_t holding `words` ([nrows, >=width] uint64), excess bits masked."""
    L = lib()
    nrows = words.shape[0]
    p = L.mzd_init(nrows, ncols)
    w = p.contents.width
    if nrows and w:
        view = mzd_words(p)
        view[:, :w] = words[:, :w]
        view[:, w - 1] &= np.uint64(p.contents.high_bitmask)
    return p


# ---------------------------------------------------------------------------------
# friendly API mirrors
# ---------------------------------------------------------------------------------
class BinVector:
    """Mirror of m4ri_rust::friendly::BinVector (binary_vector.rs:13-215): a bit
    vector stored LSB-first in 64-bit words (vob::Vob<usize> in the reference)."""

    def __init__(self, bits=()):
        self.bits = np.asarray(list(bits), dtype=np.uint8) & 1

    # constructors: binary_vector.rs:44-108
    @classmethod
    def from_bools(cls, bools):
        return cls([1 if b else 0 for b in bools])

    @classmethod
    def from_elem(cls, length, elem):
        return cls([1 if elem else 0] * length)

    @classmethod
    def from_function(cls, length, f):
        return cls([1 if f(i) else 0 for i in range(length)])

    @classmethod
    def random(cls, length, rng=None):
        rng = rng or np.random.default_rng()
        return cls(rng.integers(0, 2, size=length))

    @classmethod
    def from_bytes(cls, data):
        """MSB-first per byte (binary_vector.rs:229-237)."""
        return cls([(b >> (7 - j)) & 1 for b in data for j in range(8)])

    @classmethod
    def from_storage(cls, words, length):
        words = np.asarray(words, dtype=np.uint64)
        idx = np.arange(length)
        return cls(((words[idx // 64] >> (idx % 64).astype(np.uint64)) & np.uint64(1)).astype(np.uint8))

    def __len__(self):
        return int(self.bits.size)

    def get(self, i):
        return bool(self.bits[i]) if 0 <= i < len(self) else None

    def get_storage(self):
        words = np.zeros((len(self) + 63) // 64, dtype=np.uint64)
        for i in np.flatnonzero(self.bits):
            words[i // 64] |= np.uint64(1) << np.uint64(i % 64)
        return words

    def count_ones(self):  # binary_vector.rs:112
        return int(self.bits.sum())

    def as_matrix(self):  # binary_vector.rs:130-132: 1 x n
        return BinMatrix.new([self])

    def as_column_matrix(self):  # binary_vector.rs:135-137: n x 1
        return self.as_matrix().transposed()

    def __eq__(self, other):
        return isinstance(other, BinVector) and len(self) == len(other) and bool((self.bits == other.bits).all())

    def __add__(self, other):  # XOR, binary_vector.rs:156-192
        assert len(self) == len(other), "unequal length vectors"
        return BinVector(self.bits ^ other.bits)

    def __mul__(self, other):
        if isinstance(other, BinMatrix):  # v^T * A, binary_matrix.rs:552-563
            return (self.as_matrix() * other).as_vector()
        return bool(int((self.bits & other.bits).sum()) & 1)  # binary_vector.rs:194-215

    def __repr__(self):
        return "BinVector(" + "".join(str(int(b)) for b in self.bits) + ")"


class BinMatrix:
    """Mirror of m4ri_rust::friendly::BinMatrix (binary_matrix.rs:33-586): an owning
    wrapper around an `mzd_t*`; Drop -> mzd_free (binary_matrix.rs:41-45)."""

    # which entry point `*` uses: the reference picks it with a Cargo feature
    # (binary_matrix.rs:53-95); default = strassen = mzd_mul.
    mul_strategy = "strassen_mul"

    def __init__(self, mzd):
        if not mzd:
            raise ValueError("Can't be NULL")  # from_mzd, binary_matrix.rs:203
        self.mzd = mzd

    def __del__(self):
        try:
            if self.mzd:
                lib().mzd_free(self.mzd)
                self.mzd = None
        except Exception:
            pass

    # ---- constructors: binary_matrix.rs:99-216 ----
    @classmethod
    def zero(cls, rows, cols):
        if rows == 0 or cols == 0:
            raise ValueError("Can't create a 0 matrix")
        return cls(lib().mzd_init(rows, cols))

    @classmethod
    def new(cls, rows):
        rowlen = len(rows[0])
        return cls.from_slices([r.get_storage() for r in rows], rowlen)

    @classmethod
    def from_slices(cls, rows, rowlen):
        if len(rows) == 0 or rowlen == 0:
            raise ValueError("Can't create a 0 matrix")
        w = (rowlen + 63) // 64
        words = np.zeros((len(rows), max(w, 1)), dtype=np.uint64)
        for r, row in enumerate(rows):
            row = np.asarray(row, dtype=np.uint64)
            assert row.size * 64 >= rowlen, f"expected len {rowlen} bits but got only {row.size} blocks"
            words[r, :w] = row[:w]
        return cls(mzd_from_words(words, rowlen))

    @classmethod
    def random(cls, rows, cols):
        p = lib().mzd_init(rows, cols)
        lib().mzd_randomize(p)
        return cls(p)

    @classmethod
    def from_mzd(cls, mzd):
        return cls(mzd)

    @classmethod
    def identity(cls, rows):
        p = lib().mzd_init(rows, rows)
        lib().mzd_set_ui(p, 1)
        return cls(p)

    # ---- accessors ----
    def nrows(self):
        return int(self.mzd.contents.nrows)

    def ncols(self):
        return int(self.mzd.contents.ncols)

    def words(self):
        return mzd_words(self.mzd)

    def bit(self, row, col):  # mzd_read_bit, mzd.rs:263-269
        return bool((int(self.mzd.contents.rows[row][col // 64]) >> (col % 64)) & 1)

    def get_word(self, row, column):
        assert row < self.nrows() and column < self.ncols()
        return int(self.mzd.contents.rows[row][column])

    def count_ones(self):  # binary_matrix.rs:171-189
        assert self.nrows() == 1 or self.ncols() == 1, "only works on single row or single column matrices"
        w = self.mzd.contents.width
        return int(sum(bin(int(x)).count("1") for x in self.words()[:, :w].ravel()))

    def augmented(self, other):  # binary_matrix.rs:218-231
        assert self.nrows() == other.nrows(), "The rows need to be equal"
        return BinMatrix(lib().mzd_concat(None, self.mzd, other.mzd))

    def stacked(self, other):  # binary_matrix.rs:233-244
        return BinMatrix(lib().mzd_stack(None, self.mzd, other.mzd))

    def rank(self):  # binary_matrix.rs:246-252: echelonizes a clone
        return self.clone().echelonize()

    def echelonize(self):  # binary_matrix.rs:254-261 (in place, returns the rank)
        return int(lib().mzd_echelonize(self.mzd, 0))

    def inverted(self):  # binary_matrix.rs:263-268
        p = lib().mzd_inv_m4ri(None, self.mzd, 0)
        if not p:
            raise ValueError("Can't be NULL")  # nonnull! panics on a singular matrix
        return BinMatrix(p)

    def transposed(self):  # binary_matrix.rs:272-279
        return BinMatrix(lib().mzd_transpose(None, self.mzd))

    def clone(self):
        return BinMatrix(lib().mzd_copy(None, self.mzd))

    def as_vector(self):  # binary_matrix.rs:332-361
        if self.nrows() != 1:
            assert self.ncols() == 1, "needs to have only one column or row"
            return self.transposed().as_vector()
        w = self.mzd.contents.width
        return BinVector.from_storage(self.words()[0, :w], self.ncols())

    def get_window(self, start_row, start_col, high_row, high_col):  # binary_matrix.rs:371-392
        out = BinMatrix.zero(high_row - start_row, high_col - start_col)
        for r in range(start_row, high_row):
            for c in range(start_col, high_col):
                if self.bit(r, c):
                    out.mzd.contents.rows[r - start_row][(c - start_col) // 64] |= 1 << ((c - start_col) % 64)
        return out

    def set_window(self, start_row, start_col, other):  # binary_matrix.rs:398-413
        assert self.ncols() >= start_col + other.ncols(), "This matrix is too small!"
        assert self.nrows() >= start_row + other.nrows(), "This matrix has too few rows !"
        rows = self.mzd.contents.rows
        for r in range(other.nrows()):
            for c in range(other.ncols()):
                rr, cc = r + start_row, c + start_col
                v = int(rows[rr][cc // 64]) & ~(1 << (cc % 64))
                rows[rr][cc // 64] = v | (int(other.bit(r, c)) << (cc % 64))

    # ---- the hot path ----
    def mul_slice(self, other):  # binary_matrix.rs:416-431
        assert self.ncols() <= len(other) * 64, "Mismatched sizes"
        vec = BinMatrix.from_slices([other], self.ncols()).transposed()
        return BinMatrix(lib().mzd_mul_naive(None, self.mzd, vec.mzd))

    def __mul__(self, other):
        L = lib()
        if isinstance(other, BinVector):  # A * v^T, binary_matrix.rs:528-542
            return self.mul_slice(other.get_storage()).as_vector()
        if self.mul_strategy == "m4rm_mul":  # mul_impl!, binary_matrix.rs:53-95
            p = L.mzd_mul_m4rm(None, self.mzd, other.mzd, 0)
        elif self.mul_strategy == "naive_mul":
            p = L.mzd_mul_naive(None, self.mzd, other.mzd)
        else:
            p = L.mzd_mul(None, self.mzd, other.mzd, 0)
        if not p:
            raise RuntimeError("Multiplication failed")  # binary_matrix.rs:467-469
        return BinMatrix(p)

    def __add__(self, other):  # binary_matrix.rs:474-489
        return BinMatrix(lib().mzd_add(None, self.mzd, other.mzd))

    def __iadd__(self, other):  # binary_matrix.rs:507-526, C aliases A
        lib().mzd_add(self.mzd, self.mzd, other.mzd)
        return self

    def __eq__(self, other):  # binary_matrix.rs:434-438
        return isinstance(other, BinMatrix) and lib().mzd_equal(self.mzd, other.mzd) == 1


def solve_left(a, b):
    """Solve AX = B for X, B is overwritten with the solution; True on success
    (binary_matrix.rs:575-586)."""
    return lib().mzd_solve_left(a.mzd, b.mzd, 0, 1) == 0


# ---------------------------------------------------------------------------------
# device-resident matrices (additive API)
# ---------------------------------------------------------------------------------
class DMat:
    def __init__(self, handle, keepalive=None):
        self.h = handle
        self._keepalive = keepalive

    @classmethod
    def init(cls, nrows, ncols):
        return cls(lib().m4ri_b200_dmat_init(nrows, ncols))

    @classmethod
    def wrap(cls, dptr, nrows, ncols, rowstride, keepalive=None):
        return cls(lib().m4ri_b200_dmat_wrap(dptr, nrows, ncols, rowstride), keepalive)

    @classmethod
    def from_host(cls, mzd):
        d = cls.init(mzd.contents.nrows, mzd.contents.ncols)
        lib().m4ri_b200_upload(d.h, mzd)
        return d

    def __del__(self):
        try:
            if self.h:
                lib().m4ri_b200_dmat_free(self.h)
                self.h = None
        except Exception:
            pass

    nrows = property(lambda self: lib().m4ri_b200_dmat_nrows(self.h))
    ncols = property(lambda self: lib().m4ri_b200_dmat_ncols(self.h))
    rowstride = property(lambda self: lib().m4ri_b200_dmat_rowstride(self.h))
    ptr = property(lambda self: lib().m4ri_b200_dmat_ptr(self.h))

    def to_host(self):
        p = lib().mzd_init(self.nrows, self.ncols)
        lib().m4ri_b200_download(p, self.h)
        lib().m4ri_b200_synchronize()
        return p

    def to_words(self):
        p = self.to_host()
        try:
            return mzd_words(p).copy()
        finally:
            lib().mzd_free(p)

    def fill_random(self, seed):
        lib().m4ri_b200_dfill_random(self.h, seed)
        return self

    def checksum(self):
        return int(lib().m4ri_b200_dchecksum(self.h))


def dmul(C, A, B, clear=True, algo=ALGO_AUTO):
    lib().m4ri_b200_dmul(C.h, A.h, B.h, int(clear), algo)


def dmul_abt(C, A, Bt, clear=True, algo=ALGO_AUTO):
    lib().m4ri_b200_dmul_abt(C.h, A.h, Bt.h, int(clear), algo)
