"""ctypes loader for the CPU oracle (oracle/gf2_oracle.c).

TEST INFRASTRUCTURE ONLY. May be imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs, never by the product package.
Matrices are numpy uint64 arrays of shape [nrows, stride] (64-bit words,
LSB-first, row-major: m4ri-sys/src/mzd.rs:263-269 in the reference).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libgf2_oracle.so")
_lib = None


def build(force=False):
    src = os.path.join(_HERE, "gf2_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "all"] + (["-B"] if force else []))
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = ctypes.CDLL(_SO)
        u64p, i, u64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_uint64
        sig = {
            "oracle_fill_random": (None, [u64p, i, i, i, u64]),
            "oracle_checksum": (u64, [u64p, i, i, i]),
            "oracle_mul_naive": (None, [u64p, i, u64p, i, u64p, i, i, i, i, i]),
            "oracle_mul_m4rm": (None, [u64p, i, u64p, i, u64p, i, i, i, i, i, i]),
            "oracle_mul_strassen": (None, [u64p, i, u64p, i, u64p, i, i, i, i, i, i, i]),
            "oracle_mul_abt": (None, [u64p, i, u64p, i, u64p, i, i, i, i, i]),
            "oracle_mul_va": (None, [u64p, i, u64p, i, u64p, i, i, i, i, i]),
            "oracle_mul_v": (None, [u64p, u64p, i, u64p, i, i, i]),
            "oracle_transpose": (None, [u64p, i, u64p, i, i, i]),
            "oracle_add": (None, [u64p, i, u64p, i, u64p, i, i, i]),
            "oracle_freivalds": (i, [u64p, i, u64p, i, u64p, i, i, i, i, u64]),
            "oracle_max_threads": (i, []),
            "oracle_rref": (i, [u64p, i, i, i, i, u64p]),
            "oracle_ple": (i, [u64p, i, i, i, u64p, u64p, i]),
        }
        for name, (res, args) in sig.items():
            f = getattr(_lib, name)
            f.restype, f.argtypes = res, args
    return _lib


def width(ncols):
    return (ncols + 63) // 64


def m4ri_rowstride(ncols):
    """rowstride rule of the reference: m4ri-sys/src/mzd.rs:35-37 (paddingwidth 3)."""
    w = width(ncols)
    return w if (w < 3 or w % 2 == 0) else w + 1


def zeros(nrows, ncols, stride=None):
    return np.zeros((nrows, stride or m4ri_rowstride(ncols)), dtype=np.uint64)


def _p(a):
    assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data


def random(nrows, ncols, seed, stride=None):
    m = zeros(nrows, ncols, stride)
    lib().oracle_fill_random(_p(m), nrows, ncols, m.shape[1], seed)
    return m


def checksum(M, ncols):
    return int(lib().oracle_checksum(_p(M), M.shape[0], ncols, M.shape[1]))


def mul_naive(A, B, k, n, C=None):
    m = A.shape[0]
    clear = C is None
    C = zeros(m, n) if C is None else C
    lib().oracle_mul_naive(_p(C), C.shape[1], _p(A), A.shape[1], _p(B), B.shape[1], m, k, n, int(clear))
    return C


def mul_m4rm(A, B, k, n, C=None, threads=1):
    m = A.shape[0]
    clear = C is None
    C = zeros(m, n) if C is None else C
    lib().oracle_mul_m4rm(_p(C), C.shape[1], _p(A), A.shape[1], _p(B), B.shape[1], m, k, n, int(clear), threads)
    return C


def mul_strassen(A, B, k, n, C=None, cutoff=0, threads=0):
    """Strassen-Winograd above `cutoff` (0 = 2048) with Four-Russians leaves: the strategy behind
    the reference's default mzd_mul (m4ri-sys/src/strassen.rs:5-18)."""
    m = A.shape[0]
    clear = C is None
    C = zeros(m, n) if C is None else C
    lib().oracle_mul_strassen(_p(C), C.shape[1], _p(A), A.shape[1], _p(B), B.shape[1], m, k, n, int(clear),
                              int(cutoff), threads)
    return C


def mul_abt(A, Bt, k, C=None):
    m, n = A.shape[0], Bt.shape[0]
    clear = C is None
    C = zeros(m, n) if C is None else C
    lib().oracle_mul_abt(_p(C), C.shape[1], _p(A), A.shape[1], _p(Bt), Bt.shape[1], m, k, n, int(clear))
    return C


def mul_va(v, A, k, n, C=None):
    r = v.shape[0]
    clear = C is None
    C = zeros(r, n) if C is None else C
    lib().oracle_mul_va(_p(C), C.shape[1], _p(v), v.shape[1], _p(A), A.shape[1], r, k, n, int(clear))
    return C


def mul_v(A, x, n, y=None):
    """y = A*x with x, y packed rows (1-D uint64 arrays)."""
    m = A.shape[0]
    clear = y is None
    y = np.zeros(width(m), dtype=np.uint64) if y is None else y
    lib().oracle_mul_v(_p(y), _p(A), A.shape[1], _p(np.ascontiguousarray(x)), m, n, int(clear))
    return y


def transpose(S, ncols):
    nrows = S.shape[0]
    D = zeros(ncols, nrows)
    lib().oracle_transpose(_p(D), D.shape[1], _p(S), S.shape[1], nrows, ncols)
    return D


def add(A, B, ncols):
    C = zeros(A.shape[0], ncols)
    lib().oracle_add(_p(C), C.shape[1], _p(A), A.shape[1], _p(B), B.shape[1], A.shape[0], ncols)
    return C


def freivalds(C, A, B, k, n, seed=0xF5E1):
    return bool(lib().oracle_freivalds(_p(C), C.shape[1], _p(A), A.shape[1], _p(B), B.shape[1], A.shape[0], k, n, seed))


def max_threads():
    return int(lib().oracle_max_threads())


def get_bit(M, r, c):
    return int((int(M[r, c // 64]) >> (c % 64)) & 1)


def from_bits(bits):
    """bits: 2-D array-like of 0/1 -> packed matrix with the reference's rowstride."""
    bits = np.asarray(bits, dtype=np.uint8)
    nrows, ncols = bits.shape
    M = zeros(nrows, ncols)
    w = width(ncols)
    if nrows and w:
        padded = np.zeros((nrows, w * 64), dtype=np.uint8)
        padded[:, :ncols] = bits & 1
        M[:, :w] = np.packbits(padded, axis=1, bitorder="little").view(np.uint64).reshape(nrows, w)
    return M


def to_bits(M, ncols):
    nrows = M.shape[0]
    out = np.zeros((nrows, ncols), dtype=np.uint8)
    for c in range(ncols):
        out[:, c] = (M[:, c // 64] >> np.uint64(c % 64)) & np.uint64(1)
    return out


def rref(M, n, pivot_cols=None):
    """Reduced row echelon form of M (m x n bits): (R, rank, pivot columns). Pivots only in
    columns [0, pivot_cols)."""
    R = np.ascontiguousarray(M).copy()
    m = R.shape[0]
    piv = np.zeros(max(1, min(m, n)), dtype=np.int32)
    r = int(lib().oracle_rref(_p(R), R.shape[1], m, n, n if pivot_cols is None else pivot_cols, piv.ctypes.data))
    return R, r, piv[:r].copy()


def concat_bits(A, na, B, nb):
    """[A | B] at arbitrary bit offsets (rows of na and nb valid bits)."""
    a = np.unpackbits(np.ascontiguousarray(A[:, : width(na)]).view(np.uint8), axis=1, bitorder="little")[:, :na]
    b = np.unpackbits(np.ascontiguousarray(B[:, : width(nb)]).view(np.uint8), axis=1, bitorder="little")[:, :nb]
    return from_bits(np.concatenate([a, b], axis=1))


def inverse(A, n):
    """A^-1 (n x n) or None when A is singular: RREF of [A | I]."""
    I = np.zeros((n, width(n)), dtype=np.uint64)
    for i in range(n):
        I[i, i // 64] |= np.uint64(1) << np.uint64(i % 64)
    R, r, piv = rref(concat_bits(A, n, I, n), 2 * n, pivot_cols=n)
    if r < n:
        return None
    bits = np.unpackbits(R.view(np.uint8), axis=1, bitorder="little")[:, n:2 * n]
    return from_bits(bits)


def solve_left(A, n, B, k):
    """X (n x k) with A X = B and the free variables zero, or None when inconsistent."""
    m = A.shape[0]
    R, r, piv = rref(concat_bits(A, n, B, k), n + k, pivot_cols=n)
    bits = np.unpackbits(R.view(np.uint8), axis=1, bitorder="little")[:, n:n + k]
    if bits[r:].any():
        return None
    X = np.zeros((n, k), dtype=np.uint8)
    for t in range(r):
        X[piv[t]] = bits[t]
    return from_bits(X)


def ple(M, n, pluq=False):
    """PLE (or PLUQ) decomposition from the definition (m4ri-sys/src/ple.rs:35,65):
    (packed [L \\ E], rank, P, Q) with P, Q LAPACK-style transposition lists."""
    R = np.ascontiguousarray(M).copy()
    m = R.shape[0]
    P = np.zeros(max(1, m), dtype=np.int32)
    Q = np.zeros(max(1, n), dtype=np.int32)
    r = int(lib().oracle_ple(_p(R), R.shape[1], m, n, P.ctypes.data, Q.ctypes.data, int(pluq)))
    return R, r, P[:m].copy(), Q[:n].copy()


def ple_reconstruct(packed, n, r, P, Q, pluq=False):
    """The matrix a packed PLE / PLUQ result stands for: unpack L (m x r, unit diagonal) and E
    (r x n; for pluq U with the column swaps undone), check their shape constraints, and return
    P * L * E as packed words. Raises AssertionError when the storage contract is violated."""
    m = packed.shape[0]
    bits = to_bits(packed, n).astype(np.uint8)
    Q = [int(q) for q in Q]
    P = [int(p) for p in P]
    assert all(Q[i] < Q[i + 1] for i in range(r - 1)), "pivot columns must increase"
    assert all(Q[i] >= i for i in range(r)) and all(Q[i] == i for i in range(r, n)), "Q beyond the rank is the identity"
    assert all(P[i] >= i for i in range(m)), "transpositions point forward"
    L = np.zeros((m, max(r, 1)), dtype=np.uint8)
    E = np.zeros((max(r, 1), n), dtype=np.uint8)
    for i in range(m):
        k = min(i, r)
        L[i, :k] = bits[i, :k]
        if i < r:
            L[i, i] = 1
            start = i if pluq else Q[i]
            E[i, start:] = bits[i, start:]
            assert not bits[i, k:start].any(), f"row {i}: bits between L and the pivot column"
        else:
            assert not bits[i, k:].any(), f"row {i} (below the rank) holds more than multipliers"
    if pluq:
        for i in range(r):
            assert E[i, i] == 1, f"U[{i}][{i}] must be 1"
        for i in reversed(range(r)):  # undo the column swaps
            if Q[i] != i:
                E[:, [i, Q[i]]] = E[:, [Q[i], i]]
    for i in range(r):
        assert E[i, Q[i]] == 1 and not E[i, : Q[i]].any(), f"E row {i} must start at its pivot column"
    if r == 0:
        prod = np.zeros((m, n), dtype=np.uint8)
    else:
        prod = to_bits(mul_m4rm(from_bits(L[:, :r]), from_bits(E[:r]), r, n), n)
    for i in reversed(range(m)):  # A = P (L E): undo the row swaps
        if P[i] != i:
            prod[[i, P[i]]] = prod[[P[i], i]]
    return from_bits(prod)
