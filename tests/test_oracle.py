"""CPU tests of the oracle (oracle/gf2_oracle.c): pinned against everything the
reference's own tests hold for the path (tests/golden/reference_known_answers.json,
transcribed from m4ri-rust/src/friendly/binary_matrix.rs:661-699 and
m4ri-sys/src/mzd.rs:385-439), against the committed golden products, and its two
independent restatements (definition vs Four Russians) against each other."""
import json
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")


def identity(orc, n):
    return orc.from_bits(np.eye(n, dtype=np.uint8))


def ones_row(orc, n):
    return orc.from_bits(np.ones((1, n), dtype=np.uint8))


def test_reference_known_answers(orc):
    cases = json.load(open(os.path.join(GOLD, "reference_known_answers.json")))["cases"]
    seen = set()
    for c in cases:
        op = c["op"]
        seen.add(op)
        if op == "mul":  # binary_matrix.rs:661-670
            I = identity(orc, 8)
            for f in (orc.mul_naive, orc.mul_m4rm):
                assert (f(I, I, 8, 8) == I).all()
        elif op == "matvec":  # binary_matrix.rs:677-678
            I = identity(orc, 10)
            x = ones_row(orc, 10)[0]
            y = orc.mul_v(I, x, 10)
            assert int(y[0]) == (1 << 10) - 1
        elif op == "vecmat":  # binary_matrix.rs:680-681
            I = identity(orc, 10)
            v = ones_row(orc, 10)
            assert (orc.mul_va(v, I, 10, 10) == v).all()
            assert (orc.mul_m4rm(v, I, 10, 10) == v).all()
        elif op == "vecmat_shape":  # binary_matrix.rs:683-685
            A = orc.random(10, 3, 99)
            out = orc.mul_va(ones_row(orc, 10), A, 10, 3)
            assert out.shape[0] == 1 and orc.to_bits(out, 3).shape == (1, 3)
        elif op == "words":  # binary_matrix.rs:693-699: column j is bit j, LSB first
            I = identity(orc, 3)
            assert [[int(I[r, 0])] for r in range(3)] == c["expect_rows"]
        elif op == "read_bit_identity":  # mzd.rs:425-439
            n = c["n"]
            I = identity(orc, n)
            bits = orc.to_bits(I, n)
            assert (bits == np.eye(n, dtype=np.uint8)).all()
        elif op in ("sizeof", "from_bytes"):
            pass  # checked against the library / BinVector mirror in test_abi.py
        else:
            raise AssertionError(f"unknown known-answer op {op}")
    assert {"mul", "matvec", "vecmat", "words"} <= seen


def test_golden_products(orc):
    g = np.load(os.path.join(GOLD, "oracle_products.npz"))
    n_cases = sum(1 for k in g.files if k.startswith("shape_"))
    assert n_cases >= 14
    for idx in range(n_cases):
        m, k, n = (int(x) for x in g[f"shape_{idx}"])
        A = orc.random(m, k, 0x4D345249 ^ (idx << 32) ^ 1)
        B = orc.random(k, n, 0x4D345249 ^ (idx << 32) ^ 2)
        chk = g[f"chk_{idx}"]
        assert orc.checksum(A, k) == int(chk[0]) and orc.checksum(B, n) == int(chk[1])
        C = g[f"C_{idx}"]
        assert (orc.mul_naive(A, B, k, n) == C).all()
        assert (orc.mul_m4rm(A, B, k, n) == C).all()
        assert (orc.mul_m4rm(A, B, k, n, threads=2) == C).all()
        assert (orc.mul_strassen(A, B, k, n, cutoff=64, threads=2) == C).all()
        assert orc.checksum(C, n) == int(chk[2])
        Bt = orc.transpose(B, n)
        assert (orc.mul_abt(A, Bt, k) == C).all()
        if m <= 4:
            assert (orc.mul_va(A, B, k, n) == C).all()


SHAPES = [(1, 1, 1), (1, 64, 1), (2, 63, 65), (7, 129, 3), (65, 64, 64), (64, 65, 64),
          (64, 64, 65), (33, 31, 200), (191, 193, 129), (200, 257, 321), (5, 700, 1100)]


@pytest.mark.parametrize("m,k,n", SHAPES)
def test_m4rm_matches_definition(orc, m, k, n):
    A = orc.random(m, k, 1000 + m)
    B = orc.random(k, n, 2000 + n)
    C = orc.mul_naive(A, B, k, n)
    assert (orc.mul_m4rm(A, B, k, n) == C).all()
    # numpy integer matmul mod 2 as a third opinion
    ref = (orc.to_bits(A, k).astype(np.int64) @ orc.to_bits(B, n).astype(np.int64)) % 2
    assert (orc.to_bits(C, n) == ref).all()
    # excess bits and padding words stay zero
    w = orc.width(n)
    if n % 64:
        assert not (C[:, w - 1] >> np.uint64(n % 64)).any()
    assert not C[:, w:].any()


def test_accumulate_and_clear(orc):
    m, k, n = 70, 90, 130
    A, B = orc.random(m, k, 5), orc.random(k, n, 6)
    P = orc.mul_naive(A, B, k, n)
    C0 = orc.random(m, n, 7)
    for f in (orc.mul_naive, orc.mul_m4rm):
        C = C0.copy()
        f(A, B, k, n, C=C)
        assert (C == (C0 ^ P)).all()
        f(A, B, k, n, C=C)
        assert (C == C0).all()  # addmul twice restores C


def test_transposed_product_and_identities(orc):
    m, k, n = 50, 77, 131
    A, B = orc.random(m, k, 11), orc.random(k, n, 12)
    C = orc.mul_m4rm(A, B, k, n)
    # (A*B)^T == B^T * A^T ties the pre-transposed form to the plain one
    Ct = orc.mul_m4rm(orc.transpose(B, n), orc.transpose(A, k), k, m)
    assert (orc.transpose(C, n) == Ct).all()
    assert (orc.transpose(orc.transpose(A, k), m) == A).all()
    # distributivity: A*(B^B2) == A*B ^ A*B2
    B2 = orc.random(k, n, 13)
    lhs = orc.mul_m4rm(A, orc.add(B, B2, n), k, n)
    assert (lhs == orc.add(C, orc.mul_m4rm(A, B2, k, n), n)).all()
    # mat-vec equals column 0 of the product with [x|0...]
    x = orc.random(1, k, 14)[0]
    X = orc.transpose(x.reshape(1, -1), k)  # k x 1
    col = orc.mul_m4rm(A, X, k, 1)
    y = orc.mul_v(A, x, k)
    assert [orc.get_bit(col, i, 0) for i in range(m)] == [int(y[i // 64] >> np.uint64(i % 64)) & 1 for i in range(m)]


def test_freivalds_accepts_and_rejects(orc):
    m, k, n = 100, 120, 140
    A, B = orc.random(m, k, 21), orc.random(k, n, 22)
    C = orc.mul_m4rm(A, B, k, n)
    assert orc.freivalds(C, A, B, k, n)
    C[37, 1] ^= np.uint64(1) << np.uint64(5)
    assert not orc.freivalds(C, A, B, k, n)


def test_medium_threaded(orc):
    n = 1024
    A, B = orc.random(n, n, 31), orc.random(n, n, 32)
    C1 = orc.mul_m4rm(A, B, n, n, threads=1)
    C4 = orc.mul_m4rm(A, B, n, n, threads=4)
    assert (C1 == C4).all() and orc.freivalds(C1, A, B, n, n)


@pytest.mark.parametrize("m,k,n,cutoff", [(256, 256, 256, 64), (300, 500, 260, 64), (257, 129, 131, 64),
                                          (640, 384, 512, 64), (1000, 777, 1300, 128), (129, 128, 128, 64),
                                          (1024, 1024, 1024, 256), (100, 100, 100, 64)])
def test_strassen_strategy_matches_definition(orc, m, k, n, cutoff):
    """The Strassen-Winograd strategy (the reference's default mzd_mul, strassen.rs:5-18) against
    the definition: even cores of one and two levels, odd tails in every dimension (peeled to the
    Four Russians), accumulate, excess bits."""
    A, B = orc.random(m, k, 300 + m), orc.random(k, n, 400 + n)
    want = orc.mul_naive(A, B, k, n) if m * k * n <= 300 * 500 * 260 else orc.mul_m4rm(A, B, k, n)
    got = orc.mul_strassen(A, B, k, n, cutoff=cutoff, threads=3)
    assert (got == want).all()
    C0 = orc.random(m, n, 7)
    C = C0.copy()
    orc.mul_strassen(A, B, k, n, C=C, cutoff=cutoff, threads=2)
    assert (C == (C0 ^ want)).all()
    w = orc.width(n)
    if n % 64:
        assert not (got[:, w - 1] >> np.uint64(n % 64)).any()
    assert not got[:, w:].any()


def test_ple_restatement_satisfies_its_definition(orc):
    """oracle_ple (m4ri-sys/src/ple.rs:35,65): P * L * E reproduces the input through the storage
    contract, the rank and the pivot columns agree with the Gauss-Jordan restatement."""
    for (m, n, kind) in [(1, 1, "r"), (5, 7, "r"), (64, 64, "r"), (100, 300, "r"), (300, 100, "r"),
                         (200, 200, "low"), (130, 257, "zc"), (65, 63, "r")]:
        if kind == "r":
            A = orc.random(m, n, 3 + m)
        elif kind == "low":
            A = orc.mul_m4rm(orc.random(m, 20, 1), orc.random(20, n, 2), 20, n)
        else:
            A = orc.random(m, n, 7)
            A[:, : orc.width(n) // 2] = 0
            A[m // 2:] = A[: m - m // 2]
        _, rank, piv = orc.rref(A, n)
        for pluq in (False, True):
            R, r, P, Q = orc.ple(A, n, pluq)
            assert r == rank and list(Q[:r]) == list(piv)
            back = orc.ple_reconstruct(R, n, r, P, Q, pluq)
            assert (back[:, : orc.width(n)] == A[:, : orc.width(n)]).all(), (m, n, kind, pluq)
    # identity: no swaps, L = I, E = I
    n = 70
    I = orc.from_bits(np.eye(n, dtype=np.uint8))
    R, r, P, Q = orc.ple(I, n)
    assert r == n and list(P) == list(range(n)) and list(Q) == list(range(n)) and (R == I).all()


def test_committed_full_size_checksum_is_reproducible(orc):
    """tests/golden/full_size_checksums.json (made by make_full_size_checksums.py) holds the checksum
    of the oracle's product of bench.py's seeded operands at the benchmark sizes; bench.py holds the
    GPU result of every N to it. Recompute the n = 32768 entry here (a few seconds on all cores) with
    the OTHER oracle strategy, Four Russians without the Strassen recursion, on a row sample, and
    the full Strassen product's checksum itself."""
    import json
    import os
    import bench
    golden = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "full_size_checksums.json")))
    assert set(golden) >= {"32768", "65536"}
    for e in golden.values():
        assert (e["seed_a"], e["seed_b"]) == (bench.SEED_A, bench.SEED_B)
    n = 32768
    A, B = orc.random(n, n, bench.SEED_A), orc.random(n, n, bench.SEED_B)
    C = orc.mul_strassen(A, B, n, n, threads=0)
    assert f"{orc.checksum(C, n):016x}" == golden[str(n)]["checksum"]
    rows = np.arange(0, n, 997)  # the Strassen result against plain Four Russians on a sample of rows
    assert (orc.mul_m4rm(np.ascontiguousarray(A[rows]), B, n, n, threads=0) == C[rows]).all()
