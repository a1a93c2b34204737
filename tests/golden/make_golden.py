"""Generates tests/golden/oracle_products.npz.

The reference cannot run here (its arithmetic is the absent M4RI submodule), so
these vectors are NOT reference outputs: they are products computed from the
definition by oracle_mul_naive (bit triple loop) and cross-checked with numpy's
integer matmul mod 2 before being written. They pin the oracle and the CUDA
path against regressions; the reference-pinned answers are in
reference_known_answers.json.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

# (m, k, n): the reference's bench shapes (m4ri-rust/benches/binary_matrix.rs:30-76)
# plus odd widths / stride-padding cases.
SHAPES = [(10, 10, 10), (100, 10, 10), (100, 10, 100), (1000, 64, 1000), (1000, 10, 1000),
          (1000, 64, 1), (1, 64, 1000), (10, 1000, 10), (1, 1000, 64), (64, 1000, 1),
          (34, 128, 1), (65, 129, 193), (130, 191, 257), (3, 200, 5)]


def main():
    out = {}
    for idx, (m, k, n) in enumerate(SHAPES):
        A = oracle.random(m, k, 0x4D345249 ^ (idx << 32) ^ 1)
        B = oracle.random(k, n, 0x4D345249 ^ (idx << 32) ^ 2)
        C = oracle.mul_naive(A, B, k, n)
        a = oracle.to_bits(A, k).astype(np.int64)
        b = oracle.to_bits(B, n).astype(np.int64)
        assert ((a @ b) % 2 == oracle.to_bits(C, n)).all()
        out[f"shape_{idx}"] = np.array([m, k, n], dtype=np.int64)
        out[f"C_{idx}"] = C
        out[f"chk_{idx}"] = np.array([oracle.checksum(A, k), oracle.checksum(B, n), oracle.checksum(C, n)], dtype=np.uint64)
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "oracle_products.npz"), **out)
    print("wrote", len(SHAPES), "cases")


if __name__ == "__main__":
    main()
