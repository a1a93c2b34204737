"""Generates tests/golden/full_size_checksums.json: the checksum (oracle.checksum, the same
position-mixed XOR fold m4ri_b200_dchecksum computes) of the CPU oracle's product of bench.py's
seeded operands at the benchmark sizes. The inputs are splitmix64 streams (oracle.random), so the
result is a constant: bench.py compares the GPU result's checksum with it at every N.

    python tests/golden/make_full_size_checksums.py [n ...]     (n = 65536: ~1.5 GB, a minute)
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

SEED_A, SEED_B = 0x4D345249 ^ (2 << 32) ^ 1, 0x4D345249 ^ (2 << 32) ^ 2  # bench.py:39


def main():
    sizes = [int(x) for x in sys.argv[1:]] or [32768, 65536]
    path = os.path.join(ROOT, "tests", "golden", "full_size_checksums.json")
    out = json.load(open(path)) if os.path.exists(path) else {}
    for n in sizes:
        A, B = oracle.random(n, n, SEED_A), oracle.random(n, n, SEED_B)
        t0 = time.perf_counter()
        C = oracle.mul_strassen(A, B, n, n, threads=0)
        out[str(n)] = {"checksum": f"{oracle.checksum(C, n):016x}", "seed_a": SEED_A, "seed_b": SEED_B,
                       "by": "oracle.mul_strassen (oracle/gf2_oracle.c)", "seconds": round(time.perf_counter() - t0, 1)}
        print(n, out[str(n)], flush=True)
        del A, B, C
    json.dump(out, open(path, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
