"""The shapes of the reference's own micro-benchmarks (m4ri-rust/benches/binary_matrix.rs:27-76)
through the Python mirror of its operator API: &m1 * &m2, &m * &v, &v * &m, mul_slice.
usage: python tools/ref_bench_shapes.py"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import m4ri_rust_b200 as m4
L = m4.lib()

def med(fn, reps=40):
    for _ in range(5):
        fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        ts.append((time.perf_counter() - t0) * 1e6)
    ts.sort()
    return round(ts[len(ts) // 2], 1)

for (a, b, d) in [(10, 10, 10), (100, 10, 10), (100, 10, 100), (1000, 64, 1000), (1000, 10, 1000), (1000, 64, 1),
                  (1, 64, 1000), (10, 1000, 10)]:
    m1, m2 = L.mzd_init(a, b), L.mzd_init(b, d)
    L.mzd_randomize(m1); L.mzd_randomize(m2)
    print(json.dumps({"bench": f"matrix_multiply_{a}x{b}_{b}x{d}", "us": med(lambda: L.mzd_free(L.mzd_mul(None, m1, m2, 0)))}), flush=True)
    L.mzd_free(m1); L.mzd_free(m2)
m = L.mzd_init(64, 1000); x = L.mzd_init(1000, 1)
L.mzd_randomize(m); L.mzd_randomize(x)
print(json.dumps({"bench": "matrix_vector 64x1000 (mzd_mul_naive with a column)", "us": med(lambda: L.mzd_free(L.mzd_mul_naive(None, m, x)))}), flush=True)
m34 = L.mzd_init(34, 128); x128 = L.mzd_init(128, 1)
L.mzd_randomize(m34); L.mzd_randomize(x128)
print(json.dumps({"bench": "matrix_mul_slice 34x128", "us": med(lambda: L.mzd_free(L.mzd_mul_naive(None, m34, x128)))}), flush=True)
