"""Measured rate of the tcgen05 kind::mxf4 pipe under THIS kernel's issue pattern and clocks
(MEASURED_PEAKS.json has bf16 only). Perf-hook build only
(python m4ri-rust_b200/build.py --perf): the default tensor kernel is run at the Strassen leaf
shape with
  DBG=0   everything on                                  -> the kernel as shipped
  DBG=5   no operand expansion, no accumulator drain     -> MMA issue + the barrier protocol
  DBG=13  ... and the issuer never waits for the epilogue -> back-to-back MMAs, the pipe itself
Each mode is its own process (the hook is read once). Results are WRONG by construction for
DBG != 0; only the time is used. Output: one JSON object (-> profiles/peaks_measured.json)."""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "--one":
    sys.path.insert(0, ROOT)
    import m4ri_rust_b200 as m4
    L = m4.lib()
    n = int(sys.argv[2])
    L.m4ri_b200_set_strassen_levels(0)
    dA, dBt, dC = (m4.DMat.init(n, n) for _ in range(3))
    dA.fill_random(1); dBt.fill_random(2)
    for _ in range(3):
        m4.dmul_abt(dC, dA, dBt, clear=True, algo=m4.ALGO_MMA)
    L.m4ri_b200_synchronize()
    best = 1e9
    for _ in range(5):
        L.m4ri_b200_timing_reset()
        for _ in range(5):
            m4.dmul_abt(dC, dA, dBt, clear=True, algo=m4.ALGO_MMA)
        best = min(best, L.m4ri_b200_timing_ms() / 5)
    print(json.dumps({"ms": best}))
    sys.exit(0)
n = int(os.environ.get("N", "16384"))
out = {"fp4_peak_shape": f"{n}x{n}x{n}", "fp4_peak_kernel": "mma6_kernel<2,WIDE,FULL> (fp4s256)"}
for dbg, key in ((0, "fp4_kernel_as_shipped"), (5, "fp4_pipe_with_barrier_protocol"), (13, "fp4_pipe_back_to_back")):
    env = dict(os.environ, M4RI_B200_DBG=str(dbg), M4RI_B200_LIB=os.path.join(ROOT, "m4ri-rust_b200", "lib", "libm4ri_b200_perf.so"))
    r = subprocess.run([sys.executable, __file__, "--one", str(n)], env=env, capture_output=True, text=True, timeout=300)
    ms = json.loads(r.stdout.strip().splitlines()[-1])["ms"]
    out[key + "_ms"] = ms
    out[key + "_ops_per_s"] = 2.0 * n ** 3 / (ms * 1e-3)
out["fp4_nominal_ops_per_s"] = 32768.0 * 148 * 1.965e9
out["fp4_peak_how"] = "tools/fp4_peak.py: best of 5 x 5 launches, CUDA events around the kernel, perf-hook build; ops = 2mnk"
print(json.dumps(out))
