#!/usr/bin/env python
"""bench.py — GF(2) mzd_mul throughput on B200 (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--n SIZE] [--algo auto|m4rm|mma]

A step is one product C = A*B over GF(2) on synthetic uniform random bits.
  N = 1 : square n = 32768 (BASELINE configs[2], the size the metric is quoted on).
  N > 1 : square n = 65536 (configs[3]): C/A row-blocks sharded over the ranks, B of
          rank 0 distributed inside every timed step (the path's one exchange step;
          default: fused into the transpose with NVLink peer stores, --exchange nccl:
          NCCL broadcast in overlapped k-chunks), one process per GPU under torchrun.
`value` = effective bit-ops 2*m*k*n of the whole job / device time (operands
resident in HBM); `e2e` = the same through the reference-facing C-ABI call
mzd_mul(NULL, A, B, 0) on HOST mzd_t operands (H2D + kernels + D2H inside the
timed region). Prints ONE JSON line on rank 0.

--impl reference times the CPU arm: M4RI itself is an absent submodule of the
reference (DESIGN.md "Oracle"), so the arm is the repo's C restatement (oracle/,
kind "port") of both strategies of the reference's mul_impl! on all host threads —
Strassen-Winograd over Four-Russians leaves (the default mzd_mul; the full product
per step) and the Four Russians alone (bounded row sample); value = the faster.
"""
import argparse
import ctypes
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SEED_A, SEED_B = 0x4D345249 ^ (2 << 32) ^ 1, 0x4D345249 ^ (2 << 32) ^ 2
METRIC = "GF(2) mzd_mul effective Tbit-op/s"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "src": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "src": "fallback"}


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self._halt = index, [], set(), None, threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if not self.nv:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.02)

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def host_threads():
    """Host threads the CPU arm uses: the cores this process may run on. Passed to the oracle
    explicitly — torchrun exports OMP_NUM_THREADS=1, which would otherwise silently turn the
    N > 1 reference arm into a one-thread run."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def block_digest(words):
    """Digest of a block of result rows (valid words only), for comparing results across ranks."""
    import hashlib
    return hashlib.blake2b(np.ascontiguousarray(words).tobytes(), digest_size=16).hexdigest()


def cpu_arm(n, m_total, steps, warmup, target_s=12.0, budget_s=600.0, strassen_n=None):
    """CPU arm on the box's host cores, both strategies of the reference's `mul_impl!`
    (m4ri-rust/src/friendly/binary_matrix.rs:53-95), all host threads:
    - "strassen": Strassen-Winograd over Four-Russians leaves (the default, mzd_mul) on the
      full n x n x n workload — the recursion needs the whole square, so when (steps + warmup)
      products do not fit `budget_s` the number of STEPS shrinks (recorded), never the size;
      `strassen_n` (the cpu_baseline leg of an N > 1 line only) asks for a smaller, labelled cube;
    - "m4rm": Four Russians alone (mzd_mul_m4rm) on a bounded row sample, the first `rows` rows
      of A against the full k x n B.
    `value` is the faster of the two."""
    import oracle
    threads = host_threads()
    B = oracle.random(n, n, SEED_B)
    probe_rows = 256
    A = oracle.random(probe_rows, n, SEED_A)
    t0 = time.perf_counter()
    oracle.mul_m4rm(A, B, n, n, threads=threads)
    t_probe = time.perf_counter() - t0
    per_step = target_s / max(1, steps + warmup)
    rows = int(min(m_total, max(256, probe_rows * per_step / max(t_probe, 1e-6))))
    rows = max(256, (rows // 256) * 256)
    A = oracle.random(rows, n, SEED_A)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        oracle.mul_m4rm(A, B, n, n, threads=threads)
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    t = sum(times) / len(times)
    m4rm = {"value": 2.0 * rows * n * n / t / 1e12, "ms_per_step": t * 1e3,
            "sample": f"first {rows} of {m_total} rows of A times the full {n}x{n} B, per step"}
    # Strassen strategy: the full square product on the same seeded operands. The step count is
    # bounded by the time budget (estimated from an 8192-cube); the size is not reduced.
    ns = strassen_n or n
    s_steps, s_warm = steps, warmup
    if ns > 8192:
        Ap, Bp = oracle.random(8192, 8192, SEED_A), oracle.random(8192, 8192, SEED_B)
        oracle.mul_strassen(Ap, Bp, 8192, 8192, threads=threads)
        t0 = time.perf_counter()
        oracle.mul_strassen(Ap, Bp, 8192, 8192, threads=threads)
        rate = 2.0 * 8192 ** 3 / (time.perf_counter() - t0)
        est = 2.0 * ns ** 3 / rate
        while s_warm > 1 and (s_steps + s_warm) * est > budget_s:
            s_warm -= 1
        while s_steps > 1 and (s_steps + s_warm) * est > budget_s:
            s_steps -= 1
        del Ap, Bp
    if ns == n:
        A = oracle.random(n, n, SEED_A)
        Bs = B
    else:
        A, Bs = oracle.random(ns, ns, SEED_A), oracle.random(ns, ns, SEED_B)
    times = []
    for i in range(s_warm + s_steps):
        t0 = time.perf_counter()
        C = oracle.mul_strassen(A, Bs, ns, ns, threads=threads)
        if i >= s_warm:
            times.append(time.perf_counter() - t0)
    t = sum(times) / len(times)
    strassen = {"value": 2.0 * ns * ns * ns / t / 1e12, "ms_per_step": t * 1e3,
                "steps_timed": s_steps, "warmup_run": s_warm,
                "sample": (f"the full {n}x{n}x{n} product, per step" if ns == n else
                           f"a {ns}x{ns}x{ns} product per step (bounded sample of the {n}-cube)"),
                "checksum": f"{oracle.checksum(C, ns):016x}" if ns == n else None}
    best, name = (strassen, "strassen") if strassen["value"] >= m4rm["value"] else (m4rm, "m4rm")
    return {"value": best["value"], "unit": "Tbit-op/s", "cores": threads, "kind": "port",
            "sample": f"{name}: {best['sample']}; CPU restatement (oracle/gf2_oracle.c, OpenMP, {threads} threads) of the "
                      f"reference's two strategies, not M4RI itself (absent submodule, none on the box)",
            "ms_per_step": best["ms_per_step"], "rows": rows if name == "m4rm" else ns,
            "strategies": {"strassen": strassen, "m4rm": m4rm}}


def run_lpn(args, quick=False):
    """BASELINE configs[4] (extra bench line, not the headline): the LPN covering-code shapes.
    quick = True is the summary the default line carries under "also_measured" (product, mat-vec,
    8-vector batch only).
    (a) tall product 2^20 x 512 times 512 x 512 — compute bound, every algorithm;
    (b) mat-vec y = A*x on the resident 2^20 x 512 A, packed words in and out — HBM bound:
        algorithmic bytes = 8*(m*width + width) + ceil(m/64)*8 per product."""
    import torch
    import m4ri_rust_b200 as m4
    L = m4.lib()
    stream = torch.cuda.current_stream()
    L.m4ri_b200_set_stream(ctypes.c_void_p(stream.cuda_stream))
    m, k, n = 1 << 20, 512, 512
    dA = m4.DMat.init(m, k).fill_random(21)
    dB = m4.DMat.init(k, n).fill_random(22)
    dC = m4.DMat.init(m, n)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(steps):
            flush.fill_(1)  # operands (64 MiB + 64 MiB) fit the 126 MB L2: flush between steps
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            fn()
            b.record(stream)
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        return statistics.median(ts)  # (single flushed launches of 10-40 us kernels: the mean is outlier-driven)

    out = {"metric": "GF(2) LPN covering-code workload", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
           "data": "synthetic", "dtype": "u64", "higher_is_better": True,
           "config": {"workload": "tall 2^20x512 * 512x512 + mat-vec on the resident A", "m": m, "k": k, "n": n,
                      "l2": "L2 flushed between steps by writing a 256 MiB buffer"}}
    prod = {}
    for name in (("auto",) if quick else ("auto", "mma", "m4rm")):
        ms = timed(lambda: m4.dmul(dC, dA, dB, clear=True, algo=m4.ALGO_NAMES[name]), args.steps, args.warmup)
        ok = L.m4ri_b200_dfreivalds(dC.h, dA.h, dB.h, 0xF5E1)
        prod[name] = {"ms": ms, "tbitops": 2.0 * m * k * n / ms / 1e9, "freivalds": bool(ok),
                      "gbs_compulsory": (2 * m * 64 + k * 64) / ms / 1e6}
    out["product"] = prod
    out["value"], out["unit"], out["ms_per_step"] = prod["auto"]["tbitops"], "Tbit-op/s", prod["auto"]["ms"]
    # ---- mat-vec on the resident A: HBM bound ------------------------------------------
    pk = peaks()
    w = k // 64
    # Every launch must read its A from HBM, not from the 126 MB L2: the launches rotate over
    # four resident copies of A (4 x 64 MiB = 256 MiB > L2), timed back to back with one event
    # pair around the batch — the launch latency of a ~12 us kernel is then amortised instead of
    # being charged to a 64 MiB read. The L2-flushed single-launch time is kept beside it.
    copies = [dA] + [m4.DMat.init(m, k).fill_random(21) for _ in range(3)]

    def rotating(fn_of_A, reps=24):
        for i in range(8):
            fn_of_A(copies[i % 4])
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for i in range(reps):
            fn_of_A(copies[i % 4])
        b.record(stream)
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    def roof(bytes_alg, ms, kernel):
        return {"bound": "hbm", "achieved": bytes_alg / ms / 1e6, "peak": pk["hbm_gbs"], "unit": "GB/s",
                "frac": bytes_alg / ms / 1e6 / pk["hbm_gbs"], "traffic": None, "kernel": kernel,
                "algorithmic_bytes": bytes_alg}

    dx = m4.DMat.init(1, k).fill_random(23)
    dy = m4.DMat.init(1, m)
    ms_single = timed(lambda: L.m4ri_b200_dmul_v(dy.h, dA.h, dx.h, 1), args.steps, args.warmup)
    ms = rotating(lambda A_: L.m4ri_b200_dmul_v(dy.h, A_.h, dx.h, 1))
    bytes_alg = 8 * (m * w + w) + (m // 64) * 8
    out["matvec"] = {"ms": ms, "ms_flushed_single_launch": ms_single, "tbitops": 2.0 * m * k / ms / 1e9,
                     "timing": "24 launches back to back over 4 resident copies of A (256 MiB > L2), one event pair",
                     "roofline": roof(bytes_alg, ms, "mul_v_coal_kernel<4,2>")}
    # batched: b vectors against the resident A in one pass (m4ri_b200_dmul_v_batched)
    for bvec in ((8,) if quick else (8, 32)):
        dX = m4.DMat.init(bvec, k).fill_random(24)
        dY = m4.DMat.init(bvec, m)
        msb = rotating(lambda A_: L.m4ri_b200_dmul_v_batched(dY.h, A_.h, dX.h, 1))
        passes = (bvec + 7) // 8
        bb = passes * 8 * m * w + 8 * bvec * w + bvec * (m // 64) * 8
        out[f"matvec_batched_{bvec}"] = {"ms": msb, "tbitops": 2.0 * m * k * bvec / msb / 1e9, "vectors": bvec,
                                        "passes_over_A": passes,
                                        "roofline": roof(bb, msb, "mul_vb_coal_kernel<4,1,8>")}
    if quick:
        return out
    # the reference's own call: BinMatrix::mul_slice = mzd_mul_naive(NULL, A, x_column) on HOST
    # operands (binary_matrix.rs:416-431): 64 MiB of A cross PCIe per call, then the same kernel
    hA = L.mzd_init(m, k)
    L.m4ri_b200_download(hA, dA.h)
    L.m4ri_b200_synchronize()
    M = m4.BinMatrix(hA)
    xw = m4.mzd_words(dx.to_host())[0, :w].copy()
    for _ in range(2):
        M.mul_slice(xw)
    L.m4ri_b200_timing_reset()
    t0 = time.perf_counter()
    reps = 5
    for _ in range(reps):
        col = M.mul_slice(xw)
    e2e_ms = (time.perf_counter() - t0) * 1e3 / reps
    # the C call inside it alone (what a Rust host pays; the rest is the Python mirror's packing of
    # x into a k x 1 mzd_t and of the m one-bit rows of the result back into words)
    xcol = L.mzd_init(k, 1)
    for i in range(k):
        m4.mzd_words(xcol)[i, 0] = (int(xw[i // 64]) >> (i % 64)) & 1
    for _ in range(2):
        L.mzd_free(L.mzd_mul_naive(None, hA, xcol))
    t0 = time.perf_counter()
    for _ in range(reps):
        L.mzd_free(L.mzd_mul_naive(None, hA, xcol))
    c_ms = (time.perf_counter() - t0) * 1e3 / reps
    L.mzd_free(xcol)
    out["matvec_via_mul_slice"] = {"ms": e2e_ms, "c_call_ms": c_ms, "c_call_pcie_gbs": m * w * 8 / c_ms / 1e6,
                                   "kernel": L.m4ri_b200_timing_name().decode(),
                                   "h2d_bytes_per_call": m * w * 8, "d2h_bytes_per_call": (m // 64) * 8,
                                   "api": "BinMatrix.mul_slice -> mzd_transpose + mzd_mul_naive(NULL, A, x) on host mzd_t (binary_matrix.rs:416-431)",
                                   "pcie_gbs": m * w * 8 / e2e_ms / 1e6}
    del M
    # the same kernel on a 4x taller A (256 MiB > L2): launch latency amortised, no flush needed
    m2 = 1 << 22
    dA2 = m4.DMat.init(m2, k).fill_random(31)
    dy2 = m4.DMat.init(1, m2)
    ms2 = timed(lambda: L.m4ri_b200_dmul_v(dy2.h, dA2.h, dx.h, 1), args.steps, args.warmup)
    b2 = 8 * (m2 * w + w) + (m2 // 64) * 8
    out["matvec_2^22_rows"] = {"ms": ms2, "tbitops": 2.0 * m2 * k / ms2 / 1e9, "roofline": roof(b2, ms2, "mul_v_coal_kernel<4,2>")}
    del dA2, dy2
    # long rows: square 32768 x 32768 A (128 MiB > L2), one CTA per 32 rows
    m3 = n3 = 32768
    dA3 = m4.DMat.init(m3, n3).fill_random(41)
    dx3 = m4.DMat.init(1, n3).fill_random(42)
    dy3 = m4.DMat.init(1, m3)
    ms3 = timed(lambda: L.m4ri_b200_dmul_v(dy3.h, dA3.h, dx3.h, 1), args.steps, args.warmup)
    b3 = 8 * (m3 * (n3 // 64) + n3 // 64) + (m3 // 64) * 8
    out["matvec_32768x32768"] = {"ms": ms3, "tbitops": 2.0 * m3 * n3 / ms3 / 1e9, "roofline": roof(b3, ms3, "mul_v_long_kernel")}
    dX3 = m4.DMat.init(8, n3).fill_random(43)
    dY3 = m4.DMat.init(8, m3)
    ms3b = timed(lambda: L.m4ri_b200_dmul_v_batched(dY3.h, dA3.h, dX3.h, 1), args.steps, args.warmup)
    out["matvec_batched_8_32768x32768"] = {"ms": ms3b, "tbitops": 2.0 * m3 * n3 * 8 / ms3b / 1e9,
                                           "roofline": roof(b3 + 7 * (n3 // 8 + m3 // 8), ms3b, "mul_vb_long_kernel<8>")}
    print(json.dumps(out))
    return out


def also_measured(args):
    """Summary of the other Section-8 rows inside the default N = 1 line, so that the driver's own run
    carries them: the LPN shapes (configs[4]) and the consumers of the product (RREF and PLE of a
    random 8192 x 8192 matrix, device-resident, checked against the CPU oracle's rank, pivot
    columns and reduced form)."""
    import m4ri_rust_b200 as m4
    import oracle
    L = m4.lib()
    la = argparse.Namespace(steps=5, warmup=3)
    lpn = run_lpn(la, quick=True)
    out = {"lpn_2^20x512": {"product_ms": lpn["product"]["auto"]["ms"], "product_tbitops": lpn["product"]["auto"]["tbitops"],
                            "product_freivalds": lpn["product"]["auto"]["freivalds"],
                            "matvec_ms": lpn["matvec"]["ms"], "matvec_hbm_frac": lpn["matvec"]["roofline"]["frac"],
                            "matvec_batched_8_ms": lpn["matvec_batched_8"]["ms"],
                            "full_line": "python bench.py --workload lpn"}}
    n = 8192
    W = m4.DMat.init(n, n)
    A = W.fill_random(7).to_words()
    want, rank, piv = oracle.rref(np.ascontiguousarray(A), n)
    ts = []
    for _ in range(3):
        W.fill_random(7)
        L.m4ri_b200_synchronize()
        t0 = time.perf_counter()
        r = L.m4ri_b200_drref(W.h, 0, None, 0)
        L.m4ri_b200_synchronize()
        ts.append((time.perf_counter() - t0) * 1e3)
    got = W.to_words()
    w = n // 64
    ok_rref = bool(r == rank and (got[:, :w] == want[:, :w]).all())
    P, Q = (ctypes.c_int * n)(), (ctypes.c_int * n)()
    tp = []
    for _ in range(3):
        W.fill_random(7)
        L.m4ri_b200_synchronize()
        t0 = time.perf_counter()
        rp = L.m4ri_b200_dple(W.h, P, Q, 0)
        L.m4ri_b200_synchronize()
        tp.append((time.perf_counter() - t0) * 1e3)
    ok_ple = bool(rp == rank and list(Q)[:rp] == [int(c) for c in piv])
    try:  # the packed factors must stand for the input: P * L * E == A (what smoke() checks at 700 x 900)
        back = oracle.ple_reconstruct(np.ascontiguousarray(W.to_words()), n, rp, list(P), list(Q))
        ok_ple_identity = bool((back[:, :w] == A[:, :w]).all())
    except AssertionError:
        ok_ple_identity = False
    out["elimination_8192"] = {"rref_ms": min(ts), "ple_ms": min(tp), "rank": int(r),
                               "rref_equals_cpu_oracle": ok_rref, "ple_rank_and_pivot_columns_equal_cpu_oracle": ok_ple,
                               "ple_p_l_e_equals_a": ok_ple_identity}
    out["ok"] = bool(ok_rref and ok_ple and ok_ple_identity and lpn["product"]["auto"]["freivalds"])
    return out  # (main() prints the line first and then fails the run when ok is false)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=0)
    ap.add_argument("--algo", default="auto", choices=["auto", "m4rm", "mma", "simple"])
    ap.add_argument("--chunks", default="auto", help="k-chunks of the broadcast of B (N > 1): a count, 'ramp' = 1/8, 3/8, 1/2 of k, or 'auto'")
    ap.add_argument("--exchange", default="fused", choices=["fused", "nccl"],
                    help="N > 1: 'fused' = transpose kernel that reads the root's B and stores Bt to every rank over NVLink; "
                         "'nccl' = NCCL broadcast of B in k-chunks overlapped with the product")
    ap.add_argument("--reserve", type=int, default=8, help="SMs left to NCCL while a broadcast chunk is in flight (N > 1)")
    ap.add_argument("--workload", default="square", choices=["square", "lpn"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the also_measured summary (LPN shapes, elimination)")
    args = ap.parse_args()
    if args.chunks not in ("ramp", "ramp4", "auto"):
        args.chunks = int(args.chunks)
    os.environ.setdefault("NCCL_MAX_NCHANNELS", str(max(1, args.reserve)))  # the broadcast runs on the 8 SMs the product leaves free
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    n = args.n or (32768 if world == 1 else 65536)
    workload = f"square mzd_mul n={n} random" + (
        "" if world == 1 else f", C/A row-blocks sharded over {world} ranks, B of rank 0 distributed every step: "
        + ("exchange fused into the transpose (plan in the line's exchange_plan)"
           if args.exchange == "fused" else
           f"NCCL broadcast in k-chunks ({args.chunks}) overlapped with the product; the product leaves {args.reserve} SMs to NCCL while a chunk is in flight"))
    config = {"workload": workload, "m": n, "k": n, "n": n, "algo": args.algo,
              "l2": "operands (3 x n*n/8 bytes) exceed the 126 MB L2; no explicit flush" if n >= 32768
              else "L2 flushed between steps by writing a 256 MiB buffer"}

    if args.workload == "lpn" and args.impl != "reference":
        if rank == 0:
            run_lpn(args)
        return
    if args.impl == "reference":
        if rank != 0:
            return
        steps, warmup = args.steps, args.warmup
        r = cpu_arm(n, n, steps, warmup, target_s=30.0)
        # (threads and sample live under cpu_baseline: `config` stays word for word the product arm's)
        line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": "Tbit-op/s",
                "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": r["ms_per_step"],
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u64",
                "data": "synthetic", "config": config,
                "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "strategies")},
                "e2e": {"value": r["value"], "unit": "Tbit-op/s", "h2d_bytes_per_step": 0,
                        "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return

    import torch
    import m4ri_rust_b200 as m4
    if torch.cuda.is_available():
        torch.cuda.set_device(local_rank)  # before the library is touched: its context is this rank's GPU
    L = m4.lib()
    if m4.device_count() <= 0:
        raise SystemExit("bench.py: no CUDA device; libm4ri_b200 has no CPU fallback")
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist = dist_mod
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    stream = torch.cuda.current_stream()
    L.m4ri_b200_set_stream(ctypes.c_void_p(stream.cuda_stream))
    algo = m4.ALGO_NAMES[args.algo]

    # row-block of this rank
    from m4ri_rust_b200 import sharded
    r0, rows = sharded.row_block(n, world, rank)
    ws = n // 64  # words per row (n % 128 == 0 so device stride == width)
    fused = world > 1 and args.exchange == "fused"
    exchange_plan = None
    if fused:
        # exchange fused into the transpose: each rank transposes one k-slab of the root's B over
        # NVLink and stores it into every rank's Bt (peer stores); one product launch per step
        fm = sharded.FusedBroadcastMul(m4, n, n, n, dist, src=0)
        tA, tB, tC, dA, dB, dC = fm.tA, fm.tB, fm.tC, fm.dA, fm.dB, fm.dC
        exchange_plan = {
            "mc_root": "mc_root: the root transposes its B in one kernel whose stores go to the NVSwitch multicast mapping of the symmetric Bt buffer: one store per 16 bytes reaches every GPU (root egress |B|, no remote reads)",
            "mc_slabs": "mc_slabs: every rank transposes one k-slab of the root's B (read over NVLink) and stores it to the multicast mapping of Bt",
            "peer": "peer: every rank transposes one k-slab of the root's B (read over NVLink; root_serves=%s, pulled in %d pieces by the copy engine) and stores it into every rank's Bt with peer stores" % (fm.root_serves, fm.chunks),
        }[fm.plan]
    else:
        tA = torch.empty((rows, ws), dtype=torch.int64, device="cuda")
        tB = torch.empty((n, ws), dtype=torch.int64, device="cuda")
        tC = torch.empty((rows, ws), dtype=torch.int64, device="cuda")
        dA = m4.DMat.wrap(tA.data_ptr(), rows, n, ws, keepalive=tA)
        dB = m4.DMat.wrap(tB.data_ptr(), n, n, ws, keepalive=tB)
        dC = m4.DMat.wrap(tC.data_ptr(), rows, n, ws, keepalive=tC)
    # word (r,w) of A = splitmix64 stream index r*width+w: generate the full-matrix
    # stream positions of this rank's rows
    L.m4ri_b200_dfill_random_at(dA.h, SEED_A, r0 * ws)
    if rank == 0:
        dB.fill_random(SEED_B)
    else:
        tB.zero_()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda") if n < 32768 else None
    if fused:
        step = lambda: fm.step(algo)
    else:
        # one step = NCCL broadcast of B from rank 0 in k-chunks (N > 1), each chunk's product
        # overlapped with the next chunk's transfer (m4ri-rust_b200/sharded.py)
        nsm = m4.device_info()["sms"]
        sm = sharded.ShardedMul(n, n, n, dist=dist, src=0, chunks=args.chunks, tB=tB,
                                reserve_sms=(lambda r: L.m4ri_b200_set_sm_limit(nsm - args.reserve if r else 0)) if world > 1 and args.reserve > 0 else None,
                                chunk_product=sharded.gpu_chunk_product(m4, tA, tB, tC, rows, n, n, algo))
        step = sm.step

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    # everything slow on the host (NVML init, event creation) happens BEFORE the barrier: the
    # steps are coupled across ranks by device-side barriers, so host skew between the barrier
    # and the first timed step would be charged to that step
    sampler = ClockSampler(local_rank)
    sampler.start()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    e_all0, e_all1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    L.m4ri_b200_timing_reset()
    l0 = m4.launch_count()
    e_all0.record(stream)
    for a, b in evs:
        if flush is not None:
            flush.fill_(1)
        a.record(stream)
        step()
        b.record(stream)
    e_all1.record(stream)
    barrier()
    clocks = sampler.stop()
    launches = m4.launch_count() - l0
    step_ms = [a.elapsed_time(b) for a, b in evs]
    ms = sum(step_ms) / len(step_ms)
    timing_count = L.m4ri_b200_timing_count()
    kern_total_ms = L.m4ri_b200_timing_ms()
    timing_ops = L.m4ri_b200_timing_ops()
    kern_ms = kern_total_ms / max(1, timing_count)
    kern_name = L.m4ri_b200_timing_name().decode()

    if fused:
        dist.broadcast(tB, src=0)  # only Bt travelled inside the steps; the check below needs B itself
        torch.cuda.synchronize()
    # parity inside the run: Freivalds on this rank's block (false accept 2^-64)
    ok = L.m4ri_b200_dfreivalds(dC.h, dA.h, dB.h, 0xF5E1)
    if ok != 1:
        raise SystemExit(f"bench.py: rank {rank}: product failed the Freivalds check")
    chk = dC.checksum()

    # ---- e2e: the reference-facing call, host operands ---------------------------
    # ONE call C = mzd_mul(NULL, A, B, 0) on the FULL host operands, as m4ri-rust makes it for
    # `&a * &b` (binary_matrix.rs:465): C == NULL, the library allocates the result, the caller
    # drops it (mzd_free) — allocation and free are timed. At N > 1 rank 0 makes that one call
    # with m4ri_b200_set_devices(N): the library spreads it over the N GPUs of the box (B uploaded
    # once, slab g over GPU g's PCIe link, exchanged over NVLink); the other ranks wait on the
    # host (a store key, not an NCCL kernel that would sit on their SMs).
    e2e_ms = None
    e2e_bytes = (0, 0)
    parity_extra = {}
    n1_same_size_ms = None
    if not args.no_e2e:
        store = dist.distributed_c10d._get_default_store() if dist is not None else None
        my_digest = block_digest(dC.to_words()[:, :ws])
        if store is not None:
            store.set(f"m4ri_b200_block_{rank}", f"{r0}:{rows}:{my_digest}")
        barrier()
        if rank == 0:
            if world > 1:
                dAf = m4.DMat.init(n, n).fill_random(SEED_A)
                hA, hB = L.mzd_init(n, n), L.mzd_init(n, n)
                L.m4ri_b200_download(hA, dAf.h)
                del dAf
            else:
                hA, hB = L.mzd_init(rows, n), L.mzd_init(n, n)
                L.m4ri_b200_download(hA, dA.h)
            L.m4ri_b200_download(hB, dB.h)
            L.m4ri_b200_synchronize()
            L.m4ri_b200_set_algo(algo)
            L.m4ri_b200_set_devices(world)
            # as a host without torch calls it: the library's own stream (the steps above ran on
            # torch's current stream — the legacy default stream — for the CUDA-event timing)
            if os.environ.get("BENCH_E2E_TORCH_STREAM") != "1":
                L.m4ri_b200_use_own_stream()
            e2e_steps = max(3, min(args.steps, 5))
            for i in range(2):
                L.mzd_free(L.mzd_mul(None, hA, hB, 0))
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for i in range(e2e_steps - 1):
                L.mzd_free(L.mzd_mul(None, hA, hB, 0))  # synchronous: result words valid on return
            hC = L.mzd_mul(None, hA, hB, 0)
            e2e_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
            L.m4ri_b200_set_devices(1)
            L.m4ri_b200_set_algo(0)
            L.m4ri_b200_set_stream(ctypes.c_void_p(stream.cuda_stream))
            # The WHOLE result of the one host call against (1) the block every rank computed
            # device-resident in the timed steps (each Freivalds-checked above) and (2) the CPU
            # oracle's checksum of this seeded product (tests/golden/full_size_checksums.json)
            import oracle
            words = m4.mzd_words(hC)
            blocks = [(r0, rows, my_digest)] if store is None else [
                (int(a), int(b), c) for a, b, c in (store.get(f"m4ri_b200_block_{i}").decode().split(":") for i in range(world))]
            if sum(b for _, b, _ in blocks) != n:
                raise SystemExit("bench.py: the ranks' row-blocks do not cover the result")
            for i, (b0, brows, dig) in enumerate(blocks):
                if block_digest(words[b0:b0 + brows, :ws]) != dig:
                    raise SystemExit(f"bench.py: host-ABI result differs from rank {i}'s device-resident block (rows {b0}..{b0 + brows})")
            parity_extra["e2e_equals_every_ranks_block"] = True
            full = f"{oracle.checksum(np.ascontiguousarray(words[:, :ws]), n):016x}"
            parity_extra["checksum_all_ranks"] = full
            golden = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "tests", "golden", "full_size_checksums.json")))
            if str(n) in golden:
                parity_extra["equals_cpu_oracle_full_size"] = full == golden[str(n)]["checksum"]
                if full != golden[str(n)]["checksum"]:
                    raise SystemExit("bench.py: the full result's checksum differs from the committed CPU-oracle checksum")
            if world > 1:
                # GPUTEST only ever sees one GPU: one ragged multi-device product, fresh and accumulating,
                # against the CPU oracle, inside every N > 1 line
                L.m4ri_b200_set_devices(world)
                L.m4ri_b200_use_own_stream()
                rm, rk, rn = 5000, 20000, 9000
                rA, rB, rC0 = oracle.random(rm, rk, 21), oracle.random(rk, rn, 22), oracle.random(rm, rn, 23)
                want = oracle.mul_m4rm(rA, rB, rk, rn, threads=0)[:, : oracle.width(rn)]
                ra, rb, rc = m4.mzd_from_words(rA, rk), m4.mzd_from_words(rB, rn), m4.mzd_from_words(rC0, rn)
                fresh = L.mzd_mul(None, ra, rb, 0)
                L.mzd_addmul(rc, ra, rb, 0)
                ok_r = bool((m4.mzd_words(fresh)[:, : oracle.width(rn)] == want).all() and
                            (m4.mzd_words(rc)[:, : oracle.width(rn)] == (rC0[:, : oracle.width(rn)] ^ want)).all())
                parity_extra["ragged_multi_device_5000x20000x9000"] = ok_r
                for p_ in (ra, rb, rc, fresh):
                    L.mzd_free(p_)
                L.m4ri_b200_set_devices(1)
                L.m4ri_b200_set_stream(ctypes.c_void_p(stream.cuda_stream))
                if not ok_r:
                    raise SystemExit("bench.py: ragged multi-device product differs from the CPU oracle")
                # the same-size one-GPU step (strong-scaling denominator): the N = 1 step (m4ri_b200_dmul
                # on resident operands) at THIS n on rank 0's GPU alone, 1 warm-up + 3 timed
                dA1, dB1, dC1 = m4.DMat.init(n, n).fill_random(SEED_A), m4.DMat.init(n, n).fill_random(SEED_B), m4.DMat.init(n, n)
                m4.dmul(dC1, dA1, dB1, clear=True, algo=algo)
                ev1 = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
                ev1[0].record(stream)
                for i in range(3):
                    m4.dmul(dC1, dA1, dB1, clear=True, algo=algo)
                    ev1[i + 1].record(stream)
                torch.cuda.synchronize()
                parity_extra["n1_same_size_checksum_equal"] = f"{dC1.checksum():016x}" == full
                n1_same_size_ms = ev1[0].elapsed_time(ev1[3]) / 3
                del dA1, dB1, dC1
            e2e_bytes = (2 * n * ws * 8 if world > 1 else (rows * ws + n * ws) * 8, (n if world > 1 else rows) * ws * 8)
            for p in (hA, hB, hC):
                L.mzd_free(p)
            if store is not None:
                store.set("m4ri_b200_e2e_done", "1")
        elif store is not None:
            store.wait(["m4ri_b200_e2e_done"])

    # ---- reduce over ranks: max time ------------------------------------------------
    t = torch.tensor([ms, e2e_ms or 0.0, kern_ms], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_max, kern_ms = (float(x) for x in t.cpu())
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    total_ops = 2.0 * n * n * n
    value = total_ops / (ms * 1e-3) / 1e12
    pk = peaks()
    # executed bit-ops of the timed launches (Strassen leaves are smaller than the job and
    # there are 7^L of them) / their summed device time
    kern_ops, kern_count = timing_ops, max(1, timing_count)
    achieved = kern_ops / (kern_total_ms * 1e-3) / 1e12 if kern_total_ms > 0 else None
    if "mma" in kern_name:
        # tensor roofline. MEASURED_PEAKS.json holds bf16 only; the int8 pipe issues twice the
        # bf16 rate: 16384 int8 op/clk/SM. `peak` is that rate at the maximum SM clock (the
        # strictest denominator); the fraction of 2 x the driver-measured cuBLAS bf16 burst
        # (power-capped at ~1.3 GHz) is reported beside it.
        sm_max = (clocks.get("sm_max_mhz") or 1965) * 1e6
        fp4 = any(v in kern_name for v in ("mma4", "mma5", "mma6"))  # kind::mxf4: 64 elements per row per MMA, twice kind::i8
        per_clk = 32768.0 if fp4 else 16384.0
        peak = per_clk * 148 * sm_max / 1e12
        meas = (4.0 if fp4 else 2.0) * pk["bf16_tflops"]
        roof = {"bound": "tensor", "achieved": achieved, "peak": peak,
                "unit": "TOP/s (%s, executed 2mnk)" % ("fp4 e2m1" if fp4 else "int8"),
                "frac": achieved / peak if achieved else None, "traffic": None,
                "peak_source": "tcgen05 %s issue rate %d op/clk/SM x 148 SMs x max SM clock" % (
                    "kind::mxf4" if fp4 else "kind::i8", int(per_clk)),
                "peak_measured_scaled_bf16": meas,
                "frac_of_measured_scaled_bf16": achieved / meas if achieved else None,
                "peak_measured_source": "%d x bf16_tflops (burst) of MEASURED_PEAKS.json (%s): the dense %s rate is %dx bf16" % (
                    4 if fp4 else 2, pk["src"], "fp4" if fp4 else "int8", 4 if fp4 else 2)}
        if fp4:
            # the kind::mxf4 pipe rate measured under this kernel's own issue pattern and clocks
            # (tools/fp4_peak.py, perf-hook build: expansion, drain and epilogue waits disabled)
            try:
                pm = json.load(open(os.path.join(ROOT, "profiles", "peaks_measured.json")))
                mp = pm["fp4_pipe_back_to_back_ops_per_s"] / 1e12
                roof["peak_measured_fp4"] = mp
                roof["frac_of_measured_fp4"] = achieved / mp if achieved else None
                roof["peak_measured_fp4_source"] = "profiles/peaks_measured.json fp4_pipe_back_to_back_ops_per_s (%s, %s)" % (
                    pm.get("fp4_peak_shape"), pm.get("fp4_peak_how"))
            except (OSError, KeyError, ValueError):
                pass
    else:
        # Four-Russians gather: bound by shared-memory load bandwidth, measured by
        # m4ri-rust_b200/lib/microbench (profiles/); expressed in the metric's unit
        smem = json.load(open(os.path.join(ROOT, "profiles", "peaks_measured.json"))) \
            if os.path.exists(os.path.join(ROOT, "profiles", "peaks_measured.json")) else {}
        lds = smem.get("lds128_bytes_per_s", 128.0 * 148 * 1.965e9)
        peak = lds / 4 * 64 * 8 / 1e12  # one 32-bit table word = 32 columns x 8 k-bits x 2 ops
        roof = {"bound": "smem", "achieved": achieved, "peak": peak, "unit": "Tbit-op/s",
                "frac": achieved / peak if achieved else None, "traffic": None,
                "peak_source": "measured shared-memory load bandwidth x 512 bit-ops per 32-bit table word (t=8)"}
    # DRAM traffic of one launch of that kernel at the leaf shape, from the committed ncu capture
    try:
        per_launch = kern_ops / kern_count / 2.0  # sum of m*k*n over the products of one launch
        shape, group = None, 1
        for lv in range(4):  # the rank's row-block (n/world x n x n) after lv Strassen levels
            mm, kk = n // world >> lv, n >> lv
            for g in (1, 7):  # the fused last level runs its seven leaf products as one launch
                if abs(g * mm * kk * kk - per_launch) < 1e-6 * per_launch:
                    shape, group = (mm, kk, kk), g
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        key = "%s@%s%dx%dx%d" % ((kern_name, "7x" if group == 7 else "") + shape) if shape else None
        ent = tj.get(key) if key else None
        if shape:
            roof["launch_shape"] = ("7 products of " if group == 7 else "") + "%dx%dx%d" % shape
            roof["algorithmic_bytes"] = group * (shape[0] * shape[1] + shape[1] * shape[2] + shape[0] * shape[2]) // 8
        if ent:
            roof["traffic"] = ent["bytes"]
            roof["traffic_source"] = ent["source"]
    except Exception:
        pass
    roof["kernel"] = kern_name
    roof["kernel_ms"] = kern_ms
    roof["kernel_launches_per_step"] = timing_count / max(1, args.steps)
    roof["kernel_share_of_step"] = kern_total_ms / max(1e-9, ms * args.steps)
    line = {"metric": METRIC, "value": value, "unit": "Tbit-op/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "u64", "data": "synthetic", "config": config,
            "clocks": clocks, "gpu_launches": int(launches), "roofline": roof,
            "parity": {"freivalds_2^-64": True, "checksum_rank0": f"{chk:016x}"},
            "step_ms_rank0": [round(x, 3) for x in step_ms]}
    line["parity"].update(parity_extra)
    if n1_same_size_ms is not None:
        line["n1_same_size_ms"] = n1_same_size_ms  # one GPU, same n: ms_per_step * n_gpus / this = strong-scaling efficiency
    if exchange_plan:
        line["exchange_plan"] = exchange_plan  # beside config, not in it: both arms print one config
    if e2e_ms is not None:
        line["e2e"] = {"value": total_ops / (e2e_max * 1e-3) / 1e12, "unit": "Tbit-op/s",
                       "ms_per_step": e2e_max, "h2d_bytes_per_step": e2e_bytes[0],
                       "d2h_bytes_per_step": e2e_bytes[1],
                       "api": "ONE call C = mzd_mul(NULL, A, B, 0) on the full host mzd_t operands (pinned), result allocated by the library and freed by the caller as m4ri-rust does"
                              + ("" if world == 1 else f"; made by rank 0 with m4ri_b200_set_devices({world}): B crosses PCIe once (slab g over GPU g's link) and is exchanged over NVLink inside the library")}
    if world > 1 and not args.no_cpu_baseline:
        # N > 1 lines: a bounded, labelled sample (the full arm is `--impl reference`)
        r = cpu_arm(n, n, 1, 1, target_s=6.0, strassen_n=16384)
        line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "strategies")}
    if world == 1 and not args.no_cpu_baseline:
        r = cpu_arm(n, n, 2, 1, target_s=12.0)
        line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "strategies")}
        # the CPU Strassen leg computed the whole product from the same seeded operands: the
        # checksum of its C must equal the checksum of the C the GPU left in HBM (bit-exact parity
        # at the full benchmark size, not only Freivalds)
        same = r["strategies"]["strassen"]["checksum"] == f"{chk:016x}"
        line["parity"]["equals_cpu_oracle_full_size"] = same
        if not same:
            raise SystemExit("bench.py: GPU product differs from the CPU oracle's at full size")
    if world == 1 and not args.no_extras:
        try:
            line["also_measured"] = also_measured(args)
        except Exception as ex:  # the summary is an extra: it must never take the headline line with it
            line["also_measured"] = {"error": f"{type(ex).__name__}: {ex}"}
    # a wrong result of an extra must not be published as a timing with a nested false flag: the
    # headline line is still printed (its own parity block stands), then the run fails
    extras_ok = line.get("also_measured", {}).get("ok", True)
    line["ok"] = bool(extras_ok)
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()
    if not extras_ok:
        raise SystemExit("bench.py: an also_measured result differs from the CPU oracle (see the flags in the line)")


if __name__ == "__main__":
    main()
