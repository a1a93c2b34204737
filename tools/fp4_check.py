"""Exactness diagnostics of the FP4 tensor path (kernels_mma4.cu): the fp32 accumulators
must equal popcount(A_i & Bt_j)/4 exactly, for random rows and for the worst case (all
ones, k = 65536). Then the usual parity sweep + timing through tools/mma_variant_check.py.
(diagnostic tool: numpy/oracle are the checkers, never the thing measured)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import m4ri_rust_b200 as m4
import oracle

L = m4.lib()

def popc_matrix(A, B):
    # A [m, w] uint64, B [n, w] uint64 -> [m, n] popcount of the AND
    a8 = np.unpackbits(A.view(np.uint8), axis=1, bitorder="little").astype(np.float32)
    b8 = np.unpackbits(B.view(np.uint8), axis=1, bitorder="little").astype(np.float32)
    return a8 @ b8.T

bad = 0
for cg in (1, 2):
    for (m, k, n, mode) in [(256, 4096, 256, "random"), (200, 1000, 300, "random"), (256, 65536, 256, "ones"),
                            (128, 65536, 128, "random"), (384, 32768, 512, "ones")]:
        if mode == "ones":
            A = np.full((m, oracle.width(k)), np.uint64(0xFFFFFFFFFFFFFFFF)); B = np.full((n, oracle.width(k)), np.uint64(0xFFFFFFFFFFFFFFFF))
        else:
            A, B = oracle.random(m, k, 5 + m), oracle.random(n, k, 7 + n)
        A = np.ascontiguousarray(A[:, :oracle.width(k)]); B = np.ascontiguousarray(B[:, :oracle.width(k)])
        a, b = m4.mzd_from_words(A, k), m4.mzd_from_words(B, k)
        dA, dB, dC = m4.DMat.from_host(a), m4.DMat.from_host(b), m4.DMat.init(m, n)
        acc = torch.full((m, n), -1.0, dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        L.m4ri_b200_dmul_abt_fp4_debug(dC.h, dA.h, dB.h, acc.data_ptr(), cg)
        L.m4ri_b200_synchronize()
        got = acc.cpu().numpy()
        want = popc_matrix(m4.mzd_words(a)[:, :oracle.width(k)], m4.mzd_words(b)[:, :oracle.width(k)]) / 4.0
        ok = bool((got == want).all())
        par = ((dC.to_words()[:, :oracle.width(n)]))
        wantpar = np.packbits((want * 4).astype(np.int64) & 1, axis=1, bitorder="little")
        wantpar = np.pad(wantpar, ((0, 0), (0, (-wantpar.shape[1]) % 8))).view(np.uint64)[:, :oracle.width(n)]
        okp = bool((par == wantpar).all())
        if not (ok and okp):
            bad += 1
            d = np.argwhere(got != want)
            print(f"cg{cg} {m}x{k}x{n} {mode}: acc exact={ok} parity={okp}; mismatches {len(d)}; first {d[:4].tolist()} got {got[tuple(d[0])] if len(d) else None} want {want[tuple(d[0])] if len(d) else None} max|err| {np.abs(got-want).max()}", flush=True)
        else:
            print(f"cg{cg} {m}x{k}x{n} {mode}: accumulators exact, parity ok (max acc {got.max()})", flush=True)
print("fp4 accumulators:", "EXACT" if bad == 0 else f"{bad} FAILED", flush=True)
