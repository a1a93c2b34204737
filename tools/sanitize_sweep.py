"""Shape sweep for memory-safety and race evidence: every kernel family of the path runs on small
ragged shapes and is checked against the oracle.

  compute-sanitizer --tool memcheck|racecheck|synccheck python tools/sanitize_sweep.py [family...]
      the intended use; compute-sanitizer is CLOSED on this GPU pool (the wrapper refuses to run), so
  python tools/sanitize_sweep.py --guard --repeat 5 [family...]
      is the substitute: --guard puts every operand and result in a caller-owned buffer
      (m4ri_b200_dmat_wrap) with 4 KiB red zones before and after and two canary words past every
      row, results pre-filled with garbage, and verifies after each call that no canary byte changed
      (stray writes) and that the result equals the oracle's (stray or racy reads/writes inside the
      matrix); --repeat re-runs every case and requires identical results each time (races that
      depend on scheduling).
families: m4rm mma simple matvec elim   (default: all)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import m4ri_rust_b200 as m4
import oracle

L = m4.lib()
args = sys.argv[1:]
GUARD = "--guard" in args
REPEAT = int(args[args.index("--repeat") + 1]) if "--repeat" in args else 1
fams = [a for i, a in enumerate(args) if not a.startswith("--") and (i == 0 or args[i - 1] != "--repeat")] \
    or ["m4rm", "mma", "simple", "matvec", "elim"]
RED = 4096
guards = []
if GUARD:
    import torch
    L.m4ri_b200_set_stream(torch.cuda.current_stream().cuda_stream)


class Guarded:
    """A device matrix inside a canary-filled torch buffer."""

    def __init__(self, nrows, ncols, words=None):
        w = oracle.width(ncols)
        self.stride = (w + 1) // 2 * 2 + 2
        self.nrows, self.w = nrows, w
        self.buf = torch.full((2 * RED + max(1, nrows) * self.stride * 8,), 0xA5, dtype=torch.uint8, device="cuda")
        self.d = m4.DMat.wrap(self.buf.data_ptr() + RED, nrows, ncols, self.stride, keepalive=self.buf)
        if words is not None:
            L.m4ri_b200_upload(self.d.h, m4.mzd_from_words(words, ncols))
            L.m4ri_b200_synchronize()
        guards.append(self)

    def intact(self):
        torch.cuda.synchronize()
        b = self.buf.cpu().numpy()
        ok = (b[:RED] == 0xA5).all() and (b[RED + self.nrows * self.stride * 8:] == 0xA5).all()
        body = b[RED: RED + self.nrows * self.stride * 8].reshape(self.nrows, self.stride * 8) if self.nrows else None
        if body is not None:
            ok = ok and (body[:, (self.stride - 2) * 8:] == 0xA5).all()
        return bool(ok)


def dev(nrows, ncols, words=None):
    if GUARD:
        return Guarded(nrows, ncols, words).d
    return m4.DMat.from_host(m4.mzd_from_words(words, ncols)) if words is not None else m4.DMat.init(nrows, ncols)


def canaries_ok():
    ok = all(g.intact() for g in guards)
    guards.clear()
    return ok
SHAPES = [(1, 1, 1), (2, 63, 65), (65, 64, 64), (64, 65, 129), (191, 193, 129), (200, 257, 321),
          (128, 1024, 128), (256, 2048, 256), (300, 1100, 520), (1030, 40, 1030), (512, 4096, 512)]
bad = 0


def product(algo, name, shapes):
    global bad
    for (m, k, n) in shapes:
        A, B = oracle.random(m, k, 7 + m), oracle.random(k, n, 9 + n)
        want = oracle.mul_m4rm(A, B, k, n, threads=0)[:, : oracle.width(n)]
        dA, dB, dC = dev(m, k, A), dev(k, n, B), dev(m, n)
        ok = True
        for _ in range(REPEAT):
            m4.dmul(dC, dA, dB, clear=True, algo=algo)
            ok = ok and bool((dC.to_words()[:, : oracle.width(n)] == want).all())
            m4.dmul(dC, dA, dB, clear=False, algo=algo)
            ok = ok and not dC.to_words()[:, : oracle.width(n)].any()
        ok = ok and canaries_ok()
        bad += not ok
        print(f"{name} {m}x{k}x{n}: {'ok' if ok else 'MISMATCH'}", flush=True)


if "m4rm" in fams:
    product(m4.ALGO_M4RM, "m4rm", SHAPES)
if "mma" in fams:
    for v in ("fp4s256", "fp4s256cg1"):
        L.m4ri_b200_set_mma_variant(m4.MMA_VARIANTS[v])
        product(m4.ALGO_MMA, v, SHAPES)
    L.m4ri_b200_set_mma_variant(m4.MMA_VARIANTS["fp4s256"])
if "simple" in fams:
    product(m4.ALGO_SIMPLE, "simple", SHAPES[:8])
if "matvec" in fams:
    for (m, k) in [(1, 1), (70, 130), (1000, 512), (4096, 96), (300, 5000)]:
        for nb in (1, 3, 8, 20):
            A, X = oracle.random(m, k, 3 + m), oracle.random(nb, k, 5 + nb)
            dA, dX, dY = dev(m, k, A), dev(nb, k, X), dev(nb, m)
            At = oracle.transpose(A, k)
            want = oracle.mul_m4rm(X, At, k, m, threads=0)[:, : oracle.width(m)]
            ok = True
            for _ in range(REPEAT):
                L.m4ri_b200_dmul_v_batched(dY.h, dA.h, dX.h, 1)
                ok = ok and bool((dY.to_words()[:, : oracle.width(m)] == want).all())
            ok = ok and canaries_ok()
            bad += not ok
            print(f"matvec {m}x{k} batch {nb}: {'ok' if ok else 'MISMATCH'}", flush=True)
if "elim" in fams:
    for (m, n) in [(1, 1), (70, 130), (300, 200), (513, 1100), (1500, 1500)]:
        A = oracle.random(m, n, 11 + m)
        want, r, _ = oracle.rref(A, n)
        ok = True
        for _ in range(REPEAT):
            dW = dev(m, n, A)
            got_r = L.m4ri_b200_drref(dW.h, 0, None, 0)
            ok = ok and got_r == r and bool((dW.to_words()[:, : oracle.width(n)] == want[:, : oracle.width(n)]).all())
            ok = ok and canaries_ok()
        bad += not ok
        print(f"rref {m}x{n}: rank {got_r} {'ok' if ok else 'MISMATCH'}", flush=True)
print(f"sanitize sweep (guard={GUARD}, repeat={REPEAT}): {bad} failures", flush=True)
sys.exit(1 if bad else 0)
