"""Device transpose timing, both kernels: python tools/transpose_time.py"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, ctypes
import m4ri_rust_b200 as m4
L = m4.lib()
stream = torch.cuda.current_stream()
L.m4ri_b200_set_stream(ctypes.c_void_p(stream.cuda_stream))
for (r, c) in [(32768, 32768), (65536, 65536), (8192, 65536), (16384, 16384)]:
    S = m4.DMat.init(r, c).fill_random(1)
    D = m4.DMat.init(c, r)
    out = {"rows": r, "cols": c}
    chk = {}
    for var in ("narrow", "wide"):
        os.environ["M4RI_B200_TRANSPOSE"] = var
        for _ in range(3):
            L.m4ri_b200_dtranspose(D.h, S.h)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(10):
            L.m4ri_b200_dtranspose(D.h, S.h)
        b.record(stream)
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 10
        out[var] = {"ms": round(ms, 4), "gbs": round(2.0 * r * c / 8 / ms / 1e6, 1)}
        chk[var] = D.checksum()
    out["same"] = chk["narrow"] == chk["wide"]
    print(json.dumps(out), flush=True)
    del S, D
