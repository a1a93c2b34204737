"""Why is the same kernel slower inside bench.py? torch-allocated vs pool-allocated operands,
own stream vs torch's stream, with/without the transpose in front."""
import os, sys, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import m4ri_rust_b200 as m4
L = m4.lib()
n = 32768; ws = n // 64
def timeit(tag, dA, dBt, dC, pre=None, reps=5):
    for _ in range(2):
        if pre: pre()
        m4.dmul_abt(dC, dA, dBt, clear=True, algo=m4.ALGO_MMA)
    L.m4ri_b200_synchronize()
    L.m4ri_b200_timing_reset()
    for _ in range(reps):
        if pre: pre()
        m4.dmul_abt(dC, dA, dBt, clear=True, algo=m4.ALGO_MMA)
    ms = L.m4ri_b200_timing_ms() / L.m4ri_b200_timing_count()
    print(f"{tag}: {ms:.3f} ms", flush=True)
# 1. pool allocated, own stream
dA, dB, dBt, dC = (m4.DMat.init(n, n) for _ in range(4))
dA.fill_random(1); dB.fill_random(2); L.m4ri_b200_dtranspose(dBt.h, dB.h)
timeit("pool memory, library stream", dA, dBt, dC)
timeit("pool memory, library stream, transpose before each", dA, dBt, dC, pre=lambda: L.m4ri_b200_dtranspose(dBt.h, dB.h))
# 2. torch allocated
t = [torch.empty((n, ws), dtype=torch.int64, device="cuda") for _ in range(4)]
w = [m4.DMat.wrap(x.data_ptr(), n, n, ws, keepalive=x) for x in t]
torch.cuda.synchronize()
w[0].fill_random(1); w[1].fill_random(2); L.m4ri_b200_dtranspose(w[2].h, w[1].h)
timeit("torch memory, library stream", w[0], w[2], w[3])
print("ptrs", [hex(x.data_ptr()) for x in t], hex(dA.ptr or 0), hex(dBt.ptr or 0), hex(dC.ptr or 0))
# 3. torch stream
L.m4ri_b200_set_stream(ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
timeit("torch memory, torch stream", w[0], w[2], w[3])
timeit("pool memory, torch stream", dA, dBt, dC)
# 4. A from seed used in bench
SEED_A, SEED_B = 0x4D345249 ^ (2 << 32) ^ 1, 0x4D345249 ^ (2 << 32) ^ 2
dA.fill_random(SEED_A); dB.fill_random(SEED_B); L.m4ri_b200_dtranspose(dBt.h, dB.h)
timeit("pool memory, bench seeds", dA, dBt, dC)
