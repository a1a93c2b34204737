"""Measures the tensor-pipe denominators for path (b) on this box: cuBLASLt int8
(torch._int_mm) and bf16 dense GEMM, burst (best of 10) — library calls used only
as a yardstick, never on the product path. Prints one JSON object."""
import json
import torch

def best_ms(f, reps=10):
    for _ in range(3):
        f()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); f(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best

out = {}
for n in (8192, 16384):
    A = torch.randint(0, 2, (n, n), dtype=torch.int8, device="cuda")
    B = torch.randint(0, 2, (n, n), dtype=torch.int8, device="cuda").t().contiguous().t()
    try:
        ms = best_ms(lambda: torch._int_mm(A, B))
        out[f"int8_int_mm_n{n}_tops"] = 2.0 * n ** 3 / (ms * 1e-3) / 1e12
    except Exception as e:  # noqa
        out[f"int8_int_mm_n{n}_error"] = str(e)[:200]
    a16, b16 = A.to(torch.bfloat16), B.to(torch.bfloat16)
    ms = best_ms(lambda: torch.matmul(a16, b16))
    out[f"bf16_n{n}_tflops"] = 2.0 * n ** 3 / (ms * 1e-3) / 1e12
    try:
        a8, b8 = A.to(torch.float8_e4m3fn), B.t().contiguous().to(torch.float8_e4m3fn).t()
        one = torch.tensor(1.0, device="cuda")
        ms = best_ms(lambda: torch._scaled_mm(a8, b8, scale_a=one, scale_b=one, out_dtype=torch.bfloat16))
        out[f"fp8_scaled_mm_n{n}_tflops"] = 2.0 * n ** 3 / (ms * 1e-3) / 1e12
    except Exception as e:  # noqa
        out[f"fp8_n{n}_error"] = str(e)[:200]
    del A, B, a16, b16
print(json.dumps(out))
