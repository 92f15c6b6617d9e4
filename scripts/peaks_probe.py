"""Library GEMM rates on this GPU for bf16, fp8 (e4m3) and int8 -- the practical tensor peaks the kernels are judged by."""
import torch
def t(fn, reps=20):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
for n in (8192, 16384):
    a = torch.randn(n, n, device="cuda", dtype=torch.bfloat16); b = torch.randn(n, n, device="cuda", dtype=torch.bfloat16)
    ms = t(lambda: a @ b); print(f"bf16 {n}: {2*n**3/ms/1e9:.0f} TFLOP/s", flush=True)
    try:
        a8 = a.to(torch.float8_e4m3fn); b8 = b.to(torch.float8_e4m3fn).t().contiguous().t()
        one = torch.ones((), device="cuda")
        ms = t(lambda: torch._scaled_mm(a8, b8, scale_a=one, scale_b=one, out_dtype=torch.bfloat16)); print(f"fp8  {n}: {2*n**3/ms/1e9:.0f} TFLOP/s", flush=True)
    except Exception as e:
        print("fp8 failed:", str(e)[:200])
    try:
        ai = torch.randint(-8, 8, (n, n), device="cuda", dtype=torch.int8); bi = torch.randint(-8, 8, (n, n), device="cuda", dtype=torch.int8)
        ms = t(lambda: torch._int_mm(ai, bi)); print(f"int8 {n}: {2*n**3/ms/1e9:.0f} TOP/s", flush=True)
    except Exception as e:
        print("int8 failed:", str(e)[:200])
