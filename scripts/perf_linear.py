"""Times the projection kernels (CUDA events) at the C5 subgraph shape."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops

def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

n, p, h = 15625, 50000, 100
X = ops.synth_genotypes(n, p, "cuda")
W = torch.randn(h, p, device="cuda") * 0.01
dH = torch.randn(n, h, device="cuda")
ref_f = X.float() @ W.t()
ref_b = dH.t() @ X.float()
for prec in (2, 4):
    ms_f = timeit(lambda: ops.linear_fwd(X, W, precision=prec))
    ms_b = timeit(lambda: ops.linear_bwd_w(X, dH, precision=prec))
    ef = float((ops.linear_fwd(X, W, precision=prec) - ref_f).abs().max() / ref_f.abs().max())
    eb = float((ops.linear_bwd_w(X, dH, precision=prec) - ref_b).abs().max() / ref_b.abs().max())
    fl = 2.0 * n * p * h
    print(f"precision={prec}: fwd {ms_f:.3f} ms ({fl/ms_f/1e9:.0f} TFLOP/s alg, {n*p/ms_f/1e6:.0f} GB/s of int8) err {ef:.2e} | "
          f"bwd_w {ms_b:.3f} ms ({fl/ms_b/1e9:.0f} TFLOP/s alg) err {eb:.2e}", flush=True)
Xp = ops.pack_2bit(X)
ms_f = timeit(lambda: ops.linear_fwd(Xp, W))
ms_b = timeit(lambda: ops.linear_bwd_w(Xp, dH, p=p))
same = torch.equal(ops.linear_fwd(Xp, W), ops.linear_fwd(X, W)) and torch.equal(ops.linear_bwd_w(Xp, dH, p=p), ops.linear_bwd_w(X, dH))
print(f"2-bit packed X (precision 4): fwd {ms_f:.3f} ms | bwd_w {ms_b:.3f} ms | bit-identical to int8 X: {same}", flush=True)
