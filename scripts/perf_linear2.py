import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops
def timeit(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
n, p, h = 15625, 50000, 100
X = ops.synth_genotypes(n, p, "cuda")
W = torch.randn(h, p, device="cuda") * 0.01
dH = torch.randn(n, h, device="cuda")
out = []
for prec in (1, 2, 3):
    out.append(f"T{prec}: fwd {timeit(lambda: ops.linear_fwd(X, W, precision=prec)):.3f} bwd {timeit(lambda: ops.linear_bwd_w(X, dH, precision=prec)):.3f}")
print(os.environ.get("GNNB_ABLATE", "0"), " | ".join(out), flush=True)
