"""Times the int8-limb projection (results invalid under GNNB_LINEAR_ABLATE)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops
n, p, h = 15625, 50000, 100
X = ops.synth_genotypes(n, p, "cuda")
W = torch.randn(h, p, device="cuda") * 0.01
dH = torch.randn(n, h, device="cuda")
def t(fn, reps=10):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
print(f"ablate={os.environ.get('GNNB_LINEAR_ABLATE','0')}: fwd {t(lambda: ops.linear_fwd(X, W, precision=4)):.3f} ms  bwd_w {t(lambda: ops.linear_bwd_w(X, dH, precision=4)):.3f} ms")
