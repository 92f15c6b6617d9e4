"""Times the inner-layer (h x h) products at the C5 subgraph shape: forward, dX, weight gradient (fp32)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops
n, h = 15625, 100
A = torch.randn(n, h, device="cuda"); W = torch.randn(h, h, device="cuda") * 0.1; dH = torch.randn(n, h, device="cuda")
def t(fn, reps=20):
    """GPU time per call: the calls are replayed as one CUDA graph so that the host-side wrapper cost does not show."""
    fn(); fn(); torch.cuda.synchronize()
    side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        fn()
    torch.cuda.current_stream().wait_stream(side); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps): fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1000
def err(a, b): return float((a.double() - b).abs().max() / b.abs().max())
print(f"fwd   {t(lambda: ops.linear_fwd(A, W, precision=0)):.1f} us  err {err(ops.linear_fwd(A, W, precision=0), A.double() @ W.double().t()):.1e}")
print(f"bwd_x {t(lambda: ops.linear_bwd_x(dH, W)):.1f} us  err {err(ops.linear_bwd_x(dH, W), dH.double() @ W.double()):.1e}")
print(f"bwd_w {t(lambda: ops.linear_bwd_w(A, dH, precision=0)):.1f} us  err {err(ops.linear_bwd_w(A, dH, precision=0), dH.double().t() @ A.double()):.1e}")
