"""CUDA-event timings of the inner-layer products (15 625 / 180 000 rows x 100 x 100, float32) and a check against float64."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops


def t(fn, it=50):
    """device time per call: `it` calls captured in one CUDA graph (no host launch cost), replayed 5 times"""
    for _ in range(3): fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(it): fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (5 * it) * 1e3


torch.manual_seed(0)
for n, h, p in ((15625, 100, 100), (180000, 100, 100), (15625, 64, 100), (15000, 100, 98)):
    X = torch.randn(n, p, device="cuda"); W = torch.randn(h, p, device="cuda"); dH = torch.randn(n, h, device="cuda")
    err = lambda a, b: float((a.double() - b).abs().max() / b.abs().max())
    e = (err(ops.linear_fwd(X, W), X.double() @ W.double().t()), err(ops.linear_bwd_x(dH, W), dH.double() @ W.double()),
         err(ops.linear_bwd_w(X, dH), dH.double().t() @ X.double()))
    print(f"{n}x{p}x{h}: fwd {t(lambda: ops.linear_fwd(X, W)):.1f} us (torch {t(lambda: X @ W.t()):.1f}) | "
          f"bwd_x {t(lambda: ops.linear_bwd_x(dH, W)):.1f} us (torch {t(lambda: dH @ W):.1f}) | "
          f"bwd_w {t(lambda: ops.linear_bwd_w(X, dH)):.1f} us (torch {t(lambda: dH.t() @ X):.1f}) | max rel err {e[0]:.1e} {e[1]:.1e} {e[2]:.1e}")
