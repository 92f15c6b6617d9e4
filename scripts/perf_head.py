"""CUDA-event timings of the dense head (Linear(100 -> 1) on 15 625 rows): own kernels against torch's library path."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops
n, h, o = 15625, 100, 1
A = torch.randn(n, h, device="cuda"); W = torch.randn(o, h, device="cuda"); b = torch.randn(o, device="cuda"); dO = torch.randn(n, o, device="cuda")
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
def torch_bwd():
    return dO @ W, dO.t() @ A, dO.sum(0)
print(f"head_fwd {t(lambda: ops.head_fwd(A, W, b)):.1f} us | torch linear {t(lambda: torch.nn.functional.linear(A, W, b)):.1f} us")
print(f"head_bwd {t(lambda: ops.head_bwd(dO, A, W)):.1f} us | torch dA, dW, db {t(torch_bwd):.1f} us")
