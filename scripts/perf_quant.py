"""Projection forward at the C5 shape (15 625 x 50 000 x 100, packed genotypes) with the weight quantised by one launch
(default) and by the two-kernel path (GNNB_QUANT_FUSED=0): CUDA-graph replays, CUDA events."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops

n, p, h = 15625, 50000, 100
X = ops.pack_2bit(ops.synth_genotypes(n, p, torch.device("cuda"), row0=0, p_total=p))
W = torch.randn(h, p, device="cuda") * 0.01


def t(fn, it=20):
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


outs = {}
for mode in ("1", "0"):
    os.environ["GNNB_QUANT_FUSED"] = mode
    outs[mode] = ops.linear_fwd(X, W)
    print(f"GNNB_QUANT_FUSED={mode}: linear_fwd {t(lambda: ops.linear_fwd(X, W)):.1f} us")
print("bit-identical:", torch.equal(outs["0"], outs["1"]))
