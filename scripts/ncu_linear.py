"""Two forward and two weight-gradient projections at the C5 subgraph shape (T = 2): the target of an ncu capture."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops

n, p, h = 15625, 50000, 100
X = ops.synth_genotypes(n, p, "cuda")
W = torch.randn(h, p, device="cuda") * 0.01
dH = torch.randn(n, h, device="cuda")
for _ in range(2):
    H = ops.linear_fwd(X, W, precision=2)
for _ in range(2):
    dW = ops.linear_bwd_w(X, dH, precision=2)
torch.cuda.synchronize()
print("ok", float(H.abs().max()), float(dW.abs().max()))
