import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops
n, p, h = 15625, 50000, 100
X = ops.synth_genotypes(n, p, "cuda")
W = torch.randn(h, p, device="cuda") * 0.01
dH = torch.randn(n, h, device="cuda")
for prec in (1, 2):
    for _ in range(2):
        ops.linear_fwd(X, W, precision=prec)
    ops.linear_bwd_w(X, dH, precision=prec)
torch.cuda.synchronize()
