"""The hot kernels at the C5 subgraph shape (projection with int8 and with 2-bit packed X), twice each (first = warm-up): the target of one ncu --set full capture.
kNN (15625 x 50000, k = 32) -> graph -> projection forward / weight gradient (h = 100) -> aggregation (F = 100)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops
n, p, k, h = 15625, 50000, 32, 100
X = ops.synth_genotypes(n, p, "cuda")
for _ in range(2):
    idx, d2 = ops.knn(X, k, impl=2)
ei, ew = ops.build_edges(idx, d2, mode="distance", smoothing="laplacian", use_weights=True)
g = ops.gcn_graph(ei, ew.float(), n, True)
W = torch.randn(h, p, device="cuda") * 0.01
dH = torch.randn(n, h, device="cuda")
b = torch.randn(h, device="cuda")
for _ in range(2):
    H = ops.linear_fwd(X, W)
for _ in range(2):
    dW = ops.linear_bwd_w(X, dH)
Xp = ops.pack_2bit(X)                      # the resident format of the bench: the projection expands it on chip
for _ in range(2):
    Hp = ops.linear_fwd(Xp, W)
for _ in range(2):
    dWp = ops.linear_bwd_w(Xp, dH, p=p)
assert torch.equal(Hp, H) and torch.equal(dWp, dW)
for _ in range(2):
    out = ops.spmm(g.rowptr, g.col, g.val, H, b, hub=g.hub)
torch.cuda.synchronize()
print("ok", int(idx[0, 0]), float(H.abs().max()), float(dW.abs().max()), float(out.abs().max()))
