"""Three aggregations over one C5 subgraph's kNN Laplacian graph (15625 nodes, k = 32, F = 100): the target of an ncu capture."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops
n, p, k = 15625, 50000, 32
X = ops.synth_genotypes(n, p, "cuda")
idx, d2 = ops.knn(X, k, impl=2)
ei, ew = ops.build_edges(idx, d2, mode="distance", smoothing="laplacian", use_weights=True)
g = ops.gcn_graph(ei, ew.float(), n, True)
H = torch.randn(n, 100, device="cuda"); b = torch.randn(100, device="cuda")
for _ in range(3):
    out = ops.spmm(g.rowptr, g.col, g.val, H, b, hub=g.hub)
torch.cuda.synchronize()
print("ok", g.nnz, float(out.abs().max()))
