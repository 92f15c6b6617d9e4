"""One launch of each hot kernel at the C5 subgraph shape (15625 x 50000, k = 32, h = 100): the target of
the ncu captures kept under profiles/."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops

n, p, h, k = 15625, 50000, 100, 32
which = sys.argv[1] if len(sys.argv) > 1 else "all"
X = ops.synth_genotypes(n, p, "cuda")
W = torch.randn(h, p, device="cuda") * 0.01
dH = torch.randn(n, h, device="cuda")
torch.cuda.synchronize()
reps = 2
if which in ("all", "linear"):
    for _ in range(reps):
        ops.linear_fwd(X, W, precision=2)
        ops.linear_bwd_w(X, dH, precision=2)
if which in ("all", "knn", "spmm"):
    for _ in range(reps if which != "spmm" else 1):
        idx, d2 = ops.knn(X, k, impl=2)
if which in ("all", "spmm"):
    ei, ew = ops.build_edges(idx, d2, mode="distance", smoothing="laplacian", use_weights=True)
    g = ops.gcn_graph(ei, ew.float(), n, True)
    H = torch.randn(n, h, device="cuda")
    b = torch.randn(h, device="cuda")
    for _ in range(reps):
        out = ops.spmm(g.rowptr, g.col, g.val, H, b, hub=g.hub)
    print("nnz", g.nnz, "agg bytes", g.agg_bytes(h))
torch.cuda.synchronize()
print("done", which)
