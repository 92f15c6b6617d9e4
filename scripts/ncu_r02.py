"""Round-2 ncu target: the hot kernels of one C5 training step on the bench's own code path (2-bit packed X, 24-bit digits,
weight gradient from prepared dH) and the aggregation on a graph far larger than L2, twice each (first = warm-up).
  ncu --set full --clock-control none --import-source on -k regex:'linear_p2|spmm_planned|knn_tc2' python scripts/ncu_r02.py
The 1M-node graph is a genuine k = 32 kNN Laplacian graph of 1M synthetic individuals over 1024 markers (the search over
50 000 markers would take 40 s per ncu pass); only its aggregation is the subject here."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops
n, p, k, h = 15625, 50000, 32, 100
X = ops.synth_genotypes(n, p, "cuda")
for _ in range(2):
    idx, d2 = ops.knn(X, k, impl=2)
ei, ew = ops.build_edges(idx, d2, mode="distance", smoothing="laplacian", use_weights=True)
ew = ew.float()
g = ops.gcn_graph(ei, ew, n, True)
W = torch.randn(h, p, device="cuda") * 0.01
dH = torch.randn(n, h, device="cuda")
b = torch.randn(h, device="cuda")
Xp = ops.pack_2bit(X)
for _ in range(2):
    H = ops.linear_fwd(Xp, W)
for _ in range(2):
    dW = ops.linear_bwd_w_ranges(Xp, dH, [(0, p)], p=p)
for _ in range(2):
    out = ops.spmm(g.rowptr, g.col, g.val, H, b, hub=g.hub)
if "--big" in sys.argv:
    N = 1_000_000
    Xb = ops.synth_genotypes(N, 1024, "cuda")
    idx, d2 = ops.knn(Xb, k, impl=2)
    del Xb
    ei, ew = ops.build_edges(idx, d2, mode="distance", smoothing="laplacian", use_weights=True)
    ew = ew.float()
    gb = ops.gcn_graph(ei, ew, N, True)
    Hb = torch.randn(N, h, device="cuda")
    for _ in range(2):
        ob = ops.spmm(gb.rowptr, gb.col, gb.val, Hb, b, hub=gb.hub)
    for _ in range(2):
        ob = ops.spmm(gb.rowptr_t, gb.col_t, gb.val_t, Hb, None, hub=gb.hub_t)
    print("big graph nnz", gb.nnz, "agg_bytes", gb.agg_bytes(h), "max degree", int((gb.rowptr[1:] - gb.rowptr[:-1]).max()))
torch.cuda.synchronize()
print("ok", float(H.abs().max()), float(dW.abs().max()), float(out.abs().max()))
