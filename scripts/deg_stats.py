import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops
n, p, k = 15625, 50000, 32
X = ops.synth_genotypes(n, p, "cuda")
idx, d2 = ops.knn(X, k, impl=2)
ei, ew = ops.build_edges(idx, d2, mode="distance", smoothing="laplacian", use_weights=True)
g = ops.gcn_graph(ei, ew.float(), n, True)
deg = (g.rowptr[1:] - g.rowptr[:-1]).long()
print("nnz", g.nnz, "mean", float(deg.float().mean()), "max", int(deg.max()), "min", int(deg.min()))
for q in (0.5, 0.9, 0.99, 0.999):
    print("q", q, int(torch.quantile(deg.float(), q)))
for t in (64, 128, 256, 512, 1024, 2048, 4096):
    m = deg > t
    print(f"rows > {t}: {int(m.sum())}  nnz in them: {int(deg[m].sum())} ({100.0*int(deg[m].sum())/g.nnz:.1f}%)")
print("top10", sorted(deg.tolist())[-10:])
