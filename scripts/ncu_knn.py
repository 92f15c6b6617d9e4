"""Two tensor-core kNN passes over one C5 subgraph (15625 x 50000, k = 32): the target of an ncu capture."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops

n, p, k = 15625, 50000, 32
X = ops.synth_genotypes(n, p, "cuda")
for _ in range(2):
    idx, d2 = ops.knn(X, k, impl=2)
torch.cuda.synchronize()
print("ok", int(idx[0, 0]), float(d2[0, 0]))
