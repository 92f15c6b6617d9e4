"""ncu target: the FP4 neighbour search at the C5 subgraph shape, twice (first = warm-up)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops
X = ops.synth_genotypes(15625, 50000, "cuda")
for _ in range(2):
    idx, d2 = ops.knn(X, 32, impl=3)
torch.cuda.synchronize()
print("ok", int(idx[0, 0]))
