"""Times one tensor-core kNN pass (results may be invalid under GNNB_KNN_ABLATE)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops
n, p, k = 15625, 50000, 32
X = ops.synth_genotypes(n, p, "cuda")
ops.knn(X, k, impl=2); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3): ops.knn(X, k, impl=2)
e1.record(); torch.cuda.synchronize()
print(f"ablate={os.environ.get('GNNB_KNN_ABLATE','0')}: {e0.elapsed_time(e1)/3:.2f} ms")
