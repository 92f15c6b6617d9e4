"""Timing of the int8 and FP4 neighbour searches at a given shape (n p k)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops
n, p, k = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
X = ops.synth_genotypes(n, p, "cuda"); X4 = ops.pack_fp4(X); sq = ops.row_sqnorm(X)
d2 = torch.empty((n, k), dtype=torch.float64, device="cuda"); idx = torch.empty((n, k), dtype=torch.int32, device="cuda")
def t(fn, it=2):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it
fl = 2.0 * n * n * p
ms8 = t(lambda: ops.knn_block(X, sq, 0, X, sq, 0, k, d2, idx, 0, impl=2)); ref = idx.clone()
ms4 = t(lambda: ops.knn_block_fp4(X4, sq, 0, X4, sq, 0, p, k, d2, idx, 0))
print(f"{n} x {p} k={k} splits={os.environ.get('GNNB_KNN_SPLITS','auto')}: int8 {ms8:.2f} ms = {fl/ms8/1e9:.0f} TOP/s | fp4 {ms4:.2f} ms = {fl/ms4/1e9:.0f} TOP/s | same {torch.equal(ref, idx)}", flush=True)
