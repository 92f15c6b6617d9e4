"""FP4 neighbour search against the int8 tensor-core search: bit-equality at a few shapes, then timings at the C5 shapes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops
def t(fn, it=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it
for n, p, k in ((300, 600, 5), (1000, 4096, 10), (3534, 5003, 10), (2000, 20001, 32), (15625, 50000, 32)):
    X = ops.synth_genotypes(n, p, "cuda", row0=11)
    a_i, a_d = ops.knn(X, k, impl=2)
    b_i, b_d = ops.knn(X, k, impl=3)
    same = torch.equal(a_i, b_i) and torch.equal(a_d, b_d)
    print(f"n={n} p={p} k={k}: fp4 == int8 {same}", flush=True)
    if not same:
        bad = (a_i != b_i).any(dim=1).nonzero().flatten()
        print("  rows differing:", bad.numel(), "first", bad[:5].tolist(), a_d[bad[0]][:6].tolist(), b_d[bad[0]][:6].tolist())
        sys.exit(1)
extra = torch.randn(3534, 1, device="cuda", dtype=torch.float64)
X = ops.synth_genotypes(3534, 5003, "cuda")
a = ops.knn(X, 8, extra=extra, impl=2); b = ops.knn(X, 8, extra=extra, impl=3)
print("with a float64 side column:", torch.equal(a[0], b[0]) and torch.equal(a[1], b[1]), flush=True)
X = ops.synth_genotypes(15625, 50000, "cuda")
X4 = ops.pack_fp4(X)
sq = ops.row_sqnorm(X)
d2 = torch.empty((15625, 32), dtype=torch.float64, device="cuda"); idx = torch.empty((15625, 32), dtype=torch.int32, device="cuda")
ms8 = t(lambda: ops.knn_block(X, sq, 0, X, sq, 0, 32, d2, idx, 0, impl=2))
ms4 = t(lambda: ops.knn_block_fp4(X4, sq, 0, X4, sq, 0, 50000, 32, d2, idx, 0))
msp = t(lambda: ops.pack_fp4(X))
fl = 2.0 * 15625 * 15625 * 50000
print(f"15625 x 50000: int8 {ms8:.2f} ms = {fl / ms8 / 1e9:.0f} TOP/s | fp4 {ms4:.2f} ms = {fl / ms4 / 1e9:.0f} TOP/s | pack {msp:.3f} ms", flush=True)
