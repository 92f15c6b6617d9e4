"""Times the kNN kernels (CUDA events) on one subgraph-sized problem and a larger one."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops

def timeit(fn, reps=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

for n, p, k in ((15625, 50000, 32), (65536, 8192, 32)):
    X = ops.synth_genotypes(n, p, "cuda")
    for impl in (2, 1):
        if impl == 1 and n * n * p > 2e13: 
            continue
        ms = timeit(lambda: ops.knn(X, k, impl=impl), reps=2)
        print(f"n={n} p={p} k={k} impl={impl}: {ms:.2f} ms  {2.0*n*n*p/ms/1e9:.1f} TOP/s", flush=True)
    a = ops.knn(X, k, impl=2); 
    if n * n * p <= 2e13:
        b = ops.knn(X, k, impl=1)
        print("equal:", torch.equal(a[0], b[0]) and torch.equal(a[1], b[1]))
