import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops
def timeit(fn, reps=20):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
for n, p, k in ((15625, 50000, 32), (180000, 4096, 32)):
    X = ops.synth_genotypes(n, p, "cuda")
    idx, d2 = ops.knn(X, k, impl=2)
    ei, ew = ops.build_edges(idx, d2, mode="distance", smoothing="laplacian", use_weights=True)
    g = ops.gcn_graph(ei, ew.float(), n, True)
    H = torch.randn(n, 100, device="cuda"); b = torch.randn(100, device="cuda")
    ms0 = timeit(lambda: ops.spmm(g.rowptr, g.col, g.val, H, b))
    ms = timeit(lambda: ops.spmm(g.rowptr, g.col, g.val, H, b, hub=g.hub))
    ref = torch.sparse_csr_tensor(g.rowptr.long(), g.col.long(), g.val, (n, n)).to_dense() @ H + b if n < 20000 else None
    out = ops.spmm(g.rowptr, g.col, g.val, H, b, hub=g.hub)
    err = float((out - ref).abs().max() / ref.abs().max()) if ref is not None else -1
    print(f"n={n} nnz={g.nnz} maxdeg={int((g.rowptr[1:]-g.rowptr[:-1]).max())}: planned {ms*1000:.1f} us ({g.agg_bytes(100)/ms/1e6:.0f} GB/s alg), unplanned {ms0*1000:.1f} us, tasks {0 if g.hub is None else g.hub[0].numel()}  err {err:.1e}", flush=True)
