"""torchrun --nproc-per-node G scripts/check_knn_sharded.py : row-sharded kNN == single-GPU kNN, bit for bit."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnn_b200 import ops, parallel

rank, world, local_rank = parallel.init_distributed()
torch.cuda.set_device(local_rank)
dev = torch.device("cuda", local_rank)
n, p, k = 10000, 4096, 32
per = (n + world - 1) // world
lo, hi = min(n, rank * per), min(n, (rank + 1) * per)
X_local = ops.synth_genotypes(hi - lo, p, dev, row0=lo, p_total=p)
y = torch.arange(n, dtype=torch.float64, device=dev).mul_(0.37).sin_().reshape(-1, 1)
for extra in (None, y):
    idx, d2 = parallel.knn_sharded(X_local, lo, n, k, rank, world, None if extra is None else extra[lo:hi].contiguous())
    X_all = ops.synth_genotypes(n, p, dev)
    ridx, rd2 = ops.knn(X_all, k, extra=extra)
    ok = torch.equal(idx, ridx[lo:hi]) and torch.equal(d2, rd2[lo:hi])
    flag = torch.tensor([int(ok)], device=dev)
    if world > 1:
        torch.distributed.all_reduce(flag, op=torch.distributed.ReduceOp.MIN)
    if rank == 0:
        print(f"knn_sharded world={world} extra={'yes' if extra is not None else 'no'}: {'OK' if int(flag) else 'MISMATCH'}", flush=True)
if world > 1:
    torch.distributed.destroy_process_group()
