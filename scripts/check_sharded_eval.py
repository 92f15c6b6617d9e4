"""torchrun --nproc-per-node N scripts/check_sharded_eval.py: the row-sharded eval forward (parallel.sharded_eval) must give
what the unsharded model gives on every rank (same kernels, each output row summed by one rank in the same order)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from gnn_b200 import ops, parallel
from gnn_b200.data import GenotypeTable
from gnn_b200.loaders.load import create_graph
from gnn_b200.networks import networks as nets
from gnn_b200.trainers.util import get_args_tuple

rank, world, local = parallel.init_distributed()
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
for n, p in ((3001, 640), (20000, 2048)):
    X = ops.synth_genotypes(n, p, dev)
    g = create_graph(GenotypeTable(X, np.random.default_rng(0).normal(size=n)), n_neighbors=8, use_weights=True, mode="distance")
    g.edge_weight = g.edge_weight.float()
    torch.manual_seed(0)
    for name in ("create_network_two_no_conv_dropout", "create_network_no_conv_relu_dropout"):
        model = getattr(nets, name)(p, 1, internal_size=100, dropout=0.3).to(dev)
        parallel.broadcast_parameters(model)
        with torch.no_grad():
            a = parallel.sharded_eval(model, g, rank, world, get_args_tuple)
            model.eval()
            b = model(*get_args_tuple(g))
        torch.cuda.synchronize()
        err = float((a - b).abs().max() / b.abs().max())
        print(f"rank {rank}/{world} n={n} {name}: shape {tuple(a.shape)} max rel diff {err:.2e} bit-equal {bool(torch.equal(a, b))}", flush=True)
        assert a.shape == b.shape and err < 1e-6
print("ok")
