"""Two NCCL ranks on two GPUs: the data-parallel step with the gradient sink (range-wise overlapped all-reduce), the
row-sharded evaluation and the row-sharded neighbour search, each against its single-GPU counterpart.  Skipped on a box
with fewer than two GPUs (run it with ``gpurun --gpus 2``)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _make(rank_seed, n, p, dev):
    from gnn_b200 import ops
    from gnn_b200.data import GenotypeTable
    from gnn_b200.loaders.load import create_graph
    X = ops.synth_genotypes(n, p, dev, row0=1000 * rank_seed, p_total=p)
    y = np.random.default_rng(rank_seed).standard_normal(n)
    return create_graph(GenotypeTable(X, y), n_neighbors=6, smoothing="laplacian", use_weights=True, mode="distance",
                        algorithm="euclidean")


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import torch.distributed as dist
    from gnn_b200 import ops, parallel
    from gnn_b200.networks import networks as nets
    from gnn_b200.trainers.util import get_args_tuple
    parallel.init_distributed()
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    n, p, h = 700, 2048, 32
    graphs = [_make(s, n, p, dev) for s in range(4)]                 # every rank can build every subgraph (oracle below)
    for g in graphs:
        g.x = ops.pack_2bit(g.x)
    torch.manual_seed(0)
    model = nets.create_network_two_no_conv_dropout(p, 1, internal_size=h, dropout=0.0).to(dev)
    ref = nets.create_network_two_no_conv_dropout(p, 1, internal_size=h, dropout=0.0).to(dev)
    parallel.broadcast_parameters(model)
    ref.load_state_dict(model.state_dict())
    flat = parallel.FlatGrads(model.parameters(), sink_layer=parallel.first_genotype_layer(model), world=world, averaging="nccl", model=model)
    flat.range_markers, flat.cut = 512, "every_wave"                  # four ranges at p = 2048: staged, overlapped reduces
    opt = torch.optim.Adam(model.parameters(), lr=1e-2)
    opt_ref = torch.optim.Adam(ref.parameters(), lr=1e-2)
    lf = torch.nn.MSELoss()
    mine = parallel.owned(len(graphs), rank, world)
    for step, b in enumerate(mine):
        parallel.dp_step(model, flat, opt, graphs[b], world, lf, get_args_tuple)
        opt_ref.zero_grad()
        for r in range(world):                                        # gradient averaging over the step's subgraphs
            g = graphs[step * world + r]
            (lf(ref(*get_args_tuple(g)), g.y) / world).backward()
        for (k, a), (_, c) in zip(model.named_parameters(), ref.named_parameters()):
            err = float((a.grad - c.grad).abs().max() / c.grad.abs().max().clamp_min(1e-30))
            assert err < 1e-5, (step, k, err)
        opt_ref.step()
    for a, c in zip(model.parameters(), ref.parameters()):
        assert float((a - c).abs().max()) < 1e-5
    # gradient averaging by our own kernels over a symmetric flat buffer (NVLink peer loads / stores, and multimem through
    # the NVSwitch) against NCCL's all-reduce; all replicas must end up with the same bits
    for mode in ("nvls", "p2p"):
        torch.manual_seed(1)
        m2 = nets.create_network_two_no_conv_dropout(p, 1, internal_size=h, dropout=0.0).to(dev)
        f2 = parallel.FlatGrads(m2.parameters(), sink_layer=parallel.first_genotype_layer(m2), world=world, averaging=mode, model=m2)
        assert f2.nvls is not None and f2.averaging == mode, f2.nvls_error
        for trial in range(3):
            vals = torch.randn(f2.flat.numel(), device=dev, generator=torch.Generator(device=dev).manual_seed(10 * trial + rank))
            f2.flat.copy_(vals)
            want = vals.clone()
            dist.all_reduce(want, op=dist.ReduceOp.AVG)
            f2.allreduce_mean(world)
            assert float((f2.flat - want).abs().max()) <= 1e-6 * float(want.abs().max()), mode
            both = [torch.empty_like(f2.flat) for _ in range(world)]
            dist.all_gather(both, f2.flat.contiguous())
            assert torch.equal(both[0], both[1]), mode            # replicas hold the same bits
    # ... and a whole data-parallel step on that path (sink writes dW in place, one NVLS pass after the backward)
    ref2 = nets.create_network_two_no_conv_dropout(p, 1, internal_size=h, dropout=0.0).to(dev)
    parallel.broadcast_parameters(m2)
    ref2.load_state_dict(m2.state_dict())
    opt2 = torch.optim.Adam(m2.parameters(), lr=1e-2)
    for step, b in enumerate(mine):
        parallel.dp_step(m2, f2, opt2, graphs[b], world, lf, get_args_tuple)
        ref2.zero_grad()
        for r in range(world):
            g = graphs[step * world + r]
            (lf(ref2(*get_args_tuple(g)), g.y) / world).backward()
        for (k, a), (_, c) in zip(m2.named_parameters(), ref2.named_parameters()):
            err = float((a.grad - c.grad).abs().max() / c.grad.abs().max().clamp_min(1e-30))
            assert err < 1e-5, ("nvls", step, k, err)
        with torch.no_grad():
            for a, c in zip(m2.parameters(), ref2.parameters()):
                c.copy_(a)
    # row-sharded evaluation == the unsharded model, bit for bit
    test = graphs[3]
    whole = parallel.sharded_eval(model, test, 0, 1, get_args_tuple)
    shard = parallel.sharded_eval(model, test, rank, world, get_args_tuple)
    assert torch.equal(whole, shard)
    # row-sharded neighbour search == one GPU, bit for bit
    N = 1500
    X = ops.synth_genotypes(N, p, dev, row0=7, p_total=p)
    lo, hi = (0, 800) if rank == 0 else (800, N)
    idx, d2 = parallel.knn_sharded(X[lo:hi], lo, N, 8, rank, world)
    idx1, d21 = ops.knn(X, 8)
    assert torch.equal(idx, idx1[lo:hi]) and torch.equal(d2, d21[lo:hi])
    # the same search with the key blocks passed round a ring (double-buffered isend / irecv), with a phenotype column
    extra = torch.from_numpy(np.random.default_rng(1).standard_normal((N, 1))).to(dev)
    idx_r, d2_r = parallel.knn_ring(X[lo:hi], lo, N, 8, rank, world, extra_local=extra[lo:hi])
    idx_e, d2_e = ops.knn(X, 8, extra=extra)
    assert torch.equal(idx_r, idx_e[lo:hi]) and torch.equal(d2_r, d2_e[lo:hi])
    # rows by global id from the sharded matrix
    sizes, offs = parallel.shard_sizes(hi - lo, rank, world, dev)
    wanted = torch.randperm(N, generator=torch.Generator().manual_seed(rank))[:300]
    store = X.as_strided((N, X.stride(0)), (X.stride(0), 1))
    got = parallel.fetch_rows(store[lo:hi], sizes, offs, wanted, rank, world)
    assert torch.equal(got, store[wanted.to(dev)])
    # one graph built from row shards: same edge list as on one GPU, and the sharded evaluation accepts the shard
    from gnn_b200.data import GenotypeTable
    from gnn_b200.loaders.load import create_graph
    yv = np.random.default_rng(2).standard_normal(N)
    one = create_graph(GenotypeTable(X, yv), n_neighbors=6, smoothing="laplacian", use_weights=True, mode="distance")
    slo, shi = parallel.row_shard(N, rank, world)
    shd = parallel.create_graph_sharded(X[slo:shi], yv, 6, rank, world, use_weights=True, mode="distance")
    assert torch.equal(shd.edge_index, one.edge_index) and torch.equal(shd.edge_weight, one.edge_weight)
    assert torch.equal(shd.y, one.y) and shd.x_shard == (slo, shi)
    big = nets.create_network_two_no_conv_dropout(p, 1, internal_size=h, dropout=0.0).to(dev)
    parallel.broadcast_parameters(big)
    assert torch.equal(parallel.sharded_eval(big, shd, rank, world, get_args_tuple),
                       parallel.sharded_eval(big, one, 0, 1, get_args_tuple))
    torch.cuda.synchronize()
    dist.barrier()
    if rank == 0:
        open(out, "w").write("ok")
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_nccl_ranks(tmp_path):
    out = str(tmp_path / "ok")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    assert open(out).read() == "ok"
