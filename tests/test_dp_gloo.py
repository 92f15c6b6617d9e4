"""World-size-2 gloo test of the data-parallel host logic (no GPU): ownership of subgraph batches, the
flat-gradient all-reduce, and the row-sharded all-gather used by the evaluation pass."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gnn_b200 import parallel


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    r, w, _ = parallel.init_distributed(backend="gloo")
    assert (r, w) == (rank, world)
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(6, 4), torch.nn.ReLU(), torch.nn.Linear(4, 1))
    parallel.broadcast_parameters(model)
    flat = parallel.FlatGrads(model.parameters())
    opt = torch.optim.Adam(model.parameters(), lr=0.05)
    g = torch.Generator().manual_seed(123)
    batches = [(torch.randn(10, 6, generator=g), torch.randn(10, 1, generator=g)) for _ in range(6)]
    mine = parallel.owned(len(batches), rank, world)
    assert mine == list(range(rank, 6, world))
    losses = []
    for b in mine:                                     # one DP step per owned batch, as parallel.dp_step does
        x, y = batches[b]
        flat.zero()
        loss = torch.nn.functional.mse_loss(model(x), y)
        loss.backward()
        flat.allreduce_mean(world)
        opt.step()
        losses.append(float(loss))
    # sharded rows -> all-gather
    n = 300
    full = torch.arange(n * 3, dtype=torch.float32).reshape(n, 3)
    lo, hi = parallel.row_shard(n, rank, world, align=128)
    gathered = parallel.allgather_rows(full[lo:hi].clone(), n, rank, world, align=128)
    assert torch.equal(gathered, full)
    state = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    if rank == 0:
        torch.save({"state": state, "losses": losses}, out)
    gathered_state = [torch.zeros_like(state) for _ in range(world)]
    dist.all_gather(gathered_state, state)
    assert all(torch.equal(gathered_state[0], s) for s in gathered_state)      # replicas stay identical
    dist.destroy_process_group()


def test_dp_two_ranks_matches_gradient_averaging(tmp_path):
    world, port, out = 2, _free_port(), str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    got = torch.load(out)
    # single-process oracle of the same schedule: average the per-subgraph gradients of a step, one Adam step
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(6, 4), torch.nn.ReLU(), torch.nn.Linear(4, 1))
    opt = torch.optim.Adam(model.parameters(), lr=0.05)
    g = torch.Generator().manual_seed(123)
    batches = [(torch.randn(10, 6, generator=g), torch.randn(10, 1, generator=g)) for _ in range(6)]
    for step in range(3):
        opt.zero_grad()
        for r in range(world):
            x, y = batches[step * world + r]
            (torch.nn.functional.mse_loss(model(x), y) / world).backward()
        opt.step()
    want = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    torch.testing.assert_close(got["state"], want, rtol=1e-5, atol=1e-6)


def test_row_shard_covers_everything():
    for n in (1, 127, 128, 129, 1000, 180000):
        for world in (1, 2, 4, 8):
            spans = [parallel.row_shard(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert all((lo % 128 == 0) for lo, _ in spans if lo < n)


def _exchange_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    parallel.init_distributed(backend="gloo")
    total, ld = 11, 16
    full = (torch.arange(total * ld, dtype=torch.int32) % 127).to(torch.int8).reshape(total, ld)
    cut = [0, 4, 7, 11][:world + 1] if world == 3 else [0, 7, 11]
    lo, hi = cut[rank], cut[rank + 1]
    local = full[lo:hi].clone()
    sizes, offs = parallel.shard_sizes(hi - lo, rank, world, torch.device("cpu"))
    assert sizes == [b - a for a, b in zip(cut, cut[1:])] and offs == cut[:-1]
    # ring: every owner's block is folded exactly once, own block first, with the right contents
    seen = []

    def fold(owner, block):
        assert torch.equal(block[0], full[cut[owner]:cut[owner + 1]]) and int(block[1][0]) == owner
        seen.append(owner)

    tag = torch.full((hi - lo,), rank, dtype=torch.int64)
    parallel.ring_pass((local, tag), sizes, rank, world, fold,
                       lambda rows: (torch.empty((rows, ld), dtype=torch.int8), torch.empty(rows, dtype=torch.int64)))
    assert seen == [(rank - s) % world for s in range(world)]
    # shards of unequal length -> the whole matrix on every rank
    assert torch.equal(parallel.allgather_shards(local, sizes, rank, world), full)
    # rows by global id, any order, repeats, rows of both owners
    wanted = [10, 0, 3, 3, 8, 6, 7, 1] if rank == 0 else [2, 9, 2, 5]
    got = parallel.fetch_rows(local, sizes, offs, wanted, rank, world)
    assert torch.equal(got, full[torch.tensor(wanted)])
    dist.barrier()
    if rank == 0:
        open(out, "w").write("ok")
    dist.destroy_process_group()


def test_ring_allgather_and_row_exchange_two_ranks(tmp_path):
    out = str(tmp_path / "ok")
    mp.spawn(_exchange_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    assert open(out).read() == "ok"
