"""Data-parallel training over subsampled subgraphs: one process per GPU, NCCL over NVLink.

The reference has no distributed code; its closest analogue is ``train_batches``
(gnn/trainers/batches.py:31-68), one optimiser step per subgraph.  Here subgraphs are independent units:
rank r owns batches r, r + G, r + 2G, ...; one step = every rank runs forward/backward on its next
subgraph, the flat fp32 gradient buffer is summed with ONE all-reduce and divided by G, and every rank
applies the same Adam step.  The per-epoch evaluation on the test graph shards the expensive marker
projection by rows and all-gathers the [n, h] result; the cheap remainder is replicated.

Nothing here needs a GPU to be tested: the host logic runs under gloo in tests/test_dp_gloo.py.
"""
import os

import torch
import torch.distributed as dist


def init_distributed(backend=None):
    """(rank, world, local_rank) from the torchrun environment; initialises the process group if world > 1."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local_rank)
        dist.init_process_group(backend=backend, rank=rank, world_size=world)
    return rank, world, local_rank


def owned(n_items, rank, world):
    """Indices of the units (subgraph batches) rank owns: round robin."""
    return list(range(rank, n_items, world))


def row_shard(n_rows, rank, world, align=128):
    """[lo, hi) of the rows rank projects in the sharded evaluation; shard sizes are multiples of `align`."""
    per = (n_rows + world - 1) // world
    per = (per + align - 1) // align * align
    lo = min(n_rows, rank * per)
    return lo, min(n_rows, lo + per)


class FlatGrads:
    """All gradients of a model as views into one contiguous fp32 buffer -> a single all-reduce per step."""

    def __init__(self, params):
        self.params = [p for p in params if p.requires_grad]
        total = sum(p.numel() for p in self.params)
        dev = self.params[0].device
        self.flat = torch.zeros(total, dtype=torch.float32, device=dev)
        off = 0
        for p in self.params:
            p.grad = self.flat[off:off + p.numel()].view_as(p)
            off += p.numel()

    def zero(self):
        self.flat.zero_()

    def allreduce_mean(self, world, group=None):
        if world > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)
            self.flat.div_(world)


def broadcast_parameters(model, src=0):
    if dist.is_initialized() and dist.get_world_size() > 1:
        for p in model.parameters():
            dist.broadcast(p.data, src=src)
        for b in model.buffers():
            dist.broadcast(b.data, src=src)


def allgather_rows(local, n_rows, rank, world, align=128):
    """Concatenate row shards [lo_r, hi_r) of an [n_rows, F] matrix held one per rank."""
    if world == 1:
        return local
    per = row_shard(n_rows, 0, world, align)[1]
    pad = torch.zeros((per, local.shape[1]), dtype=local.dtype, device=local.device)
    pad[:local.shape[0]] = local
    out = torch.empty((world * per, local.shape[1]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, pad)
    return out[:n_rows]


def dp_step(model, flat, optimizer, batch, world, loss_func, args_of):
    """One data-parallel optimiser step on this rank's subgraph; returns the local loss (0-d tensor)."""
    model.train()
    flat.zero()
    out = model(*args_of(batch))
    loss = loss_func(out, batch.y)
    loss.backward()
    flat.allreduce_mean(world)
    optimizer.step()
    return loss.detach()


_shard_plans = {}


def _row_plan(g, lo, hi):
    """The CSR rows [lo, hi) of a cached GcnGraph as their own row pointer (offsets stay absolute) + hub-row plan."""
    from . import ops
    key = (g.rowptr.data_ptr(), lo, hi)
    plan = _shard_plans.get(key)
    if plan is None:
        rp = g.rowptr[lo:hi + 1].contiguous()
        nnz = int(rp[-1] - rp[0]) if hi > lo else 0
        plan = _shard_plans[key] = (rp, ops.hub_plan(rp) if hi > lo else None, nnz, g.rowptr)   # g.rowptr: keeps the key's address alive
    return plan[0], plan[1], plan[2]


@torch.no_grad()
def sharded_eval(model, data, rank, world, args_of):
    """Eval-mode forward on the test graph with every GCN layer sharded by node rows: rank r projects and aggregates the
    rows [lo_r, hi_r) only (each output row is one rank's complete, deterministic sum) and the [n, h] activations are
    all-gathered where the next aggregation needs every neighbour's row.  Row-wise modules run on the local rows."""
    from . import ops
    from .networks.gcn_conv import GCNConv
    model.eval()
    args = args_of(data)
    first = model.nns[0]
    if world == 1 or not isinstance(first, GCNConv):
        return model(*args)
    x, edge_index = args[0], args[1]
    edge_weight = args[2] if len(args) > 2 else None
    n = x.shape[0]
    lo, hi = row_shard(n, rank, world)
    cur = x[lo:hi]
    for m, f in zip(model.nns, model.fns):
        if f is not None:
            cur = f(cur)
        elif isinstance(m, GCNConv):
            h = allgather_rows(ops.linear_fwd(cur, m.lin.weight), n, rank, world)
            g = ops.gcn_graph(edge_index, edge_weight, n, m.add_self_loops)
            rp, hub, nnz = _row_plan(g, lo, hi)
            cur = ops.spmm(rp, g.col, g.val, h, m.bias, n=hi - lo, hub=hub, nnz=nnz)
        else:
            cur = m(cur)
    return allgather_rows(cur, n, rank, world)


def knn_sharded(X_local, row0, n_total, k, rank, world, extra_local=None):
    """Row-sharded kNN graph build: rank r owns query rows [row0, row0 + n_local) of the n_total
    individuals.  Key blocks are passed round by NCCL broadcast, one owner at a time, and folded into the
    owner-local running lists with gnnb_knn_i8(merge=1); no neighbour list is ever reduced across ranks.
    Returns (idx int32 [n_local, k] of GLOBAL row ids, d2 float64 [n_local, k]) -- bit-identical to ops.knn on the
    whole matrix, whatever the number of ranks."""
    from . import ops
    dev = X_local.device
    n_local, p = X_local.shape
    sq_local = ops.row_sqnorm(X_local)
    best_d2 = torch.empty((n_local, k), dtype=torch.float64, device=dev)
    best_idx = torch.empty((n_local, k), dtype=torch.int32, device=dev)
    if world == 1:
        ops.knn_block(X_local, sq_local, row0, X_local, sq_local, row0, k, best_d2, best_idx, 0, extra_local, extra_local)
        return best_idx, best_d2
    sizes = torch.zeros(world, dtype=torch.int64, device=dev)
    sizes[rank] = n_local
    dist.all_reduce(sizes)
    offs = torch.cumsum(sizes, 0) - sizes
    sizes, offs = sizes.tolist(), offs.tolist()
    ld = X_local.stride(0)
    n_extra = 0 if extra_local is None else extra_local.shape[1]
    for step, owner in enumerate(range(world)):
        nk = int(sizes[owner])
        if owner == rank:
            keys, sqk, ek = X_local, sq_local, extra_local
        else:
            buf = torch.empty((nk, ld), dtype=torch.int8, device=dev)
            keys, sqk = buf[:, :p], torch.empty(nk, dtype=torch.int64, device=dev)
            ek = torch.empty((nk, n_extra), dtype=torch.float64, device=dev) if n_extra else None
        # broadcast the padded storage so that the receiver keeps the 16-byte row stride TMA needs
        store = X_local.as_strided((n_local, ld), (ld, 1)) if owner == rank else buf
        dist.broadcast(store, src=owner)
        dist.broadcast(sqk, src=owner)
        if n_extra:
            dist.broadcast(ek, src=owner)
        ops.knn_block(X_local, sq_local, row0, keys, sqk, int(offs[owner]), k, best_d2, best_idx, int(step > 0),
                      extra_local, ek)
    return best_idx, best_d2


# ----------------------------------------------------------------------------------------------- CUDA-graphed steps
def make_capturable(optimizer):
    """True if ``optimizer.step()`` may be captured in a CUDA graph.  torch's Adam-family optimisers keep their step
    counter on the host unless ``capturable`` is set; a captured host-side counter would silently freeze the bias
    correction, so: groups that already say capturable=True are fine, an optimiser that has not stepped yet is switched
    (the update rule is unchanged), anything else is refused."""
    groups = getattr(optimizer, "param_groups", None)
    if not groups or any("capturable" not in g for g in groups):
        return False
    if all(g["capturable"] for g in groups):
        return True
    if len(optimizer.state) > 0:
        return False
    for g in groups:
        g["capturable"] = True
    return True


class GraphedStep:
    """A fixed piece of device work -- typically one epoch body / optimiser step on one subgraph (ensemble2.py:51-88,
    batches.py:34-61) -- that is run eagerly ``warmup`` times on a side stream, then captured into a CUDA graph and
    replayed.  The ~100 small launches of a step on a 10^2..10^4-node subgraph are host-bound when issued one by one
    ("small-graph regime"); a replay costs one launch.

    ``fn`` must do the same device work on every call, must not synchronise with the host, and must leave its results in
    tensors it returns (they belong to the graph's memory pool and are overwritten by the next replay -- copy what must
    survive).  Every call performs the work exactly once, whether eager, capturing or replaying.  If capture fails the
    object falls back to eager execution for good and keeps the reason in ``.error``."""

    _side = {}

    def __init__(self, fn, warmup=2, pool=None, enabled=True):
        self.fn, self.warmup, self.pool = fn, max(1, int(warmup)), pool
        self.enabled = bool(enabled)
        self.calls, self.graph, self.out, self.error = 0, None, None, None

    @classmethod
    def side_stream(cls):
        dev = torch.cuda.current_device()
        s = cls._side.get(dev)
        if s is None:
            s = cls._side[dev] = torch.cuda.Stream()
        return s

    def _eager_on_side_stream(self):
        cur, side = torch.cuda.current_stream(), self.side_stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            out = self.fn()
        cur.wait_stream(side)
        return out

    def __call__(self):
        self.calls += 1
        if self.graph is not None:
            self.graph.replay()
            return self.out
        if not self.enabled or self.error is not None:
            return self.fn()
        if self.calls <= self.warmup:
            return self._eager_on_side_stream()         # allocator, autograd and NCCL warm-up outside capture
        try:
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, pool=self.pool):
                out = self.fn()
            self.graph, self.out = graph, out
        except Exception as exc:                        # noqa: BLE001 -- eager execution is always available
            self.error = f"{type(exc).__name__}: {str(exc)[:300]}"
            torch.cuda.synchronize()
            return self.fn()
        self.graph.replay()                             # capture records, replay executes: the step happens once
        return self.out

    @property
    def captured(self):
        return self.graph is not None
