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
    """All gradients of a model as views into one contiguous fp32 buffer.

    The marker projection of the first GCN layer owns almost all of it (h x p floats, 20 MB at C5).  Given that layer
    (``sink_layer``: a GCNConv fed with integer genotypes), its weight and bias come FIRST in the buffer and the layer gets
    a *gradient sink*: its backward writes dW and db straight into their slices -- no zero-fill of and no accumulate pass
    over 20 MB per step -- and, on several GPUs, starts the all-reduce of every finished marker range while the GEMM of
    the next range runs (``ops.linear_bwd_w_ranges``).  Everything else is the small tail.  Without a sink layer the
    buffer is zeroed and reduced in one piece."""

    def __init__(self, params, sink_layer=None, world=1, group=None, averaging="auto", model=None):
        params = [p for p in params if p.requires_grad]
        head = []
        if sink_layer is not None and getattr(sink_layer, "bias", None) is not None:
            head = [sink_layer.lin.weight, sink_layer.bias]
        # With ``model``: every other GCNConv and dense head of it (with a bias, all parameters trainable) also writes its
        # gradients in place -- its backward hands autograd None, so a step has no zero-fill of and no accumulate pass over
        # them either (seven small launches per step at C5).  The first backward of such a layer after zero() overwrites,
        # later ones in the same step accumulate as usual; a marked layer that takes no part in a step would keep the
        # previous step's gradient, so pass ``model`` only when every layer runs in every step (the GCN factories' do).
        self.inplace_layers = []
        owned = list(head)
        if model is not None:
            from . import ops
            from .networks.gcn_conv import GCNConv
            from .networks.head import Linear
            for m in model.modules():
                if m is sink_layer or not isinstance(m, (GCNConv, Linear)) or getattr(m, "bias", None) is None:
                    continue
                if isinstance(m, Linear) and not ops.head_supported(m.in_features, m.out_features):
                    continue                                      # runs through torch's Linear: autograd accumulates
                mine = [m.lin.weight if isinstance(m, GCNConv) else m.weight, m.bias]
                if all(q.requires_grad and any(q is r for r in params) for q in mine) and \
                        not any(q is r for q in mine for r in owned):
                    self.inplace_layers.append(m)
                    owned += mine
        params = owned + [p for p in params if all(p is not q for q in owned)]
        self.params = params
        total = sum(p.numel() for p in self.params)
        dev = self.params[0].device
        self.nvls, self.nvls_error = None, None
        # "p2p" | "nvls" (both: our kernels over symmetric memory) | "nccl".  Measured on 20 MB (barrier + kernel + barrier
        # against NCCL's all-reduce): 2 GPUs 48 / 70 / 62 us, 8 GPUs 81 / 66 / 106 us -- peer loads win while a rank has few
        # peers, the in-switch reduction once it would have to read seven of them
        self.averaging = ("p2p" if int(world) < 8 else "nvls") if averaging == "auto" else averaging
        if (int(world) > 1 and self.averaging in ("p2p", "nvls") and dev.type == "cuda" and dist.is_initialized()
                and dist.get_backend(group) == "nccl"):
            self._try_symmetric(total, dev, int(world), group)
        if self.nvls is None:
            self.flat = torch.zeros(total, dtype=torch.float32, device=dev)
        off = 0
        for p in self.params:
            p.grad = self.flat[off:off + p.numel()].view_as(p)
            off += p.numel()
        self.world, self.group, self.pending, self._stage = int(world), group, [], {}
        self.has_sink = bool(head)
        self.range_markers = None           # markers per wave of the weight-gradient kernel; None = ask the library
        self.cut = "last_wave"              # see _ranges
        self.tail = self.flat[sum(q.numel() for q in owned):]          # what autograd still accumulates into
        if head:
            self.w_numel = head[0].numel()
            sink_layer.grad_sink = self
        for m in self.inplace_layers:
            m.grad_inplace, m._grad_written = True, False

    # ---- averaging through the NVSwitch -------------------------------------------------------------------------------
    def _try_symmetric(self, total, dev, world, group):
        """Put the flat buffer into ONE symmetric allocation (every rank maps every replica, plus a multicast object over
        all of them), so that the step's gradient averaging is a single pass of our own kernel (csrc/comm.cu) between two
        cross-GPU barriers instead of a library all-reduce: "p2p" reads and writes the replicas over NVLink directly,
        "nvls" lets the switch add them in flight (multimem).  Falls back to NCCL (reason kept in ``nvls_error``) where
        symmetric memory is unavailable.  Collective: every rank must construct it."""
        try:
            import torch.distributed._symmetric_memory as symm
            pg = dist.group.WORLD if group is None else group
            padded = (total + 4 * world - 1) // (4 * world) * (4 * world)
            buf = symm.empty(padded, dtype=torch.float32, device=dev)
            handle = symm.rendezvous(buf, pg.group_name)
            usable = len(handle.buffer_ptrs) == world and all(int(p) != 0 for p in handle.buffer_ptrs)
            if self.averaging == "nvls":
                usable = usable and bool(handle.has_multicast_support) and int(handle.multicast_ptr) != 0
            ok = torch.tensor([1 if usable else 0], device=dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
            if not int(ok):
                raise RuntimeError("symmetric memory (peer mappings / multicast) is not usable on this box")
            buf.zero_()
            off = int(getattr(handle, "offset", 0) or 0)
            import ctypes
            peers = (ctypes.c_uint64 * world)(*[int(p) + off for p in handle.buffer_ptrs])
            if int(peers[dist.get_rank(group)]) != buf.data_ptr():
                raise RuntimeError("symmetric buffer: the local mapping is not where the tensor lives")
            mc = int(handle.multicast_ptr) + off if handle.has_multicast_support else 0
            self.nvls = (handle, buf, mc, padded, dist.get_rank(group), peers)
            self.flat = buf[:total]
        except Exception as exc:                                  # noqa: BLE001 -- NCCL is always available
            self.nvls, self.nvls_error = None, f"{type(exc).__name__}: {str(exc)[:200]}"

    def _nvls_mean(self):
        import ctypes
        from . import ops
        handle, buf, mc_ptr, padded, rank, peers = self.nvls
        handle.barrier(channel=0)                                 # every rank's gradients are written
        with torch.cuda.device(buf.device):
            if self.averaging == "nvls":
                ops._call("gnnb_allreduce_mean_multimem", ctypes.c_void_p(mc_ptr), padded, rank, self.world, 32, ops._stream())
            else:
                ops._call("gnnb_allreduce_mean_p2p", peers, padded, rank, self.world, 64, ops._stream())
        handle.barrier(channel=1)                                 # every slice has been delivered everywhere

    # ---- called by the sink layer's backward (networks/gcn_conv.py) -------------------------------------------------
    def write(self, layer, x, d_h):
        """dW = d_h^T x of the sink layer written in place (the layer's backward has already put db into ``bias.grad``);
        all-reduces started range by range."""
        from . import ops
        w = layer.lin.weight
        many = self.world > 1 and self.nvls is None               # NVLS path: one pass over the whole buffer after the backward
        if many:                 # db and the tail are final here: every other layer's backward ran before this one
            self._start(self.flat[self.w_numel:])
        h, p = w.shape
        fmt = 2 if x.dtype == torch.uint8 else 1
        ranges = self._ranges(p, fmt, x.device) if many else [(0, p)]
        if many and len(ranges) > 1:
            # a marker range is a column block of dW [h, p] -- not contiguous, and a collective wants one piece of memory:
            # the GEMM of a range writes into its own contiguous staging block, the all-reduce runs on that block while
            # the next range is computed, and finish() copies the averaged block into the gradient (one pass over dW)
            outs = [self._staging(h, r) for r in ranges]

            def after(c0, c1):
                self._start(self._stage[(h, c0, c1)], back=(w.grad, c0, c1))
            ops.linear_bwd_w_ranges(x, d_h, ranges, p=p, out=outs, after_range=after)
        else:
            ops.linear_bwd_w_ranges(x, d_h, ranges, p=p, out=w.grad)
            if many:
                self._start(self.flat[:self.w_numel])

    def _ranges(self, p, fmt, device):
        """Where the markers are cut for the overlapped all-reduce.  The weight-gradient GEMM runs in waves of clusters
        (``ops.bwd_w_wave_markers`` markers each); a cut costs a launch, a pipeline fill and a drain (~10 us), and what stays
        exposed in any case is the all-reduce of the LAST range.  ``cut``: "last_wave" (default) = all full waves but the
        last one | the rest: one cut, the first all-reduce hides behind the last wave; "every_wave" = one range per wave;
        "none" = one GEMM, one all-reduce after it."""
        from . import ops
        wave = self.range_markers or ops.bwd_w_wave_markers(fmt, device)
        if self.cut == "none" or wave <= 0 or p <= wave:
            return [(0, p)]
        per_wave = ops.marker_ranges(p, None, fmt, wave=wave)
        if self.cut == "every_wave" or len(per_wave) <= 2:
            return per_wave
        return [(0, per_wave[-1][0]), per_wave[-1]]

    def _staging(self, h, r):
        key = (h, r[0], r[1])
        buf = self._stage.get(key)
        if buf is None:
            buf = self._stage[key] = torch.empty((h, r[1] - r[0]), dtype=torch.float32, device=self.flat.device)
        return buf

    def _start(self, buf, back=None):
        avg = dist.get_backend(self.group) == "nccl"             # gloo has no AVG: sum, then divide in finish()
        work = dist.all_reduce(buf, op=dist.ReduceOp.AVG if avg else dist.ReduceOp.SUM, group=self.group, async_op=True)
        self.pending.append((work, buf, back, avg))

    def finish(self):
        """Wait for every all-reduce started by ``write`` and copy staged ranges into the gradient."""
        for work, buf, back, avg in self.pending:
            work.wait()
            if not avg:
                buf.div_(self.world)
            if back is not None:
                grad, c0, c1 = back
                grad[:, c0:c1].copy_(buf)
        self.pending = []

    # ---- per step ----------------------------------------------------------------------------------------------------
    def zero(self):
        if self.tail.numel():
            self.tail.zero_()                    # the slices of the sink / in-place layers are overwritten, not accumulated into
        for m in self.inplace_layers:
            m._grad_written = False

    def allreduce_mean(self, world, group=None):
        if world <= 1:
            return
        if self.nvls is not None:
            self._nvls_mean()
            return
        if self.pending:                         # the sink started everything during the backward pass
            self.finish()
            return
        if dist.get_backend(group) == "nccl":
            dist.all_reduce(self.flat, op=dist.ReduceOp.AVG, group=group)
        else:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)
            self.flat.div_(world)


def first_genotype_layer(model):
    """The GCNConv that reads the genotype matrix (the first module of the reference's GCN factories), or None."""
    from .networks.gcn_conv import GCNConv
    nns = getattr(model, "nns", None)
    if nns is not None and len(nns) > 0 and isinstance(nns[0], GCNConv):
        return nns[0]
    return None


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
    shard = getattr(data, "x_shard", None)               # create_graph_sharded: x holds this rank's rows only
    n = data.y.shape[0] if shard is not None else x.shape[0]
    lo, hi = row_shard(n, rank, world)
    if shard is not None and tuple(shard) != (lo, hi):
        raise ValueError(f"sharded_eval: x holds rows {tuple(shard)}, this rank evaluates rows {(lo, hi)}")
    cur = x if shard is not None else x[lo:hi]
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


# ----------------------------------------------------------------------------------------------- sharded graph build
def shard_sizes(n_local, rank, world, device):
    """(sizes, offsets) of the row shards of all ranks, as Python lists."""
    sizes = torch.zeros(world, dtype=torch.int64, device=device)
    sizes[rank] = n_local
    if world > 1:
        dist.all_reduce(sizes)
    offs = torch.cumsum(sizes, 0) - sizes
    return [int(v) for v in sizes.tolist()], [int(v) for v in offs.tolist()]


def ring_pass(local, sizes, rank, world, fold, make_buffers):
    """Pass every rank's block once round the ring (rank r sends to r + 1, receives from r - 1), double-buffered:
    ``fold(owner, tensors)`` is called for the block of every owner, starting with the rank's own, while the next block
    is already in flight (the transfers run on the backend's stream, the fold on the current one).
    ``local``: tuple of tensors of this rank's block; ``make_buffers(n_rows)``: a matching tuple of empty tensors."""
    cur, owner = tuple(local), rank
    spare = [None, None]
    for s in range(world):
        reqs, nxt, nxt_owner = [], None, None
        if s + 1 < world:
            nxt_owner = (rank - s - 1) % world
            slot = s & 1
            if spare[slot] is None or spare[slot][0].shape[0] != sizes[nxt_owner]:
                spare[slot] = tuple(make_buffers(sizes[nxt_owner]))
            nxt = spare[slot]
            ops_ = [dist.P2POp(dist.isend, t, (rank + 1) % world) for t in cur]
            ops_ += [dist.P2POp(dist.irecv, t, (rank - 1) % world) for t in nxt]
            reqs = dist.batch_isend_irecv(ops_)
        fold(owner, cur)
        for r in reqs:
            r.wait()
        if nxt is not None:
            cur, owner = nxt, nxt_owner


def warm_up_collectives(rank, world, device):
    """Create the communicator and its point-to-point / all-to-all connections before anything is timed: NCCL builds them
    lazily on first use (seconds at 8 ranks), which would otherwise be charged to the first section that communicates."""
    if world == 1:
        return
    t = torch.zeros(world * 4, dtype=torch.float32, device=device)
    dist.all_reduce(t)
    out = torch.empty(world * world * 4, dtype=torch.float32, device=device)
    dist.all_gather_into_tensor(out, t)
    dist.all_to_all_single(torch.empty_like(t), t)
    ring_pass((t,), [t.shape[0]] * world, rank, world, lambda owner, block: None,
              lambda rows: (torch.empty(rows, dtype=torch.float32, device=device),))
    torch.cuda.synchronize()
    dist.barrier()


def knn_ring(X_local, row0, n_total, k, rank, world, extra_local=None, impl=None):
    """Row-sharded kNN graph build with the key blocks passed round a ring: rank r owns the query rows
    [row0, row0 + n_local); in step s it folds the block that started s ranks behind it into its running lists
    (gnnb_knn_i8, merge = 1) while the next block arrives.  No neighbour list is ever reduced across ranks, and the lists
    are bit-identical to ``ops.knn`` on the whole matrix for any number of ranks.
    The received blocks land in this GPU's own memory before any kernel reads them.
    Returns (idx int32 [n_local, k] of GLOBAL row ids, d2 float64 [n_local, k])."""
    from . import ops
    dev = X_local.device
    n_local, p = X_local.shape
    ld = X_local.stride(0)
    sq_local = ops.row_sqnorm(X_local)
    best_d2 = torch.empty((n_local, k), dtype=torch.float64, device=dev)
    best_idx = torch.empty((n_local, k), dtype=torch.int32, device=dev)
    n_extra = 0 if extra_local is None else extra_local.shape[1]
    sizes, offs = shard_sizes(n_local, rank, world, dev)
    assert sum(sizes) == n_total and offs[rank] == row0
    qb = ops.config["knn_query_block"]
    qb = -(-n_local // max(1, -(-n_local // qb)))         # equal query blocks: none of them tiny (the FP4 kernel wants >= 129 rows)
    state = {"first": True}
    # codes 0 / 1 / 2 on every rank: the blocks travel and are searched in their e2m1-packed form (half the bytes on the
    # links, FP4 tensor tiles); otherwise as int8 codes
    X4 = None
    if impl in (None, 0, 3) and ops.config["knn_fp4"] and ops.fp4_worthwhile(n_local, p):
        X4 = ops.pack_fp4(X_local)
        if world > 1:
            ok = torch.tensor([0 if X4 is None else 1], device=dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if not int(ok):
                X4 = None
    if X4 is not None and min(sizes) < 129:
        X4 = None

    def fold(owner, block):
        ek = block[2] if n_extra else None
        for q0 in range(0, n_local, qb):
            q1 = min(n_local, q0 + qb)
            eq = None if n_extra == 0 else extra_local[q0:q1]
            if X4 is not None:
                ops.knn_block_fp4(X4[q0:q1], sq_local[q0:q1], row0 + q0, block[0], block[1], offs[owner], p, k,
                                  best_d2[q0:q1], best_idx[q0:q1], 0 if state["first"] else 1, eq, ek)
            else:
                ops.knn_block(X_local[q0:q1], sq_local[q0:q1], row0 + q0, block[0][:, :p], block[1], offs[owner], k,
                              best_d2[q0:q1], best_idx[q0:q1], 0 if state["first"] else 1, eq, ek, impl)
        state["first"] = False

    width = X4.shape[1] if X4 is not None else ld

    def make_buffers(rows):
        out = [torch.empty((rows, width), dtype=torch.uint8 if X4 is not None else torch.int8, device=dev),
               torch.empty(rows, dtype=torch.int64, device=dev)]
        if n_extra:
            out.append(torch.empty((rows, n_extra), dtype=torch.float64, device=dev))
        return out

    # the padded storage travels: receivers keep the 16-byte row stride TMA needs
    local = [X4 if X4 is not None else X_local.as_strided((n_local, ld), (ld, 1)), sq_local]
    if n_extra:
        local.append(extra_local.contiguous())
    ring_pass(local, sizes, rank, world, fold, make_buffers)
    return best_idx, best_d2


def allgather_shards(local, sizes, rank, world):
    """Concatenate per-rank row shards of unequal length (``sizes``) on every rank."""
    if world == 1:
        return local
    per = max(sizes)
    pad = torch.zeros((per,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[:local.shape[0]] = local
    out = torch.empty((world * per,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, pad)
    return torch.cat([out[r * per:r * per + sizes[r]] for r in range(world)])


def fetch_rows(store_local, sizes, offs, wanted, rank, world):
    """Rows ``wanted`` (global ids, any order, repeats allowed) of a row-sharded matrix, as one [len(wanted), ld] tensor in
    the order asked for: one all-to-all of the requested ids, one all-to-all of the rows (df.iloc[batch_nodes] of
    util.py:68 when the frame is spread over the GPUs).  ``store_local``: this rank's shard [sizes[rank], ld]."""
    dev = store_local.device
    wanted = torch.as_tensor(wanted, dtype=torch.int64, device=dev)
    if world == 1:
        return store_local[wanted]
    uniq, inverse = torch.unique(wanted, sorted=True, return_inverse=True)
    bounds = torch.tensor(offs[1:] + [offs[-1] + sizes[-1]], dtype=torch.int64, device=dev)
    owner = torch.searchsorted(bounds, uniq, right=True)
    ask = torch.bincount(owner, minlength=world)
    give = torch.empty_like(ask)
    dist.all_to_all_single(give, ask)
    ask_l, give_l = [int(v) for v in ask.tolist()], [int(v) for v in give.tolist()]
    ids = torch.empty(sum(give_l), dtype=torch.int64, device=dev)
    dist.all_to_all_single(ids, uniq, output_split_sizes=give_l, input_split_sizes=ask_l)
    rows_out = store_local[ids - offs[rank]]
    rows_in = torch.empty((int(uniq.numel()), store_local.shape[1]), dtype=store_local.dtype, device=dev)
    dist.all_to_all_single(rows_in, rows_out, output_split_sizes=ask_l, input_split_sizes=give_l)
    return rows_in[inverse]


def create_graph_sharded(X_shard, y_all, n_neighbors, rank, world, smoothing="laplacian", use_weights=False,
                         mode="connectivity", remove_mean=True):
    """``create_graph`` (load.py:21-81) for ONE graph whose node rows are sharded over the ranks by ``row_shard``: the
    O(n^2 p) neighbour search runs as a ring over the shards, the k x n neighbour lists (12 bytes per entry) are
    all-gathered, and every rank runs the O(n k) similarity / symmetrisation / Laplacian pass on the full lists -- the same
    bytes a reverse-edge exchange would move, no distributed sort, and an edge list that is bit-identical to the one-GPU
    build by construction (the global d_max and max L are then local reductions).  Returns a Data whose ``x`` holds this
    rank's rows only (``x_shard = (lo, hi)``), as ``sharded_eval`` wants it."""
    import numpy as np
    from . import ops
    from .data import Data
    n = int(len(y_all))
    lo, hi = row_shard(n, rank, world)
    assert X_shard.shape[0] == hi - lo
    idx_l, d2_l = knn_ring(X_shard, lo, n, n_neighbors, rank, world)
    sizes = [row_shard(n, r, world)[1] - row_shard(n, r, world)[0] for r in range(world)]
    idx, d2 = allgather_shards(idx_l, sizes, rank, world), allgather_shards(d2_l, sizes, rank, world)
    ei, ew = ops.build_edges(idx, d2, mode=mode, smoothing=smoothing, use_weights=use_weights)
    yv = np.asarray(y_all, dtype=np.float64)
    if remove_mean:
        yv = yv - yv.mean()
    y = torch.from_numpy(yv.astype(np.float32).reshape(-1, 1)).to(X_shard.device)
    data = Data(x=X_shard, y=y, edge_index=ei)
    if use_weights:
        data.edge_weight = ew
    data.x_shard = (lo, hi)
    return data


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
