"""Tensor-level wrappers over the C ABI (include/gnnb.h).  torch is plumbing here: device memory,
streams, autograd bookkeeping.  Every function launches hand-written sm_100a kernels from
libgnn_b200.so and raises if the library or a B200 is missing -- there is no fallback path.
"""
import ctypes
import os
import weakref
from collections import OrderedDict

import torch

from . import _lib

# knobs: 0 = CUDA-core reference tiles, 1..3 = tcgen05 with that many bf16 terms, "auto"
config = {
    "linear_precision": "auto",
    "knn_impl": 0,              # 0 = pick, 1 = dp4a tiles, 2 = tcgen05 int8 tiles, 3 = tcgen05 FP4 tiles (codes 0..2)
    "knn_fp4": os.environ.get("GNNB_KNN_FP4", "1") != "0",      # "pick" prefers the FP4 tiles where they apply
    "knn_query_block": 1 << 18,
    "fork": os.environ.get("GNNB_FORK", "1") != "0",    # independent pieces of a backward pass on side streams (ops.Fork)
    # a train-mode torch.nn.Dropout directly in front of the dense head is applied inside the head's kernels (networks.Sequential)
    "fused_dropout": os.environ.get("GNNB_FUSED_DROPOUT", "1") != "0",
}

_launches = 0                   # C-ABI calls that enqueue kernels (bench.py reports it)
_dropout_states = 0


def launches():
    return _launches


def _p(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _need_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise _lib.GnnbError("gnn_b200 operators need CUDA tensors (no CPU fallback)")


_ws_cache = {}


def _ws(nbytes, device):
    """Scratch for one C-ABI call.  One grow-only buffer per (device, stream): calls on a stream are ordered, so the
    next call may reuse it; this keeps the allocator (and its lock) off the per-step path."""
    nbytes = max(int(nbytes), 256)
    key = (device.index if device.index is not None else torch.cuda.current_device(), torch.cuda.current_stream(device).cuda_stream)
    buf = _ws_cache.get(key)
    if buf is None or buf.numel() < nbytes or torch.cuda.is_current_stream_capturing():
        buf = torch.empty(nbytes, dtype=torch.uint8, device=device)
        if not torch.cuda.is_current_stream_capturing():
            _ws_cache[key] = buf
    return buf


_wsb_cache = {}


def _ws_bytes(name, *args):
    key = (name,) + args
    v = _wsb_cache.get(key)
    if v is None:
        v = _wsb_cache[key] = _lib.workspace_bytes(name, *args)
    return v


class Profiler:
    """CUDA-event timing of every C-ABI call on the launching stream, with the algorithmic work each call
    declares (bench.py derives the roofline numbers from this; events cost ~1 us per call)."""

    def __init__(self):
        self.records = []            # (name, work_dict, start_event, end_event)

    def summary(self, by_shape=False):
        """name -> totals; with ``by_shape`` the key is ``name[shape]`` for calls that declared one (projection, aggregation,
        neighbour search), so that e.g. the 50 000-wide tensor-core projection and the 100-wide inner products are apart."""
        torch.cuda.synchronize()
        out = {}
        for name, work, e0, e1 in self.records:
            key = f"{name}[{work['shape']}]" if by_shape and "shape" in work else name
            slot = out.setdefault(key, {"calls": 0, "ms": 0.0, "flops": 0.0, "bytes": 0.0})
            slot["calls"] += 1
            slot["ms"] += e0.elapsed_time(e1)
            slot["flops"] += work.get("flops", 0.0)
            slot["bytes"] += work.get("bytes", 0.0)
        return out


profiler = None                 # set to a Profiler() to time calls
_work = {}                      # algorithmic work of the next call (set by the wrappers below)


def _call(name, *args):
    global _launches, _work
    _launches += 1
    if profiler is None:
        _work = {}
        return _lib.call(name, *args)
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    rc = _lib.call(name, *args)
    e1.record()
    profiler.records.append((name, _work, e0, e1))
    _work = {}
    return rc


def _declare(**work):
    global _work
    _work = work


# --------------------------------------------------------------------------------------- staging
def synth_genotypes(n, p, device, row0=0, p_total=None, seed=20240601, seed_freq=20240602, out=None):
    """int8 [n, p] block of the stateless synthetic genotype matrix."""
    p_total = p if p_total is None else p_total
    # rows padded to 16 bytes, like every genotype matrix of this package (TMA needs 16-byte strides)
    X = out if out is not None else torch.zeros((n, (p + 15) // 16 * 16), dtype=torch.int8, device=device)[:, :p]
    _need_cuda(X)
    with torch.cuda.device(X.device):
        _call("gnnb_synth_genotypes", _p(X), n, p, X.stride(0), row0, p_total, seed, seed_freq, _stream())
    return X


def gather_rows(X, rows, ld_align=16):
    """X[rows] for int8 genotypes (df.iloc[batch_nodes], util.py:68); row stride padded to 16 bytes."""
    _need_cuda(X, rows)
    rows = rows.to(torch.int64).contiguous()
    p = X.shape[1]
    ld = (p + ld_align - 1) // ld_align * ld_align
    buf = torch.zeros((rows.numel(), ld), dtype=torch.int8, device=X.device)
    with torch.cuda.device(X.device):
        _call("gnnb_gather_rows_i8", _p(X), X.stride(0), _p(rows), rows.numel(), p, _p(buf), ld, _stream())
    return buf[:, :p]


def packed_stride(p):
    """Row stride in bytes of 2-bit packed markers: four codes per byte, padded to 16 bytes."""
    return ((p + 3) // 4 + 15) // 16 * 16


def pack_2bit_host(X):
    """numpy int8 codes 0..3 [n, p] -> uint8 [n, packed_stride(p)] (code j in bits 2*(j%4) of byte j//4)."""
    import numpy as np
    X = np.asarray(X)
    n, p = X.shape
    if X.size and (X.min() < 0 or X.max() > 3):
        raise ValueError("2-bit packing needs genotype codes in 0..3")
    q = (p + 3) // 4
    padded = np.zeros((n, q * 4), dtype=np.uint8)
    padded[:, :p] = X
    v = padded.reshape(n, q, 4)
    out = np.zeros((n, packed_stride(p)), dtype=np.uint8)
    out[:, :q] = v[:, :, 0] | (v[:, :, 1] << 2) | (v[:, :, 2] << 4) | (v[:, :, 3] << 6)
    return out


def unpack_2bit(P, p, out=None):
    """uint8 packed markers on the GPU -> int8 codes [n, p] (row stride padded to 16 bytes)."""
    _need_cuda(P)
    if P.dtype != torch.uint8 or P.dim() != 2 or P.stride(1) != 1:
        raise _lib.GnnbError("packed markers must be a 2-D uint8 tensor with unit column stride")
    n = P.shape[0]
    if out is None:
        ld = (p + 15) // 16 * 16
        out = torch.empty((n, ld), dtype=torch.int8, device=P.device)[:, :p]
    with torch.cuda.device(P.device):
        _declare(bytes=float(n * ((p + 3) // 4) + n * p))
        _call("gnnb_unpack_2bit", _p(P), P.stride(0), n, p, _p(out), out.stride(0), _stream())
    return out


def pack_2bit(X):
    """int8 codes on the GPU -> packed uint8 [n, packed_stride(p)]; raises if a code is outside 0..3 (synchronises)."""
    _rows_ok(X)
    n, p = X.shape
    out = torch.zeros((n, packed_stride(p)), dtype=torch.uint8, device=X.device)
    bad = torch.zeros(1, dtype=torch.int32, device=X.device)
    with torch.cuda.device(X.device):
        _call("gnnb_pack_2bit", _p(X), X.stride(0), n, p, _p(out), out.stride(0), _p(bad), _stream())
    if int(bad.item()):
        raise ValueError("2-bit packing needs genotype codes in 0..3")
    return out


def base3_stride(p):
    """Row stride in bytes of base-3 packed markers: five codes per byte, padded to 16 bytes."""
    return ((p + 4) // 5 + 15) // 16 * 16


def pack_base3(X):
    """int8 codes 0..2 on the GPU -> base-3 packed uint8 [n, base3_stride(p)] (five codes per byte: the host / PCIe form).
    Set-up only (index arithmetic in torch); raises if a code is outside 0..2 (synchronises)."""
    _rows_ok(X)
    n, p = X.shape
    if n and (int(X.min()) < 0 or int(X.max()) > 2):
        raise ValueError("base-3 packing needs genotype codes in 0..2")
    q = (p + 4) // 5
    out = torch.zeros((n, base3_stride(p)), dtype=torch.uint8, device=X.device)
    weights = torch.tensor([1, 3, 9, 27, 81], dtype=torch.int32, device=X.device)
    for r0 in range(0, n, 8192):
        blk = torch.zeros((min(n, r0 + 8192) - r0, q * 5), dtype=torch.int32, device=X.device)
        blk[:, :p] = X[r0:r0 + 8192]
        out[r0:r0 + 8192, :q] = (blk.view(-1, q, 5) * weights).sum(dim=2).to(torch.uint8)
    return out


def unpack_base3(P3, p, out=None):
    """base-3 packed markers on the GPU -> 2-bit packed uint8 [n, packed_stride(p)] (what the projection kernels read)."""
    _need_cuda(P3)
    if P3.dtype != torch.uint8 or P3.dim() != 2 or P3.stride(1) != 1:
        raise _lib.GnnbError("base-3 packed markers must be a 2-D uint8 tensor with unit column stride")
    n = P3.shape[0]
    if out is None:
        out = torch.empty((n, packed_stride(p)), dtype=torch.uint8, device=P3.device)
        if packed_stride(p) > (p + 3) // 4:
            out[:, (p + 3) // 4:] = 0
    with torch.cuda.device(P3.device):
        _declare(bytes=float(n * ((p + 4) // 5) + n * ((p + 3) // 4)))
        _call("gnnb_base3_to_2bit", _p(P3), P3.stride(0), n, p, _p(out), out.stride(0), None, _stream())
    return out


def i8_to_f32(X):
    _need_cuda(X)
    out = torch.empty(X.shape, dtype=torch.float32, device=X.device)
    with torch.cuda.device(X.device):
        _call("gnnb_i8_to_f32", _p(X), X.stride(0), X.shape[0], X.shape[1], _p(out), out.stride(0), _stream())
    return out


def row_sqnorm(X):
    _need_cuda(X)
    sq = torch.empty(X.shape[0], dtype=torch.int64, device=X.device)
    with torch.cuda.device(X.device):
        _call("gnnb_row_sqnorm_i8", _p(X), X.stride(0), X.shape[0], X.shape[1], _p(sq), _stream())
    return sq


def _rows_ok(X):
    if X.dtype != torch.int8 or X.dim() != 2 or X.stride(1) != 1:
        raise _lib.GnnbError("genotypes must be a 2-D int8 tensor with unit column stride")


# --------------------------------------------------------------------------------------- K1
def knn_block(Q, sq_q, q_off, K, sq_k, k_off, k, best_d2, best_idx, merge, extra_q=None, extra_k=None, impl=None):
    """Fold the key block K (global rows k_off..) into the running lists of the query block Q."""
    _rows_ok(Q)
    _rows_ok(K)
    _need_cuda(Q, K, best_d2, best_idx)
    impl = config["knn_impl"] if impl is None else impl
    n_extra = 0 if extra_q is None else extra_q.shape[1]
    with torch.cuda.device(Q.device):
        nbytes = _ws_bytes("gnnb_knn_workspace_bytes", Q.shape[0], K.shape[0], Q.shape[1], k, impl)
        ws = _ws(nbytes, Q.device)
        _declare(flops=2.0 * Q.shape[0] * K.shape[0] * Q.shape[1], shape=f"{Q.shape[0]}x{K.shape[0]}x{Q.shape[1]}")
        _call("gnnb_knn_i8", _p(Q), Q.stride(0), _p(sq_q), Q.shape[0], q_off,
              _p(K), K.stride(0), _p(sq_k), K.shape[0], k_off, Q.shape[1],
              _p(extra_q), _p(extra_k), n_extra, k, _p(best_d2), _p(best_idx), int(merge), impl,
              _p(ws), ws.numel(), _stream())


def fp4_stride(p):
    """Row stride in bytes of e2m1-packed markers: two codes per byte, padded to 16 bytes."""
    return ((p + 1) // 2 + 15) // 16 * 16


def pack_fp4(X):
    """int8 codes 0..2 -> e2m1-packed uint8 [n, fp4_stride(p)] (the operand form of the FP4 neighbour search), or None if a
    code is outside 0..2 (one host read of a flag)."""
    _rows_ok(X)
    n, p = X.shape
    out = torch.empty((n, fp4_stride(p)), dtype=torch.uint8, device=X.device)
    bad = torch.zeros(1, dtype=torch.int32, device=X.device)
    with torch.cuda.device(X.device):
        _declare(bytes=float(n * p + n * fp4_stride(p)))
        _call("gnnb_pack_fp4", _p(X), X.stride(0), n, p, _p(out), out.stride(0), _p(bad), _stream())
    return None if int(bad.item()) else out


def knn_block_fp4(Q4, sq_q, q_off, K4, sq_k, k_off, p, k, best_d2, best_idx, merge, extra_q=None, extra_k=None):
    """``knn_block`` on e2m1-packed operands (``pack_fp4``): tcgen05 kind::mxf4 tiles, bit-identical lists."""
    _need_cuda(Q4, K4, best_d2, best_idx)
    n_extra = 0 if extra_q is None else extra_q.shape[1]
    with torch.cuda.device(Q4.device):
        nbytes = _ws_bytes("gnnb_knn_workspace_bytes", Q4.shape[0], K4.shape[0], p, k, 2)
        ws = _ws(nbytes, Q4.device)
        _declare(flops=2.0 * Q4.shape[0] * K4.shape[0] * p, shape=f"{Q4.shape[0]}x{K4.shape[0]}x{p},fp4")
        _call("gnnb_knn_fp4", _p(Q4), Q4.stride(0), _p(sq_q), Q4.shape[0], q_off,
              _p(K4), K4.stride(0), _p(sq_k), K4.shape[0], k_off, p,
              _p(extra_q), _p(extra_k), n_extra, k, _p(best_d2), _p(best_idx), int(merge), _p(ws), ws.numel(), _stream())


def fp4_worthwhile(n_q, p):
    """The FP4 search needs the CTA-pair kernel (>= 129 query rows) and pays for its packing pass on long rows."""
    return n_q >= 256 and p >= 512


def knn(X, k, extra=None, impl=None):
    """Canonical k nearest neighbours of every row of int8 X (sklearn call at load.py:52-59, util.py:52-57).

    extra: optional float64 [n, m] side columns (the phenotype column the partition graph keeps).
    Returns (idx int32 [n, k], d2 float64 [n, k]) in ascending (d2, index) order.
    """
    _rows_ok(X)
    n = X.shape[0]
    if not 0 < k < n:
        raise ValueError(f"Expected n_neighbors <= n_samples_fit, but n_neighbors = {k + 1}, n_samples_fit = {n}")
    if extra is not None:
        extra = extra.to(device=X.device, dtype=torch.float64).reshape(n, -1).contiguous()
    sq = row_sqnorm(X)
    best_d2 = torch.empty((n, k), dtype=torch.float64, device=X.device)
    best_idx = torch.empty((n, k), dtype=torch.int32, device=X.device)
    qb = config["knn_query_block"]
    impl = config["knn_impl"] if impl is None else impl
    X4 = None
    if impl == 3 or (impl == 0 and config["knn_fp4"] and fp4_worthwhile(n, X.shape[1])):
        X4 = pack_fp4(X)                                   # None when a code is outside 0..2: the int8 tiles take any code
        if X4 is None and impl == 3:
            raise _lib.GnnbError("knn impl 3 (FP4 tensor path) needs genotype codes in 0..2")
    for q0 in range(0, n, qb):
        q1 = min(n, q0 + qb)
        eq = None if extra is None else extra[q0:q1]
        if X4 is not None and q1 - q0 >= 129:
            knn_block_fp4(X4[q0:q1], sq[q0:q1], q0, X4, sq, 0, X.shape[1], k, best_d2[q0:q1], best_idx[q0:q1], 0, eq, extra)
        else:
            knn_block(X[q0:q1], sq[q0:q1], q0, X, sq, 0, k, best_d2[q0:q1], best_idx[q0:q1], 0, eq, extra, 0 if impl == 3 else impl)
    return best_idx, best_d2


def insert_self(best_idx, best_d2, q_off=0):
    """include_self=True (load.py:28,57): fold (0, i) into self-excluded lists, in place."""
    _need_cuda(best_idx, best_d2)
    n, k = best_idx.shape
    assert best_idx.is_contiguous() and best_d2.is_contiguous()
    with torch.cuda.device(best_idx.device):
        _call("gnnb_knn_insert_self", _p(best_d2), _p(best_idx), n, k, q_off, _stream())
    return best_idx, best_d2


def knn_including_self(search, n, k):
    """k smallest (d2, j) over all j, self included, from a self-excluding search ``search(kk) -> (idx, d2)``."""
    if not 0 < k <= n:
        raise ValueError(f"Expected n_neighbors <= n_samples_fit, but n_neighbors = {k}, n_samples_fit = {n}")
    kk = min(k, n - 1)
    if kk == 0:                                                  # one row: its only neighbour is itself
        dev = torch.device("cuda", torch.cuda.current_device())
        idx = torch.full((n, k), -1, dtype=torch.int32, device=dev)
        d2 = torch.full((n, k), float("inf"), dtype=torch.float64, device=dev)
        return insert_self(idx, d2)
    idx, d2 = search(kk)
    if kk < k:
        idx = torch.cat([idx, torch.full((n, k - kk), -1, dtype=torch.int32, device=idx.device)], dim=1).contiguous()
        d2 = torch.cat([d2, torch.full((n, k - kk), float("inf"), dtype=torch.float64, device=d2.device)], dim=1).contiguous()
    return insert_self(idx, d2)


def scale_columns(X, want64=True, want32=True):
    """``scaled=True`` of load.py:45-48 on int8 codes: (x - mean) / std per column with pandas' arithmetic, NaN -> 0.
    Returns (float64 [n, p] or None, float32 [n, p] or None)."""
    _rows_ok(X)
    n, p = X.shape
    o64 = torch.empty((n, p), dtype=torch.float64, device=X.device) if want64 else None
    ld32 = (p + 3) // 4 * 4
    o32 = torch.empty((n, ld32), dtype=torch.float32, device=X.device)[:, :p] if want32 else None
    with torch.cuda.device(X.device):
        _declare(bytes=3.0 * n * p + (8.0 if want64 else 0) * n * p + (4.0 if want32 else 0) * n * p)
        _call("gnnb_scale_columns_i8", _p(X), X.stride(0), n, p, _p(o64) if want64 else None, p,
              _p(o32) if want32 else None, ld32, _stream())
    return o64, o32


def knn_float(X, k):
    """Canonical kNN for float64 features [n, p] (scaled / transformed markers).  Same output contract as ``knn``."""
    _need_cuda(X)
    X = X.to(torch.float64).contiguous()
    n, p = X.shape
    if not 0 < k < n:
        raise ValueError(f"Expected n_neighbors <= n_samples_fit, but n_neighbors = {k + 1}, n_samples_fit = {n}")
    best_d2 = torch.empty((n, k), dtype=torch.float64, device=X.device)
    best_idx = torch.empty((n, k), dtype=torch.int32, device=X.device)
    with torch.cuda.device(X.device):
        ws = _ws(_ws_bytes("gnnb_knn_workspace_bytes", n, n, p, k, 1), X.device)
        _declare(flops=3.0 * n * n * p)
        _call("gnnb_knn_f64", _p(X), X.stride(0), n, 0, _p(X), X.stride(0), n, 0, p, k, _p(best_d2), _p(best_idx), 0,
              _p(ws), ws.numel(), _stream())
    return best_idx, best_d2


# --------------------------------------------------------------------------------------- K2 / K3
def build_edges(idx, d2, mode="connectivity", smoothing="laplacian", use_weights=False):
    """Edge list of load.py:61-80 from canonical neighbour lists.
    Returns (edge_index int64 [2, E], edge_weight float64 [E] or None)."""
    _need_cuda(idx, d2)
    idx = idx.contiguous()
    d2 = d2.contiguous()
    n, k = idx.shape
    m = 1 if mode == "distance" else 0
    dev = idx.device
    with torch.cuda.device(dev):
        if smoothing == "laplacian":
            ws = _ws(_ws_bytes("gnnb_graph_workspace_bytes", n, k), dev)
            n_edges = torch.zeros(1, dtype=torch.int64, device=dev)
            _call("gnnb_graph_count", _p(idx), _p(d2), n, k, m, _p(n_edges), _p(ws), ws.numel(), _stream())
            E = int(n_edges.item())                       # the one host sync of the graph build
            ei = torch.empty((2, E), dtype=torch.int64, device=dev)
            ew = torch.empty(E, dtype=torch.float64, device=dev)
            _call("gnnb_graph_fill", n, k, int(bool(use_weights)), E, _p(ei), _p(ew), _p(ws), ws.numel(), _stream())
        else:
            E = n * k
            ws = _ws(256, dev)
            ei = torch.empty((2, E), dtype=torch.int64, device=dev)
            ew = torch.empty(E, dtype=torch.float64, device=dev)
            _call("gnnb_graph_knn_edges", _p(idx), _p(d2), n, k, m, int(bool(use_weights)), _p(ei), _p(ew),
                  _p(ws), ws.numel(), _stream())
    return ei, (ew if use_weights else None)


# --------------------------------------------------------------------------------------- K4
class GcnGraph:
    """GCN-normalised adjacency in CSR, by target (forward) and by source (backward)."""
    __slots__ = ("n", "nnz", "rowptr", "col", "val", "rowptr_t", "col_t", "val_t", "hub", "hub_t", "nbytes")

    def agg_bytes(self, F, s=4):
        """Algorithmic bytes of one aggregation launch."""
        return self.nnz * (F * s + 8) + self.n * F * s + (self.n + 1) * 4 + self.n * 4 + F * 4


_graph_cache = OrderedDict()
_GRAPH_CACHE_MAX = 256                      # entries; the byte bound below is the one that matters for large graphs
_GRAPH_CACHE_MAX_BYTES = 48 << 30


def _graph_bytes(g):
    total = 0
    for name in ("rowptr", "col", "val", "rowptr_t", "col_t", "val_t"):
        t = getattr(g, name)
        total += t.numel() * t.element_size()
    for plan in (g.hub, g.hub_t):
        if plan is not None:
            total += sum(t.numel() * t.element_size() for t in plan)
    return total


def _evict_dead(key):
    _graph_cache.pop(key, None)


def gcn_graph(edge_index, edge_weight, n, add_self_loops=True):
    """gcn_norm + CSR of PyG's GCNConv, built once per (edge list, weights) and cached:
    the reference recomputes it on every forward (cached=False); the result is identical.

    The cache is keyed on the IDENTITY (and version counter) of the two tensors and holds them only weakly: an entry
    lives exactly as long as the tensors it was built from, so a temporary (``edge_weight.float()`` made per call) can
    neither produce a stale hit through a reused address nor pin its CSR in device memory.  Callers that want hits pass
    the same tensor objects again (trainers.util.get_args_tuple keeps the float32 weights on the Data object)."""
    _need_cuda(edge_index, edge_weight)
    if edge_index.dtype != torch.int64 or edge_index.dim() != 2 or edge_index.shape[0] != 2:
        raise _lib.GnnbError("edge_index must be int64 [2, E]")
    key = (id(edge_index), edge_index._version, edge_index.shape[1],
           None if edge_weight is None else (id(edge_weight), edge_weight._version),
           int(n), bool(add_self_loops))
    hit = _graph_cache.get(key)
    if hit is not None:
        g, ei_ref, ew_ref = hit
        if ei_ref() is edge_index and (edge_weight is None or (ew_ref is not None and ew_ref() is edge_weight)):
            _graph_cache.move_to_end(key)
            return g
        del _graph_cache[key]                   # an id reused by another tensor before the weakref callback ran
    ei = edge_index.contiguous()
    ew = None if edge_weight is None else edge_weight.to(torch.float32).contiguous()
    E = ei.shape[1]
    dev = ei.device
    g = GcnGraph()
    with torch.cuda.device(dev):
        ws = _ws(_ws_bytes("gnnb_csr_workspace_bytes", n, E), dev)
        nnz_t = torch.zeros(1, dtype=torch.int64, device=dev)
        _call("gnnb_csr_count", _p(ei), E, n, int(add_self_loops), _p(nnz_t), _p(ws), ws.numel(), _stream())
        nnz = int(nnz_t.item())
        g.n, g.nnz = int(n), nnz
        g.rowptr = torch.empty(n + 1, dtype=torch.int32, device=dev)
        g.rowptr_t = torch.empty(n + 1, dtype=torch.int32, device=dev)
        g.col = torch.empty(max(nnz, 1), dtype=torch.int32, device=dev)
        g.col_t = torch.empty(max(nnz, 1), dtype=torch.int32, device=dev)
        g.val = torch.empty(max(nnz, 1), dtype=torch.float32, device=dev)
        g.val_t = torch.empty(max(nnz, 1), dtype=torch.float32, device=dev)
        _call("gnnb_csr_build", _p(ei), _p(ew), E, n, int(add_self_loops), nnz,
              _p(g.rowptr), _p(g.col), _p(g.val), _p(g.rowptr_t), _p(g.col_t), _p(g.val_t),
              _p(ws), ws.numel(), _stream())
    g.hub = hub_plan(g.rowptr)              # hub rows cut into tasks (forward CSR and its transpose)
    g.hub_t = hub_plan(g.rowptr_t)
    g.nbytes = _graph_bytes(g)
    drop = lambda _ref, key=key: _evict_dead(key)       # noqa: E731 -- the entry dies with either tensor
    _graph_cache[key] = (g, weakref.ref(edge_index, drop), None if edge_weight is None else weakref.ref(edge_weight, drop))
    total = sum(v[0].nbytes for v in _graph_cache.values())
    while len(_graph_cache) > 1 and (len(_graph_cache) > _GRAPH_CACHE_MAX or total > _GRAPH_CACHE_MAX_BYTES):
        _, (old, _a, _b) = _graph_cache.popitem(last=False)
        total -= old.nbytes
    return g


def hub_plan(rowptr):
    """Task list for rows longer than gnnb_spmm_long_row() (see include/gnnb.h): (task_row, task_chunk, task_first,
    task_count) as int32 tensors, or None when the graph has no such row.  Index bookkeeping only -- done once per graph."""
    chunk = int(_lib.lib().gnnb_spmm_long_row())
    deg = (rowptr[1:] - rowptr[:-1]).to(torch.int64)
    long_rows = torch.nonzero(deg > chunk).flatten()
    if long_rows.numel() == 0:
        return None
    nch = (deg[long_rows] + chunk - 1) // chunk
    first = torch.cumsum(nch, 0) - nch
    task_row = torch.repeat_interleave(long_rows, nch)
    task_first = torch.repeat_interleave(first, nch)
    task_count = torch.repeat_interleave(nch, nch)
    task_chunk = torch.arange(task_row.numel(), device=rowptr.device) - task_first
    return tuple(t.to(torch.int32).contiguous() for t in (task_row, task_chunk, task_first, task_count))


def clear_graph_cache():
    _graph_cache.clear()


# --------------------------------------------------------------------------------------- K6 / K7
def spmm(rowptr, col, val, H, bias=None, n=None, hub=None, nnz=None):
    """out[i] = sum_e val[e] * H[col[e]] (+ bias) over CSR row i.  ``hub``: optional task plan from ``hub_plan``.
    ``rowptr`` may be a slice of a larger row pointer (its offsets stay absolute): the rows of one shard; ``nnz`` then
    states the shard's entry count for the work accounting."""
    _need_cuda(rowptr, H)
    if H.dtype != torch.float32 or H.stride(1) != 1:
        H = H.to(torch.float32).contiguous()
    n = rowptr.numel() - 1 if n is None else n
    out = torch.empty((n, H.shape[1]), dtype=torch.float32, device=H.device)
    nnz = int(col.numel()) if nnz is None else int(nnz)
    F = int(H.shape[1])
    _declare(bytes=float(nnz * (F * 4 + 8) + n * F * 4 + (n + 1) * 4 + n * 4 + F * 4),     # AGG_BYTES, SURVEY.md 8d
             shape=f"n{n},F{F}")
    with torch.cuda.device(H.device):
        if hub is not None:
            nt = hub[0].numel()
            ws = _ws(_ws_bytes("gnnb_spmm_workspace_bytes", nt, F), H.device)
            _call("gnnb_spmm", _p(rowptr), _p(col), _p(val), n, _p(H), H.stride(0), F, _p(bias), _p(out), out.stride(0),
                  _p(hub[0]), _p(hub[1]), _p(hub[2]), _p(hub[3]), nt, _p(ws), ws.numel(), _stream())
        else:
            _call("gnnb_spmm", _p(rowptr), _p(col), _p(val), n, _p(H), H.stride(0), F, _p(bias), _p(out), out.stride(0),
                  None, None, None, None, 0, None, 0, _stream())
    return out


# --------------------------------------------------------------------------------------- K5
def _precision(X, fmt, n, p, h):
    """The projection path for this call.  ``config["linear_precision"]``: an explicit 0..4 is passed through (the library
    refuses shapes that precision does not take); "auto" = int8 tensor cores with 24-bit fixed-point digits (4) for
    integer genotypes wherever that kernel takes the shape, else three bf16 terms (3, ~fp32-exact on integer codes) for
    int8 codes, exact fp32 tiles (0) for float features."""
    prec = config["linear_precision"]
    if fmt == 0:
        return 0                          # float features (inner layers, scaled=True) always run on the exact fp32 tiles
    if prec != "auto":
        return int(prec)
    if _lib.lib().gnnb_linear_supported(_p(X), fmt, X.stride(0), n, p, h, 4):
        return 4
    return 3 if fmt == 1 else 4           # packed X outside the int8 kernel's shapes: let the library refuse it loudly


def _x_operand(X, p_packed):
    """(tensor, x_format, p) of a projection input: float32 -> 0, int8 codes -> 1, uint8 = 2-bit packed codes -> 2
    (then the marker count cannot be read off the tensor and comes from the weight's shape)."""
    if X.dtype == torch.uint8:
        if p_packed is None or X.shape[1] * 4 < p_packed:
            raise _lib.GnnbError("2-bit packed genotypes: the row is shorter than the marker count")
        fmt, p = 2, int(p_packed)
    elif X.dtype == torch.int8:
        fmt, p = 1, X.shape[1]
    else:
        if X.dtype != torch.float32:
            X = X.to(torch.float32)
        fmt, p = 0, X.shape[1]
    if X.stride(1) != 1:
        X = X.contiguous()
    return X, fmt, p


def linear_fwd(X, W, precision=None):
    """H = X . W^T with X int8 genotype codes, 2-bit packed codes (uint8, see pack_2bit) or float32; W float32 [h, p]."""
    _need_cuda(X, W)
    W = W.contiguous()
    h = W.shape[0]
    X, fmt, p = _x_operand(X, W.shape[1])
    if p != W.shape[1]:
        raise _lib.GnnbError(f"linear_fwd: x has {p} columns, the weight {W.shape[1]}")
    n = X.shape[0]
    prec = _precision(X, fmt, n, p, h) if precision is None else precision
    H = torch.empty((n, h), dtype=torch.float32, device=X.device)
    with torch.cuda.device(X.device):
        ws = _ws(_ws_bytes("gnnb_linear_workspace_bytes", n, p, h, prec), X.device)
        _declare(flops=2.0 * n * p * h, bytes=float(n * X.shape[1] * X.element_size() + h * p * 4 + n * h * 4),
                 shape=f"{n}x{p}x{h},fmt{fmt},prec{prec}")
        _call("gnnb_linear_fwd", _p(X), fmt, X.stride(0), n, p, _p(W), W.stride(0), h,
              _p(H), H.stride(0), prec, _p(ws), ws.numel(), _stream())
    return H


def linear_bwd_w(X, dH, precision=None, p=None, out=None):
    """dW [h, p] = dH^T . X  (``p``: marker count, needed when X is 2-bit packed; ``out``: optional float32 [h, p] with unit
    column stride to write into, e.g. the weight's slice of a flat gradient buffer)"""
    _need_cuda(X, dH)
    dH = dH.contiguous()
    h = dH.shape[1]
    X, fmt, p = _x_operand(X, p)
    n = X.shape[0]
    if dH.shape[0] != n:
        raise _lib.GnnbError(f"linear_bwd_w: x has {n} rows, dH {dH.shape[0]}")
    prec = _precision(X, fmt, n, p, h) if precision is None else precision
    dW = _grad_out(out, h, p, X.device)
    with torch.cuda.device(X.device):
        ws = _ws(_ws_bytes("gnnb_linear_workspace_bytes", n, p, h, prec), X.device)
        _declare(flops=2.0 * n * p * h, bytes=float(n * X.shape[1] * X.element_size() + h * p * 4 + n * h * 4),
                 shape=f"{n}x{p}x{h},fmt{fmt},prec{prec}")
        _call("gnnb_linear_bwd_w", _p(X), fmt, X.stride(0), n, p, _p(dH), dH.stride(0), h,
              _p(dW), dW.stride(0), prec, _p(ws), ws.numel(), _stream())
    return dW


def _grad_out(out, h, p, device):
    if out is None:
        return torch.empty((h, p), dtype=torch.float32, device=device)
    if out.shape != (h, p) or out.dtype != torch.float32 or out.stride(1) != 1 or out.device != device:
        raise _lib.GnnbError("linear_bwd_w: `out` must be a float32 [h, p] tensor with unit column stride on x's device")
    return out


def marker_ranges(p, n_ranges=None, fmt=1, wave=0):
    """Cut the markers into ranges whose starts are 16-byte aligned in the genotype matrix (a multiple of 64 codes when
    packed, 16 when int8) and a multiple of 512 markers (the M tiles of one cluster of the weight-gradient kernel).
    With ``wave`` (markers per full wave of clusters, ``bwd_w_wave_markers``) and no ``n_ranges`` every range but the
    last is exactly one wave: cutting there costs no tensor-core time, cutting finer idles SMs at the end of each range."""
    if n_ranges is None:
        step = wave if wave > 0 else p
    else:
        step = (p + n_ranges - 1) // n_ranges
    step = max(512, (step + 511) // 512 * 512)
    return [(c0, min(p, c0 + step)) for c0 in range(0, p, step)]


def linear_bwd_w_ranges(X, dH, ranges, p=None, out=None, after_range=None):
    """dW = dH^T . X computed one marker range at a time from ONE quantisation of dH (gnnb_linear_prepare_dh +
    gnnb_linear_bwd_w_prepared); ``after_range(c0, c1)`` is called after the kernels of each range have been enqueued --
    the data-parallel step starts that range's all-reduce there, so it overlaps the next range's GEMM.
    ``out``: None, one float32 [h, p] tensor (the ranges are its column blocks), or a list with one contiguous
    [h, c1 - c0] tensor per range.  Same bits as ``linear_bwd_w``.  Needs the int8-digit path (int8 or packed X); other
    inputs fall back to one ``linear_bwd_w`` over all markers."""
    _need_cuda(X, dH)
    dH = dH.contiguous()
    h = dH.shape[1]
    X, fmt, p = _x_operand(X, p)
    n = X.shape[0]
    per_range = isinstance(out, (list, tuple))
    if (fmt == 0 or config["linear_precision"] not in ("auto", 4, "4")
            or not _lib.lib().gnnb_linear_supported(_p(X), fmt, X.stride(0), n, p, h, 4)):
        dW = linear_bwd_w(X, dH, p=p, out=None if per_range else out)
        for i, (c0, c1) in enumerate(ranges):
            if per_range:
                out[i].copy_(dW[:, c0:c1])
            if after_range is not None:
                after_range(c0, c1)
        return out if per_range else dW
    dW = None if per_range else _grad_out(out, h, p, X.device)
    per_code = 4 if fmt == 2 else 1
    with torch.cuda.device(X.device):
        ws = _ws(_ws_bytes("gnnb_linear_workspace_bytes", n, p, h, 4), X.device)
        _call("gnnb_linear_prepare_dh", fmt, n, _p(dH), dH.stride(0), h, _p(ws), ws.numel(), _stream())
        for i, (c0, c1) in enumerate(ranges):
            if c0 % (16 * per_code) != 0:
                raise _lib.GnnbError("linear_bwd_w_ranges: range starts must be 16-byte aligned in the genotype matrix")
            xs = ctypes.c_void_p(X.data_ptr() + c0 // per_code)
            if per_range:
                dst = _grad_out(out[i], h, c1 - c0, X.device)
                ds, ldd = _p(dst), dst.stride(0)
            else:
                ds, ldd = ctypes.c_void_p(dW.data_ptr() + 4 * c0), dW.stride(0)
            _declare(flops=2.0 * n * (c1 - c0) * h, bytes=float(n * (c1 - c0) // per_code + h * (c1 - c0) * 4),
                     shape=f"{n}x{c1 - c0}x{h},fmt{fmt},prec4")
            _call("gnnb_linear_bwd_w_prepared", xs, fmt, X.stride(0), n, c1 - c0, h, ds, ldd, _p(ws), ws.numel(), _stream())
            if after_range is not None:
                after_range(c0, c1)
    return out if per_range else dW


_wave_cache = {}


def bwd_w_wave_markers(fmt, device):
    """Markers one full wave of the weight-gradient kernel's clusters covers on ``device`` (0 if unknown)."""
    key = (fmt, device.index)
    if key not in _wave_cache:
        with torch.cuda.device(device):
            _wave_cache[key] = int(_lib.lib().gnnb_linear_bwd_w_wave_markers(fmt))
    return _wave_cache[key]


def linear_bwd_x(dH, W):
    """dX [n, p] = dH . W"""
    _need_cuda(dH, W)
    dH = dH.contiguous()
    W = W.contiguous()
    n, h = dH.shape
    p = W.shape[1]
    dX = torch.empty((n, p), dtype=torch.float32, device=dH.device)
    _declare(flops=2.0 * n * p * h, shape=f"{n}x{p}x{h}")
    with torch.cuda.device(dH.device):
        _call("gnnb_linear_bwd_x", _p(dH), dH.stride(0), n, h, _p(W), W.stride(0), p, _p(dX), dX.stride(0), _stream())
    return dX


def colsum(A, out=None):
    _need_cuda(A)
    A = A.contiguous()
    n, F = A.shape
    if out is None:
        out = torch.empty(F, dtype=torch.float32, device=A.device)
    elif out.shape != (F,) or out.dtype != torch.float32 or not out.is_contiguous():
        raise _lib.GnnbError("colsum: `out` must be a contiguous float32 [F] tensor")
    with torch.cuda.device(A.device):
        ws = _ws(_ws_bytes("gnnb_colsum_workspace_bytes", F), A.device)
        _call("gnnb_colsum", _p(A), A.stride(0), n, F, _p(out), _p(ws), ws.numel(), _stream())
    return out


# --------------------------------------------------------------------------------------- K8: dense head
def head_supported(h, out):
    return bool(_lib.lib().gnnb_head_supported(int(h), int(out)))


def dropout_state(device, seed=None):
    """Generator state of the fused dropout (``head_fwd(..., drop=(p, state))``): int64 [4] = {seed, call counter, 0, 0} on
    ``device``.  The seed defaults to a function of ``torch.initial_seed()`` and of how many states were made before, so
    ``torch.manual_seed`` makes runs repeatable."""
    global _dropout_states
    if seed is None:                # torch's seed (not a draw: the caller's random stream stays where it is) + a per-state stride
        _dropout_states += 1
        seed = (int(torch.initial_seed()) + 0x9E3779B97F4A7C15 * _dropout_states) & (2 ** 62 - 1)
    return torch.tensor([seed, 0, 0, 0], dtype=torch.int64, device=device)


def head_fwd(A, W, bias=None, drop=None):
    """out = A . W^T + bias for a Linear(h -> out) with a handful of outputs (see include/gnnb.h).
    ``drop = (p, state)``: train-mode dropout with probability p applied to A on the fly (``state`` from ``dropout_state``,
    advanced on the device); returns ``(out, used)`` then, ``used`` being what ``head_bwd(drop=(p, used))`` needs."""
    _need_cuda(A, W)
    A = A if (A.dtype == torch.float32 and A.stride(1) == 1) else A.float().contiguous()
    W = W.contiguous()
    n, h = A.shape
    out = torch.empty((n, W.shape[0]), dtype=torch.float32, device=A.device)
    with torch.cuda.device(A.device):
        if drop is None:
            _call("gnnb_head_fwd", _p(A), A.stride(0), n, h, _p(W), W.stride(0), W.shape[0], _p(bias), _p(out), out.stride(0), _stream())
            return out
        p, state = drop
        if state.dtype != torch.int64 or state.numel() != 4 or state.device != A.device or not state.is_contiguous():
            raise _lib.GnnbError("head_fwd: the dropout state is an int64 [4] tensor on the inputs' device (ops.dropout_state)")
        used = torch.empty(2, dtype=torch.int64, device=A.device)
        _call("gnnb_head_dropout_fwd", _p(A), A.stride(0), n, h, _p(W), W.stride(0), W.shape[0], _p(bias), _p(out), out.stride(0),
              float(p), _p(state), _p(used), _stream())
    return out, used


def head_bwd(d_out, A, W, need_dA=True, need_db=True, out_w=None, out_b=None, drop=None):
    """(dA or None, dW, db or None) of ``head_fwd``; ``out_w`` [o, h] / ``out_b`` [o]: contiguous float32 tensors to write
    dW / db into (e.g. the layer's slices of a flat gradient buffer); ``drop = (p, used)`` from the dropout forward."""
    _need_cuda(d_out, A, W)
    A = A if (A.dtype == torch.float32 and A.stride(1) == 1) else A.float().contiguous()
    d_out = d_out.contiguous()
    W = W.contiguous()
    n, h = A.shape
    o = W.shape[0]
    dA = torch.empty((n, h), dtype=torch.float32, device=A.device) if need_dA else None
    for t, shape in ((out_w, (o, h)), (out_b, (o,))):
        if t is not None and (tuple(t.shape) != shape or t.dtype != torch.float32 or not t.is_contiguous() or t.device != A.device):
            raise _lib.GnnbError(f"head_bwd: an output tensor must be contiguous float32 {shape} on the inputs' device")
    dW = out_w if out_w is not None else torch.empty((o, h), dtype=torch.float32, device=A.device)
    db = (out_b if out_b is not None else torch.empty(o, dtype=torch.float32, device=A.device)) if need_db else None
    with torch.cuda.device(A.device):
        ws = _ws(_ws_bytes("gnnb_head_workspace_bytes", n, h, o), A.device)
        if drop is None:
            _call("gnnb_head_bwd", _p(d_out), d_out.stride(0), _p(A), A.stride(0), n, h, _p(W), W.stride(0), o,
                  _p(dA), h, _p(dW), h, _p(db), _p(ws), ws.numel(), _stream())
        else:
            _call("gnnb_head_dropout_bwd", _p(d_out), d_out.stride(0), _p(A), A.stride(0), n, h, _p(W), W.stride(0), o,
                  _p(dA), h, _p(dW), h, _p(db), float(drop[0]), _p(drop[1]), _p(ws), ws.numel(), _stream())
    return dA, dW, db


# --------------------------------------------------------------------------------------- subsampling
def first_occurrence(idx):
    """first_pos[v] = first position of node v in the flattened neighbour lists."""
    _need_cuda(idx)
    S = idx.contiguous().view(-1)
    n = idx.shape[0]
    fp = torch.empty(n, dtype=torch.int64, device=idx.device)
    with torch.cuda.device(idx.device):
        _call("gnnb_first_occurrence", _p(S), S.numel(), n, _p(fp), _stream())
    return fp


def partition_walk(order, n_nodes, sizes, taken=None):
    """Host cursor walk of naive_partition over all batches (util.py:15-45, 63-68).
    ``taken``: optional uint8 [n_nodes] numpy array, the `traversed` set, updated in place."""
    import numpy as np
    order = np.ascontiguousarray(order, dtype=np.int64)
    sizes = np.ascontiguousarray(sizes, dtype=np.int64)
    out = np.empty(n_nodes + 1, dtype=np.int64)
    lens = np.zeros(len(sizes), dtype=np.int64)
    if taken is not None:
        assert taken.dtype == np.uint8 and taken.size == n_nodes and taken.flags["C_CONTIGUOUS"]
    rc = _lib.lib().gnnb_partition_walk_host(
        order.ctypes.data_as(ctypes.c_void_p), order.size, n_nodes,
        sizes.ctypes.data_as(ctypes.c_void_p), sizes.size,
        None if taken is None else taken.ctypes.data_as(ctypes.c_void_p),
        out.ctypes.data_as(ctypes.c_void_p), lens.ctypes.data_as(ctypes.c_void_p))
    if rc < 0:
        _lib.check(int(rc), "gnnb_partition_walk_host")
    res, pos = [], 0
    for ln in lens:
        res.append(out[pos:pos + int(ln)].copy())
        pos += int(ln)
    return res


# --------------------------------------------------------------------------------------- independent work side by side
_side_streams = {}


class Fork:
    """Run an independent piece of a backward pass on a side stream while the main stream carries on (e.g. the bias
    gradient's column sums next to the aggregation of the same dOut: each fills a fraction of the GPU).  Usage:

        job = Fork(lambda: colsum(d_out), reads=(d_out,))     # enqueued on the side stream, after everything so far
        d_h = spmm(...)                                        # main stream, concurrently
        d_bias = job.join()                                    # main stream waits; the result is safe to use on it

    Works eagerly and under CUDA-graph capture (the fork and join become graph edges).  ``slot`` picks one of a few side
    streams per device so that two forks can be in flight at once."""

    def __init__(self, fn, reads=(), slot=0):
        if not config["fork"]:                                   # A/B switch: run in line on the main stream
            self.side, self.out = None, fn()
            return
        self.main = torch.cuda.current_stream()
        dev = self.main.device_index
        key = (dev, slot)
        side = _side_streams.get(key)
        if side is None:
            side = _side_streams[key] = torch.cuda.Stream(device=dev)
        self.side = side
        side.wait_stream(self.main)
        with torch.cuda.stream(side):
            self.out = fn()
        for t in reads:
            if t is not None:
                t.record_stream(side)

    def join(self):
        if self.side is None:
            return self.out
        self.main.wait_stream(self.side)
        outs = self.out if isinstance(self.out, (tuple, list)) else (self.out,)
        for t in outs:
            if torch.is_tensor(t):
                t.record_stream(self.main)
        return self.out
