"""Graph-aware subsampling -- the surface of the reference's gnn/loaders/util.py.

``split_dataset_graph`` builds the global kNN graph on the device (all columns, phenotype included,
util.py:52-57), finds the first occurrence of every node in the neighbour stream with a scatter-min
kernel and walks the batches on the host in O(n + n k).  The output -- the row order of
every batch -- is identical to the reference's ``naive_partition`` under the canonical neighbour order.
"""
import numpy as np
import torch

from .. import ops
from ..data import GenotypeTable


def _first_occurrence_order(stream):
    """Nodes that occur in ``stream`` by ascending first occurrence (host, numpy)."""
    stream = np.asarray(stream, dtype=np.int64)
    if stream.size == 0:
        return stream
    _, first = np.unique(stream, return_index=True)
    return stream[np.sort(first)]


def naive_partition(edges, size, bidirectional=True, traversed=None):
    """Same contract as the reference's ``naive_partition`` (util.py:15-45), including the effect of the
    shared adjacency list of util.py:7: the scan order is the edge stream itself.

    edges: sequence of (source, target) pairs over nodes 0..n-1; returns the picked nodes in pick order.
    """
    traversed = set() if traversed is None else set(int(v) for v in traversed)
    e = np.asarray(edges, dtype=np.int64).reshape(-1, 2)
    if e.size == 0:
        return []
    n_nodes = len(set(e[:, 0].tolist()) | set(e[:, 1].tolist()))
    # util.py:8-11: edges[x].append(y) (and edges[y].append(x)) all land in ONE list
    stream = np.stack([e[:, 1], e[:, 0]], axis=1).ravel() if bidirectional else e[:, 1]
    if int(e.max()) >= n_nodes:
        raise IndexError("list index out of range")     # what the reference raises for non-contiguous ids
    taken = np.zeros(n_nodes, dtype=np.uint8)
    for v in traversed:
        if 0 <= v < n_nodes:
            taken[v] = 1
    out = ops.partition_walk(_first_occurrence_order(stream), n_nodes, [int(size)], taken)
    return [int(v) for v in out[0]]


def _sizes(n_rows, batches):
    if isinstance(batches, (int, np.integer)):
        return [n_rows // int(batches)] * int(batches)
    return [int(b) for b in batches]


def partition_order(idx):
    """Nodes in the order ``naive_partition`` meets them: by first occurrence in the neighbour stream (the flattened
    canonical lists).  Device scatter-min (gnnb_first_occurrence), one D2H copy of n int64, host argsort."""
    first_pos = ops.first_occurrence(idx).cpu().numpy()
    seen = np.nonzero(first_pos != np.iinfo(np.int64).max)[0]
    return seen[np.argsort(first_pos[seen], kind="stable")]


def split_indices_graph(table, batches, neighbours=5):
    """Row indices of every batch of ``split_dataset_graph`` for a device-resident table."""
    n = len(table)
    cache = table.__dict__.setdefault("_partition_order", {})
    order = cache.get(neighbours)
    if order is None:       # the reference rebuilds this graph for every split of the same frame (load.py:203, :238)
        extra = torch.from_numpy(np.array(table.value, dtype=np.float64).reshape(-1, 1))
        idx, _ = ops.knn(table.markers, neighbours, extra=extra)
        order = cache[neighbours] = partition_order(idx)
    return ops.partition_walk(order, n, _sizes(n, batches))


def split_dataset_graph(df, batches, neighbours=5, metric='euclidian', partition=naive_partition):
    """util.py:48-73.  ``df`` is a pandas DataFrame (column 0 = phenotype) or a GenotypeTable; the
    batches come back as the same kind of object (``df.iloc[batch_nodes]``)."""
    if metric not in ("euclidian", "euclidean", "minkowski", "l2"):
        raise NotImplementedError(f"metric {metric!r}: only the Euclidean distance is implemented")
    table = df if isinstance(df, GenotypeTable) else GenotypeTable.from_frame(df)
    if partition is naive_partition:
        parts = split_indices_graph(table, batches, neighbours)
    else:
        # any other partitioner gets what the reference gives it: the (row, col) edge list (util.py:59-68)
        extra = torch.from_numpy(np.array(table.value, dtype=np.float64).reshape(-1, 1))
        idx, _ = ops.knn(table.markers, neighbours, extra=extra)
        nbr = idx.cpu().numpy()
        edges = [(i, int(c)) for i in range(nbr.shape[0]) for c in nbr[i]]
        parts, traversed = [], set()
        for b in _sizes(len(table), batches):
            nodes = partition(edges, b, bidirectional=False, traversed=set(traversed))
            traversed |= set(nodes)
            parts.append(np.asarray(nodes, dtype=np.int64))
    return tuple(df.iloc[np.asarray(p, dtype=np.int64)] for p in parts)


def split_dataset(df, batches, **_):
    """util.py:76-84: contiguous row slices."""
    start, out = 0, []
    for b in _sizes(len(df), batches):
        out.append(df.iloc[start:start + b, :])
        start += b
    return tuple(out)
