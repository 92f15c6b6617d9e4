"""Stand-in for the torch_geometric 2.1.0 symbols the reference imports.

TEST INFRASTRUCTURE (see oracle/__init__.py).  torch_geometric is pinned by the
reference (requirements.txt:2, torch-geometric==2.1.0) but is not vendored under
/root/reference and not installed in this image, so its published behaviour is
restated here.  Call sites in the reference that define what is needed:

* ``Data`` / ``InMemoryDataset``      gnn/loaders/load.py:6, 72, 84-126, 283-284
* ``from_scipy_sparse_matrix``        gnn/loaders/load.py:70,77; gnn/loaders/util.py:59
* ``GCNConv``                         gnn/networks/networks.py:5, 142-866
* ``DenseGCNConv``                    imported at networks.py:6, not on the hot path

Semantics restated (PyG 2.1.0):
``GCNConv(in, out, improved=False, cached=False, add_self_loops=True,
normalize=True, bias=True)`` holds ``lin = Linear(in, out, bias=False)`` with
Glorot-uniform weights and a zero ``bias``; ``forward`` runs ``gcn_norm``
(add the remaining self loops with weight 1 while keeping the weight of self
loops that already exist, in-degree by target, symmetric D^-1/2 A D^-1/2
scaling with inf -> 0), then ``x @ W^T``, then a sum over incoming edges in
edge order, then ``+ bias``.
"""
import copy
import math
import os

import torch


# --------------------------------------------------------------------------- data
class Data:
    """Attribute bag with the few conveniences the reference relies on."""

    _DEFAULT_NONE = ("x", "y", "edge_index", "edge_weight", "edge_attr", "pos")

    def __init__(self, x=None, edge_index=None, edge_attr=None, y=None, pos=None, **kwargs):
        for key, val in (("x", x), ("edge_index", edge_index), ("edge_attr", edge_attr),
                         ("y", y), ("pos", pos)):
            if val is not None:
                self.__dict__[key] = val
        for key, val in kwargs.items():
            self.__dict__[key] = val

    def __getattr__(self, key):
        # only reached when the attribute is absent
        if key in Data._DEFAULT_NONE:
            return None
        raise AttributeError(f"'Data' object has no attribute '{key}'")

    @property
    def keys(self):
        return [k for k, v in self.__dict__.items() if v is not None]

    @property
    def num_nodes(self):
        return None if self.x is None else self.x.shape[0]

    @property
    def num_features(self):
        return 0 if self.x is None else (1 if self.x.dim() == 1 else self.x.shape[1])

    num_node_features = num_features

    def to(self, device, *_, **__):
        seen = {}

        def move(obj):
            if id(obj) in seen:
                return seen[id(obj)]
            if isinstance(obj, Data):
                seen[id(obj)] = obj
                for k, v in list(obj.__dict__.items()):
                    obj.__dict__[k] = move(v)
                return obj
            if torch.is_tensor(obj):
                return obj.to(device)
            return obj

        return move(self)

    def __repr__(self):
        parts = []
        for k, v in self.__dict__.items():
            if torch.is_tensor(v):
                parts.append(f"{k}={list(v.shape)}")
            elif isinstance(v, Data):
                parts.append(f"{k}=Data(...)")
            else:
                parts.append(f"{k}={v!r}")
        return "Data(" + ", ".join(parts) + ")"


class _Collated:
    """What ``collate`` hands back for more than one graph: an opaque holder."""

    def __init__(self, items):
        self.items = list(items)


class InMemoryDataset:
    """Runs ``process()`` once per processed file name, then serves the list.

    Mirrors the life cycle the reference depends on (load.py:123-126):
    ``__init__(root)`` -> ``process()`` iff the processed file is missing ->
    subclass loads ``self.data, self.slices`` from ``processed_paths[0]``.
    """

    def __init__(self, root=None, transform=None, pre_transform=None, pre_filter=None):
        self.root = root or "."
        self._indices = None
        os.makedirs(self.processed_dir, exist_ok=True)
        if not all(os.path.exists(p) for p in self.processed_paths):
            self.process()

    # -- paths
    @property
    def raw_dir(self):
        return os.path.join(self.root, "raw")

    @property
    def processed_dir(self):
        return os.path.join(self.root, "processed")

    @property
    def raw_file_names(self):
        return []

    @property
    def processed_file_names(self):
        return []

    @property
    def processed_paths(self):
        names = self.processed_file_names
        if isinstance(names, str):
            names = [names]
        return [os.path.join(self.processed_dir, n) for n in names]

    def download(self):
        pass

    def process(self):
        raise NotImplementedError

    # -- collate / access
    @staticmethod
    def collate(data_list):
        if len(data_list) == 1:
            return data_list[0], None
        return _Collated(data_list), {"n": len(data_list)}

    def _items(self):
        data = self.data
        items = data.items if isinstance(data, _Collated) else [data]
        if self._indices is not None:
            items = [items[i] for i in self._indices]
        return items

    def __len__(self):
        return len(self._items())

    def __getitem__(self, idx):
        items = self._items()
        if isinstance(idx, slice):
            sub = copy.copy(self)
            base = list(range(len(items))) if self._indices is None else list(self._indices)
            sub._indices = base[idx]
            return sub
        return copy.copy(items[idx])

    def __iter__(self):
        for i in range(len(self)):
            yield self[i]

    @property
    def num_features(self):
        return self._items()[0].num_features

    num_node_features = num_features


# --------------------------------------------------------------------------- utils
def from_scipy_sparse_matrix(A):
    """COO view of a scipy sparse matrix: (int64 [2, E], data [E]) in COO order."""
    A = A.tocoo()
    row = torch.from_numpy(A.row).to(torch.long)
    col = torch.from_numpy(A.col).to(torch.long)
    edge_index = torch.stack([row, col], dim=0)
    edge_weight = torch.from_numpy(A.data)
    return edge_index, edge_weight


def add_remaining_self_loops(edge_index, edge_weight, fill_value, num_nodes):
    """Drop existing self loops from the edge list, append one loop per node;
    a node that already had a loop keeps that loop's weight, others get fill."""
    row, col = edge_index[0], edge_index[1]
    mask = row != col
    loop_index = torch.arange(num_nodes, dtype=row.dtype, device=row.device)
    loop_index = loop_index.unsqueeze(0).repeat(2, 1)
    if edge_weight is not None:
        loop_weight = edge_weight.new_full((num_nodes,), fill_value)
        inv = ~mask
        loop_weight[row[inv]] = edge_weight[inv]
        edge_weight = torch.cat([edge_weight[mask], loop_weight], dim=0)
    edge_index = torch.cat([edge_index[:, mask], loop_index], dim=1)
    return edge_index, edge_weight


def gcn_norm(edge_index, edge_weight, num_nodes, add_self_loops=True, dtype=torch.float32):
    if edge_weight is None:
        edge_weight = torch.ones((edge_index.size(1),), dtype=dtype, device=edge_index.device)
    if add_self_loops:
        edge_index, edge_weight = add_remaining_self_loops(edge_index, edge_weight, 1.0, num_nodes)
    row, col = edge_index[0], edge_index[1]
    deg = torch.zeros(num_nodes, dtype=edge_weight.dtype, device=edge_weight.device)
    deg.index_add_(0, col, edge_weight)
    dis = deg.pow(-0.5)
    dis.masked_fill_(dis == float("inf"), 0)
    return edge_index, dis[row] * edge_weight * dis[col]


class _Lin(torch.nn.Module):
    """PyG ``Linear(in, out, bias=False, weight_initializer='glorot')``."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.weight = torch.nn.Parameter(torch.empty(out_channels, in_channels))
        self.reset_parameters()

    def reset_parameters(self):
        a = math.sqrt(6.0 / (self.in_channels + self.out_channels))
        with torch.no_grad():
            self.weight.uniform_(-a, a)

    def forward(self, x):
        return torch.nn.functional.linear(x, self.weight)


class GCNConv(torch.nn.Module):
    def __init__(self, in_channels, out_channels, improved=False, cached=False,
                 add_self_loops=True, normalize=True, bias=True, **_):
        super().__init__()
        assert not improved and not cached and normalize
        self.in_channels, self.out_channels = in_channels, out_channels
        self.add_self_loops = add_self_loops
        self.lin = _Lin(in_channels, out_channels)
        if bias:
            self.bias = torch.nn.Parameter(torch.zeros(out_channels))
        else:
            self.register_parameter("bias", None)

    def reset_parameters(self):
        self.lin.reset_parameters()
        if self.bias is not None:
            with torch.no_grad():
                self.bias.zero_()

    def forward(self, x, edge_index, edge_weight=None):
        n = x.size(0)
        edge_index, w = gcn_norm(edge_index, edge_weight, n, self.add_self_loops, x.dtype)
        h = self.lin(x)
        src, dst = edge_index[0], edge_index[1]
        out = torch.zeros(n, h.size(1), dtype=h.dtype, device=h.device)
        out.index_add_(0, dst, h[src] * w.unsqueeze(1))
        if self.bias is not None:
            out = out + self.bias
        return out


class DenseGCNConv(torch.nn.Module):  # imported by the reference, never on the hot path
    def __init__(self, *_, **__):
        super().__init__()
        raise NotImplementedError("DenseGCNConv is outside the hot path (SURVEY.md 2.1 #15)")
