"""Graph containers with the surface of torch_geometric's ``Data`` / ``InMemoryDataset`` that the
reference relies on (gnn/loaders/load.py:6, 72, 84-126; gnn/main.py:193-197): attribute access,
``x / y / edge_index / edge_weight`` defaulting to None, nested ``test`` / ``valid`` graphs, ``.to()``,
``len`` / indexing / slicing / iteration and ``num_features`` on the dataset.
"""
import copy

import torch


class Data:
    _OPTIONAL = ("x", "y", "edge_index", "edge_weight", "edge_attr", "pos")

    def __init__(self, x=None, edge_index=None, edge_attr=None, y=None, pos=None, **kwargs):
        for key, val in (("x", x), ("edge_index", edge_index), ("edge_attr", edge_attr), ("y", y), ("pos", pos)):
            if val is not None:
                setattr(self, key, val)
        for key, val in kwargs.items():
            setattr(self, key, val)

    def __getattr__(self, key):          # only called for attributes that are not set
        if key in Data._OPTIONAL:
            return None
        raise AttributeError(f"'Data' object has no attribute '{key}'")

    @property
    def keys(self):
        return [k for k, v in vars(self).items() if v is not None]

    @property
    def num_nodes(self):
        return None if self.x is None else int(self.x.shape[0])

    @property
    def num_features(self):
        if self.x is None:
            return 0
        return 1 if self.x.dim() == 1 else int(self.x.shape[1])

    num_node_features = num_features

    def to(self, device, *_, **__):
        """Moves every tensor (and nested graph) in place and returns self, like PyG."""
        done = set()

        def move(obj):
            if isinstance(obj, Data):
                if id(obj) not in done:
                    done.add(id(obj))
                    for k, v in list(vars(obj).items()):
                        object.__setattr__(obj, k, move(v))
                return obj
            return obj.to(device) if torch.is_tensor(obj) else obj

        return move(self)

    def cuda(self, device=None):
        return self.to("cuda" if device is None else device)

    def cpu(self):
        return self.to("cpu")

    def __repr__(self):
        parts = []
        for k, v in vars(self).items():
            if torch.is_tensor(v):
                parts.append(f"{k}={list(v.shape)}")
            elif isinstance(v, Data):
                parts.append(f"{k}=Data(...)")
            else:
                parts.append(f"{k}={v!r}")
        return "Data(" + ", ".join(parts) + ")"


class GraphDataset:
    """A list of graphs behaving like the reference's dataset object: ``dataset.data`` is the first
    graph (so ``hasattr(dataset.data, 'edge_index')`` works, main.py:190), ``dataset.valid`` mirrors
    ``data.valid`` when there is a single graph (load.py:125-126)."""

    def __init__(self, graphs):
        self._graphs = list(graphs)
        if len(self._graphs) == 1 and hasattr(self._graphs[0], "valid"):
            self.valid = self._graphs[0].valid

    @property
    def data(self):
        return self._graphs[0]

    def __len__(self):
        return len(self._graphs)

    def __getitem__(self, i):
        if isinstance(i, slice):
            return GraphDataset(self._graphs[i])
        return copy.copy(self._graphs[i])

    def __iter__(self):
        for i in range(len(self)):
            yield self[i]

    @property
    def num_features(self):
        return self._graphs[0].num_features

    num_node_features = num_features

    def to(self, device):
        for g in self._graphs:
            g.to(device)
        return self


class GenotypeTable:
    """The reference's DataFrame (column 'value' + marker columns, load.py:182-187) kept where the
    kernels need it: markers as int8 codes on the GPU (row stride padded to 16 bytes), phenotypes as
    float64 on the host.  ``.iloc[rows]`` gathers rows on the device (util.py:68, :82).

    Markers must be integer codes in [0, 127] (0/1/2 genotypes); anything else is refused -- the
    float feature transforms of load.py:38-48 are not part of this path yet.
    """

    def __init__(self, markers, value, index=None, int_labels=False):
        import numpy as np
        assert markers.dtype == torch.int8 and markers.dim() == 2 and markers.is_cuda
        self.markers = markers
        self.int_labels = int_labels        # marker columns labelled 1..P as integers (load.py:189-191), else '1'..'p'
        self.value = np.asarray(value, dtype=np.float64).reshape(-1)
        self.index = np.arange(len(self.value)) if index is None else np.asarray(index)
        assert self.markers.shape[0] == self.value.shape[0]

    # -- construction
    @staticmethod
    def from_numpy(X, y, device="cuda"):
        import numpy as np
        X = np.asarray(X)
        if X.dtype != np.int8:
            Xi = np.rint(X).astype(np.int8)
            if not np.array_equal(Xi, X) or (X.size and (X.min() < 0 or X.max() > 127)):
                raise NotImplementedError("a GenotypeTable holds integer genotype codes in [0, 127]; standardised "
                                          "features are produced on the device by scaled=True (load.py:45-48)")
            X = Xi
        n, p = X.shape
        if not torch.cuda.is_available():
            raise RuntimeError("gnn_b200 needs a CUDA device (no CPU fallback)")
        from . import ops
        if X.size and X.min() >= 0 and X.max() <= 3:
            # 0/1/2 genotypes cross PCIe as 2-bit codes (a quarter of the bytes) and are expanded on the device
            host = torch.from_numpy(ops.pack_2bit_host(X)).pin_memory()
            return GenotypeTable(ops.unpack_2bit(host.to(device, non_blocking=True), p), y)
        ld = (p + 15) // 16 * 16
        host = torch.zeros((n, ld), dtype=torch.int8).pin_memory()
        host[:, :p] = torch.from_numpy(X)
        dev = host.to(device, non_blocking=True)
        return GenotypeTable(dev[:, :p], y)

    @staticmethod
    def from_frame(df, device="cuda"):
        import numpy as np
        vals = df.values
        t = GenotypeTable.from_numpy(vals[:, 1:], vals[:, 0], device=device)._with_index(df.index.values)
        t.int_labels = list(df.columns[1:]) == list(range(1, df.shape[1])) and all(isinstance(c, (int, np.integer)) for c in df.columns[1:])
        return t

    def _with_index(self, index):
        import numpy as np
        self.index = np.asarray(index)
        return self

    # -- frame-like surface
    def __len__(self):
        return self.markers.shape[0]

    @property
    def shape(self):
        return (self.markers.shape[0], self.markers.shape[1] + 1)

    @property
    def iloc(self):
        return _ILoc(self)

    def take(self, rows):
        import numpy as np
        from . import ops
        rows = np.asarray(rows, dtype=np.int64)
        if rows.size and (rows.min() < 0 or rows.max() >= len(self)):
            raise IndexError("positional indexers are out-of-bounds")
        r = torch.from_numpy(rows).to(self.markers.device)
        return GenotypeTable(ops.gather_rows(self.markers, r), self.value[rows], self.index[rows], self.int_labels)


class _ILoc:
    def __init__(self, table):
        self.t = table

    def __getitem__(self, key):
        import numpy as np
        if isinstance(key, tuple):          # df.iloc[start:stop, :]
            key = key[0]
        n = len(self.t)
        if isinstance(key, slice) and key.step in (None, 1):
            a, b, _ = key.indices(n)                       # contiguous rows: a view, no copy (util.py:82)
            b = max(a, b)
            return GenotypeTable(self.t.markers[a:b], self.t.value[a:b], self.t.index[a:b], self.t.int_labels)
        if isinstance(key, slice):
            rows = np.arange(n)[key]
        else:
            rows = np.asarray(key, dtype=np.int64)
        return self.t.take(rows)
