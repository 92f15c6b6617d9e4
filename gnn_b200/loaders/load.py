"""Graph construction -- the surface of the reference's gnn/loaders/load.py.

``create_graph`` keeps the reference's signature and returns the same ``Data(x, y, edge_index
[, edge_weight])``, but the O(n^2 p) neighbour search, the similarity / symmetrisation / Laplacian edge
list and the weight map run as CUDA kernels (gnn_b200/csrc/knn*.cu, graph.cu) and the tensors stay on
the GPU.  Differences a caller can see, all deliberate:
  * ``x`` is the int8 genotype matrix (exact for 0/1/2 codes), not a float32 copy -- ``x.float()`` gives
    the reference's tensor; GCNConv consumes int8 directly (``scaled=True`` gives float32 as in the reference);
  * neighbour ties are ordered by (distance, index);
  * metrics other than the Euclidean one are refused instead of silently ignored.
"""
import os

import numpy as np
import torch

from .. import ops
from ..data import Data, GenotypeTable, GraphDataset
from .util import split_dataset


def hash_dict(d):
    return "".join(f"{k}:{getattr(v, '__name__', v)};" for k, v in d.items())


_EUCLIDEAN = ("minkowski", "euclidean", "euclidian", "l2")


def hot_markers(table):
    """``hot=True`` exactly as load.py:38-43 executes it.  The loops address columns by the INTEGER labels 1..P-1, so
    they only run on a frame whose marker columns were renamed by ``bits`` (load.py:189-191); the first loop overwrites
    columns 1..P-1 with ``== 1`` indicators, and the second loop then compares those 0/1 indicators with 2: column P and
    the P-2 appended columns are all zero.  Result: [ (x[:, :P-1] == 1) | zeros(P-1) ], 2P-2 columns."""
    if not table.int_labels:
        raise KeyError(1)                                      # what pandas raises at load.py:41 for the labels '1'..'p'
    X = table.markers
    n, P = X.shape
    if P < 2:
        return X
    ld = (2 * P - 2 + 15) // 16 * 16
    out = torch.zeros((n, ld), dtype=torch.int8, device=X.device)
    out[:, :P - 1] = (X[:, :P - 1] == 1).to(torch.int8)
    return out[:, :2 * P - 2]


def node_features(markers, scaled):
    """(x handed to the network, search(k) -> canonical lists).  load.py:45-59."""
    if scaled:
        x64, x32 = ops.scale_columns(markers)                  # load.py:45-48 (float64 for the search, float32 for x)
        return x32, (lambda kk: ops.knn_float(x64, kk))
    return markers, (lambda kk: ops.knn(markers, kk))


def graph_from_markers(markers, yv, n_neighbors, smoothing, use_weights, scaled, include_self, mode, remove_mean):
    """load.py:33-81 after the frame has been taken apart: markers int8 [n, p] on the GPU, yv float64 on the host."""
    if remove_mean:
        yv = yv - yv.mean()                                    # load.py:35-36 (float64, then float32 at :51)
    dev = markers.device
    y = torch.from_numpy(np.ascontiguousarray(yv, dtype=np.float64).astype(np.float32).reshape(-1, 1)).to(dev)
    x, search = node_features(markers, scaled)
    n = markers.shape[0]
    if include_self:
        idx, d2 = ops.knn_including_self(search, n, n_neighbors)       # load.py:57
    else:
        idx, d2 = search(n_neighbors)                          # load.py:52-59
    edge_index, edge_weight = ops.build_edges(idx, d2, mode=mode, smoothing=smoothing, use_weights=use_weights)
    data = Data(x=x, y=y, edge_index=edge_index)
    if use_weights:
        data.edge_weight = edge_weight
    return data


def check_metric(algorithm, mode):
    name = algorithm if isinstance(algorithm, str) else getattr(algorithm, "__name__", str(algorithm))
    if name not in _EUCLIDEAN:
        raise NotImplementedError(f"metric {name!r}: only the Euclidean distance is implemented")
    if mode not in ("connectivity", "distance"):
        raise ValueError('Unsupported mode, must be one of "connectivity", or "distance" but got "%s" instead' % mode)


def create_graph(df, n_neighbors=1, smoothing="laplacian", use_weights=False, hot=False, remove_mean=True,
                 scaled=False, include_self=False, mode="connectivity", algorithm="minkowski"):
    """load.py:21-81 on the GPU.  ``df``: pandas DataFrame (column 'value' + markers) or GenotypeTable."""
    check_metric(algorithm, mode)
    table = df if isinstance(df, GenotypeTable) else GenotypeTable.from_frame(df)
    markers = hot_markers(table) if hot else table.markers
    return graph_from_markers(markers, table.value, n_neighbors, smoothing, use_weights, scaled, include_self, mode,
                              remove_mean)


class ProcessedDataset(GraphDataset):
    """A dataset built by ``process()`` and kept under ``<root>/processed/<processed_file_names[0]>`` like the
    reference's InMemoryDataset subclasses (load.py:123-124, 283-284; load_ensembles2.py:35, 296-297)."""

    root = "."
    device = "cuda"

    def _build(self, cache):
        graphs = self._load_cache() if cache else None
        if graphs is None:
            graphs = self.process()
            if cache:
                self._save_cache(graphs)
        GraphDataset.__init__(self, graphs)

    # -- processed/<name>.pt cache (load.py:123-124, 283-284): same file name, plain tensors inside
    @property
    def processed_paths(self):
        return [os.path.join(self.root, "processed", n) for n in self.processed_file_names]

    @staticmethod
    def _pack(g):
        out = {k: v.detach().cpu() for k, v in vars(g).items() if torch.is_tensor(v)}
        out["x_stride"] = int(g.x.stride(0))
        return out

    @staticmethod
    def _unpack(d, device):
        x = d["x"]
        buf = torch.zeros((x.shape[0], d.get("x_stride", x.shape[1])), dtype=x.dtype)
        buf[:, :x.shape[1]] = x
        g = Data(x=buf.to(device)[:, :x.shape[1]], y=d["y"].to(device), edge_index=d["edge_index"].to(device))
        if "edge_weight" in d:
            g.edge_weight = d["edge_weight"].to(device)
        return g

    def _save_cache(self, graphs):
        """One entry per graph; the shared test graph and the validation graph are stored once and re-linked on load."""
        os.makedirs(os.path.join(self.root, "processed"), exist_ok=True)
        pool, refs = [], {}

        def ref(g):
            if id(g) not in refs:
                refs[id(g)] = len(pool)
                pool.append(self._pack(g))
            return refs[id(g)]

        items = []
        for g in graphs:
            items.append({"self": ref(g), "test": ref(g.test) if hasattr(g, "test") else None,
                          "valid": ref(g.valid) if hasattr(g, "valid") else None})
        torch.save({"format": "gnn_b200-1", "pool": pool, "items": items}, self.processed_paths[0])

    def _load_cache(self):
        path = self.processed_paths[0]
        if not os.path.exists(path):
            return None
        try:
            blob = torch.load(path, map_location="cpu")
        except Exception:
            return None
        if not isinstance(blob, dict) or blob.get("format") != "gnn_b200-1":
            return None               # a file written by the reference itself: rebuild rather than guess its layout
        pool = [self._unpack(d, self.device) for d in blob["pool"]]
        graphs = []
        for it in blob["items"]:
            g = pool[it["self"]]
            if it["test"] is not None:
                g.test = pool[it["test"]]
            if it["valid"] is not None:
                g.valid = pool[it["valid"]]
            graphs.append(g)
        return graphs



class PlainGraph(ProcessedDataset):
    """load.py:84-284: CSV -> train/test/validation split -> one graph per subsampled batch (or one
    training graph), each carrying ``.test`` (and ``.valid``).  Same constructor arguments."""

    def __init__(self, root, bits, raw_filename, num_neighbours=1, smoothing="laplacian", mode="connectivity",
                 use_validation=False, split=None, use_weights=False, algorithm="minkowski", validation_size=0.1,
                 include_self=False, hot=False, scaled=False, remove_mean=True, batches=None,
                 split_algorithm=split_dataset, split_algorithm_params=None, device="cuda", cache=True):
        self.root = root
        self.use_validation = use_validation
        self.num_neighbours = num_neighbours
        self.smoothing = smoothing
        self.mode = mode
        self.raw_filename = raw_filename
        self.bits = bits
        self.split = split if split is not None else 0.8
        self.use_weights = use_weights
        self.algorithm = algorithm
        self.include_self = include_self
        self.validation_size = validation_size
        self.hot = hot
        self.scaled = scaled
        self.remove_mean = remove_mean
        self.batches = batches
        self.split_algorithm = split_algorithm
        self.split_algorithm_params = split_algorithm_params if split_algorithm_params is not None else {}
        self.device = device
        self._build(cache)

    @property
    def raw_file_names(self):
        return [self.raw_filename]

    @property
    def processed_file_names(self):
        """The cache key of load.py:132-155 (kept so that runs can be matched to the reference's files)."""
        bits = "whole" if self.bits is None else "bits"
        filename = self.raw_filename.replace("/", ":")
        algorithm = self.algorithm if isinstance(self.algorithm, str) else self.algorithm.__name__
        return [f"data-{filename}-{self.split}-{self.use_validation}-{self.num_neighbours}-{self.smoothing}-"
                f"{self.mode}-{self.use_weights}-{algorithm}-{self.validation_size}-{self.include_self}-{self.hot}-"
                f"{self.scaled}-{self.remove_mean}-{self.batches}-{self.split_algorithm.__name__}-"
                f"{hash_dict(self.split_algorithm_params)}-{bits}.pt"]

    def _read(self, filename):
        import pandas as pd
        with open(filename) as fp:
            column_count = len(fp.readline().split(","))
        labels = ["value"] + [str(i + 1) for i in range(column_count - 1)]
        df = pd.read_csv(filename, names=labels)                # load.py:182-187
        if self.bits is not None:
            df = df[["value"] + self.bits]
            df.columns = ["value"] + list(range(1, len(self.bits) + 1))      # load.py:189-191
        return df

    def _graph(self, table):
        return create_graph(table, n_neighbors=self.num_neighbours, smoothing=self.smoothing,
                            use_weights=self.use_weights, hot=self.hot, remove_mean=self.remove_mean,
                            scaled=self.scaled, include_self=self.include_self, mode=self.mode,
                            algorithm=self.algorithm)

    def process(self):
        data_list = []
        for filename in self.raw_file_names:
            df = self._read(filename)
            whole = GenotypeTable.from_frame(df, device=self.device)       # one H2D copy of the genotypes
            n = len(whole)
            split = int(n * self.split) if self.split < 1 else self.split  # load.py:193-201
            valid = (self.validation_size if isinstance(self.validation_size, int)
                     else int(self.validation_size * (n - split)))
            valid = valid if self.use_validation else 0
            t_train, t_test, t_valid = self.split_algorithm(
                whole, (split, n - valid - split, n - valid),
                neighbours=self.num_neighbours, metric=self.algorithm, **self.split_algorithm_params)
            test_data = self._graph(t_test)
            assert test_data.edge_index.shape[0] > 0
            valid_data = None
            if self.use_validation:
                valid_data = self._graph(t_valid)
                valid_data.test = valid_data
            if self.batches is not None:
                for batch in self.split_algorithm(whole, self.batches, neighbours=self.num_neighbours,
                                                  metric=self.algorithm, **self.split_algorithm_params):
                    data = self._graph(batch)                   # per-batch kNN rebuilt inside the batch (load.py:245-257)
                    assert data.edge_index.shape[0] > 0
                    data.test = test_data
                    data_list.append(data)
                if self.use_validation:
                    data_list.append(valid_data)
            else:
                data = self._graph(t_train)
                data.test = test_data
                if self.use_validation:
                    data.valid = valid_data
                assert data.edge_index.shape[0] > 0
                data_list.append(data)
        return data_list


def load_data(filename, bits=None, num_neighbours=1, smoothing="laplacian", mode="connectivity", use_validation=False,
              split=None, use_weights=False, algorithm="minkowski", validation_size=0.1, hot=False, scaled=False,
              remove_mean=True, batches=None, split_algorithm=split_dataset, split_algorithm_params=None, **_):
    """load.py:287-325."""
    return PlainGraph(".", bits=bits, raw_filename=filename, num_neighbours=num_neighbours, smoothing=smoothing,
                      mode=mode, use_validation=use_validation, split=split, use_weights=use_weights,
                      algorithm=algorithm, validation_size=validation_size, hot=hot, scaled=scaled,
                      remove_mean=remove_mean, batches=batches, split_algorithm=split_algorithm,
                      split_algorithm_params=split_algorithm_params)
