"""Import the UNMODIFIED reference (/root/reference) on top of oracle.pyg_shim.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Works only where /root/reference
is mounted, i.e. in the build container -- never on the GPU box.  Used by
oracle/make_golden.py and by the CPU tests that re-check the restatement
against the live reference when it is present.

Nothing of the reference is copied: its modules are imported from where they
lie.  Two third-party names are substituted, both documented in DESIGN.md:

* ``torch_geometric``  -> oracle.pyg_shim (the wheel is absent from the image);
* ``kneighbors_graph`` -> the canonical (d^2, index) tie order of SURVEY.md A.2,
  installed into ``gnn.loaders.load`` / ``gnn.loaders.util`` only when
  ``canonical_knn=True`` (sklearn's own tie order depends on its thread count,
  so it cannot be a bit-exact target).
"""
import importlib
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("GNN_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "gnn", "loaders", "load.py"))


def _install_shim():
    from . import pyg_shim
    if "torch_geometric" in sys.modules and not getattr(sys.modules["torch_geometric"], "_is_oracle_shim", False):
        return  # a real torch_geometric is installed: use it
    tg = types.ModuleType("torch_geometric")
    tg._is_oracle_shim = True
    tg.__version__ = "2.1.0-shim"
    data = types.ModuleType("torch_geometric.data")
    data.Data, data.InMemoryDataset = pyg_shim.Data, pyg_shim.InMemoryDataset
    utils = types.ModuleType("torch_geometric.utils")
    utils.from_scipy_sparse_matrix = pyg_shim.from_scipy_sparse_matrix
    utils.add_remaining_self_loops = pyg_shim.add_remaining_self_loops

    def _no_networkx(*_a, **_k):            # imported by load_ensembles2.py:10, only used in commented-out code
        raise NotImplementedError("from_networkx is not on the path")
    utils.from_networkx = _no_networkx
    nn = types.ModuleType("torch_geometric.nn")
    nn.GCNConv = pyg_shim.GCNConv
    dense = types.ModuleType("torch_geometric.nn.dense")
    dense.DenseGCNConv = pyg_shim.DenseGCNConv
    nn.dense, nn.DenseGCNConv = dense, pyg_shim.DenseGCNConv
    tg.data, tg.utils, tg.nn = data, utils, nn
    sys.modules.update({
        "torch_geometric": tg, "torch_geometric.data": data, "torch_geometric.utils": utils,
        "torch_geometric.nn": nn, "torch_geometric.nn.dense": dense,
    })
    # the reference calls bare torch.load() on files it pickled itself (load.py:124), written for
    # torch 1.11; torch >= 2.6 defaults to weights_only=True, so allow-list the shim's own classes
    try:
        torch_ser = importlib.import_module("torch.serialization")
        torch_ser.add_safe_globals([pyg_shim.Data, pyg_shim._Collated])
    except Exception:
        pass
    for name in ("seaborn", "matplotlib", "matplotlib.pyplot"):      # plotting imports of load_ensembles2.py:11-12
        if name not in sys.modules:
            try:
                importlib.import_module(name)
            except Exception:
                sys.modules[name] = types.ModuleType(name)
    if not hasattr(sys.modules["matplotlib"], "pyplot"):
        sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    if "dcor" not in sys.modules:
        try:
            importlib.import_module("dcor")
        except Exception:
            # the package is absent from the image: plain.py:76-79 gets the restated V-statistic (restate.distance_correlation)
            from . import restate
            dcor = types.ModuleType("dcor")
            dcor.distance_correlation = restate.distance_correlation
            sys.modules["dcor"] = dcor


def load(canonical_knn=True):
    """Return a namespace with the reference's hot-path modules."""
    if not available():
        raise RuntimeError(f"reference not found under {REFERENCE_ROOT}")
    _install_shim()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    ns = types.SimpleNamespace()
    ns.load = importlib.import_module("gnn.loaders.load")
    ns.util = importlib.import_module("gnn.loaders.util")
    ns.networks = importlib.import_module("gnn.networks.networks")
    ns.sequential = importlib.import_module("gnn.networks.sequential")
    ns.trainers_util = importlib.import_module("gnn.trainers.util")
    ns.ensemble2 = importlib.import_module("gnn.trainers.ensemble2")
    ns.batches = importlib.import_module("gnn.trainers.batches")
    ns.plain = importlib.import_module("gnn.trainers.plain")
    ns.load_ensembles2 = importlib.import_module("gnn.loaders.load_ensembles2")
    if canonical_knn:
        from .restate import canonical_kneighbors_graph as knn_graph
    else:
        from sklearn.neighbors import kneighbors_graph as knn_graph
    ns.load.kneighbors_graph = knn_graph
    ns.util.kneighbors_graph = knn_graph
    ns.load_ensembles2.kneighbors_graph = knn_graph
    return ns
