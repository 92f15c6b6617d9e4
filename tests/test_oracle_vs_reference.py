"""Live cross-check of the oracle (and the host-side drop-ins) against the reference itself.  Only runs
where /root/reference is mounted (the build container); the GPU box relies on tests/golden/."""
import numpy as np
import pytest

from oracle import ref_import, restate

pytestmark = pytest.mark.skipif(not ref_import.available(), reason="/root/reference is not mounted here")


@pytest.fixture(scope="module")
def ref():
    return ref_import.load(canonical_knn=True)


def test_naive_partition_drop_in_equals_reference(ref):
    from gnn_b200.loaders.util import naive_partition
    rng = np.random.default_rng(3)
    for _ in range(60):
        n = int(rng.integers(3, 20))
        k = int(rng.integers(1, min(3, n - 1) + 1))
        nbr = np.stack([rng.choice(np.delete(np.arange(n), i), size=k, replace=False) for i in range(n)])
        edges = [(i, int(c)) for i in range(n) for c in nbr[i]]
        for bidir in (False, True):
            size = int(rng.integers(1, n + 2))
            trav = set(int(v) for v in rng.choice(n, size=int(rng.integers(0, n)), replace=False))
            want = [int(v) for v in ref.util.naive_partition(edges, size, bidirectional=bidir, traversed=set(trav))]
            got = naive_partition(edges, size, bidirectional=bidir, traversed=set(trav))
            assert got == want, (edges, size, bidir, trav)


def test_create_graph_on_synthetic_genotypes(ref):
    import pandas as pd
    from oracle import synth
    X = synth.genotypes(150, 97)
    y = synth.phenotypes(150, 97)[:, 0]
    df = pd.DataFrame(np.column_stack([y, X.astype(np.float64)]), columns=["value"] + [str(i + 1) for i in range(97)])
    for mode in ("distance", "connectivity"):
        d = ref.load.create_graph(df.copy(), n_neighbors=4, smoothing="laplacian", use_weights=True, mode=mode,
                                  algorithm="euclidean")
        mine = restate.create_graph(X, y, 4, "laplacian", True, True, mode)
        assert np.array_equal(d.edge_index.numpy(), mine["edge_index"])
        assert np.array_equal(d.edge_weight.numpy(), mine["edge_weight"])


def test_sklearn_agrees_on_tie_free_rows():
    from oracle import synth
    from sklearn.neighbors import kneighbors_graph
    X = synth.genotypes(300, 501).astype(np.float64)
    k = 7
    idx, d2 = restate.canonical_knn(X, k + 1)
    g = kneighbors_graph(X, k, mode="distance", include_self=False).tocsr()
    sk_i, sk_d = g.indices.reshape(300, k), g.data.reshape(300, k)
    assert np.array_equal(np.sort(sk_d, axis=1), np.sqrt(d2[:, :k]))
    tie_free = d2[:, k - 1] != d2[:, k]
    same = np.array([set(a) == set(b) for a, b in zip(sk_i, idx[:, :k])])
    assert tie_free.sum() > 100 and np.all(same[tie_free])
