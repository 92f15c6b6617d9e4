"""Feature transforms and the seeded-subsample loader on the B200:
``scaled`` / ``include_self`` / ``hot`` of load.py:38-48,:57 and load_ensembles2.py:53-297 against the fixtures
frozen from the reference (tests/golden/wheat_features.npz, wheat_ensemble.npz) and against the oracle on seeded
inputs.  Indices, neighbour lists and the scaled features are bit-exact; float64 weights to 1e-12 (exp)."""
import os

import numpy as np
import pytest
import torch

from gnn_b200 import ops
from gnn_b200.data import GenotypeTable
from gnn_b200.loaders.load import create_graph, load_data
from gnn_b200.loaders.load_ensembles2 import EnsembleGraph, load_data_ensembles2
from oracle import restate
from oracle.make_golden import digest

pytestmark = pytest.mark.gpu


def dev(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
    return t if dtype is None else t.to(dtype)


def cmp_edges(data, ref_ei, ref_ew=None):
    torch.cuda.synchronize()
    ei = data.edge_index.cpu().numpy()
    assert ei.shape == ref_ei.shape, (ei.shape, ref_ei.shape)
    assert np.array_equal(ei, ref_ei.astype(np.int64))
    if ref_ew is not None:
        np.testing.assert_allclose(data.edge_weight.cpu().numpy(), ref_ew, rtol=1e-12, atol=1e-15)


# ------------------------------------------------------------------------------------------ scaling
@pytest.mark.parametrize("n", [1, 2, 7, 8, 9, 64, 127, 128, 129, 255, 256, 257, 300, 1000, 1031, 4099])
def test_scale_columns_bit_exact(n):
    """(x - mean) / std with pandas' summation order: every tree shape of numpy's pairwise sum."""
    rng = np.random.default_rng(n)
    p = 37
    X = rng.integers(0, 3, size=(n, p)).astype(np.int8)
    X[:, 3] = 1                                         # constant column: 0 / 0 -> NaN -> 0
    if n > 2:
        X[:, 5] = 0
        X[n // 2, 5] = 2                                # a single carrier
    want = restate.scale_features(X)
    x64, x32 = ops.scale_columns(dev(X))
    torch.cuda.synchronize()
    got = x64.cpu().numpy()
    assert np.array_equal(got, want), np.abs(got - want).max()
    assert np.array_equal(x32.cpu().numpy(), want.astype(np.float32))
    assert x32.stride(0) % 4 == 0


def test_scale_columns_wheat(wheat, golden):
    X, _ = wheat
    g = golden("wheat_features.npz")
    _, x32 = ops.scale_columns(dev(X[:479]), want64=False)
    torch.cuda.synchronize()
    assert np.array_equal(x32.cpu().numpy(), g["scaled_train_x"])


# ------------------------------------------------------------------------------------------ float kNN
@pytest.mark.parametrize("n,p,k", [(50, 7, 3), (333, 130, 9), (1500, 64, 32), (700, 1000, 64)])
def test_knn_float_vs_oracle(n, p, k):
    rng = np.random.default_rng(n + p)
    X = rng.normal(size=(n, p))
    X[n // 3] = X[1]                                    # exact duplicates: ties at distance 0 and elsewhere
    X[n // 2] = X[1]
    want_i, want_d = restate.canonical_knn(np.zeros((n, 0)), k, extra=X)
    idx, d2 = ops.knn_float(dev(X), k)
    torch.cuda.synchronize()
    assert np.array_equal(d2.cpu().numpy(), want_d)
    assert np.array_equal(idx.cpu().numpy().astype(np.int64), want_i)


def test_knn_float_wheat_golden(wheat, golden):
    X, _ = wheat
    g = golden("wheat_features.npz")
    x64, _ = ops.scale_columns(dev(X[:479]), want32=False)
    idx, d2 = ops.knn_float(x64, 5)
    torch.cuda.synchronize()
    assert np.array_equal(idx.cpu().numpy(), g["scaled_train_idx"])
    assert np.array_equal(d2.cpu().numpy(), g["scaled_train_d2"])


@pytest.mark.parametrize("n,k", [(1, 1), (5, 5), (40, 3), (300, 64)])
def test_include_self_lists(n, k):
    rng = np.random.default_rng(n * 7 + k)
    X = rng.integers(0, 3, size=(n, 24)).astype(np.int8)
    if n >= 40:
        X[10] = X[4]; X[20] = X[4]; X[2] = X[4]          # duplicates on both sides of the row index
    want_i, want_d = restate.canonical_knn(X, k, include_self=True)
    idx, d2 = ops.knn_including_self(lambda kk: ops.knn(dev(X), kk), n, k)
    torch.cuda.synchronize()
    assert np.array_equal(idx.cpu().numpy().astype(np.int64), want_i)
    assert np.array_equal(d2.cpu().numpy(), want_d)
    with pytest.raises(ValueError):
        ops.knn_including_self(lambda kk: ops.knn(dev(X), kk), n, n + 1)


# ------------------------------------------------------------------------------------------ create_graph
@pytest.mark.parametrize("mode", ["distance", "connectivity"])
@pytest.mark.parametrize("smoothing", ["laplacian", "none"])
def test_create_graph_feature_goldens(wheat, golden, mode, smoothing):
    X, y = wheat
    g = golden("wheat_features.npz")
    train, test = slice(0, 479), slice(479, 587)
    cases = [
        (f"scaled_train_k5_{mode}_{smoothing}", train, 5, dict(scaled=True)),
        (f"self_train_k5_{mode}_{smoothing}", train, 5, dict(include_self=True)),
        (f"scaledself_test_k3_{mode}_{smoothing}", test, 3, dict(scaled=True, include_self=True)),
    ]
    for tag, rows, k, kw in cases:
        d = create_graph(GenotypeTable.from_numpy(X[rows], y[rows]), n_neighbors=k, smoothing=smoothing,
                         use_weights=True, mode=mode, algorithm="euclidean", **kw)
        cmp_edges(d, g[tag + "_ei"], g[tag + "_ew"])
        x = d.x.float().cpu().numpy()
        assert digest(x) == g[tag + "_xdigest"], tag
        assert d.x.dtype == (torch.float32 if kw.get("scaled") else torch.int8)


def test_constant_columns_and_duplicates(wheat, golden):
    import pandas as pd
    X, y = wheat
    g = golden("wheat_features.npz")
    Xd = X[:80].copy()
    Xd[:, g["const_cols"][0]] = 1
    Xd[:, g["const_cols"][1]] = 0
    r = g["const_dup_rows"]
    Xd[r[1]] = Xd[r[0]]; Xd[r[2]] = Xd[r[0]]
    fr = pd.DataFrame(np.column_stack([y[:80], Xd.astype(np.float64)]),
                      columns=["value"] + [str(i + 1) for i in range(Xd.shape[1])])
    d = create_graph(fr, n_neighbors=4, smoothing="laplacian", use_weights=True, mode="distance", scaled=True,
                     include_self=True)
    cmp_edges(d, g["const_k4_distance_laplacian_ei"], g["const_k4_distance_laplacian_ew"])
    assert digest(d.x.cpu().numpy()) == g["const_k4_distance_laplacian_xdigest"]
    d = create_graph(fr, n_neighbors=4, smoothing="laplacian", use_weights=True, mode="connectivity", include_self=True)
    cmp_edges(d, g["const_k4_connectivity_laplacian_ei"], g["const_k4_connectivity_laplacian_ew"])


def test_hot_golden_and_keyerror(wheat, golden):
    import pandas as pd
    X, y = wheat
    g = golden("wheat_features.npz")
    cols = g["hot_bits"] - 1
    fb = pd.DataFrame(np.column_stack([y, X[:, cols].astype(np.float64)]),
                      columns=["value"] + list(range(1, len(cols) + 1)))
    d = create_graph(fb.iloc[:479], n_neighbors=5, smoothing="laplacian", use_weights=True, mode="distance", hot=True)
    cmp_edges(d, g["hot_train_k5_distance_laplacian_ei"], g["hot_train_k5_distance_laplacian_ew"])
    assert np.array_equal(d.x.cpu().numpy(), g["hot_train_x"])
    assert d.x.shape[1] == 2 * len(cols) - 2
    d = create_graph(fb.iloc[479:587], n_neighbors=3, smoothing="none", use_weights=True, mode="distance", hot=True,
                     scaled=True)
    cmp_edges(d, g["hotscaled_test_k3_distance_none_ei"], g["hotscaled_test_k3_distance_none_ew"])
    assert digest(d.x.cpu().numpy()) == g["hotscaled_test_k3_distance_none_xdigest"]
    with pytest.raises(KeyError):                        # string labels '1'..'p': load.py:41 raises KeyError(1)
        create_graph(GenotypeTable.from_numpy(X[:100], y[:100]), n_neighbors=3, hot=True)


def test_scaled_graph_trains(wheat):
    """float32 node features go through the same projection / aggregation kernels as int8 ones."""
    from gnn_b200.networks.networks import create_network_two_no_conv_dropout
    X, y = wheat
    d = create_graph(GenotypeTable.from_numpy(X[:479], y[:479]), n_neighbors=5, use_weights=True, mode="distance", scaled=True)
    o = restate.create_graph(X[:479], y[:479], n_neighbors=5, use_weights=True, mode="distance", scaled=True)
    torch.manual_seed(0)
    ref = restate.make_model("create_network_two_no_conv_dropout", X.shape[1], 1, internal_size=32, dropout=0.0)
    m = create_network_two_no_conv_dropout(X.shape[1], 1, internal_size=32, dropout=0.0).cuda()
    m.load_state_dict(ref.state_dict())
    out = m(d.x, d.edge_index, d.edge_weight.float())
    want = ref(*restate.as_tensors(o))
    err = (out.cpu() - want).abs().max().item() / want.abs().max().item()
    assert err < 1e-4, err
    loss = torch.nn.functional.mse_loss(out, d.y)
    loss.backward()
    lref = torch.nn.functional.mse_loss(want, torch.from_numpy(o["y"]))
    lref.backward()
    for (k, a), (_, b) in zip(m.named_parameters(), ref.named_parameters()):
        e = (a.grad.cpu() - b.grad).abs().max().item() / max(b.grad.abs().max().item(), 1e-12)
        assert e < 1e-4, (k, e)


# ------------------------------------------------------------------------------------------ ensemble loader
ENSEMBLE_CASES = {
    "a": dict(num_neighbours=3, split=3, use_weights=True, use_validation=True, mode="distance"),
    "b": dict(num_neighbours=2, split=2, use_weights=True, scaled=False, add_full=True),
    "c": dict(num_neighbours=2, split=2, use_weights=True, scaled=False, add_full=True, full_split=0.8),
    "d": dict(num_neighbours=2, split=2, use_weights=False, separate_sets=False, train_size=300),
    "e": dict(num_neighbours=2, split=2, use_weights=True, separate_sets=False, full_split=400, train_size=0.6,
              include_self=True),
}


@pytest.mark.parametrize("name", sorted(ENSEMBLE_CASES))
def test_ensemble_loader_golden(wheat_csv, golden, tmp_path, monkeypatch, name):
    monkeypatch.chdir(tmp_path)
    g = golden("wheat_ensemble.npz")
    kw = ENSEMBLE_CASES[name]
    np.random.seed(int(g[f"{name}_seed"]))
    ds = load_data_ensembles2(wheat_csv, **kw)
    assert len(ds) == int(g[f"{name}_n"])
    for i in range(len(ds)):
        for sub, d in (("", ds[i]), ("_test", ds[i].test)):
            key = f"{name}_{i}{sub}"
            cmp_edges(d, g[key + "_ei"], g[key + "_ew"] if kw.get("use_weights") else None)
            assert np.array_equal(d.y.cpu().numpy(), g[key + "_y"]), key
            assert digest(d.x.float().cpu().numpy()) == g[key + "_xdigest"], key
    if kw.get("use_validation"):
        v = ds[len(ds) - 1]                              # items are shallow copies, like dataset[i] in the reference
        assert v.test.x.data_ptr() == v.x.data_ptr() and v.test.edge_index.data_ptr() == v.edge_index.data_ptr()
    # second construction comes from processed/<reference file name> and is identical
    files = os.listdir(tmp_path / "processed")
    assert len(files) == 1 and files[0].startswith("data-ensemble2-") and files[0].endswith("-whole.pt")
    again = load_data_ensembles2(wheat_csv, **kw)
    assert len(again) == len(ds)
    for a, b in zip(again, ds):
        assert torch.equal(a.x, b.x) and torch.equal(a.edge_index, b.edge_index) and torch.equal(a.test.y, b.test.y)


def test_ensemble_loader_hot_raises(wheat_csv, golden, tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    with pytest.raises(KeyError) as e:
        EnsembleGraph(".", None, wheat_csv, split=1, hot=True, scaled=False)
    assert e.value.args[0] == int(golden("wheat_ensemble.npz")["hot_keyerror"])


def test_plain_loader_scaled(wheat_csv, wheat, tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    X, y = wheat
    ds = load_data(wheat_csv, num_neighbours=4, mode="distance", use_weights=True, scaled=True)
    d = ds[0]
    o = restate.create_graph(X[:479], y[:479], n_neighbors=4, use_weights=True, mode="distance", scaled=True)
    cmp_edges(d, o["edge_index"], o["edge_weight"])
    assert np.array_equal(d.x.cpu().numpy(), o["x"])
    t = restate.create_graph(X[479:], y[479:], n_neighbors=4, use_weights=True, mode="distance", scaled=True)
    cmp_edges(d.test, t["edge_index"], t["edge_weight"])
