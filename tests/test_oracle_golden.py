"""The oracle restatement against the fixtures frozen from the reference's own run (oracle/make_golden.py)."""
import numpy as np
import pytest
import torch

from oracle import restate, synth


def test_wheat_shape(wheat):
    X, y = wheat
    assert X.shape == (599, 1279) and set(np.unique(X)) <= {0, 1}
    assert restate.split_sizes(599, 0.8, True, 0.1) == (479, 108, 587)          # load.py:193-209


def test_knn_golden_and_prefix_property(wheat, golden):
    X, y = wheat
    g = golden("wheat_knn.npz")
    idx, d2 = restate.canonical_knn(X, 32)
    assert np.array_equal(idx, g["idx_all"]) and np.array_equal(d2, g["d2_all"])
    for k in (1, 5, 10):                       # canonical order => shorter lists are prefixes
        a, b = restate.canonical_knn(X, k)
        assert np.array_equal(a, idx[:, :k]) and np.array_equal(b, d2[:, :k])
    it, dt = restate.canonical_knn(X[:479], 32)
    assert np.array_equal(it, g["idx_train"]) and np.array_equal(dt, g["d2_train"])
    ip, dp = restate.canonical_knn(X, 10, extra=y[:, None])
    assert np.array_equal(ip, g["idx_with_pheno"]) and np.array_equal(dp, g["d2_with_pheno"])
    # lists are sorted by (d2, j) and never contain the row itself
    assert np.all(np.diff(d2, axis=1) >= 0)
    ties = np.diff(d2, axis=1) == 0
    assert np.all(np.diff(idx, axis=1)[ties] > 0)
    assert not np.any(idx == np.arange(599)[:, None])


@pytest.mark.parametrize("k", [3, 5, 10])
@pytest.mark.parametrize("mode", ["distance", "connectivity"])
@pytest.mark.parametrize("smoothing", ["laplacian", "none"])
def test_graph_golden(wheat, golden, k, mode, smoothing):
    X, y = wheat
    g = golden("wheat_graph.npz")
    for rows, tag in ((slice(0, 479), "train"), (slice(479, 587), "test")):
        mine = restate.create_graph(X[rows], y[rows], k, smoothing, True, True, mode)
        key = f"{tag}_k{k}_{mode}_{smoothing}"
        assert np.array_equal(mine["edge_index"], g[key + "_ei"])
        assert np.array_equal(mine["edge_weight"], g[key + "_ew"])
        if smoothing == "laplacian":           # sorted off-diagonal block, then the diagonal
            n = X[rows].shape[0]
            ei = mine["edge_index"]
            assert np.array_equal(ei[:, -n:], np.stack([np.arange(n)] * 2))
            off = ei[:, :-n]
            assert np.all(off[0] != off[1])
            lin = off[0] * n + off[1]
            assert np.all(np.diff(lin) > 0)


def test_graph_duplicates_golden(wheat, golden):
    X, y = wheat
    g = golden("wheat_graph.npz")
    Xd = X[:60].copy()
    Xd[7] = Xd[3]; Xd[11] = Xd[3]; Xd[40] = Xd[41]
    mine = restate.create_graph(Xd, y[:60], 2, "laplacian", True, True, "distance")
    assert np.array_equal(mine["edge_index"], g["dup_k2_distance_laplacian_ei"])
    assert np.array_equal(mine["edge_weight"], g["dup_k2_distance_laplacian_ew"])


def test_partition_golden(wheat, golden):
    X, y = wheat
    g = golden("wheat_partition.npz")
    table = np.column_stack([y, X.astype(np.float64)])
    for k in (3, 5, 10):
        for batches, tag in ((3, "b3"), (5, "b5"), (8, "b8"), (10, "b10"), ((479, 108, 587), "split")):
            mine = restate.split_indices_graph(table, batches, neighbours=k)
            for i, m in enumerate(mine):
                assert np.array_equal(np.asarray(m), g[f"k{k}_{tag}_{i}"]), (k, tag, i)


def test_partition_small_golden(golden):
    g = golden("partition_small.npz")
    for c in range(40):
        nbr, szs = g[f"c{c}_nbr"], [int(s) for s in g[f"c{c}_sizes"]]
        S = restate.neighbour_stream(nbr)
        fast = restate.partition_linear(S, nbr.shape[0], szs)
        assert [v for b in fast for v in b] == list(g[f"c{c}_flat"])
        assert [len(b) for b in fast] == list(g[f"c{c}_lens"])
        trav, slow = set(), []
        for s in szs:
            got = restate.partition_scan(S, nbr.shape[0], s, trav)
            trav |= set(got)
            slow.append(got)
        assert slow == fast


def test_model_golden(wheat, golden):
    X, y = wheat
    g = golden("wheat_model.npz")
    graph = restate.create_graph(X[:479], y[:479], 5, "laplacian", True, True, "distance")
    args = restate.as_tensors(graph)
    yt = torch.as_tensor(graph["y"])
    tags = sorted({k.split("__")[0] for k in g.files})
    assert len(tags) == 5
    for tag in tags:
        name, asl = tag.rsplit("_asl", 1)
        m = restate.make_model(name, 1279, 1, internal_size=32, dropout=0.0, add_self_loops=bool(int(asl)))
        m.load_state_dict({k.split("__w__")[1]: torch.from_numpy(g[k]) for k in g.files if k.startswith(tag + "__w__")})
        m.train()
        out = m(*args)
        loss = torch.nn.MSELoss()(out, yt)
        loss.backward()
        np.testing.assert_allclose(out.detach().numpy(), g[tag + "__out"], rtol=1e-5, atol=1e-6)
        assert abs(float(loss) - float(g[tag + "__loss"])) < 1e-6
        for kk, p in m.named_parameters():
            np.testing.assert_allclose(p.grad.numpy(), g[f"{tag}__g__{kk}"], rtol=1e-4, atol=1e-6)


def test_train_trajectory_golden(wheat, golden):
    X, y = wheat
    g = golden("wheat_train.npz")
    te = g["test_rows"]
    g_test = restate.create_graph(X[te], y[te], 5, "laplacian", True, True, "distance")
    for b in range(3):
        rows = g[f"b{b}_rows"]
        gb = restate.create_graph(X[rows], y[rows], 5, "laplacian", True, True, "distance")
        m = restate.make_model("create_network_two_no_conv_dropout", 1279, 1, internal_size=int(g["hidden"]), dropout=0.0)
        m.load_state_dict({k.split("__w0__")[1]: torch.from_numpy(g[k]) for k in g.files if k.startswith(f"b{b}__w0__")})
        opt = torch.optim.Adam(m.parameters(), lr=float(g["lr"]))
        tl, vl = [], []
        for _ in range(int(g["epochs"])):
            a, c, out_test = restate.epoch_step(m, opt, gb, g_test)
            tl.append(a)
            vl.append(c)
        np.testing.assert_allclose(tl, g[f"b{b}_train_loss"], rtol=1e-5)
        np.testing.assert_allclose(vl, g[f"b{b}_test_loss"], rtol=1e-5)
        assert restate.pearson(out_test.numpy(), g[f"b{b}_pred_test"]) > 0.999999


def test_synth_is_stateless_and_in_range():
    a = synth.genotypes(50, 300)
    assert a.dtype == np.int8 and set(np.unique(a)) <= {0, 1, 2}
    b = synth.genotypes(20, 300, row0=30)
    assert np.array_equal(a[30:], b)
    c = synth.genotypes(50, 7, cols=[3, 10, 11, 200, 201, 202, 299], p_total=300)
    assert np.array_equal(c, a[:, [3, 10, 11, 200, 201, 202, 299]])
    # pinned values: any change to the generator changes every benchmark input
    assert synth.splitmix64(np.uint64(0)) == np.uint64(0xE220A8397B1DCDAF)
    assert int(a.astype(np.int64).sum()) == int(synth.genotypes(50, 300).astype(np.int64).sum())
    f = a.mean(axis=0) / 2
    assert 0.0 <= f.min() and f.max() <= 0.75
    y = synth.phenotypes(200, 300, traits=3)
    assert y.shape == (200, 3) and np.allclose(y.mean(axis=0), 0)


# ---- feature transforms and the ensemble loader (fixtures frozen from the reference by oracle/make_golden.py) ----
def test_feature_graphs_golden(wheat, golden):
    from oracle.make_golden import digest
    X, y = wheat
    g = golden("wheat_features.npz")
    train, test = slice(0, 479), slice(479, 587)
    for mode, smoothing in (("distance", "laplacian"), ("connectivity", "none")):
        for tag, rows, k, kw in ((f"scaled_train_k5_{mode}_{smoothing}", train, 5, dict(scaled=True)),
                                 (f"self_train_k5_{mode}_{smoothing}", train, 5, dict(include_self=True)),
                                 (f"scaledself_test_k3_{mode}_{smoothing}", test, 3, dict(scaled=True, include_self=True))):
            m = restate.create_graph(X[rows].astype(np.float64), y[rows], n_neighbors=k, smoothing=smoothing,
                                     use_weights=True, mode=mode, **kw)
            assert np.array_equal(m["edge_index"], g[tag + "_ei"]), tag
            assert np.array_equal(m["edge_weight"], g[tag + "_ew"]), tag
            assert digest(m["x"]) == g[tag + "_xdigest"], tag
    cols = g["hot_bits"] - 1
    m = restate.create_graph(X[train][:, cols].astype(np.float64), y[train], n_neighbors=5, use_weights=True,
                             mode="distance", hot=True, int_labels=True)
    assert np.array_equal(m["edge_index"], g["hot_train_k5_distance_laplacian_ei"])
    assert np.array_equal(m["x"].astype(np.int8), g["hot_train_x"])
    with pytest.raises(KeyError):
        restate.create_graph(X[:50].astype(np.float64), y[:50], n_neighbors=2, hot=True)


def test_pairwise_sum_restatement_matches_pandas_std():
    """The summation order the CUDA scaling kernel reproduces (numpy's pairwise sum) is what pandas' std uses."""
    def pw(a):
        n = len(a)
        if n < 8:
            r = 0.0
            for v in a:
                r = r + v
            return r
        if n <= 128:
            r = [a[i] for i in range(8)]
            i = 8
            while i < n - (n % 8):
                for j in range(8):
                    r[j] = r[j] + a[i + j]
                i += 8
            res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]))
            while i < n:
                res = res + a[i]
                i += 1
            return res
        n2 = n // 2
        n2 -= n2 % 8
        return pw(a[:n2]) + pw(a[n2:])

    rng = np.random.default_rng(3)
    for n in (5, 8, 100, 129, 300, 1031):
        X = rng.integers(0, 3, size=(n, 6)).astype(np.float64)
        want = restate.scale_features(X)
        for c in range(X.shape[1]):
            d = X[:, c] - X[:, c].sum() / n
            avg = pw(list(d)) / n
            sd = np.sqrt(pw(list((avg - d) ** 2)) / (n - 1))
            assert np.array_equal(np.nan_to_num(d / sd, nan=0.0) if sd == 0 else d / sd, want[:, c]), (n, c)


@pytest.mark.parametrize("name,kw", [
    ("a", dict(num_neighbours=3, split=3, use_weights=True, use_validation=True, mode="distance")),
    ("e", dict(num_neighbours=2, split=2, use_weights=True, separate_sets=False, full_split=400, train_size=0.6,
               include_self=True)),
])
def test_ensemble_loader_golden(wheat, golden, name, kw):
    from oracle.make_golden import digest
    X, y = wheat
    g = golden("wheat_ensemble.npz")
    np.random.seed(int(g[f"{name}_seed"]))
    graphs = restate.ensemble_graphs(X.astype(np.float64), y, **kw)
    assert len(graphs) == int(g[f"{name}_n"])
    for i, m in enumerate(graphs):
        for sub, d in (("", m), ("_test", m["test"])):
            key = f"{name}_{i}{sub}"
            assert np.array_equal(d["edge_index"], g[key + "_ei"]), key
            assert np.array_equal(d["edge_weight"], g[key + "_ew"]), key
            assert np.array_equal(d["y"], g[key + "_y"]) and digest(d["x"]) == g[key + "_xdigest"], key


def test_distance_correlation_restatement(golden):
    """restate.distance_correlation (dcor's biased V-statistic, plain.py:76-79): frozen values and the defining properties."""
    from oracle import restate
    g = golden("wheat_trainers.npz")
    for name in ("linear", "noise", "square", "const"):
        assert restate.distance_correlation(g[f"dcor_{name}_a"], g[f"dcor_{name}_b"]) == pytest.approx(float(g[f"dcor_{name}"]), abs=1e-12)
    a = g["dcor_noise_a"]
    assert restate.distance_correlation(a, -3.0 * a + 2.0) == pytest.approx(1.0, abs=1e-12)       # invariant under affine maps
    assert 0.0 < float(g["dcor_noise"]) < 0.3 < float(g["dcor_square"]) < 1.0
    assert restate.distance_correlation(a, g["dcor_noise_b"]) == pytest.approx(restate.distance_correlation(g["dcor_noise_b"], a), abs=1e-15)
    assert np.all(np.isfinite(g["plain_dcors"])) and np.all((g["plain_dcors"] >= 0) & (g["plain_dcors"] <= 1))
