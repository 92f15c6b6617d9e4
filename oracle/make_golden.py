"""Freeze tests/golden/*.npz from the reference's own Python -- TEST INFRASTRUCTURE.

Run in the build container (needs /root/reference):

    python -m oracle.make_golden

For every fixture the UNMODIFIED reference code (imported by oracle.ref_import,
with torch_geometric shimmed and the canonical kNN tie order installed) produces
the value, the restatement in oracle.restate is asserted to reproduce it, and
only then is it written.  The reference ships no golden vectors of its own
(SURVEY.md section 4), so these files are what pins parity.

Fixtures
  wheat.npz            the one dataset the reference ships (csv-data/WHEAT_combined.csv),
                       bit-packed markers + float64 phenotypes: the INPUT of every WHEAT case
  wheat_knn.npz        canonical neighbour lists, k = 32 (shorter k are prefixes), whole file and the
                       479-row training split; cross-checked against sklearn itself
  wheat_graph.npz      create_graph(load.py:21-81) edge_index / edge_weight for several settings
  wheat_partition.npz  split_dataset_graph(util.py:48-73) row indices
  wheat_model.npz      seeded forward / backward of the GCN factories
  wheat_train.npz      loss trajectories of train_ensemble2 phase 1 (ensemble2.py:39-124)
  partition_small.npz  naive_partition(util.py:15-45) on small random neighbour streams
  wheat_features.npz   create_graph with scaled / include_self / hot (load.py:38-48, :57)
  wheat_ensemble.npz   the seeded-subsample loader (load_ensembles2.py:53-297): rows, labels, edges, feature digests
  wheat_trainers.npz   trajectories of ``train`` (plain.py) and ``train_batches`` (batches.py), distance correlations

``python -m oracle.make_golden features`` regenerates only wheat_features / wheat_ensemble, ``... trainers`` only the last.
"""
import contextlib
import io
import os
import re
import sys
import tempfile

import numpy as np
import pandas as pd
import torch

from . import ref_import, restate

GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
WHEAT_CSV = os.path.join(ref_import.REFERENCE_ROOT, "csv-data", "WHEAT_combined.csv")


def wheat_frame():
    with open(WHEAT_CSV) as fp:
        ncol = len(fp.readline().split(","))
    return pd.read_csv(WHEAT_CSV, names=["value"] + [str(i + 1) for i in range(ncol - 1)])   # load.py:182-187


def save(name, **arrays):
    path = os.path.join(GOLDEN, name)
    np.savez_compressed(path, **arrays)
    print(f"wrote {path}  ({os.path.getsize(path) / 1024:.0f} KiB)")


def check_against_sklearn(X, k, idx, d2):
    """Canonical lists vs sklearn proper: distances bit-equal; sets equal where no tie straddles rank k."""
    from sklearn.neighbors import kneighbors_graph
    g = kneighbors_graph(X, k, mode="distance", include_self=False).tocsr()
    n = X.shape[0]
    sk_d = g.data.reshape(n, k)
    sk_i = g.indices.reshape(n, k)
    assert np.array_equal(np.sort(sk_d, axis=1), np.sqrt(d2)), "sorted distance vectors differ from sklearn"
    kk_idx, kk_d2 = restate.canonical_knn(X, min(k + 1, n - 1))
    tie_free = kk_d2[:, k - 1] != kk_d2[:, -1] if kk_d2.shape[1] > k else np.ones(n, bool)
    same = np.array([set(a) == set(b) for a, b in zip(sk_i, idx)])
    assert np.all(same[tie_free]), "neighbour sets differ from sklearn on tie-free rows"
    return int(tie_free.sum()), int(same.sum())


def digest(a):
    """Bit-exact fingerprint of an array (dtype, shape and bytes)."""
    import hashlib
    a = np.ascontiguousarray(a)
    h = hashlib.sha256(f"{a.dtype.str}{a.shape}".encode() + a.tobytes()).digest()[:8]
    return np.frombuffer(h, dtype=np.uint64)[0]


def features_and_ensemble(ref, df):
    """wheat_features.npz and wheat_ensemble.npz."""
    X = df.iloc[:, 1:].values
    y = df["value"].values
    out = {}

    def case(tag, rows, k, smoothing, mode, frame=None, int_labels=False, **kw):
        sub = (df if frame is None else frame).iloc[rows].copy()
        Xs = sub.iloc[:, 1:].values
        d = ref.load.create_graph(sub, n_neighbors=k, smoothing=smoothing, use_weights=True, mode=mode,
                                  algorithm="euclidean", **kw)
        mine = restate.create_graph(Xs, y[rows], n_neighbors=k, smoothing=smoothing, use_weights=True, mode=mode,
                                    int_labels=int_labels, **kw)
        assert np.array_equal(d.edge_index.numpy(), mine["edge_index"]), tag
        assert np.array_equal(d.edge_weight.numpy(), mine["edge_weight"]), tag
        assert np.array_equal(d.x.numpy(), mine["x"]) and np.array_equal(d.y.numpy(), mine["y"]), tag
        out[tag + "_ei"] = mine["edge_index"].astype(np.int32)
        out[tag + "_ew"] = mine["edge_weight"]
        out[tag + "_xdigest"] = digest(mine["x"])
        return mine

    train, test = slice(0, 479), slice(479, 587)
    for mode in ("distance", "connectivity"):
        for smoothing in ("laplacian", "none"):
            m = case(f"scaled_train_k5_{mode}_{smoothing}", train, 5, smoothing, mode, scaled=True)
            case(f"self_train_k5_{mode}_{smoothing}", train, 5, smoothing, mode, include_self=True)
            case(f"scaledself_test_k3_{mode}_{smoothing}", test, 3, smoothing, mode, scaled=True, include_self=True)
    out["scaled_train_x"] = m["x"]
    out["scaled_train_idx"] = m["knn_idx"].astype(np.int32)
    out["scaled_train_d2"] = m["knn_d2"]
    # constant columns: std = 0 -> NaN -> 0 (load.py:48); duplicates under include_self
    Xd = X[:80].copy()
    Xd[:, 5] = 1.0
    Xd[:, 17] = 0.0
    Xd[9] = Xd[2]
    Xd[30] = Xd[2]
    fr = pd.DataFrame(np.column_stack([y[:80], Xd]), columns=list(df.columns))
    out["const_cols"] = np.array([5, 17], dtype=np.int32)
    out["const_dup_rows"] = np.array([2, 9, 30], dtype=np.int32)
    case("const_k4_distance_laplacian", slice(0, 80), 4, "laplacian", "distance", frame=fr, scaled=True, include_self=True)
    case("const_k4_connectivity_laplacian", slice(0, 80), 4, "laplacian", "connectivity", frame=fr, include_self=True)
    # hot needs the integer labels that ``bits`` installs (load.py:189-191)
    bits = [str(i) for i in range(3, 120, 2)]
    fb = df[["value"] + bits].copy()
    fb.columns = ["value"] + list(range(1, len(bits) + 1))
    out["hot_bits"] = np.asarray([int(b) for b in bits], dtype=np.int32)
    mh = case("hot_train_k5_distance_laplacian", train, 5, "laplacian", "distance", frame=fb, int_labels=True, hot=True)
    out["hot_train_x"] = mh["x"].astype(np.int8)
    case("hotscaled_test_k3_distance_none", test, 3, "none", "distance", frame=fb, int_labels=True, hot=True, scaled=True)
    try:
        ref.load.create_graph(df.iloc[train].copy(), n_neighbors=3, hot=True)
        raise AssertionError("hot=True on string labels was expected to fail")
    except KeyError as e:
        assert e.args == (1,)
    save("wheat_features.npz", **out)

    # ---- ensemble loader -----------------------------------------------------------------------
    ens = {}
    cases = {
        "a": dict(seed=11, num_neighbours=3, split=3, use_weights=True, use_validation=True, mode="distance"),
        "b": dict(seed=5, num_neighbours=2, split=2, use_weights=True, scaled=False, add_full=True),
        "c": dict(seed=5, num_neighbours=2, split=2, use_weights=True, scaled=False, add_full=True, full_split=0.8),
        "d": dict(seed=7, num_neighbours=2, split=2, use_weights=False, separate_sets=False, train_size=300),
        "e": dict(seed=7, num_neighbours=2, split=2, use_weights=True, separate_sets=False, full_split=400,
                  train_size=0.6, include_self=True),
    }
    cwd = os.getcwd()
    for name, kw in cases.items():
        kw = dict(kw)
        seed = kw.pop("seed")
        with tempfile.TemporaryDirectory() as tmp:
            os.chdir(tmp)
            try:
                np.random.seed(seed)
                with contextlib.redirect_stdout(io.StringIO()):
                    ds = ref.load_ensembles2.load_data_ensembles2(WHEAT_CSV, **kw)
            finally:
                os.chdir(cwd)
        np.random.seed(seed)
        mine = restate.ensemble_graphs(X, y, **kw)
        np.random.seed(seed)
        rows, valid = restate.ensemble_rows(len(y), kw["split"], kw.get("train_size", 0.7), kw.get("add_full", False),
                                            kw.get("full_split"), kw.get("use_validation", False), 0.1,
                                            kw.get("separate_sets", True))
        assert len(ds) == len(mine)
        ens[f"{name}_n"] = np.int64(len(ds))
        ens[f"{name}_seed"] = np.int64(seed)
        for i in range(len(ds)):
            for sub, a, b in (("", ds[i], mine[i]), ("_test", ds[i].test, mine[i]["test"])):
                assert np.array_equal(a.x.numpy(), b["x"]) and np.array_equal(a.y.numpy(), b["y"]), (name, i, sub)
                assert np.array_equal(a.edge_index.numpy(), b["edge_index"]), (name, i, sub)
                if kw.get("use_weights"):
                    assert np.array_equal(a.edge_weight.numpy(), b["edge_weight"]), (name, i, sub)
                    ens[f"{name}_{i}{sub}_ew"] = b["edge_weight"]
                ens[f"{name}_{i}{sub}_ei"] = b["edge_index"].astype(np.int32)
                ens[f"{name}_{i}{sub}_y"] = b["y"]
                ens[f"{name}_{i}{sub}_xdigest"] = digest(b["x"])
        for i, (tr, te) in enumerate(rows):
            ens[f"{name}_{i}_rows"] = tr.astype(np.int32)
            ens[f"{name}_{i}_test_rows"] = te.astype(np.int32)
        if valid is not None:
            ens[f"{name}_valid_rows"] = valid.astype(np.int32)
    try:
        with tempfile.TemporaryDirectory() as tmp:
            os.chdir(tmp)
            try:
                with contextlib.redirect_stdout(io.StringIO()):
                    ref.load_ensembles2.load_data_ensembles2(WHEAT_CSV, split=1, hot=True, scaled=False)
            finally:
                os.chdir(cwd)
        raise AssertionError("hot=True was expected to fail in the ensemble loader")
    except KeyError as e:
        ens["hot_keyerror"] = np.int64(e.args[0])
    save("wheat_ensemble.npz", **ens)


def trainers_golden(ref, df):
    """wheat_trainers.npz: trajectories of the reference's ``train`` (plain.py:8-114) and ``train_batches``
    (batches.py:7-81) through its own loader, and the distance-correlation values ``train`` reports."""
    X = df.iloc[:, 1:].values
    y = df["value"].values
    H, EPOCHS, LR_ = 32, 5, 0.0020594745443455593
    out = {"hidden": np.int64(H), "epochs": np.int64(EPOCHS), "lr": np.float64(LR_)}
    cwd = os.getcwd()

    def opt(model):
        return torch.optim.Adam(model.parameters(), lr=LR_)

    with tempfile.TemporaryDirectory() as tmp:
        os.chdir(tmp)
        try:
            os.makedirs("output")
            # ---- train: one training graph + test graph (batches=None), weights used (plain.py:38-41)
            params = dict(num_neighbours=5, smoothing="laplacian", mode="distance", use_validation=False, split=0.8,
                          use_weights=True, algorithm="euclidean")
            with contextlib.redirect_stdout(io.StringIO()):
                ds = ref.load.load_data(WHEAT_CSV, **params)
            data = ds[0]
            torch.manual_seed(0)
            m_ref = ref.networks.create_network_two_no_conv_dropout(X.shape[1], 1, internal_size=H, dropout=0.0)
            w0 = {k: v.clone().numpy() for k, v in m_ref.state_dict().items()}
            buf = io.StringIO()
            with contextlib.redirect_stdout(buf):
                res = ref.plain.train(m_ref, opt(m_ref), data, epochs=EPOCHS, save_loss=False)["basics"]
            printed = re.findall(r"Epoch: (\d+), Loss: ([-\d.]+), Test Loss: ([-\d.]+)", buf.getvalue())
            assert len(printed) == EPOCHS
            n = len(df)
            tr_rows, te_rows = np.arange(0, int(n * 0.8)), np.arange(int(n * 0.8), n)      # split_dataset (util.py:76-84)
            g = restate.create_graph(X[tr_rows], y[tr_rows], 5, "laplacian", True, True, "distance")
            g_test = restate.create_graph(X[te_rows], y[te_rows], 5, "laplacian", True, True, "distance")
            assert np.array_equal(data.edge_index.numpy(), g["edge_index"])
            assert np.array_equal(data.test.edge_index.numpy(), g_test["edge_index"])
            torch.manual_seed(0)
            m = restate.make_model("create_network_two_no_conv_dropout", X.shape[1], 1, internal_size=H, dropout=0.0)
            tl, vl, cors, dcors, out_test = restate.train_plain(m, opt(m), g, g_test, EPOCHS)
            np.testing.assert_allclose(vl, res["losses"], rtol=1e-5)
            np.testing.assert_allclose(cors, res["cors"], rtol=1e-6, atol=1e-7)
            np.testing.assert_allclose(dcors, res["dcors"], rtol=1e-6, atol=1e-7)
            for ep in range(EPOCHS):
                assert abs(tl[ep] - float(printed[ep][1])) < 2e-5
            out.update(plain_train_rows=tr_rows.astype(np.int32), plain_test_rows=te_rows.astype(np.int32),
                       plain_train_loss=np.asarray(tl), plain_test_loss=np.asarray(res["losses"], dtype=np.float64),
                       plain_cors=np.asarray(res["cors"], dtype=np.float64), plain_dcors=np.asarray(res["dcors"], dtype=np.float64),
                       plain_pred_test=out_test.numpy())
            for kk, v in w0.items():
                out[f"plain__w0__{kk}"] = v

            # ---- train_batches: three partitioned subgraphs, one step each per epoch, NO edge weights (batches.py:37)
            params = dict(num_neighbours=5, smoothing="laplacian", mode="distance", use_validation=False, split=0.8,
                          use_weights=True, algorithm="euclidean", batches=3,
                          split_algorithm=ref.util.split_dataset_graph,
                          split_algorithm_params={"partition": ref.util.naive_partition})
            with contextlib.redirect_stdout(io.StringIO()):
                ds = ref.load.load_data(WHEAT_CSV, **params)
            data = [d for d in ds]
            assert len(data) == 3
            torch.manual_seed(1)
            m_ref = ref.networks.create_network_two_no_conv_dropout(X.shape[1], 1, internal_size=H, dropout=0.0)
            w0 = {k: v.clone().numpy() for k, v in m_ref.state_dict().items()}
            buf = io.StringIO()
            with contextlib.redirect_stdout(buf):
                losses, last_epoch = ref.batches.train_batches(m_ref, opt(m_ref), data, epochs=EPOCHS, save_loss=False)
            printed = re.findall(r"Epoch: (\d+):(\d+), Loss: ([-\d.]+), Test Loss: ([-\d.]+)", buf.getvalue())
            assert len(printed) == 3 * EPOCHS and last_epoch == EPOCHS - 1
            sizes = restate.split_sizes(n, 0.8, False, 0.1)
            _, te, _ = restate.split_indices_graph(df, sizes, neighbours=5)
            part = restate.split_indices_graph(df, 3, neighbours=5)
            g_test = restate.create_graph(X[te], y[te], 5, "laplacian", True, True, "distance")
            graphs = []
            for b, rows in enumerate(part):
                gb = restate.create_graph(X[rows], y[rows], 5, "laplacian", True, True, "distance")
                assert np.array_equal(data[b].edge_index.numpy(), gb["edge_index"])
                gb["test"] = g_test
                graphs.append(gb)
                out[f"batches_rows_{b}"] = np.asarray(rows, dtype=np.int32)
            assert np.array_equal(data[0].test.edge_index.numpy(), g_test["edge_index"])
            torch.manual_seed(1)
            m = restate.make_model("create_network_two_no_conv_dropout", X.shape[1], 1, internal_size=H, dropout=0.0)
            r_losses, r_train, steps, out_test = restate.train_batches(m, opt(m), graphs, EPOCHS)
            np.testing.assert_allclose(r_losses, losses, rtol=1e-5)
            for i, (a, c) in enumerate(steps):
                assert abs(a - float(printed[i][2])) < 2e-5 and abs(c - float(printed[i][3])) < 2e-5, (i, a, c, printed[i])
            out.update(batches_test_rows=np.asarray(te, dtype=np.int32), batches_losses=np.asarray(losses, dtype=np.float64),
                       batches_train_losses=np.asarray(r_train), batches_steps=np.asarray(steps),
                       batches_pred_test=out_test.numpy())
            for kk, v in w0.items():
                out[f"batches__w0__{kk}"] = v
        finally:
            os.chdir(cwd)
    # distance correlation on its own: a few vectors with known structure
    rng = np.random.default_rng(3)
    a = rng.standard_normal(200)
    cases = {"linear": (a, 2.0 * a + 1.0), "noise": (a, rng.standard_normal(200)), "square": (a, a * a),
             "const": (a, np.ones(200))}
    for name, (u, v) in cases.items():
        out[f"dcor_{name}_a"], out[f"dcor_{name}_b"] = u, v
        out[f"dcor_{name}"] = np.float64(restate.distance_correlation(u, v))
    assert abs(out["dcor_linear"] - 1.0) < 1e-12 and out["dcor_const"] == 0.0
    save("wheat_trainers.npz", **out)


def main(argv=None):
    argv = sys.argv[1:] if argv is None else argv
    os.makedirs(GOLDEN, exist_ok=True)
    ref = ref_import.load(canonical_knn=True)
    df = wheat_frame()
    if argv and argv[0] == "trainers":
        trainers_golden(ref, df)
        return 0
    features_and_ensemble(ref, df)
    if argv and argv[0] == "features":
        return 0
    trainers_golden(ref, df)
    N = len(df)
    X = df.iloc[:, 1:].values
    y = df["value"].values
    assert set(np.unique(X)) <= {0.0, 1.0}
    save("wheat.npz", x_bits=np.packbits(X.astype(np.uint8), axis=1), n=N, p=X.shape[1], y=y)

    sizes = restate.split_sizes(N, 0.8, True, 0.1)
    assert sizes == (479, 108, 587)

    # ---- kNN ---------------------------------------------------------------------------------
    idx_all, d2_all = restate.canonical_knn(X, 32)
    idx_tr, d2_tr = restate.canonical_knn(X[:479], 32)
    for k in (1, 3, 5, 8, 10, 32):
        a, b = restate.canonical_knn(X, k)
        assert np.array_equal(a, idx_all[:, :k]) and np.array_equal(b, d2_all[:, :k])
        free, same = check_against_sklearn(X, k, a, b)
        print(f"k={k}: tie-free rows {free}/{N}, rows with sklearn's neighbour set {same}/{N}")
    # the partition graph includes the phenotype column (util.py:52-53)
    idx_full, d2_full = restate.canonical_knn(X, 10, extra=y[:, None])
    save("wheat_knn.npz", idx_all=idx_all.astype(np.int32), d2_all=d2_all.astype(np.int32),
         idx_train=idx_tr.astype(np.int32), d2_train=d2_tr.astype(np.int32),
         idx_with_pheno=idx_full.astype(np.int32), d2_with_pheno=d2_full)

    # ---- create_graph: reference vs restatement ----------------------------------------------
    graphs = {}
    for rows, tag in ((slice(0, 479), "train"), (slice(479, 587), "test")):
        for k in (3, 5, 10):
            for mode in ("distance", "connectivity"):
                for smoothing in ("laplacian", "none"):
                    sub = df.iloc[rows].copy()
                    d = ref.load.create_graph(sub, n_neighbors=k, smoothing=smoothing, use_weights=True,
                                              mode=mode, algorithm="euclidean")
                    mine = restate.create_graph(X[rows], y[rows], n_neighbors=k, smoothing=smoothing,
                                                use_weights=True, mode=mode)
                    ei, ew = d.edge_index.numpy(), d.edge_weight.numpy()
                    assert np.array_equal(ei, mine["edge_index"]), (tag, k, mode, smoothing)
                    assert np.array_equal(ew, mine["edge_weight"]), (tag, k, mode, smoothing, np.abs(ew - mine["edge_weight"]).max())
                    assert np.array_equal(d.x.numpy(), mine["x"]) and np.array_equal(d.y.numpy(), mine["y"])
                    key = f"{tag}_k{k}_{mode}_{smoothing}"
                    graphs[key + "_ei"] = ei.astype(np.int32)
                    graphs[key + "_ew"] = ew
    # duplicates: zero distances are dropped by .nonzero() (load.py:64)
    Xd = X[:60].copy()
    Xd[7] = Xd[3]
    Xd[11] = Xd[3]
    Xd[40] = Xd[41]
    yd = y[:60]
    sub = pd.DataFrame(np.column_stack([yd, Xd]), columns=["value"] + [str(i + 1) for i in range(Xd.shape[1])])
    d = ref.load.create_graph(sub.copy(), n_neighbors=2, smoothing="laplacian", use_weights=True, mode="distance",
                              algorithm="euclidean")
    mine = restate.create_graph(Xd, yd, n_neighbors=2, smoothing="laplacian", use_weights=True, mode="distance")
    assert np.array_equal(d.edge_index.numpy(), mine["edge_index"])
    assert np.array_equal(d.edge_weight.numpy(), mine["edge_weight"])
    graphs["dup_rows"] = np.array([3, 7, 11, 40, 41], dtype=np.int32)
    graphs["dup_k2_distance_laplacian_ei"] = mine["edge_index"].astype(np.int32)
    graphs["dup_k2_distance_laplacian_ew"] = mine["edge_weight"]
    save("wheat_graph.npz", **graphs)

    # ---- partition ---------------------------------------------------------------------------
    parts = {}
    for k in (3, 5, 10):
        for batches in (3, 5, 8, 10, sizes):
            got = ref.util.split_dataset_graph(df, batches, neighbours=k, metric="euclidean",
                                               partition=ref.util.naive_partition)
            want = [list(map(int, g.index.values)) for g in got]
            mine = restate.split_indices_graph(df, batches, neighbours=k)
            assert [list(map(int, m)) for m in mine] == want, (k, batches)
            tag = "split" if isinstance(batches, tuple) else f"b{batches}"
            for i, w in enumerate(want):
                parts[f"k{k}_{tag}_{i}"] = np.asarray(w, dtype=np.int32)
    save("wheat_partition.npz", **parts)

    rng = np.random.default_rng(7)
    small = {}
    for case in range(40):
        n = int(rng.integers(3, 30))
        k = int(rng.integers(1, min(4, n - 1) + 1))
        nbr = np.stack([rng.choice(np.delete(np.arange(n), i), size=k, replace=False) for i in range(n)])
        if case % 4 == 0:      # a few nodes that nobody points to, several pointing at node 0
            nbr[:, 0] = np.where(np.arange(n) == 0, 1, 0)
        S = restate.neighbour_stream(nbr)
        edges = [(i, int(c)) for i in range(n) for c in nbr[i]]
        szs = [int(s) for s in rng.integers(1, max(2, n // 2) + 1, size=int(rng.integers(1, 5)))]
        trav, want = set(), []
        for s in szs:
            got = [int(v) for v in ref.util.naive_partition(edges, s, bidirectional=False, traversed=set(trav))]
            trav |= set(got)
            want.append(got)
        assert restate.partition_linear(S, n, szs) == want, (case, want, restate.partition_linear(S, n, szs))
        t2, slow = set(), []
        for s in szs:
            got = restate.partition_scan(S, n, s, t2)
            t2 |= set(got)
            slow.append(got)
        assert slow == want, (case, slow, want)
        small[f"c{case}_nbr"] = nbr.astype(np.int32)
        small[f"c{case}_sizes"] = np.asarray(szs, dtype=np.int32)
        small[f"c{case}_flat"] = np.asarray([v for b in want for v in b], dtype=np.int32)
        small[f"c{case}_lens"] = np.asarray([len(b) for b in want], dtype=np.int32)
    save("partition_small.npz", **small)

    # ---- model forward / backward ------------------------------------------------------------
    H = 32
    train_g = restate.create_graph(X[:479], y[:479], n_neighbors=5, smoothing="laplacian", use_weights=True,
                                   mode="distance")
    model_out = {}
    for name in ("create_network_two_no_conv_dropout", "create_network_no_conv_relu_dropout",
                 "create_network_two_relu", "create_network_no_conv"):
        for asl in (True, False):
            if name != "create_network_two_no_conv_dropout" and not asl:
                continue
            torch.manual_seed(0)
            kw = {"internal_size": H}
            if "dropout" in name:
                kw["dropout"] = 0.0
            if name in ("create_network_two_no_conv_dropout", "create_network_two_relu", "create_network_no_conv"):
                kw["add_self_loops"] = asl
            m_ref = getattr(ref.networks, name)(X.shape[1], 1, **kw)
            torch.manual_seed(0)
            m_mine = restate.make_model(name, X.shape[1], 1, internal_size=H, dropout=0.0, add_self_loops=asl)
            sd_ref, sd_mine = m_ref.state_dict(), m_mine.state_dict()
            assert list(sd_ref.keys()) == list(sd_mine.keys()), (list(sd_ref.keys()), list(sd_mine.keys()))
            for kk in sd_ref:
                assert torch.equal(sd_ref[kk], sd_mine[kk]), kk
            args = restate.as_tensors(train_g)
            yt = torch.as_tensor(train_g["y"])
            outs = []
            for m in (m_ref, m_mine):
                m.train()
                m.zero_grad()
                out = m(*args)
                loss = torch.nn.MSELoss()(out, yt)
                loss.backward()
                outs.append((out.detach(), float(loss), {k: p.grad.clone() for k, p in m.named_parameters()}))
            assert torch.equal(outs[0][0], outs[1][0]) and outs[0][1] == outs[1][1]
            for kk in outs[0][2]:   # torch-CPU index_add_/gather backward is not run-to-run bit-stable
                assert torch.allclose(outs[0][2][kk], outs[1][2][kk], rtol=1e-5, atol=1e-7), kk
            # the tuple call form must give the same result (sequential.py:100-101)
            assert torch.equal(m_ref(args), outs[0][0])
            tag = f"{name}_asl{int(asl)}"
            for kk, v in sd_ref.items():
                model_out[f"{tag}__w__{kk}"] = v.numpy()
            model_out[f"{tag}__out"] = outs[0][0].numpy()
            model_out[f"{tag}__loss"] = np.float64(outs[0][1])
            for kk, v in outs[0][2].items():
                model_out[f"{tag}__g__{kk}"] = v.numpy()
    save("wheat_model.npz", **model_out)

    # ---- training trajectories through the reference's loader + trainer ------------------------
    cwd = os.getcwd()
    traj = {}
    with tempfile.TemporaryDirectory() as tmp:
        os.chdir(tmp)
        try:
            os.makedirs("output")
            os.makedirs("model")
            params = dict(num_neighbours=5, smoothing="laplacian", mode="distance", use_validation=True,
                          split=0.8, use_weights=True, algorithm="euclidean", batches=3,
                          split_algorithm=ref.util.split_dataset_graph,
                          split_algorithm_params={"partition": ref.util.naive_partition})
            with contextlib.redirect_stdout(io.StringIO()):
                ds = ref.load.load_data(WHEAT_CSV, **params)
            data = [g for g in ds]
            assert len(data) == 4
            EPOCHS = 6

            def creator():
                return ref.networks.create_network_two_no_conv_dropout(X.shape[1], 1, internal_size=H, dropout=0.0)

            def opt(model):
                return torch.optim.Adam(model.parameters(), lr=0.0020594745443455593)

            torch.manual_seed(0)
            buf = io.StringIO()
            with contextlib.redirect_stdout(buf):
                res = ref.ensemble2.train_ensemble2(creator, opt, data, epochs=EPOCHS, aggregate_epochs=2,
                                                    save_loss=False, print_epochs=True)
            printed = re.findall(r"Batch: (\d+), Epoch: (\d+), Loss: ([-\d.]+), Test Loss: ([-\d.]+)", buf.getvalue())
            assert len(printed) == 3 * EPOCHS

            # restatement: same seed, same order of model creation
            part = restate.split_indices_graph(df, 3, neighbours=5)
            tr, te, va = restate.split_indices_graph(df, sizes, neighbours=5)
            g_test = restate.create_graph(X[te], y[te], 5, "laplacian", True, True, "distance")
            g_valid = restate.create_graph(X[va], y[va], 5, "laplacian", True, True, "distance")
            assert np.array_equal(data[0].test.edge_index.numpy(), g_test["edge_index"])
            assert np.array_equal(data[-1].edge_index.numpy(), g_valid["edge_index"])
            torch.manual_seed(0)
            for b, rows in enumerate(part):
                g = restate.create_graph(X[rows], y[rows], 5, "laplacian", True, True, "distance")
                assert np.array_equal(data[b].edge_index.numpy(), g["edge_index"])
                assert np.array_equal(data[b].edge_weight.numpy(), g["edge_weight"])
                assert np.array_equal(data[b].y.numpy(), g["y"])
                m = restate.make_model("create_network_two_no_conv_dropout", X.shape[1], 1, internal_size=H, dropout=0.0)
                w0 = {k: v.clone().numpy() for k, v in m.state_dict().items()}
                o = opt(m)
                tl, vl = [], []
                for ep in range(EPOCHS):
                    a, c, out_test = restate.epoch_step(m, o, g, g_test)
                    tl.append(a)
                    vl.append(c)
                    pa, pc = printed[b * EPOCHS + ep][2:]
                    assert abs(a - float(pa)) < 2e-5 and abs(c - float(pc)) < 2e-5, (b, ep, a, pa, c, pc)
                assert abs(min(vl) - res["basics"][f"min_loss_{b}"]) < 1e-6 * abs(min(vl))
                traj[f"b{b}_rows"] = np.asarray(rows, dtype=np.int32)
                traj[f"b{b}_train_loss"] = np.asarray(tl)
                traj[f"b{b}_test_loss"] = np.asarray(vl)
                traj[f"b{b}_pred_test"] = out_test.numpy()
                for kk, v in w0.items():
                    traj[f"b{b}__w0__{kk}"] = v
                traj[f"b{b}_pred_train"] = m.eval()(*restate.as_tensors(g)).detach().numpy()
            for key, val in res["basics"].items():          # final metrics of ensemble2.py:214-256
                if isinstance(val, (int, float, np.floating, np.integer)):
                    traj[f"ens_{key}"] = np.float64(val)
            traj["ens_cors"] = np.asarray(res["basics"]["cors"], dtype=np.float64)
            traj["ens_losses"] = np.asarray(res["basics"]["losses"], dtype=np.float64)
            traj["test_rows"] = np.asarray(te, dtype=np.int32)
            traj["valid_rows"] = np.asarray(va, dtype=np.int32)
            traj["lr"] = np.float64(0.0020594745443455593)
            traj["epochs"] = np.int64(EPOCHS)
            traj["hidden"] = np.int64(H)
        finally:
            os.chdir(cwd)
    save("wheat_train.npz", **traj)
    print("golden fixtures regenerated; restatement == reference on every case")


if __name__ == "__main__":
    sys.exit(main())
