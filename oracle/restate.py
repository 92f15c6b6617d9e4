"""numpy / torch-CPU restatement of the reference hot path -- TEST INFRASTRUCTURE.

See oracle/__init__.py for the rules.  Each function cites the reference lines
it follows (paths relative to /root/reference).  ``oracle/make_golden.py`` runs
the reference itself next to these functions and asserts equality before it
freezes tests/golden/*.npz, so the restatement is pinned to the reference's own
output, not to my reading of it.

Canonical neighbour order: the k nearest of row i are the k
smallest pairs (d2_ij, j), j != i, in lexicographic order.  d2 is evaluated
exactly: the integer-coded columns contribute an exact integer, every other
column c contributes fl(fl(x_ic - x_jc)^2) accumulated left to right in
float64 on top of it.
"""
import math

import numpy as np
import scipy.sparse as sp
import torch

from .pyg_shim import GCNConv


# ----------------------------------------------------------------------------- kNN
MAX_SIDE_COLUMNS = 4      # integer-coded markers plus at most this many float columns (the phenotype) use the exact path


def _split_columns(X):
    """Integer-coded columns (exact path) vs the rest (float64 path)."""
    X = np.asarray(X, dtype=np.float64)
    is_int = np.all((X == np.rint(X)) & (np.abs(X) <= 127), axis=0)
    return X[:, is_int], X[:, ~is_int]


def canonical_knn(X_int, k, extra=None, block=1024, include_self=False):
    """k nearest neighbours of every row under squared Euclidean distance with
    the canonical (d2, j) order.  Stands in for the sklearn call at
    gnn/loaders/load.py:52-59 and gnn/loaders/util.py:52-57.

    X_int : [n, p] integer-valued array (any dtype), exact path
    extra : optional [n, m] float64 columns added as sum_c (x_ic - x_jc)^2
    returns idx int64 [n, k], d2 float64 [n, k]
    """
    Xf = np.ascontiguousarray(X_int, dtype=np.float64)      # products/sums of small ints are exact in f64
    n = Xf.shape[0]
    if include_self:                                            # load.py:28,57: the row competes like any other
        if not 0 < k <= n:
            raise ValueError(f"Expected n_neighbors <= n_samples_fit, but n_neighbors = {k}, n_samples_fit = {n}")
    elif not 0 < k < n:
        raise ValueError(f"Expected n_neighbors <= n_samples_fit, but n_neighbors = {k + 1}, n_samples_fit = {n}")
    sq = np.einsum("ij,ij->i", Xf, Xf)
    idx = np.empty((n, k), dtype=np.int64)
    d2o = np.empty((n, k), dtype=np.float64)
    if extra is not None:
        extra = np.asarray(extra, dtype=np.float64).reshape(n, -1)
    for r0 in range(0, n, block):
        r1 = min(n, r0 + block)
        d2 = sq[r0:r1, None] + sq[None, :] - 2.0 * (Xf[r0:r1] @ Xf.T)   # exact integers
        if extra is not None:
            for c in range(extra.shape[1]):
                diff = extra[r0:r1, c][:, None] - extra[None, :, c]
                d2 = d2 + diff * diff
        if not include_self:
            d2[np.arange(r1 - r0), np.arange(r0, r1)] = np.inf          # never the row itself
        order = np.argsort(d2, axis=1, kind="stable")[:, :k]             # stable => ties by ascending j
        idx[r0:r1] = order
        d2o[r0:r1] = np.take_along_axis(d2, order, axis=1)
    return idx, d2o


def canonical_kneighbors_graph(X, n_neighbors, mode="connectivity", metric="minkowski",
                               p=2, metric_params=None, include_self=False, n_jobs=None):
    """Drop-in for ``sklearn.neighbors.kneighbors_graph`` as the reference calls it
    (load.py:52-59; util.py:52-57): CSR with exactly k entries per row in
    ascending (distance, index) order; data = distances or ones."""
    if metric not in ("minkowski", "euclidean", "euclidian", "l2") or p != 2:
        raise NotImplementedError("oracle covers the Euclidean call of the reference")
    A = X.values if hasattr(X, "values") else np.asarray(X)
    xi, xf = _split_columns(A)
    if xf.shape[1] > MAX_SIDE_COLUMNS:      # a float feature matrix (scaled=True): every column, left to right, in float64
        xi, xf = np.zeros((A.shape[0], 0)), np.asarray(A, dtype=np.float64)
    idx, d2 = canonical_knn(xi, n_neighbors, extra=xf if xf.shape[1] else None, include_self=include_self)
    n = idx.shape[0]
    data = np.sqrt(d2).ravel() if mode == "distance" else np.ones(n * n_neighbors)
    indptr = np.arange(0, n * n_neighbors + 1, n_neighbors)
    return sp.csr_matrix((data, idx.ravel().astype(np.int32), indptr.astype(np.int32)), shape=(n, n))


# ----------------------------------------------------------------------------- graph build
def graph_edges(idx, d2, mode="connectivity", smoothing="laplacian", use_weights=False):
    """Edge list the loader hands to the model, from canonical neighbour lists.

    Follows gnn/loaders/load.py:61-80.
    laplacian: zero-distance neighbours vanish (``.nonzero()``, :64), distances are
    divided by the largest one (:65), s = exp(-(d/dmax)^2/2) (:67), W = (S+S^T)/2
    (:68), L = D - W as COO with the off-diagonal block in (row, col) order and the
    n diagonal entries appended (:69-70); weights 1 - |L| / float32(max L) (:74-75).
    otherwise: the directed kNN CSR itself in row-major order (:77-80).
    Returns (edge_index int64 [2, E], edge_weight float64 [E] or None).
    """
    n, k = idx.shape
    dist = np.sqrt(d2) if mode == "distance" else np.ones_like(d2)
    rows = np.repeat(np.arange(n, dtype=np.int64), k)
    cols = idx.ravel().astype(np.int64)
    vals = dist.ravel()
    if smoothing != "laplacian":
        edge_index = np.stack([rows, cols])
        if not use_weights:
            return edge_index, None
        mx = np.float64(np.float32(vals.max()))
        return edge_index, 1.0 - np.abs(vals) / mx
    keep = vals != 0
    rows, cols, vals = rows[keep], cols[keep], vals[keep]
    normalized = vals / vals.max()
    s = np.exp(-normalized ** 2 / 2.0 * 1 ** 2)
    S = sp.csr_matrix((s, (rows, cols)), shape=(n, n))
    W = (0.5 * (S + S.T)).tocsr()
    W.sort_indices()
    deg = np.asarray(W.sum(axis=0)).ravel() - W.diagonal()
    Wc = W.tocoo()
    off = Wc.row != Wc.col                  # include_self leaves W[i, i]; the Laplacian's setdiag replaces it
    e_row = np.concatenate([Wc.row[off].astype(np.int64), np.arange(n, dtype=np.int64)])
    e_col = np.concatenate([Wc.col[off].astype(np.int64), np.arange(n, dtype=np.int64)])
    e_val = np.concatenate([-Wc.data[off], deg])
    edge_index = np.stack([e_row, e_col])
    if not use_weights:
        return edge_index, None
    mx = np.float64(np.float32(e_val.max()))
    return edge_index, 1.0 - np.abs(e_val) / mx


def hot_features(X, int_labels):
    """load.py:38-43 as it executes: the loops address columns by the integer labels 1..P-1, which exist only after
    ``bits`` renamed them (load.py:189-191; KeyError otherwise).  Loop 1 overwrites columns 1..P-1 with ``== 1``; loop 2
    compares the overwritten 0/1 columns with 2, so column P and the P-2 appended columns are zero."""
    if not int_labels:
        raise KeyError(1)
    X = np.asarray(X)
    n, P = X.shape
    if P < 2:
        return X
    return np.concatenate([(X[:, :P - 1] == 1).astype(np.int64), np.zeros((n, P - 1), dtype=np.int64)], axis=1)


def scale_features(X):
    """load.py:45-48 with the same pandas calls (mean, std with ddof 1, fillna)."""
    import pandas as pd
    df = pd.DataFrame(np.asarray(X, dtype=np.float64))
    df -= df.mean()
    df /= df.std()
    return df.fillna(0).values


def create_graph(X, y, n_neighbors=1, smoothing="laplacian", use_weights=False, remove_mean=True,
                 mode="connectivity", hot=False, scaled=False, include_self=False, int_labels=False):
    """gnn/loaders/load.py:21-81:
    x float32 [n, p], y float32 [n, 1] (centred in float64 first, :33-36, :50-51),
    edge_index / edge_weight from ``graph_edges``."""
    X = np.asarray(X)
    y = np.asarray(y, dtype=np.float64).reshape(-1)
    if remove_mean:
        y = y - y.mean()
    if hot:
        X = hot_features(X, int_labels)
    if scaled:
        X = scale_features(X)
        idx, d2 = canonical_knn(np.zeros((X.shape[0], 0)), n_neighbors, extra=X, include_self=include_self)
    else:
        idx, d2 = canonical_knn(X, n_neighbors, include_self=include_self)
    edge_index, edge_weight = graph_edges(idx, d2, mode=mode, smoothing=smoothing, use_weights=use_weights)
    return {
        "x": X.astype(np.float32), "y": y.astype(np.float32).reshape(-1, 1),
        "edge_index": edge_index, "edge_weight": edge_weight, "knn_idx": idx, "knn_d2": d2,
    }


# ----------------------------------------------------------------------------- seeded subsampling (ensemble loader)
def ensemble_rows(n_rows, split, train_size=0.7, add_full=False, full_split=None, use_validation=False,
                  validation_size=0.1, separate_sets=True):
    """Which file rows every graph of load_ensembles2.py:74-132 is made of, as positions 0..n_rows-1.

    Follows the frame manipulations literally on a one-column frame holding the row position:
    validation rows are the FIRST ``valid_size`` rows of the already truncated frame (:82-83), the unseeded
    ``df_whole.sample(frac)`` (:99) draws from numpy's global generator, and ``whole_test`` (:100) is chosen by index
    LABEL against the re-numbered 0..m-1 labels of ``whole_train``, i.e. it is every row whose label is >= m, sampled or not.
    Returns (list of (train_rows, test_rows), valid_rows or None)."""
    import pandas as pd
    df_whole = pd.DataFrame({"pos": np.arange(n_rows)})
    df_valid = None
    if use_validation:
        valid_size = validation_size if isinstance(validation_size, int) else int(df_whole.shape[0] * validation_size)
        df_whole = df_whole[valid_size:]
        df_valid = df_whole[:valid_size]
    full_split = int(full_split * len(df_whole)) if isinstance(full_split, float) else full_split
    frac = train_size if train_size < 1 else train_size / df_whole.shape[0]
    if separate_sets:
        whole_train = df_whole.sample(frac=frac).reset_index(drop=True)
        whole_test = df_whole.loc[~df_whole.index.isin(whole_train.index)].reset_index(drop=True)
    else:
        whole_train = whole_test = df_whole
    out = []
    for s_ in range(split + (1 if add_full else 0)):
        if s_ == split and add_full:
            if full_split is None:
                df_train = df_whole.sample(frac=frac, random_state=s_).reset_index(drop=True)
                df_test = df_whole.copy().reset_index(drop=True)
            else:
                df_train = df_whole[:full_split].reset_index(drop=True)
                df_test = df_whole[full_split:].reset_index(drop=True)
        elif separate_sets:
            df_train = whole_train.sample(frac=0.5, random_state=s_).reset_index(drop=True)
            df_test = whole_test
        elif full_split is None:
            df_train = df_whole.sample(frac=frac, random_state=s_).reset_index(drop=True)
            df_test = df_whole.loc[~df_whole.index.isin(df_train.index)].reset_index(drop=True)
        else:
            less = df_whole[:full_split]
            frac = train_size if train_size < 1 else train_size / less.shape[0]
            df_train = less.sample(frac=frac, random_state=s_).reset_index(drop=True)
            df_test = less.loc[~less.index.isin(df_train.index)].reset_index(drop=True)
        out.append((df_train["pos"].values.copy(), df_test["pos"].values.copy()))
    return out, (None if df_valid is None else df_valid["pos"].values.copy())


def ensemble_graphs(X, y, split, num_neighbours=1, smoothing="laplacian", mode="connectivity", scaled=True,
                    remove_mean=True, add_full=False, train_size=0.7, use_weights=False, use_validation=False,
                    validation_size=0.1, full_split=None, include_self=False, separate_sets=True):
    """load_ensembles2.py:53-297 (hot=False; hot=True dies with a KeyError at :139 once ``i`` passes the marker count):
    one dict per graph as ``create_graph`` returns it, with ``test`` nested; the validation graph last."""
    rows, valid = ensemble_rows(len(y), split, train_size, add_full, full_split, use_validation, validation_size,
                                separate_sets)
    kw = dict(n_neighbors=num_neighbours, smoothing=smoothing, use_weights=use_weights, remove_mean=remove_mean,
              mode=mode, scaled=scaled, include_self=include_self)
    out = []
    for tr, te in rows:
        g = create_graph(X[tr], y[tr], **kw)
        g["test"] = create_graph(X[te], y[te], **kw)
        out.append(g)
    if valid is not None:
        g = create_graph(X[valid], y[valid], **kw)
        g["test"] = g
        out.append(g)
    return out


# ----------------------------------------------------------------------------- subsampling
def neighbour_stream(idx):
    """S of SURVEY.md A.4: the neighbour column of the directed kNN edge list in
    CSR row-major order (util.py:59-61 + the aliasing of util.py:7)."""
    return np.asarray(idx, dtype=np.int64).ravel()


def partition_scan(S, n_nodes, size, traversed):
    """One ``naive_partition(edges, size, bidirectional=False, traversed)`` call
    (gnn/loaders/util.py:15-45) written the slow, obvious way: rescan S from its
    head each pass.  ``traversed`` is a set and is not modified."""
    traversed = set(traversed)
    free = [v for v in range(n_nodes) if v not in traversed]
    if not free:
        return []
    out = [free[0]]
    traversed.add(free[0])
    pending = len(S) > 0            # util.py:28 starts from edges[seed] == S
    while len(out) < size and len(traversed) < n_nodes and pending:
        took, full = 0, False
        for v in S:
            v = int(v)
            if v in traversed:
                continue
            traversed.add(v)
            out.append(v)
            took += 1
            if len(out) >= size:
                full = True
                break
        # util.py:38 queues one more copy of S per accepted node, except for the one that hit `size`
        queued = took - 1 if full else took
        if queued <= 0:
            pending = False
            if len(traversed) < n_nodes:          # util.py:39-43: one forced pick, then rescan
                v = min(set(range(n_nodes)) - traversed)
                traversed.add(v)
                out.append(v)
                pending = len(S) > 0
    return out


def partition_linear(S, n_nodes, sizes):
    """All batches of ``split_dataset_graph`` (util.py:63-68) in O(n + |S|):
    a monotone cursor over the first-occurrence order of S and a monotone cursor
    over node ids reproduce the repeated rescans."""
    S = np.asarray(S, dtype=np.int64)
    _, first = np.unique(S, return_index=True)
    order = S[np.sort(first)]
    traversed = np.zeros(n_nodes, dtype=bool)
    cur, lo, left = 0, 0, n_nodes
    batches = []

    def take_min():
        nonlocal lo, left
        while traversed[lo]:
            lo += 1
        traversed[lo] = True
        left -= 1
        return lo

    for size in sizes:
        if left == 0:
            batches.append([])
            continue
        out = [take_min()]
        if size > 1 and len(S) > 0:
            took = 0
            while len(out) < size and cur < len(order):
                v = int(order[cur])
                cur += 1
                if traversed[v]:
                    continue
                traversed[v] = True
                left -= 1
                out.append(v)
                took += 1
            if len(out) >= size:
                if took == 1 and left > 0:       # the size == 2 quirk of util.py:36-43
                    out.append(take_min())
            else:
                while len(out) < size and left > 0:
                    out.append(take_min())
        batches.append(out)
    return batches


def batch_sizes(n_rows, batches):
    """util.py:49-50 / :77-78: an int b means b batches of n // b rows."""
    if isinstance(batches, (int, np.integer)):
        return [n_rows // int(batches)] * int(batches)
    return list(batches)


def split_indices_graph(table, batches, neighbours=5):
    """Row indices per batch for ``split_dataset_graph`` (util.py:48-73).  ``table``
    is the whole frame, phenotype column included (util.py:52-53 passes X=df)."""
    A = table.values if hasattr(table, "values") else np.asarray(table)
    xi, xf = _split_columns(A)
    idx, _ = canonical_knn(xi, neighbours, extra=xf if xf.shape[1] else None)
    return partition_linear(neighbour_stream(idx), A.shape[0], batch_sizes(A.shape[0], batches))


def split_indices_plain(n_rows, batches):
    """Contiguous slices of ``split_dataset`` (util.py:76-84)."""
    out, start = [], 0
    for b in batch_sizes(n_rows, batches):
        out.append(list(range(min(start, n_rows), min(start + b, n_rows))))
        start += b
    return out


def split_sizes(n_rows, split=None, use_validation=False, validation_size=0.1):
    """The three sizes PlainGraph.process passes to the split algorithm (load.py:193-209)."""
    split = 0.8 if split is None else split
    train = int(n_rows * split) if split < 1 else split
    valid = validation_size if isinstance(validation_size, int) else int(validation_size * (n_rows - train))
    valid = valid if use_validation else 0
    return (train, n_rows - valid - train, n_rows - valid)


# ----------------------------------------------------------------------------- model
class OracleSequential(torch.nn.Module):
    """The GCN factories of gnn/networks/networks.py, evaluated in order.

    ``nns`` mirrors the reference's ModuleList (gnn/networks/sequential.py:72-84)
    so state_dict keys are identical: nns.{i}.lin.weight, nns.{i}.bias, ...
    """

    def __init__(self, modules):
        super().__init__()
        self.nns = torch.nn.ModuleList(modules)

    def forward(self, x, edge_index, edge_weight=None):
        for m in self.nns:
            x = m(x, edge_index, edge_weight) if isinstance(m, GCNConv) else m(x)
        return x


_FACTORIES = {
    # name: (layout, which GCNConv layers receive the add_self_loops argument)
    # C = GCNConv, R = ReLU, D = Dropout, L = Linear head
    "create_network_no_conv": ("CL", (True,)),                          # networks.py:337-345
    "create_network_no_relu": ("CL", (False,)),                         # networks.py:162-170
    "create_network_no_conv_relu": ("CRL", (False,)),                   # networks.py:151-160
    "create_network_no_conv_dropout": ("CDL", (True,)),                 # networks.py:254-263
    "create_network_no_conv_relu_dropout": ("CRDL", (False,)),          # networks.py:173-184
    "create_network_two_no_conv": ("CCL", (True, True)),                # networks.py:348-357
    "create_network_two_no_conv_dropout": ("CCDL", (True, True)),       # networks.py:266-276 (shipped, main.py:509)
    "create_network_two_no_conv_relu": ("CCRL", (True, False)),         # networks.py:208-219
    "create_network_two_no_conv_relu_dropout": ("CCRDL", (True, False)),  # networks.py:187-205
    "create_network_two_relu": ("CRCRL", (True, False)),                # networks.py:221-233
}


def make_model(name, inp_size, out_size, internal_size=-1, dropout=0.5, add_self_loops=True):
    if internal_size <= 0:
        internal_size = inp_size
    layout, takes_flag = _FACTORIES[name]
    mods, width, conv_no = [], inp_size, 0
    for ch in layout:
        if ch == "C":
            asl = add_self_loops if takes_flag[conv_no] else True
            mods.append(GCNConv(width, internal_size, add_self_loops=asl))
            width, conv_no = internal_size, conv_no + 1
        elif ch == "R":
            mods.append(torch.nn.ReLU())
        elif ch == "D":
            mods.append(torch.nn.Dropout(dropout))
        elif ch == "L":
            mods.append(torch.nn.Linear(width, out_size))
    return OracleSequential(mods)


def as_tensors(g):
    """(x float32, edge_index int64, edge_weight float32 | None) -- trainers/util.py:16-24."""
    x = torch.as_tensor(g["x"], dtype=torch.float32)
    ei = torch.as_tensor(g["edge_index"], dtype=torch.long)
    ew = None if g.get("edge_weight") is None else torch.as_tensor(g["edge_weight"]).float()
    return x, ei, ew


def epoch_step(model, optimizer, batch, test, l1_lambda=None):
    """One epoch body of train_ensemble2 phase 1 (gnn/trainers/ensemble2.py:51-88; the
    same body is plain.py:35-86 and batches.py:34-61): train-mode forward on the
    subgraph, eval-mode forward on the test graph, MSE (+ optional L1,
    trainers/util.py:4-13), backward, optimiser step.  Returns (train_loss,
    test_loss, out_test)."""
    loss_func = torch.nn.MSELoss()
    model.train()
    optimizer.zero_grad()
    out = model(*as_tensors(batch))
    model.eval()
    out_test = model(*as_tensors(test))
    loss = loss_func(out, torch.as_tensor(batch["y"], dtype=torch.float32))
    if l1_lambda is not None:
        loss = loss + l1_lambda * torch.abs(torch.cat([p.view(-1) for p in model.parameters()])).sum()
    loss.backward()
    optimizer.step()
    test_loss = float(loss_func(out_test, torch.as_tensor(test["y"], dtype=torch.float32)))
    return float(loss), test_loss, out_test.detach()


def pearson(a, b):
    a = np.asarray(a, dtype=np.float64).ravel()
    b = np.asarray(b, dtype=np.float64).ravel()
    a = a - a.mean()
    b = b - b.mean()
    den = math.sqrt(float(a @ a) * float(b @ b))
    return float(a @ b) / den if den > 0 else float("nan")


def distance_correlation(a, b):
    """``dcor.distance_correlation(x, y)`` as gnn/trainers/plain.py:76-79 calls it (two 1-D float arrays of equal length).
    The package (dcor, un-pinned in requirements.txt; absent from the image) computes the BIASED sample statistic of
    Szekely, Rizzo and Bakirov (2007): pairwise distances a_ij = |x_i - x_j|, double centring
    A_ij = a_ij - mean_i - mean_j + mean, dCov^2 = mean(A * B), dCor = sqrt(dCov^2 / sqrt(dVar_x^2 dVar_y^2)), and 0 when
    either variance vanishes.  Restated from the published definition; the reference has no fixture for it."""
    a = np.asarray(a, dtype=np.float64).reshape(-1)
    b = np.asarray(b, dtype=np.float64).reshape(-1)

    def centred(v):
        d = np.abs(v[:, None] - v[None, :])
        return d - d.mean(axis=0, keepdims=True) - d.mean(axis=1, keepdims=True) + d.mean()

    A, B = centred(a), centred(b)
    dcov2 = max(float((A * B).mean()), 0.0)
    dvar = math.sqrt(float((A * A).mean()) * float((B * B).mean()))
    return math.sqrt(dcov2 / dvar) if dvar > 0 else 0.0


def train_plain(model, optimizer, graph, test, epochs):
    """The loop of gnn/trainers/plain.py:35-100 (no ``range`` attribute, non-empty test graph): per epoch a train-mode
    forward WITH edge weights, an eval-mode forward on the test graph, Pearson r (:71-74) and distance correlation
    (:76-79) of the test predictions, MSE, backward, step.  Returns (train_losses, test_losses, cors, dcors, last test
    predictions)."""
    tl, vl, cors, dcors, out_test = [], [], [], [], None
    yt = np.asarray(test["y"], dtype=np.float32).reshape(-1)
    for _ in range(epochs):
        a, c, out_test = epoch_step(model, optimizer, graph, test)
        tl.append(a)
        vl.append(c)
        pred = out_test.numpy().reshape(-1)
        cors.append(float(np.corrcoef(pred, yt)[0][1]))
        dcors.append(distance_correlation(pred, yt))
    return tl, vl, cors, dcors, out_test


def train_batches(model, optimizer, graphs, epochs):
    """The loop of gnn/trainers/batches.py:31-68: one optimiser step per subgraph, the model called WITHOUT edge
    weights (:37, :44), per-epoch averages of the train and the test loss.  ``graphs`` = list of dicts with ``test``
    nested.  Returns (test-loss averages, train-loss averages, per-step (train, test) pairs, last test predictions)."""
    loss_func = torch.nn.MSELoss()
    losses, train_losses, steps, out_test = [], [], [], None
    for _ in range(epochs):
        losses.append(0)
        train_losses.append(0)
        for g in graphs:
            x, ei, _ = as_tensors(g)
            xt, eit, _ = as_tensors(g["test"])
            model.train()
            optimizer.zero_grad()
            out = model(x, ei)
            model.eval()
            out_test = model(xt, eit)
            loss = loss_func(out, torch.as_tensor(g["y"], dtype=torch.float32))
            loss.backward()
            optimizer.step()
            test_loss = float(loss_func(out_test, torch.as_tensor(g["test"]["y"], dtype=torch.float32)))
            losses[-1] += test_loss / len(graphs)
            train_losses[-1] += float(loss) / len(graphs)
            steps.append((float(loss), test_loss))
    return losses, train_losses, steps, out_test.detach()
