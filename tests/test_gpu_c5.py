"""Parity at the shapes and on the code paths bench.py quotes (BASELINE.json config C5): one subgraph of
15 625 individuals x 50 000 SNPs, the 180 000-node test graph, k = 32, hidden 100.

What runs here is what the benchmark runs -- the 2-bit packed projection in clusters of four (linear_p2_kernel), the
tcgen05 quad-cluster neighbour search (knn_tc2_kernel, int8 tiles and the FP4 tiles the graph build uses), the planned aggregation with hub rows of
tens of thousands of entries -- compared with independent float64 / exact-integer arithmetic on the same inputs and,
for the whole training step, with the oracle's torch-CPU restatement (oracle.restate.epoch_step, ensemble2.py:51-88).
Bars as everywhere: bit-exact indices and integer distances, 1e-4 relative for fp32 activations / gradients (the
int8-digit projection is held to its own 3e-6 per feature), Pearson r >= 0.9999.
Needs about 25 GB of device memory and 10 GB of host memory; about a minute on a B200.
"""
import numpy as np
import pytest
import torch

from gnn_b200 import ops
from gnn_b200.data import GenotypeTable
from gnn_b200.loaders.load import create_graph
from gnn_b200.networks import networks as nets
from oracle import restate

pytestmark = pytest.mark.gpu
DEV = "cuda"
N_SUB, N_TEST, P, K, HID = 15_625, 180_000, 50_000, 32, 100
ROW0_TEST = 800_000                         # bench.py: test rows of the 1M frame start at int(0.8 * N)


def rel_err(a, b):
    a, b = a.detach().double(), b.detach().double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.fixture(scope="module")
def sub():
    """Subgraph 0 of the benchmark: int8 codes, their 2-bit packing, phenotypes."""
    X = ops.synth_genotypes(N_SUB, P, DEV, row0=0, p_total=P)
    return X, ops.pack_2bit(X)


@pytest.fixture(scope="module")
def big():
    X = ops.synth_genotypes(N_TEST, P, DEV, row0=ROW0_TEST, p_total=P)
    return X


def f64_rows_times_wt(X, rows, W):
    """float64 X[rows] . W^T in row blocks (exact products of small integers with float32 weights, float64 sums)."""
    Wd = W.double().t().contiguous()
    out = torch.empty((rows.numel(), W.shape[0]), dtype=torch.float64, device=X.device)
    for r0 in range(0, rows.numel(), 2048):
        r = rows[r0:r0 + 2048]
        out[r0:r0 + r.numel()] = X[r].double() @ Wd
    return out


# ------------------------------------------------------------------------------------------------- (a) projection
def test_packed_projection_equals_int8_and_float64_at_the_subgraph_shape(sub):
    X, Xp = sub
    g = torch.Generator(device="cpu").manual_seed(5)
    a = (6.0 / (P + HID)) ** 0.5                                         # Glorot range of GCNConv.lin (gcn_conv.py)
    W = ((torch.rand(HID, P, generator=g) * 2 - 1) * a).to(DEV)
    dH = (torch.randn(N_SUB, HID, generator=g) * 1e-3).to(DEV)
    fwd_p, fwd_i = ops.linear_fwd(Xp, W), ops.linear_fwd(X, W)
    bwd_p, bwd_i = ops.linear_bwd_w(Xp, dH, p=P), ops.linear_bwd_w(X, dH)
    torch.cuda.synchronize()
    assert torch.equal(fwd_p, fwd_i), "packed forward differs from the int8 forward"
    assert torch.equal(bwd_p, bwd_i), "packed weight gradient differs from the int8 one"
    ref_f = f64_rows_times_wt(X, torch.arange(N_SUB, device=DEV), W)
    ref_b = torch.zeros((HID, P), dtype=torch.float64, device=DEV)
    for r0 in range(0, N_SUB, 2048):
        ref_b += dH[r0:r0 + 2048].double().t() @ X[r0:r0 + 2048].double()
    for f in range(HID):
        assert rel_err(fwd_p[:, f], ref_f[:, f]) < 3e-6, ("fwd", f)
        assert rel_err(bwd_p[f], ref_b[f]) < 3e-6, ("bwd", f)
    # every row and every marker is written (the last partial tile included)
    assert torch.isfinite(fwd_p).all() and torch.isfinite(bwd_p).all()


def test_packed_projection_on_the_test_graph_rows(big):
    Xp = ops.pack_2bit(big)
    g = torch.Generator(device="cpu").manual_seed(6)
    W = ((torch.rand(HID, P, generator=g) * 2 - 1) * 0.011).to(DEV)
    H = ops.linear_fwd(Xp, W)
    rows = torch.randperm(N_TEST, generator=g)[:4096].sort().values.to(DEV)
    rows[0], rows[-1] = 0, N_TEST - 1                                       # first and last row of the matrix
    ref = f64_rows_times_wt(big, rows, W)
    for f in range(HID):
        assert rel_err(H[rows][:, f], ref[:, f]) < 3e-6, f
    del Xp


# ------------------------------------------------------------------------------------------------- (b) neighbour search
def brute_force_lists(X, sq, rows, k):
    """Canonical (d2, j) lists of the sampled rows by exact arithmetic: float64 Gram blocks (exact: |dot| <= 4 p), a
    single sortable int64 key d2 * 2^20 + j, self excluded."""
    n = X.shape[0]
    keys = torch.empty((rows.numel(), n), dtype=torch.int64, device=X.device)
    Q = X[rows].double()
    for c0 in range(0, n, 16384):
        c1 = min(n, c0 + 16384)
        dot = (Q @ X[c0:c1].double().t()).to(torch.int64)
        d2 = sq[rows][:, None] + sq[None, c0:c1] - 2 * dot
        keys[:, c0:c1] = d2 * (1 << 20) + torch.arange(c0, c1, device=X.device)[None, :]
    keys[torch.arange(rows.numel(), device=X.device), rows] = torch.iinfo(torch.int64).max
    best = torch.topk(keys, k, dim=1, largest=False, sorted=True).values
    return (best & ((1 << 20) - 1)).to(torch.int32), (best >> 20).double()


def sample_rows(n, seed, count=64):
    g = torch.Generator(device="cpu").manual_seed(seed)
    rows = torch.randperm(n, generator=g)[:count].sort().values
    rows[0], rows[-1] = 0, n - 1
    return rows.to(DEV)


def test_tensor_core_knn_at_the_subgraph_shape(sub):
    X, _ = sub
    idx2, d2 = ops.knn(X, K, impl=2)
    torch.cuda.synchronize()
    rows = sample_rows(N_SUB, 7)
    bi, bd = brute_force_lists(X, ops.row_sqnorm(X), rows, K)
    assert torch.equal(idx2[rows], bi), "neighbour lists differ from the exact brute force"
    assert torch.equal(d2[rows], bd)
    idx1, d1 = ops.knn(X, K, impl=1)                                        # CUDA-core dp4a tiles on the full matrix
    assert torch.equal(idx2, idx1) and torch.equal(d2, d1)
    idx4, d4 = ops.knn(X, K, impl=3)                                        # FP4 tensor tiles: what the bench's graph build runs
    assert torch.equal(idx2, idx4) and torch.equal(d2, d4)
    # size-independent properties: ascending (d2, j), no self, all slots filled
    assert bool((d2[:, 1:] >= d2[:, :-1]).all())
    tie = d2[:, 1:] == d2[:, :-1]
    assert bool((idx2[:, 1:][tie] > idx2[:, :-1][tie]).all())
    assert bool((idx2 != torch.arange(N_SUB, device=DEV, dtype=torch.int32)[:, None]).all()) and bool((idx2 >= 0).all())


def test_tensor_core_knn_at_the_test_graph_shape(big):
    idx, d2 = ops.knn(big, K, impl=2)
    torch.cuda.synchronize()
    rows = sample_rows(N_TEST, 8)
    bi, bd = brute_force_lists(big, ops.row_sqnorm(big), rows, K)
    assert torch.equal(idx[rows], bi) and torch.equal(d2[rows], bd)
    assert bool((d2[:, 1:] >= d2[:, :-1]).all()) and bool((idx >= 0).all())
    idx4, d4 = ops.knn(big, K, impl=3)                                      # FP4 tensor tiles on all 180 000 rows
    assert torch.equal(idx, idx4) and torch.equal(d2, d4)


# ------------------------------------------------------------------------------------------------- (c) aggregation
def test_planned_aggregation_on_the_test_graph(big):
    """GCNConv's normalised aggregation on the 180 000-node kNN graph (hub rows with > 10 000 entries) against a float64
    restatement of gcn_norm + scatter-add evaluated with index_add_ on the same edge list."""
    y = np.zeros(N_TEST)
    data = create_graph(GenotypeTable(big, y), n_neighbors=K, smoothing="laplacian", use_weights=True, mode="distance",
                        algorithm="euclidean")
    ei, ew = data.edge_index, data.edge_weight.float()
    n = N_TEST
    g = ops.gcn_graph(ei, ew, n, True)
    deg_max = int((g.rowptr[1:] - g.rowptr[:-1]).max())
    assert g.hub is not None and deg_max > 2048, deg_max
    # float64 gcn_norm: drop explicit self loops, append one of weight 1 per node unless one was there (keep its weight)
    src, dst = ei[0], ei[1]
    loop = src == dst
    w64 = ew.double()
    self_w = torch.ones(n, dtype=torch.float64, device=DEV)
    self_w[src[loop]] = w64[loop]
    s2 = torch.cat([src[~loop], torch.arange(n, device=DEV)])
    d2_ = torch.cat([dst[~loop], torch.arange(n, device=DEV)])
    w2 = torch.cat([w64[~loop], self_w])
    deg = torch.zeros(n, dtype=torch.float64, device=DEV).index_add_(0, d2_, w2)
    dis = deg.pow(-0.5)
    dis[torch.isinf(dis)] = 0
    norm = dis[s2] * w2 * dis[d2_]
    assert g.nnz == s2.numel()
    gen = torch.Generator(device="cpu").manual_seed(9)
    H = torch.randn(n, HID, generator=gen).to(DEV)
    b = torch.randn(HID, generator=gen).to(DEV)

    def ref_agg(a_idx, b_idx):                                               # out[a] += norm * H[b]
        out = torch.zeros((n, HID), dtype=torch.float64, device=DEV)
        for e0 in range(0, s2.numel(), 1 << 21):
            sl = slice(e0, e0 + (1 << 21))
            out.index_add_(0, a_idx[sl], norm[sl, None] * H[b_idx[sl]].double())
        return out

    # fp32 sums over rows of up to tens of thousands of entries: the bar is the path's 1e-4 (measured 1e-6 .. 3e-5)
    fwd = ops.spmm(g.rowptr, g.col, g.val, H, b, n=n, hub=g.hub)
    assert rel_err(fwd, ref_agg(d2_, s2) + b.double()) < 1e-4
    bwd = ops.spmm(g.rowptr_t, g.col_t, g.val_t, H, None, n=n, hub=g.hub_t)
    assert rel_err(bwd, ref_agg(s2, d2_)) < 1e-4
    assert torch.equal(ops.spmm(g.rowptr, g.col, g.val, H, b, n=n, hub=g.hub), fwd)      # deterministic
    ops.clear_graph_cache()


# ------------------------------------------------------------------------------------------------- (d) one training step
def test_one_training_step_of_subgraph_0_against_the_oracle(sub):
    """The benchmark's step on its own inputs -- packed genotypes, k = 32 Laplacian graph, two GCN layers, MSE, backward --
    against oracle.restate.epoch_step on the CPU with the same weights and the same edge list (the list itself is
    checked against brute force above and against the oracle at every smaller shape)."""
    from bench import phenotypes
    X, Xp = sub
    y = phenotypes(np.arange(N_SUB), P)
    data = create_graph(GenotypeTable(X, y), n_neighbors=K, smoothing="laplacian", use_weights=True, mode="distance",
                        algorithm="euclidean")
    ref_graph = {"x": X.cpu().numpy().astype(np.float32), "y": data.y.cpu().numpy(),
                 "edge_index": data.edge_index.cpu().numpy(), "edge_weight": data.edge_weight.cpu().numpy()}
    torch.manual_seed(0)
    m_ref = restate.make_model("create_network_two_no_conv_dropout", P, 1, internal_size=HID, dropout=0.0)
    m = nets.create_network_two_no_conv_dropout(P, 1, internal_size=HID, dropout=0.0)
    m.load_state_dict(m_ref.state_dict())
    m.to(DEV)
    opt_ref = torch.optim.Adam(m_ref.parameters(), lr=2e-3)
    l_ref, _, out_ref = restate.epoch_step(m_ref, opt_ref, ref_graph, ref_graph)
    m.train()
    out = m(Xp, data.edge_index, data.edge_weight.float())
    loss = torch.nn.MSELoss()(out, data.y)
    loss.backward()
    torch.cuda.synchronize()
    assert abs(float(loss) - l_ref) <= 1e-4 * abs(l_ref), (float(loss), l_ref)
    grads_ref = {k: p.grad for k, p in m_ref.named_parameters()}
    for name, prm in m.named_parameters():
        assert rel_err(prm.grad.cpu(), grads_ref[name]) < 1e-4, name
    # out_ref is the eval-mode forward of the PRE-step weights on the same graph = the train-mode forward (dropout 0)
    assert rel_err(out.cpu(), out_ref) < 1e-4
    assert restate.pearson(out.detach().cpu().numpy(), out_ref.numpy()) >= 0.9999
    ops.clear_graph_cache()
