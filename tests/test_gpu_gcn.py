"""K4-K7 on the B200: GCN normalisation + CSR, aggregation, projection and their gradients against the
oracle's torch-CPU restatement of PyG's GCNConv.  Tolerance 1e-4 relative (north_star)."""
import numpy as np
import pytest
import torch

from gnn_b200 import ops
from gnn_b200.networks import networks as nets
from gnn_b200.networks.gcn_conv import GCNConv
from oracle import pyg_shim, restate, synth

pytestmark = pytest.mark.gpu
DEV = "cuda"


def rel_err(a, b):
    a = a.detach().cpu().double()
    b = b.detach().cpu().double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def random_graph(n, e, seed, loops=True, weights=True):
    rng = np.random.default_rng(seed)
    ei = rng.integers(0, n, size=(2, e))
    if loops:
        ei[:, : max(1, e // 20)] = rng.integers(0, n, size=max(1, e // 20))      # some explicit self loops (+ duplicates)
    ew = rng.random(e).astype(np.float32) if weights else None
    return torch.from_numpy(ei), (None if ew is None else torch.from_numpy(ew))


def dense_from_csr(rowptr, col, val, n):
    A = torch.zeros(n, n, dtype=torch.float64)
    rp = rowptr.cpu().numpy()
    c = col.cpu().numpy()
    v = val.cpu().double()
    for i in range(n):
        for q in range(rp[i], rp[i + 1]):
            A[i, c[q]] += v[q]
    return A


@pytest.mark.parametrize("asl", [True, False])
@pytest.mark.parametrize("weights", [True, False])
def test_csr_build_matches_gcn_norm(asl, weights):
    n, e = 97, 900
    ei, ew = random_graph(n, e, 1, weights=weights)
    g = ops.gcn_graph(ei.to(DEV), None if ew is None else ew.to(DEV), n, asl)
    torch.cuda.synchronize()
    ei2, w2 = pyg_shim.gcn_norm(ei, ew, n, asl)
    A_ref = torch.zeros(n, n, dtype=torch.float64)
    A_ref.index_put_((ei2[1], ei2[0]), w2.double(), accumulate=True)              # [target, source]
    A = dense_from_csr(g.rowptr, g.col, g.val, n)
    At = dense_from_csr(g.rowptr_t, g.col_t, g.val_t, n)
    assert g.nnz == ei2.shape[1] == int(g.rowptr[-1])
    torch.testing.assert_close(A, A_ref, rtol=1e-6, atol=1e-7)
    torch.testing.assert_close(At, A_ref.t(), rtol=1e-6, atol=1e-7)
    # order inside a row = edge order, self loop last
    rp, col = g.rowptr.cpu().numpy(), g.col.cpu().numpy()
    src, dst = ei2[0].numpy(), ei2[1].numpy()
    for i in (0, 13, n - 1):
        assert list(col[rp[i]:rp[i + 1]]) == list(src[dst == i])
    assert ops.gcn_graph(ei.to(DEV), None, n, asl) is not None


def test_graph_cache_hits_same_tensors():
    ei, ew = random_graph(50, 300, 2)
    ei, ew = ei.to(DEV), ew.to(DEV)
    a = ops.gcn_graph(ei, ew, 50, True)
    assert ops.gcn_graph(ei, ew, 50, True) is a
    assert ops.gcn_graph(ei, ew, 50, False) is not a
    ew.mul_(2.0)                                              # in-place change bumps the version counter
    assert ops.gcn_graph(ei, ew, 50, True) is not a


def test_float64_weights_through_get_args_tuple_build_one_graph():
    """The loaders hand out float64 edge weights and the trainers call get_args_tuple every epoch (plain.py:44): the float32
    copy is made once per weight tensor, so consecutive forwards hit the gcn_norm / CSR cache, and an entry dies with
    the tensors it was built from instead of pinning its CSR in device memory."""
    import gc
    from gnn_b200.data import Data
    from gnn_b200.trainers.util import get_args_tuple
    ops.clear_graph_cache()
    ei, ew = random_graph(60, 400, 4)
    d = Data(x=torch.randint(0, 3, (60, 16), dtype=torch.int8, device=DEV), edge_index=ei.to(DEV), y=torch.zeros(60, 1, device=DEV))
    d.edge_weight = ew.double().to(DEV)
    conv = GCNConv(16, 8).to(DEV)
    outs = [conv(*get_args_tuple(d)) for _ in range(3)]
    assert len(ops._graph_cache) == 1
    assert torch.equal(outs[0], outs[2])
    assert get_args_tuple(d)[2] is get_args_tuple(d)[2] and get_args_tuple(d)[2].dtype == torch.float32
    d.edge_weight.mul_(0.5)                                   # in-place change: new float32 copy, new graph
    assert not torch.equal(conv(*get_args_tuple(d)), outs[0])
    before = len(ops._graph_cache)
    tmp = conv(d.x, d.edge_index, d.edge_weight.float())     # a caller's own temporary: its entry dies with it
    gc.collect()
    assert len(ops._graph_cache) == before and tmp.shape == outs[0].shape
    del d, tmp, outs
    gc.collect()
    assert len(ops._graph_cache) == 0


def test_linear_first_factories_take_int8_genotypes():
    """create_network_linear / create_network_test put a stock Linear on x (networks.py:130-134, 295-306); x is int8 here."""
    n, p = 50, 24
    x = torch.randint(0, 3, (n, p), dtype=torch.int8, device=DEV)
    ei, _ = random_graph(n, 200, 5, weights=False)
    ei = ei.to(DEV)
    for factory, kw in ((nets.create_network_linear, {}), (nets.create_network_test, {"internal_size": 8, "dropout": 0.0})):
        torch.manual_seed(0)
        m = factory(p, 1, **kw).to(DEV)
        out = m(x, ei)
        assert out.shape == (n, 1) and rel_err(out, m(x.float(), ei)) < 1e-6
        out.sum().backward()
    ens = nets.create_ensemble_creator(nets.create_network_linear, nets.create_network_linear)(p, 1).to(DEV)
    ens.nns[0].to(DEV)                                        # the sub-models are unregistered on purpose (networks.py:103)
    assert ens(x, ei).shape == (n, 1)


@pytest.mark.parametrize("F", [1, 3, 32, 100, 128, 132, 260])
def test_spmm_against_dense(F):
    n = 300
    ei, ew = random_graph(n, 4000, 3)
    g = ops.gcn_graph(ei.to(DEV), ew.to(DEV), n, True)
    H = torch.randn(n, F, device=DEV)
    b = torch.randn(F, device=DEV)
    A = dense_from_csr(g.rowptr, g.col, g.val, n).to(DEV)
    out = ops.spmm(g.rowptr, g.col, g.val, H, b)
    ref = (A @ H.double() + b.double()).float()
    assert rel_err(out, ref) < 1e-5
    assert rel_err(ops.spmm(g.rowptr, g.col, g.val, H, b, hub=g.hub), ref) < 1e-5
    out_t = ops.spmm(g.rowptr_t, g.col_t, g.val_t, H, None)
    assert rel_err(out_t, (A.t() @ H.double()).float()) < 1e-5
    # linearity (size-independent property)
    H2 = torch.randn(n, F, device=DEV)
    lhs = ops.spmm(g.rowptr, g.col, g.val, H + 2 * H2, None)
    rhs = ops.spmm(g.rowptr, g.col, g.val, H, None) + 2 * ops.spmm(g.rowptr, g.col, g.val, H2, None)
    assert rel_err(lhs, rhs) < 1e-5


@pytest.mark.parametrize("n,p,h", [(70, 33, 5), (479, 1279, 100), (300, 2048, 32), (1000, 100, 100), (5, 7, 3)])
@pytest.mark.parametrize("int8_x", [True, False])
def test_linear_fp32_path(n, p, h, int8_x):
    X = synth.genotypes(n, p, seed=n)
    Xd = torch.from_numpy(X).to(DEV)
    if not int8_x:
        Xd = Xd.float() + 0.25 * torch.randn(n, p, device=DEV)
    W = torch.randn(h, p, device=DEV) * 0.05
    dH = torch.randn(n, h, device=DEV)
    Xf = Xd.double()
    assert rel_err(ops.linear_fwd(Xd, W, precision=0), Xf @ W.double().t()) < 1e-5
    assert rel_err(ops.linear_bwd_w(Xd, dH, precision=0), dH.double().t() @ Xf) < 1e-5
    assert rel_err(ops.linear_bwd_x(dH, W), dH.double() @ W.double()) < 1e-5


@pytest.mark.parametrize("n,p,h", [(128, 64, 16), (479, 1279, 100), (1000, 4096, 100), (300, 2050, 128), (130, 200, 7), (4000, 512, 32)])
@pytest.mark.parametrize("terms,tol", [(3, 3e-5), (2, 6e-5), (1, 8e-3)])
def test_linear_tensor_core_path(n, p, h, terms, tol):
    """tcgen05 projection with the fp32 operand split into `terms` bf16 terms, against float64."""
    X = synth.genotypes(n, p, seed=n + p)
    buf = torch.zeros((n, (p + 15) // 16 * 16), dtype=torch.int8, device=DEV)
    buf[:, :p] = torch.from_numpy(X).to(DEV)
    Xd = buf[:, :p]
    g = torch.Generator(device="cpu").manual_seed(n)
    W = (torch.randn(h, p, generator=g) * 0.05).to(DEV)
    dH = torch.randn(n, h, generator=g).to(DEV)
    Xf = Xd.double()
    fwd = ops.linear_fwd(Xd, W, precision=terms)
    bwd = ops.linear_bwd_w(Xd, dH, precision=terms)
    torch.cuda.synchronize()
    assert rel_err(fwd, Xf @ W.double().t()) < tol
    assert rel_err(bwd, dH.double().t() @ Xf) < tol
    # general codes up to 127 are exact in bf16 as well
    Xg = torch.randint(0, 128, (n, p), dtype=torch.int8, device=DEV)
    buf[:, :p] = Xg
    assert rel_err(ops.linear_fwd(buf[:, :p], W, precision=3), Xg.double() @ W.double().t()) < 3e-5
    assert rel_err(ops.linear_bwd_w(buf[:, :p], dH, precision=3), dH.double().t() @ Xg.double()) < 3e-5


@pytest.mark.parametrize("n,p,h", [(128, 128, 16), (479, 1279, 100), (1000, 4096, 100), (129, 200, 7), (4000, 512, 32),
                                   (700, 130, 104), (15625 // 4, 5000, 100)])
def test_linear_int8_limb_path(n, p, h):
    """precision 4: the float operand as three int8 digits per feature (21 bits), exact int32 tensor-core sums, one rounding."""
    X = synth.genotypes(n, p, seed=n + p)
    buf = torch.zeros((n, (p + 15) // 16 * 16), dtype=torch.int8, device=DEV)
    buf[:, :p] = torch.from_numpy(X).to(DEV)
    Xd = buf[:, :p]
    g = torch.Generator(device="cpu").manual_seed(n + 1)
    W = (torch.randn(h, p, generator=g) * 0.05).to(DEV)
    W[0] *= 1e-3                                         # a feature on a very different scale: the scale is per feature
    W[-1, ::2] = 0.0
    dH = torch.randn(n, h, generator=g).to(DEV)
    dH[:, 0] *= 1e4
    Xf = Xd.double()
    fwd = ops.linear_fwd(Xd, W, precision=4)
    bwd = ops.linear_bwd_w(Xd, dH, precision=4)
    torch.cuda.synchronize()
    ref_f, ref_b = Xf @ W.double().t(), dH.double().t() @ Xf
    # per output feature: the error is relative to that feature's own magnitude
    for f in range(h):
        assert rel_err(fwd[:, f], ref_f[:, f]) < 3e-6, ("fwd", f)
        assert rel_err(bwd[f], ref_b[f]) < 3e-6, ("bwd", f)
    # shapes the int8 kernels do not take are refused when asked for by number, and run as three bf16 terms under "auto"
    for h2 in (105, 120):                                # h > 104: three digit blocks no longer fit two MMAs
        W2 = torch.randn(h2, p, device=DEV) * 0.05
        with pytest.raises(Exception, match="precision 4"):
            ops.linear_fwd(Xd, W2, precision=4)
        assert rel_err(ops.linear_fwd(Xd, W2), Xf @ W2.double().t()) < 3e-5
    # any int8 code, negative ones included
    Xg = torch.randint(-128, 128, (n, p), dtype=torch.int8, device=DEV)
    buf[:, :p] = Xg
    assert rel_err(ops.linear_fwd(buf[:, :p], W, precision=4), Xg.double() @ W.double().t()) < 3e-6
    assert rel_err(ops.linear_bwd_w(buf[:, :p], dH, precision=4), dH.double().t() @ Xg.double()) < 3e-6


@pytest.mark.parametrize("spread", [1e2, 1e3, 1e4])
def test_linear_int8_digits_with_outliers_inside_a_feature(spread):
    """The fixed-point scale is per output feature, so what matters is the spread INSIDE a feature: a few rows of dH (hub
    nodes) or a few markers of W far above the rest.  24 bits relative to the feature's maximum keep every output within
    the path's 1e-4 of float64 (max-abs over max-abs per feature, the metric of every other test here) and within 1e-5 of
    what the exact-fp32 tiles (precision 0) reach on the same inputs."""
    n, p, h = 2048, 4096, 100
    X = torch.from_numpy(synth.genotypes(n, p, seed=11)).to(DEV)
    g = torch.Generator(device="cpu").manual_seed(int(spread))
    W = torch.randn(h, p, generator=g) * 0.01
    dH = torch.randn(n, h, generator=g) * 1e-3
    for f in range(h):                                    # outliers: 3 entries per feature at `spread` x the rms
        W[f, torch.randint(0, p, (3,), generator=g)] *= spread
        dH[torch.randint(0, n, (3,), generator=g), f] *= spread
    W, dH, Xf = W.to(DEV), dH.to(DEV), X.double()
    ref_f, ref_b = Xf @ W.double().t(), dH.double().t() @ Xf
    fwd, bwd = ops.linear_fwd(X, W, precision=4), ops.linear_bwd_w(X, dH, precision=4)
    fwd0, bwd0 = ops.linear_fwd(X, W, precision=0), ops.linear_bwd_w(X, dH, precision=0)
    worst = 0.0
    for f in range(h):
        ef, eb = rel_err(fwd[:, f], ref_f[:, f]), rel_err(bwd[f], ref_b[f])
        worst = max(worst, ef, eb)
        assert ef < 1e-4 and eb < 1e-4, (f, ef, eb)
        assert ef < rel_err(fwd0[:, f], ref_f[:, f]) + 1e-5 and eb < rel_err(bwd0[f], ref_b[f]) + 1e-5, f
    # measured 2.6e-6 .. 3.8e-6 whatever the spread: the error is 2^-24 of the feature's maximum per element, so it is
    # constant relative to the feature's largest outputs (relative to its TYPICAL outputs it grows like max / rms)
    assert worst < 1e-5, worst


def test_linear_int8_digit_sums_cannot_overflow():
    """int32 digit sums: |sum| <= K * cmax * dmax.  Any-int8 X (cmax = 128) gets the 24-bit form below K = 131 072, the
    21-bit form below 262 144 and is refused from there on; packed codes (cmax = 3) keep 24 bits far beyond."""
    n, h = 128, 8
    for p, tol in ((131_072 - 128, 1e-6), (131_072 + 128, 3e-6)):
        X = torch.randint(-128, 128, (n, p), dtype=torch.int8, device=DEV)
        X[:, ::3] = -128                                      # worst case for the accumulator: long runs of -128 * digit
        W = torch.ones(h, p, device=DEV)
        W[1] = -1.0
        W[2, ::2] = 0.5
        out = ops.linear_fwd(X, W, precision=4)
        ref = X.double() @ W.double().t()
        for f in range(h):
            assert rel_err(out[:, f], ref[:, f]) < tol, (p, f)
    X = torch.zeros((n, 262_144), dtype=torch.int8, device=DEV)
    with pytest.raises(Exception, match="precision 4"):
        ops.linear_fwd(X, torch.ones(h, 262_144, device=DEV), precision=4)
    assert rel_err(ops.linear_fwd(X + 1, torch.ones(h, 262_144, device=DEV)), torch.full((n, h), 262_144.0)) < 1e-6   # auto -> 3
    Xp = ops.pack_2bit(torch.full((n, 262_144), 3, dtype=torch.int8, device=DEV))
    out = ops.linear_fwd(Xp, torch.ones(h, 262_144, device=DEV))
    assert torch.equal(out, torch.full((n, h), 3.0 * 262_144, device=DEV))


@pytest.mark.parametrize("bad", [float("nan"), float("inf")])
def test_linear_int8_digits_propagate_nan_and_inf(bad):
    """A diverged run must stay visible: NaN / Inf in W or dH give NaN in every output of that feature (the reference's
    fp32 matmul gives NaN or Inf there), for int8 and for packed X; the other features are untouched."""
    n, p, h = 256, 512, 16
    X = torch.from_numpy(synth.genotypes(n, p, seed=5)).to(DEV)
    Xp = ops.pack_2bit(X)
    W = torch.randn(h, p, device=DEV)
    dH = torch.randn(n, h, device=DEV)
    clean_f, clean_b = ops.linear_fwd(X, W), ops.linear_bwd_w(X, dH)
    W[3, 77] = bad
    dH[100, 5] = bad
    for x in (X, Xp):
        fwd, bwd = ops.linear_fwd(x, W), ops.linear_bwd_w(x, dH, p=p)
        assert torch.isnan(fwd[:, 3]).all() and torch.isnan(bwd[5]).all()
        keep_f = [f for f in range(h) if f != 3]
        keep_b = [f for f in range(h) if f != 5]
        assert torch.equal(fwd[:, keep_f], clean_f[:, keep_f]) and torch.equal(bwd[keep_b], clean_b[keep_b])


def test_weight_gradient_by_marker_ranges_and_into_a_given_buffer():
    """ops.linear_bwd_w_ranges (one quantisation of dH, one GEMM per marker range, a callback after each range) and the
    ``out=`` form of linear_bwd_w give the bits of the plain call, for int8 and packed X."""
    n, p, h = 1000, 5003, 100
    buf = torch.zeros((n, (p + 15) // 16 * 16), dtype=torch.int8, device=DEV)
    buf[:, :p] = torch.from_numpy(synth.genotypes(n, p, seed=2)).to(DEV)
    X = buf[:, :p]
    Xp = ops.pack_2bit(X)
    dH = torch.randn(n, h, device=DEV)
    want = ops.linear_bwd_w(X, dH)
    flat = torch.full((h * p + 7,), 9.0, device=DEV)
    view = flat[3:3 + h * p].view(h, p)
    assert ops.linear_bwd_w(X, dH, out=view) is view and torch.equal(view, want) and float(flat[0]) == 9.0
    for x, fmt in ((X, 1), (Xp, 2)):
        ranges = ops.marker_ranges(p, 4, fmt)
        assert ops.marker_ranges(p, None, fmt, wave=1024) == [(c, min(p, c + 1024)) for c in range(0, p, 1024)]
        assert ranges[0][0] == 0 and ranges[-1][1] == p and all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
        seen = []
        view.fill_(-1.0)
        got = ops.linear_bwd_w_ranges(x, dH, ranges, p=p, out=view, after_range=lambda c0, c1: seen.append((c0, c1)))
        assert seen == ranges and torch.equal(got, want)
        outs = [torch.empty((h, c1 - c0), device=DEV) for c0, c1 in ranges]      # one contiguous block per range
        ops.linear_bwd_w_ranges(x, dH, ranges, p=p, out=outs)
        assert torch.equal(torch.cat(outs, dim=1), want)
    assert ops.bwd_w_wave_markers(2, X.device) % 512 == 0 and ops.bwd_w_wave_markers(2, X.device) >= 512 * 16
    # float features take the fallback: one plain call, one callback
    seen = []
    xf = torch.randn(n, 64, device=DEV)
    assert rel_err(ops.linear_bwd_w_ranges(xf, dH, [(0, 64)], after_range=lambda a, b: seen.append((a, b))),
                   dH.double().t() @ xf.double()) < 1e-5 and seen == [(0, 64)]


@pytest.mark.parametrize("packed", [False, True])
def test_gradient_sink_writes_the_first_layer_gradients_in_place(packed):
    """parallel.FlatGrads with a sink layer: the first GCNConv's backward writes dW / db into the flat buffer itself (no
    zero-fill, no accumulate pass); two steps in a row give the bits of plain autograd, i.e. nothing stale is added."""
    from gnn_b200 import parallel
    n, p, h = 600, 1024, 32
    X = torch.from_numpy(synth.genotypes(n, p, seed=4)).to(DEV)
    x = ops.pack_2bit(X) if packed else X
    ei, ew = random_graph(n, 5000, 6)
    ei, ew = ei.to(DEV), ew.to(DEV)
    y = torch.randn(n, 1, device=DEV)
    torch.manual_seed(0)
    plain = nets.create_network_two_no_conv_dropout(p, 1, internal_size=h, dropout=0.0).to(DEV)
    sunk = nets.create_network_two_no_conv_dropout(p, 1, internal_size=h, dropout=0.0).to(DEV)
    sunk.load_state_dict(plain.state_dict())
    layer = parallel.first_genotype_layer(sunk)
    assert layer is sunk.nns[0]
    flat = parallel.FlatGrads(sunk.parameters(), sink_layer=layer)
    assert flat.has_sink and flat.params[0] is layer.lin.weight and flat.params[1] is layer.bias
    flat.flat.fill_(123.0)                                   # stale contents everywhere: the step must not see them
    for step in range(2):
        plain.zero_grad()
        torch.nn.functional.mse_loss(plain(x, ei, ew), y).backward()
        flat.zero()
        torch.nn.functional.mse_loss(sunk(x, ei, ew), y).backward()
        flat.allreduce_mean(1)
        for (k, a), (_, b) in zip(plain.named_parameters(), sunk.named_parameters()):
            assert torch.equal(a.grad, b.grad), (step, k)
            assert b.grad.data_ptr() >= flat.flat.data_ptr() and b.grad.data_ptr() < flat.flat.data_ptr() + 4 * flat.flat.numel()
    # float features through the same layer keep the contract (overwrite, not accumulate)
    flat.zero()
    torch.nn.functional.mse_loss(sunk(X.float(), ei, ew), y).backward()
    assert rel_err(layer.lin.weight.grad, plain.nns[0].lin.weight.grad) < 1e-5


@pytest.mark.parametrize("p,h", [(50000, 100), (4096, 104), (1000, 7), (131, 33)])
def test_one_launch_weight_quantiser_is_bit_identical(p, h):
    """quantize_rows_fused_kernel (maximum, scale and digits of a weight row in one pass through shared memory) against the
    two-kernel path it replaces (GNNB_QUANT_FUSED=0): the projection's output is the same bits, also with an outlier column,
    an all-zero feature and a NaN."""
    import os
    n = 300
    wide = torch.zeros((n, (p + 15) // 16 * 16), dtype=torch.int8, device=DEV)
    wide[:, :p] = torch.from_numpy(synth.genotypes(n, p, seed=p)).to(DEV)
    X = wide[:, :p]                                           # rows start on 16-byte boundaries whatever p is
    torch.manual_seed(p + h)
    W = torch.randn(h, p, device=DEV) * 0.02
    W[0, p // 2] = 40.0                                       # outlier: the feature's scale is set by one element
    W[h - 1] = 0.0                                            # all-zero feature
    for nan in (False, True):
        if nan:
            W[1, 3] = float("nan")
        got = {}
        try:
            for mode in ("1", "0"):
                os.environ["GNNB_QUANT_FUSED"] = mode
                got[mode] = ops.linear_fwd(X, W, precision=4)
        finally:
            os.environ.pop("GNNB_QUANT_FUSED", None)
        assert torch.equal(got["1"].isnan(), got["0"].isnan()) and bool(got["1"][:, 1].isnan().all()) == nan
        assert torch.equal(torch.nan_to_num(got["1"]), torch.nan_to_num(got["0"]))
    assert rel_err(torch.nan_to_num(got["1"][:, 2:]), (X.double() @ W.double().t())[:, 2:]) < 1e-5


@pytest.mark.parametrize("n,h,o,p", [(15625, 100, 1, 0.4012), (300, 128, 3, 0.5), (1000, 260, 2, 0.1)])
def test_head_with_fused_dropout(n, h, o, p):
    """gnnb_head_dropout_fwd / _bwd: the mask is never stored, so recover it from dA (W = 1, dOut = 1 gives dA = m) and check
    that the forward pass, dW and db are those of the explicitly masked activations; keep rate, scaling, a new mask per call
    (the counter advances on the device, also under CUDA-graph replay), the same mask again from the same (seed, counter)."""
    torch.manual_seed(n)
    A = torch.randn(n, h, device=DEV)
    W = torch.randn(o, h, device=DEV)
    b = torch.randn(o, device=DEV)
    state = ops.dropout_state(DEV, seed=1234)
    ones_w, ones_d = torch.ones(o, h, device=DEV), torch.ones(n, o, device=DEV)
    out, used = ops.head_fwd(A, W, b, drop=(p, state))
    assert used.tolist() == [1234, 0] and state.tolist() == [1234, 1, 0, 0]
    mask = ops.head_bwd(ones_d, A, ones_w, drop=(p, used))[0] / o               # dA = m * sum_o 1 * 1
    scale = 1.0 / (1.0 - p)
    kept = mask != 0
    assert torch.allclose(mask[kept], torch.full_like(mask[kept], scale), rtol=1e-6)
    rate = float(kept.double().mean())
    assert abs(rate - (1 - p)) < 5 * (p * (1 - p) / (n * h)) ** 0.5, rate
    assert abs(float(kept.double().mean(0).std())) < 3 * (p * (1 - p) / n) ** 0.5          # no column is favoured
    ref = (A * mask).double() @ W.double().t() + b.double()
    assert rel_err(out, ref) < 1e-5
    dO = torch.randn(n, o, device=DEV)
    dA, dW, db = ops.head_bwd(dO, A, W, drop=(p, used))
    assert rel_err(dA, (dO.double() @ W.double()) * mask.double()) < 1e-5
    assert rel_err(dW, dO.double().t() @ (A * mask).double()) < 1e-5
    assert rel_err(db, dO.double().sum(0)) < 1e-5
    # the next call draws another mask; the same (seed, counter) the same one
    out2, used2 = ops.head_fwd(A, W, b, drop=(p, state))
    assert used2.tolist() == [1234, 1] and not torch.equal(out2, out)
    again = ops.dropout_state(DEV, seed=1234)
    assert torch.equal(ops.head_fwd(A, W, b, drop=(p, again))[0], out)
    # replayed as a CUDA graph: every replay advances the counter on the device
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        ops.head_fwd(A, W, b, drop=(p, state))
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        got, _ = ops.head_fwd(A, W, b, drop=(p, state))
    before = int(state[1])
    g.replay(); first = got.clone()
    g.replay(); second = got.clone()
    assert int(state[1]) == before + 2 and not torch.equal(first, second)


def test_sequential_fuses_train_mode_dropout_into_the_head():
    """Dropout -> Linear at the end of the GCN factories: in train mode the head's kernels apply the dropout (no dropout kernel
    of torch runs), in eval mode and with ``ops.config["fused_dropout"] = False`` nothing changes; the gradients reaching the
    layers in front are those of the masked head."""
    n, p, h = 500, 256, 64
    X = torch.from_numpy(synth.genotypes(n, p, seed=2)).to(DEV)
    ei, ew = random_graph(n, 4000, 3)
    ei, ew = ei.to(DEV), ew.to(DEV)
    torch.manual_seed(5)
    net = nets.create_network_two_no_conv_dropout(p, 1, internal_size=h, dropout=0.5).to(DEV)
    net.eval()
    quiet = net(X, ei, ew)
    net.train()
    calls = []
    real = torch.nn.functional.dropout
    torch.nn.functional.dropout = lambda *a, **k: (calls.append(1), real(*a, **k))[1]
    try:
        noisy = net(X, ei, ew)
        assert not calls and not torch.equal(noisy, quiet)
        assert net.nns[-1]._drop_state is not None and int(net.nns[-1]._drop_state[1]) == 1
        noisy.sum().backward()
        assert all(q.grad is not None and bool(torch.isfinite(q.grad).all()) for q in net.parameters())
        # E[dropout(x)] = x: the mean over many masks approaches the eval output
        with torch.no_grad():
            mean = torch.stack([net(X, ei, ew) for _ in range(200)]).mean(0)
        assert rel_err(mean, quiet) < 0.1
        ops.config["fused_dropout"] = False
        net(X, ei, ew)
        assert calls
    finally:
        torch.nn.functional.dropout = real
        ops.config["fused_dropout"] = True
    net.eval()
    assert torch.equal(net(X, ei, ew), quiet)


@pytest.mark.parametrize("factory", ["create_network_two_no_conv_dropout", "create_network_test", "create_network_two_relu"])
def test_flat_gradients_written_in_place_by_every_layer(factory):
    """parallel.FlatGrads(model=...): the inner GCNConv layers and the dense head write dW / db into their slices of the flat
    buffer too, so a step zero-fills nothing and autograd accumulates nothing; the bits are those of plain autograd over
    two steps that start from stale buffer contents.  A layer applied twice in one step still accumulates."""
    from gnn_b200 import parallel
    n, p, h = 700, 512, 40
    X = torch.from_numpy(synth.genotypes(n, p, seed=9)).to(DEV)
    ei, ew = random_graph(n, 6000, 8)
    ei, ew = ei.to(DEV), ew.to(DEV)
    y = torch.randn(n, 1, device=DEV)
    torch.manual_seed(1)
    plain = getattr(nets, factory)(p, 1, internal_size=h, dropout=0.0).to(DEV)
    mine = getattr(nets, factory)(p, 1, internal_size=h, dropout=0.0).to(DEV)
    mine.load_state_dict(plain.state_dict())
    flat = parallel.FlatGrads(mine.parameters(), sink_layer=parallel.first_genotype_layer(mine), model=mine)
    marked = [m for m in mine.modules() if getattr(m, "grad_inplace", False)]
    assert marked and flat.inplace_layers == marked
    covered = sum(q.numel() for m in marked for q in m.parameters()) + (flat.w_numel + h if flat.has_sink else 0)
    assert flat.tail.numel() == flat.flat.numel() - covered
    flat.flat.fill_(-7.0)
    flat.tail.zero_()                                        # (what zero() keeps clean; everything else is stale)
    for step in range(2):
        plain.zero_grad()
        torch.nn.functional.mse_loss(plain(X, ei, ew), y).backward()
        flat.zero()
        torch.nn.functional.mse_loss(mine(X, ei, ew), y).backward()
        flat.allreduce_mean(1)
        for (k, a), (_, b) in zip(plain.named_parameters(), mine.named_parameters()):
            assert torch.equal(a.grad, b.grad), (factory, step, k)
    # the same model applied twice in one step: the second backward of every marked layer accumulates
    plain.zero_grad()
    (torch.nn.functional.mse_loss(plain(X, ei, ew), y) + torch.nn.functional.mse_loss(plain(X, ei, ew), -y)).backward()
    flat.zero()
    (torch.nn.functional.mse_loss(mine(X, ei, ew), y) + torch.nn.functional.mse_loss(mine(X, ei, ew), -y)).backward()
    for (k, a), (_, b) in zip(plain.named_parameters(), mine.named_parameters()):
        if flat.has_sink and any(b is q for q in flat.params[:2]):
            continue                                         # the sink layer itself overwrites: one use per step
        assert rel_err(b.grad, a.grad) < 1e-5, (factory, k)


@pytest.mark.parametrize("n,h,o", [(15625, 100, 1), (1000, 100, 3), (77, 128, 4), (500, 300, 1), (64, 1024, 1), (10, 7, 2), (1, 100, 1)])
def test_head_linear_kernels_against_torch(n, h, o):
    """networks.head.Linear (the dense head of every GCN factory): one kernel forward, one backward, against torch's own
    Linear on the same parameters -- outputs and all three gradients, a strided input, reproducible sums."""
    from gnn_b200.networks.head import Linear
    torch.manual_seed(n + h)
    mine = Linear(h, o).to(DEV)
    ref = torch.nn.Linear(h, o).to(DEV)
    ref.load_state_dict(mine.state_dict())
    assert ops.head_supported(h, o)
    base = torch.randn(n, h + 4, device=DEV)
    x1 = base[:, :h].clone().requires_grad_(True)
    x2 = base[:, :h].clone().requires_grad_(True)
    wgt = torch.randn(n, o, device=DEV)
    y1, y2 = mine(x1), ref(x2)
    assert rel_err(y1, y2) < 1e-5
    (y1 * wgt).sum().backward()
    (y2 * wgt).sum().backward()
    assert rel_err(x1.grad, x2.grad) < 1e-5
    assert rel_err(mine.weight.grad, ref.weight.grad) < 1e-5 and rel_err(mine.bias.grad, ref.bias.grad) < 1e-5
    # strided rows (a column slice of a wider matrix), no input gradient, same bits on a second run
    g1 = mine.weight.grad.clone()
    mine.zero_grad()
    (mine(base[:, :h]) * wgt).sum().backward()
    assert torch.equal(mine.weight.grad, g1)
    no_bias = Linear(h, o, bias=False).to(DEV)
    assert rel_err(no_bias(base[:, :h]), base[:, :h] @ no_bias.weight.t()) < 1e-5


def test_head_linear_falls_back_outside_its_shapes():
    from gnn_b200.networks.head import Linear
    for h, o in ((2000, 1), (100, 8), (600, 3)):
        assert not ops.head_supported(h, o)
        m = Linear(h, o).to(DEV)
        x = torch.randn(33, h, device=DEV)
        assert rel_err(m(x), torch.nn.functional.linear(x, m.weight, m.bias)) < 1e-6
    m = Linear(100, 1).to(DEV)
    assert m(torch.randn(4, 5, 100, device=DEV)).shape == (4, 5, 1)           # 3-D input: torch's path
    assert Linear(100, 1)(torch.randn(6, 100)).shape == (6, 1)                 # CPU tensors: torch's path


@pytest.mark.parametrize("F", [4, 100, 256])
def test_spmm_hub_rows_planned_and_unplanned(F):
    """A graph with hubs (rows of 300 .. 5000 entries): the planned path (tasks + last-arriver reduction) and the
    CTA-per-row path both match a dense product, and repeated calls are bit-identical (deterministic)."""
    n = 6000
    rng = np.random.default_rng(0)
    src = [rng.integers(0, n, size=40 * n)]
    dst = [rng.integers(0, n, size=40 * n)]
    for hub, deg in ((5, 300), (17, 1000), (4000, 5000), (n - 1, 257)):
        src.append(rng.integers(0, n, size=deg))
        dst.append(np.full(deg, hub))
    ei = torch.from_numpy(np.stack([np.concatenate(src), np.concatenate(dst)])).to(DEV)
    ew = torch.rand(ei.shape[1], device=DEV)
    g = ops.gcn_graph(ei, ew, n, True)
    assert g.hub is not None and int(g.hub[0].numel()) >= 2 + 4 + 20 + 2
    H = torch.randn(n, F, device=DEV)
    b = torch.randn(F, device=DEV)
    ref = torch.zeros(n, F, dtype=torch.float64, device=DEV)
    rp = g.rowptr.long()
    rows = torch.repeat_interleave(torch.arange(n, device=DEV), rp[1:] - rp[:-1])
    ref.index_add_(0, rows, g.val[: g.nnz].double()[:, None] * H.double()[g.col[: g.nnz].long()])
    ref += b.double()
    a1 = ops.spmm(g.rowptr, g.col, g.val, H, b, hub=g.hub)
    a2 = ops.spmm(g.rowptr, g.col, g.val, H, b, hub=None)
    assert rel_err(a1, ref) < 1e-5 and rel_err(a2, ref) < 1e-5
    for _ in range(3):
        assert torch.equal(ops.spmm(g.rowptr, g.col, g.val, H, b, hub=g.hub), a1)


def oracle_and_mine(name, inp, hidden, asl=True, seed=0):
    torch.manual_seed(seed)
    ref = restate.make_model(name, inp, 1, internal_size=hidden, dropout=0.0, add_self_loops=asl)
    kw = {"internal_size": hidden}
    if "dropout" in name:
        kw["dropout"] = 0.0
    if name in ("create_network_two_no_conv_dropout", "create_network_two_relu", "create_network_no_conv",
                "create_network_two_no_conv", "create_network_no_conv_dropout", "create_network_two_no_conv_relu",
                "create_network_two_no_conv_relu_dropout"):
        kw["add_self_loops"] = asl
    mine = getattr(nets, name)(inp, 1, **kw)
    mine.load_state_dict(ref.state_dict())                    # weights always come from the oracle model
    return ref, mine.to(DEV)


def to_dev(g):
    x = torch.from_numpy(np.ascontiguousarray(g["x"]).astype(np.int8)).to(DEV)
    ei = torch.from_numpy(g["edge_index"]).to(DEV)
    ew = None if g["edge_weight"] is None else torch.from_numpy(g["edge_weight"]).to(DEV)
    return x, ei, ew, torch.from_numpy(g["y"]).to(DEV)


@pytest.mark.parametrize("name,asl", [("create_network_two_no_conv_dropout", True), ("create_network_two_no_conv_dropout", False),
                                      ("create_network_no_conv_relu_dropout", True), ("create_network_two_relu", True),
                                      ("create_network_no_conv", True)])
def test_model_forward_backward_golden(wheat, golden, name, asl):
    """Seeded forward/backward of the reference factories on the WHEAT training graph (golden from the
    reference's own run): activations, loss and every gradient within 1e-4 relative."""
    X, y = wheat
    gold = golden("wheat_model.npz")
    tag = f"{name}_asl{int(asl)}"
    graph = restate.create_graph(X[:479], y[:479], 5, "laplacian", True, True, "distance")
    x, ei, ew, yt = to_dev(graph)
    _, mine = oracle_and_mine(name, 1279, 32, asl)
    mine.load_state_dict({k.split("__w__")[1]: torch.from_numpy(gold[k]) for k in gold.files if k.startswith(tag + "__w__")})
    mine.train()
    out = mine(x, ei, ew.float())
    loss = torch.nn.MSELoss()(out, yt)
    loss.backward()
    assert rel_err(out, torch.from_numpy(gold[tag + "__out"])) < 1e-4
    assert abs(float(loss) - float(gold[tag + "__loss"])) < 1e-4 * float(gold[tag + "__loss"])
    for kk, prm in mine.named_parameters():
        assert rel_err(prm.grad, torch.from_numpy(gold[f"{tag}__g__{kk}"])) < 1e-4, kk
    # call forms of sequential.py:100-101 and the no-weights form of batches.py:37
    assert torch.equal(mine((x, ei, ew.float())), mine(x, ei, ew.float()))
    assert mine((x, ei)).shape == out.shape
    # float32 x (the reference's own dtype) gives the same activations as the int8 codes
    assert rel_err(mine(x.float(), ei, ew.float()), out) < 1e-5


def test_inner_layer_input_gradient_and_raw_knn_graph():
    """dx through an inner GCNConv, and a non-symmetric edge list (smoothing='none', load.py:76-80) where the
    backward needs the transposed CSR."""
    n, p, k = 400, 64, 6
    X = synth.genotypes(n, p, seed=9)
    y = synth.phenotypes(n, p)[:, 0]
    g = restate.create_graph(X, y, k, "none", True, True, "distance")
    x, ei, ew, yt = to_dev(g)
    torch.manual_seed(1)
    ref = pyg_shim.GCNConv(p, 16)
    mine = GCNConv(p, 16).to(DEV)
    mine.load_state_dict(ref.state_dict())
    xr = torch.from_numpy(g["x"]).clone().requires_grad_(True)
    xm = x.float().clone().requires_grad_(True)
    o_ref = ref(xr, torch.from_numpy(g["edge_index"]), torch.from_numpy(g["edge_weight"]).float())
    o_mine = mine(xm, ei, ew.float())
    wgt = torch.randn(n, 16)
    (o_ref * wgt).sum().backward()
    (o_mine * wgt.to(DEV)).sum().backward()
    assert rel_err(o_mine, o_ref) < 1e-4
    assert rel_err(xm.grad, xr.grad) < 1e-4
    assert rel_err(mine.lin.weight.grad, ref.lin.weight.grad) < 1e-4
    assert rel_err(mine.bias.grad, ref.bias.grad) < 1e-4


def test_training_trajectory_golden(wheat, golden):
    """Six epochs of train_ensemble2's epoch body (ensemble2.py:51-88) per subgraph: losses within 1e-4
    relative of the reference's trajectory, predictions with Pearson r >= 0.9999."""
    from gnn_b200.data import GenotypeTable
    from gnn_b200.loaders.load import create_graph
    from gnn_b200.trainers.util import get_args_tuple
    X, y = wheat
    g = golden("wheat_train.npz")
    H, E, lr = int(g["hidden"]), int(g["epochs"]), float(g["lr"])
    te = g["test_rows"]
    test = create_graph(GenotypeTable.from_numpy(X[te], y[te]), 5, "laplacian", True, mode="distance")
    for b in range(3):
        rows = g[f"b{b}_rows"]
        batch = create_graph(GenotypeTable.from_numpy(X[rows], y[rows]), 5, "laplacian", True, mode="distance")
        model = nets.create_network_two_no_conv_dropout(1279, 1, internal_size=H, dropout=0.0)
        model.load_state_dict({k.split("__w0__")[1]: torch.from_numpy(g[k]) for k in g.files if k.startswith(f"b{b}__w0__")})
        model.to(DEV)
        opt = torch.optim.Adam(model.parameters(), lr=lr)
        lf = torch.nn.MSELoss()
        tl, vl = [], []
        for _ in range(E):
            model.train()
            opt.zero_grad()
            out = model(*get_args_tuple(batch))
            model.eval()
            out_test = model(*get_args_tuple(batch.test if hasattr(batch, "test") else test))
            loss = lf(out, batch.y)
            loss.backward()
            opt.step()
            tl.append(float(loss))
            vl.append(float(lf(out_test, test.y)))
        np.testing.assert_allclose(tl, g[f"b{b}_train_loss"], rtol=1e-4)
        np.testing.assert_allclose(vl, g[f"b{b}_test_loss"], rtol=1e-4)
        assert restate.pearson(out_test.detach().cpu().numpy(), g[f"b{b}_pred_test"]) >= 0.9999
        model.eval()
        pred_train = model(*get_args_tuple(batch)).detach().cpu().numpy()
        assert restate.pearson(pred_train, g[f"b{b}_pred_train"]) >= 0.9999
        assert rel_err(torch.from_numpy(pred_train), torch.from_numpy(g[f"b{b}_pred_train"])) < 1e-3


@pytest.mark.parametrize("n,p,h", [(128, 128, 16), (479, 1279, 100), (1000, 4096, 100), (129, 200, 7), (3000, 5003, 32)])
def test_linear_packed_genotypes(n, p, h):
    """2-bit packed genotypes (uint8, four codes per byte) through the projection: the converter warps must hand the MMA
    exactly the int8 tile, so the results are BIT-IDENTICAL to the int8 path (same exact integer sums, same rounding)."""
    X = synth.genotypes(n, p, seed=3 * n + p)
    buf = torch.zeros((n, (p + 15) // 16 * 16), dtype=torch.int8, device=DEV)
    buf[:, :p] = torch.from_numpy(X).to(DEV)
    Xd = buf[:, :p]
    Xp = ops.pack_2bit(Xd)
    assert Xp.dtype == torch.uint8 and Xp.shape == (n, ops.packed_stride(p))
    g = torch.Generator(device="cpu").manual_seed(n)
    W = (torch.randn(h, p, generator=g) * 0.05).to(DEV)
    dH = torch.randn(n, h, generator=g).to(DEV)
    assert torch.equal(ops.linear_fwd(Xp, W), ops.linear_fwd(Xd, W))
    assert torch.equal(ops.linear_bwd_w(Xp, dH, p=p), ops.linear_bwd_w(Xd, dH))
    # through the operator: same outputs and gradients with packed node features
    ei = torch.stack([torch.arange(n, device=DEV), torch.roll(torch.arange(n, device=DEV), 1)])
    torch.manual_seed(0)
    conv = GCNConv(p, h).to(DEV)
    out_a = conv(Xd, ei)
    out_b = conv(Xp, ei)
    assert torch.equal(out_a, out_b)
    (out_b * dH).sum().backward()
    gb = conv.lin.weight.grad.clone()
    conv.zero_grad()
    (conv(Xd, ei) * dH).sum().backward()
    assert torch.equal(gb, conv.lin.weight.grad)
    # shapes the packed kernel does not take are refused, not silently converted
    with pytest.raises(Exception):
        ops.linear_fwd(Xp, torch.randn(120, p, device=DEV))
