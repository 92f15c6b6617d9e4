"""Host-side logic of the drop-in surface (no GPU): Sequential, the factories' state_dict layout, the
graph containers, and the C++ partition walk against the oracle and the golden fixtures."""
import numpy as np
import pytest
import torch
from hypothesis import given, settings, strategies as st

from gnn_b200 import ops
from gnn_b200.data import Data, GraphDataset
from gnn_b200.loaders.util import naive_partition, split_dataset
from gnn_b200.networks import networks as nets
from gnn_b200.networks.sequential import Sequential, parse_desc
from oracle import restate


# ------------------------------------------------------------------------------- Sequential (sequential.py:65-151)
def test_sequential_headers_optional_args_and_call_forms():
    calls = []

    class Probe(torch.nn.Module):
        def forward(self, *args):
            calls.append(len(args))
            return args[0] + 1

    m = Sequential("x, edge_index, edge_weights?", [
        (Probe(), "x, edge_index, edge_weights -> x"),
        torch.nn.Identity(),
        (lambda x: (x, x * 2), "x -> a, b"),
        (lambda a, b: a + b, "a, b -> x"),
    ])
    x, ei, ew = torch.ones(3, 2), torch.zeros(2, 0), torch.ones(0)
    out3 = m(x, ei, ew)
    out2 = m(x, ei)                       # optional trailing argument absent
    out_t = m((x, ei, ew))                # tuple call form (sequential.py:100-101)
    assert calls == [3, 2, 3]
    assert torch.equal(out3, (x + 1) * 3) and torch.equal(out2, out3) and torch.equal(out_t, out3)
    with pytest.raises(AssertionError):
        m(x, ei, ew, ew)
    assert parse_desc("x, edge_index -> x") == (["x", "edge_index"], ["x"])
    assert m.get(1).__class__ is torch.nn.Identity and callable(m.get(2))
    m.replace(1, torch.nn.ReLU())
    assert isinstance(m.nns[1], torch.nn.ReLU) and m.fns[1] is None
    m.replace(1, lambda x: x)
    assert m.nns[1] is None and callable(m.fns[1])
    assert "Sequential(" in repr(m)


def test_sequential_l1_and_reset():
    m = nets.create_network_two_no_conv_dropout(6, 1, internal_size=4, dropout=0.0)
    loss = torch.zeros(())
    total = sum(p.abs().sum() for p in m.parameters())
    assert torch.allclose(m.l1_regularize(loss, 0.5), 0.5 * total)
    before = m.nns[0].lin.weight.clone()
    m.reset_parameters()
    assert not torch.equal(before, m.nns[0].lin.weight)
    assert torch.count_nonzero(m.nns[0].bias) == 0


@pytest.mark.parametrize("name", sorted(restate._FACTORIES))
def test_factories_match_the_oracle_layout(name):
    """Same module order, shapes and state_dict keys as the reference factories (pinned via the oracle)."""
    mine = getattr(nets, name)(12, 2, internal_size=5)
    ref = restate.make_model(name, 12, 2, internal_size=5)
    a, b = mine.state_dict(), ref.state_dict()
    assert list(a.keys()) == list(b.keys())
    assert all(a[k].shape == b[k].shape for k in a)
    kinds = lambda seq: [type(m).__name__ for m in seq]
    assert kinds(mine.nns) == kinds(ref.nns)
    mine.load_state_dict(b)               # reference-style checkpoints load
    # glorot bound of PyG's Linear(bias=False): |w| <= sqrt(6 / (in + out))
    w = mine.nns[0].lin.weight
    assert float(w.abs().max()) <= (6.0 / (12 + 5)) ** 0.5 + 1e-7


def test_factories_add_self_loops_routing():
    m = nets.create_network_two_no_conv_relu_dropout(8, 1, internal_size=4, add_self_loops=False)
    assert m.nns[0].add_self_loops is False and m.nns[1].add_self_loops is True      # networks.py:199-200
    m = nets.create_network_two_no_conv_dropout(8, 1, internal_size=4, add_self_loops=False)
    assert m.nns[0].add_self_loops is False and m.nns[1].add_self_loops is False     # networks.py:271-272
    m = nets.create_network_no_conv(8, 1)
    assert m.nns[0].out_channels == 8                                               # internal_size <= 0 -> inp_size


def test_ensemble_registers_only_the_aggregator():
    subs = [nets.create_network_no_conv(6, 1, internal_size=3) for _ in range(2)]
    e = nets.Ensemble(subs, in_size=3)
    assert [k for k, _ in e.named_parameters()] == ["aggregator.weight", "aggregator.bias"]   # networks.py:103
    assert e.aggregator.in_features == 6
    e.eval()
    assert not subs[0].training
    e.train()
    assert subs[0].training


# ------------------------------------------------------------------------------- containers
def test_data_and_dataset_surface():
    d = Data(x=torch.zeros(4, 3), y=torch.zeros(4, 1), edge_index=torch.zeros(2, 5, dtype=torch.long))
    assert d.edge_weight is None and hasattr(d, "edge_weight") and not hasattr(d, "range")
    d.test = Data(x=torch.zeros(2, 3))
    assert d.num_features == 3 and d.num_nodes == 4 and "edge_index" in d.keys
    assert d.to("cpu") is d and d.test.x.device.type == "cpu"
    ds = GraphDataset([d, Data(x=torch.ones(1, 3))])
    assert len(ds) == 2 and ds.num_features == 3 and ds.data is d
    assert len(ds[:-1]) == 1 and ds[-1].x.sum() == 3 and [g.num_nodes for g in ds] == [4, 1]
    single = Data(x=torch.zeros(1, 2))
    single.valid = Data(x=torch.zeros(1, 2))
    assert GraphDataset([single]).valid is single.valid


def test_split_dataset_slices_like_the_reference():
    import pandas as pd
    df = pd.DataFrame(np.arange(40).reshape(10, 4))
    a, b, c = split_dataset(df, (6, 3, 9))
    assert list(a.index) == list(range(6)) and list(b.index) == [6, 7, 8] and list(c.index) == [9]    # clipped tail
    parts = split_dataset(df, 3)
    assert [len(p) for p in parts] == [3, 3, 3]                                                        # N mod b dropped
    assert [list(map(int, p)) for p in restate.split_indices_plain(10, (6, 3, 9))] == [list(a.index), list(b.index), list(c.index)]


# ------------------------------------------------------------------------------- partition walk (C++, host)
def test_partition_walk_golden_small(golden):
    g = golden("partition_small.npz")
    for c in range(40):
        nbr, szs = g[f"c{c}_nbr"], [int(s) for s in g[f"c{c}_sizes"]]
        n = nbr.shape[0]
        S = nbr.ravel().astype(np.int64)
        _, first = np.unique(S, return_index=True)
        got = ops.partition_walk(S[np.sort(first)], n, szs)
        assert [int(v) for b in got for v in b] == list(g[f"c{c}_flat"]), c
        assert [len(b) for b in got] == list(g[f"c{c}_lens"]), c


def test_partition_walk_wheat_golden(wheat, golden):
    X, y = wheat
    gk, gp = golden("wheat_knn.npz"), golden("wheat_partition.npz")
    S = gk["idx_with_pheno"].astype(np.int64).ravel()          # k = 10 neighbour stream of the whole frame
    _, first = np.unique(S, return_index=True)
    order = S[np.sort(first)]
    for batches, tag in ((3, "b3"), (5, "b5"), (8, "b8"), (10, "b10"), ((479, 108, 587), "split")):
        sizes = [599 // batches] * batches if isinstance(batches, int) else list(batches)
        got = ops.partition_walk(order, 599, sizes)
        for i, b in enumerate(got):
            assert np.array_equal(b, gp[f"k10_{tag}_{i}"]), (tag, i)


@st.composite
def streams(draw):
    n = draw(st.integers(2, 24))
    k = draw(st.integers(1, min(4, n - 1)))
    rows = []
    for i in range(n):
        others = [v for v in range(n) if v != i]
        rows.append(draw(st.permutations(others))[:k])
    sizes = draw(st.lists(st.integers(0, n + 2), min_size=1, max_size=5))
    return np.asarray(rows, dtype=np.int64), sizes


@settings(max_examples=150, deadline=None)
@given(streams())
def test_naive_partition_equals_the_slow_restatement(case):
    """The drop-in naive_partition (C++ walk) vs the literal rescan restatement of util.py:15-45, with the
    traversed set carried across calls as split_dataset_graph does (util.py:63-68)."""
    nbr, sizes = case
    n = nbr.shape[0]
    S = restate.neighbour_stream(nbr)
    edges = [(i, int(c)) for i in range(n) for c in nbr[i]]
    t1, t2 = set(), set()
    for s in sizes:
        want = restate.partition_scan(S, n, s, t1)
        got = naive_partition(edges, s, bidirectional=False, traversed=set(t2))
        assert got == want, (sizes, s, got, want)
        t1 |= set(want)
        t2 |= set(got)
    assert restate.partition_linear(S, n, sizes) == [restate.partition_scan(S, n, s, set()) if i == 0 else b
                                                      for i, (s, b) in enumerate(zip(sizes, restate.partition_linear(S, n, sizes)))]


def test_naive_partition_bidirectional_stream():
    # util.py:8-11 with bidirectional=True appends target then source to the one shared list
    edges = [(0, 3), (1, 2), (2, 0)]
    assert naive_partition(edges, 3) == [0, 3, 2]          # S = [3, 0, 2, 1, 0, 2]
    assert naive_partition(edges, 4, traversed={0}) == [1, 3, 2]
    assert naive_partition([], 3) == []


@pytest.mark.parametrize("name,kw", [
    ("a", dict(split=3, use_validation=True)),
    ("b", dict(split=2, add_full=True)),
    ("c", dict(split=2, add_full=True, full_split=0.8)),
    ("d", dict(split=2, separate_sets=False, train_size=300)),
    ("e", dict(split=2, separate_sets=False, full_split=400, train_size=0.6)),
])
def test_ensemble_rows_golden(golden, name, kw):
    """load_ensembles2.py:80-132: the host-side row draws equal the reference's (frozen rows) and the oracle's."""
    from gnn_b200.loaders.load_ensembles2 import ensemble_rows
    g = golden("wheat_ensemble.npz")
    seed = int(g[f"{name}_seed"])
    np.random.seed(seed)
    rows, valid = ensemble_rows(599, **kw)
    np.random.seed(seed)
    o_rows, o_valid = restate.ensemble_rows(599, **kw)
    for i, (tr, te) in enumerate(rows):
        assert np.array_equal(tr, g[f"{name}_{i}_rows"]) and np.array_equal(te, g[f"{name}_{i}_test_rows"])
        assert np.array_equal(tr, o_rows[i][0]) and np.array_equal(te, o_rows[i][1])
    if valid is None:
        assert o_valid is None
    else:
        assert np.array_equal(valid, g[f"{name}_valid_rows"]) and np.array_equal(valid, o_valid)
        assert valid[0] == len(valid)          # the validation rows come from the already truncated frame (:82-83)


def test_ensemble_rows_sizes_property():
    from gnn_b200.loaders.load_ensembles2 import ensemble_rows
    for n in (10, 57, 200):
        for split in (1, 4):
            np.random.seed(n)
            rows, valid = ensemble_rows(n, split, train_size=0.7)
            m = round(0.7 * n)
            assert len(rows) == split and valid is None
            for tr, te in rows:
                assert len(tr) == round(0.5 * m) and len(set(tr.tolist())) == len(tr)
                assert np.array_equal(te, np.arange(m, n))


def test_pack_2bit_host_layout():
    """Host packer of the PCIe / HBM genotype format: code j of a row in bits 2*(j%4) of byte j//4, rows padded to 16 bytes."""
    rng = np.random.default_rng(0)
    for n, p in ((1, 1), (3, 5), (7, 64), (5, 130)):
        X = rng.integers(0, 4, size=(n, p)).astype(np.int8)
        P = ops.pack_2bit_host(X)
        assert P.dtype == np.uint8 and P.shape == (n, ops.packed_stride(p)) and ops.packed_stride(p) % 16 == 0
        back = np.stack([(P[:, j // 4] >> (2 * (j % 4))) & 3 for j in range(p)], axis=1)
        assert np.array_equal(back, X)
        assert not P[:, (p + 3) // 4:].any()                       # padding bytes are zero
        if p % 4:
            assert not (P[:, p // 4] >> (2 * (p % 4))).any()       # padding codes inside the last byte are zero
    with pytest.raises(ValueError):
        ops.pack_2bit_host(np.array([[0, 4]], dtype=np.int8))


def test_flat_grads_marks_the_layers_that_write_in_place():
    """parallel.FlatGrads(model=...): the sink layer's parameters come first, then those of the other GCNConv layers and of
    a dense head our kernels take, then what autograd still accumulates into (here: nothing, or a wide Linear)."""
    from gnn_b200 import parallel
    from gnn_b200.networks.gcn_conv import GCNConv
    net = nets.create_network_two_no_conv_dropout(300, 1, internal_size=20, dropout=0.5)
    first = parallel.first_genotype_layer(net)
    flat = parallel.FlatGrads(net.parameters(), sink_layer=first, model=net)
    assert first.grad_sink is flat and not first.grad_inplace
    assert flat.inplace_layers == [net.nns[1], net.nns[3]] and all(m.grad_inplace for m in flat.inplace_layers)
    assert flat.params[0] is first.lin.weight and flat.params[1] is first.bias and flat.tail.numel() == 0
    assert sum(p.numel() for p in flat.params) == flat.flat.numel() == sum(p.numel() for p in net.parameters())
    for p in net.parameters():                                     # every gradient is a view into the flat buffer
        assert flat.flat.data_ptr() <= p.grad.data_ptr() < flat.flat.data_ptr() + 4 * flat.flat.numel()
    for m in flat.inplace_layers:
        m._grad_written = True
    flat.zero()
    assert not any(m._grad_written for m in flat.inplace_layers)
    # a wide first Linear (create_network_test) stays with autograd: its gradients form the zeroed tail
    net = nets.create_network_test(300, 1, internal_size=20)
    flat = parallel.FlatGrads(net.parameters(), sink_layer=parallel.first_genotype_layer(net), model=net)
    assert not flat.has_sink and not net.nns[0].grad_inplace
    assert [type(m) for m in flat.inplace_layers] == [GCNConv, GCNConv, type(net.nns[-1])]
    assert flat.tail.numel() == net.nns[0].weight.numel() + net.nns[0].bias.numel()
    flat.tail.fill_(3.0)
    flat.zero()
    assert float(flat.tail.abs().max()) == 0.0
    # without `model` nothing is marked and everything is zeroed
    net = nets.create_network_two_no_conv_dropout(300, 1, internal_size=20)
    flat = parallel.FlatGrads(net.parameters())
    assert flat.inplace_layers == [] and flat.tail.numel() == flat.flat.numel()


def test_dropout_in_front_of_the_head_stays_with_torch_on_the_host():
    """The fused Dropout -> head path needs CUDA float32 activations; CPU tensors (and eval mode) go through torch's modules
    unchanged, and a head marked for in-place gradients clears its slices itself on that path."""
    from gnn_b200 import parallel
    from gnn_b200.networks.head import Linear
    torch.manual_seed(0)
    net = Sequential("x", [(Linear(12, 6), "x -> x"), torch.nn.Dropout(0.5), Linear(6, 1)])
    x = torch.randn(40, 12)
    assert not net._dropout_feeds_head(1, [x])
    net.eval()
    quiet = net(x)
    assert torch.equal(quiet, net.nns[2](net.nns[0](x)))
    net.train()
    assert not torch.equal(net(x), quiet)                           # torch's dropout ran
    flat = parallel.FlatGrads(net.parameters(), model=net)
    assert flat.inplace_layers == [net.nns[2]]                       # 6 -> 1 is a head our kernels take; 12 -> 6 is not
    flat.flat.fill_(5.0)
    flat.zero()
    net.eval()
    net(x).sum().backward()
    ref = Sequential("x", [(Linear(12, 6), "x -> x"), torch.nn.Dropout(0.5), Linear(6, 1)])
    ref.load_state_dict(net.state_dict())
    ref.eval()
    ref(x).sum().backward()
    for a, b in zip(net.parameters(), ref.parameters()):
        assert torch.allclose(a.grad, b.grad)
