"""The trainer drop-ins on the B200 against the reference's own run (golden): loader -> partition ->
per-batch graphs -> train_ensemble2 phase 1 + 2 -> final metrics; plus train / train_batches smoke parity."""
import os

import numpy as np
import pytest
import torch

from gnn_b200.loaders.load import load_data
from gnn_b200.loaders.util import naive_partition, split_dataset_graph
from gnn_b200.networks import networks as nets
from gnn_b200.trainers.batches import train_batches
from gnn_b200.trainers.ensemble2 import train_ensemble2
from gnn_b200.trainers.plain import train

pytestmark = pytest.mark.gpu
LR = 0.0020594745443455593


@pytest.fixture(scope="module")
def wheat_batches(wheat_csv):
    ds = load_data(wheat_csv, num_neighbours=5, smoothing="laplacian", mode="distance", use_validation=True, split=0.8,
                   use_weights=True, algorithm="euclidean", batches=3, split_algorithm=split_dataset_graph,
                   split_algorithm_params={"partition": naive_partition})
    return ds


def test_processed_cache_round_trip(wheat_csv, tmp_path, monkeypatch):
    """load.py:123-124, 283-284: the dataset is cached under processed/<name built from every loader argument>.pt."""
    monkeypatch.chdir(tmp_path)
    kw = dict(num_neighbours=4, smoothing="laplacian", mode="distance", use_validation=True, split=0.8, use_weights=True,
              algorithm="euclidean", batches=3, split_algorithm=split_dataset_graph, split_algorithm_params={"partition": naive_partition})
    a = load_data(wheat_csv, **kw)
    files = os.listdir(tmp_path / "processed")
    assert len(files) == 1 and files[0].startswith("data-") and "split_dataset_graph-partition:naive_partition;" in files[0]
    b = load_data(wheat_csv, **kw)                       # served from the cache
    assert len(a) == len(b) == 4
    for ga, gb in zip(a, b):
        for key in ("x", "y", "edge_index", "edge_weight"):
            assert torch.equal(getattr(ga, key), getattr(gb, key))
        assert torch.equal(ga.test.edge_index, gb.test.edge_index) and gb.x.stride(0) % 16 == 0
    gl = [g for g in b]
    assert gl[0].test.x.data_ptr() == gl[1].test.x.data_ptr()       # the shared test graph is stored once
    assert torch.equal(gl[-1].test.x, gl[-1].x)                      # valid_data.test = valid_data (load.py:236)


def test_loader_structure(wheat_batches, golden):
    g = golden("wheat_train.npz")
    ds = wheat_batches
    assert len(ds) == 4 and ds.num_features == 1279                      # 3 batches + the validation graph
    data = [d for d in ds]
    for b in range(3):
        assert data[b].x.shape == (199, 1279) and data[b].test.x.shape == (108, 1279)
    assert data[-1].x.shape == (12, 1279) and data[-1].test is not None
    assert data[0].edge_weight.dtype == torch.float64 and data[0].edge_index.dtype == torch.int64


def test_train_ensemble2_against_reference(wheat_batches, golden):
    g = golden("wheat_train.npz")
    data = [d for d in wheat_batches]
    H, E = int(g["hidden"]), int(g["epochs"])
    torch.manual_seed(0)                                                  # same draw order as the reference run
    res = train_ensemble2(lambda: nets.create_network_two_no_conv_dropout(1279, 1, internal_size=H, dropout=0.0),
                          lambda m: torch.optim.Adam(m.parameters(), lr=LR), data, epochs=E, aggregate_epochs=2,
                          save_loss=False, print_epochs=False)["basics"]
    for b in range(3):
        assert abs(res[f"min_loss_{b}"] - float(g[f"ens_min_loss_{b}"])) <= 1e-4 * float(g[f"ens_min_loss_{b}"])
        assert res[f"min_epoch_{b}"] == int(g[f"ens_min_epoch_{b}"])
    np.testing.assert_allclose(res["losses"], g["ens_losses"], rtol=1e-4)
    for key in ("best_loss", "aggr_loss", "min_loss_validation"):
        assert abs(res[key] - float(g[f"ens_{key}"])) <= 2e-4 * abs(float(g[f"ens_{key}"])), key
    for key in ("cor_best", "cor_valid"):
        assert abs(res[key] - float(g[f"ens_{key}"])) < 1e-3, key
    np.testing.assert_allclose(res["cors"], g["ens_cors"], atol=1e-3)
    assert abs(res["heritability"]) < 1e-5
    assert res["epoch"] == E and isinstance(res["aggr_model"], nets.Ensemble)
    assert [k for k, _ in res["aggr_model"].named_parameters()] == ["aggregator.weight", "aggregator.bias"]


def test_train_and_train_batches_run_and_learn(wheat_csv):
    ds = load_data(wheat_csv, num_neighbours=5, smoothing="laplacian", mode="distance", use_validation=False, split=0.8,
                   use_weights=True, algorithm="euclidean")
    assert len(ds) == 1
    data = ds[0]
    torch.manual_seed(1)
    out = train(lambda: nets.create_network_no_conv_relu_dropout(1279, 1, internal_size=16, dropout=0.0),
                lambda m: torch.optim.Adam(m.parameters(), lr=0.01), data, epochs=12, save_loss=False)["basics"]
    assert len(out["losses"]) == 12 and len(out["cors"]) == 12 and len(out["dcors"]) == 12
    assert all(np.isfinite(out["losses"])) and 0 <= out["dcors"][-1] <= 1
    ds = load_data(wheat_csv, num_neighbours=3, use_validation=False, split=0.8, algorithm="euclidean", batches=4,
                   split_algorithm=split_dataset_graph, split_algorithm_params={"partition": naive_partition})
    assert len(ds) == 4
    torch.manual_seed(2)
    model = nets.create_network_two_no_conv_relu(1279, 1, internal_size=16)
    opt = torch.optim.Adam(model.parameters(), lr=0.003)
    losses, epoch = train_batches(model, opt, [d for d in ds], epochs=5, save_loss=False, print_steps=False)
    assert epoch == 4 and len(losses) == 5 and all(np.isfinite(losses))


def _golden_model(g, prefix, hidden):
    model = nets.create_network_two_no_conv_dropout(1279, 1, internal_size=hidden, dropout=0.0)
    model.load_state_dict({k.split("__w0__")[1]: torch.from_numpy(g[k]) for k in g.files if k.startswith(prefix + "__w0__")})
    return model


@pytest.mark.parametrize("graphs", [False, True])
def test_train_against_the_reference_trajectory(wheat_csv, golden, graphs, capsys):
    """gnn/trainers/plain.py:8-114 run by the reference on WHEAT (golden): per-epoch train / test losses, Pearson r and
    distance correlation of the test predictions, eager and CUDA-graph replayed."""
    from oracle import restate
    g = golden("wheat_trainers.npz")
    E, H, lr = int(g["epochs"]), int(g["hidden"]), float(g["lr"])
    ds = load_data(wheat_csv, num_neighbours=5, smoothing="laplacian", mode="distance", use_validation=False, split=0.8,
                   use_weights=True, algorithm="euclidean")
    data = ds[0]
    model = _golden_model(g, "plain", H)
    res = train(model, torch.optim.Adam(model.parameters(), lr=lr), data, epochs=E, save_loss=False, cuda_graphs=graphs)["basics"]
    np.testing.assert_allclose(res["losses"], g["plain_test_loss"], rtol=1e-4)
    np.testing.assert_allclose(res["cors"], g["plain_cors"], atol=1e-4)
    np.testing.assert_allclose(res["dcors"], g["plain_dcors"], atol=1e-4)
    printed = [float(line.split("Loss: ")[1].split(",")[0]) for line in capsys.readouterr().out.splitlines() if line.startswith("Epoch:")]
    np.testing.assert_allclose(printed, g["plain_train_loss"], rtol=1e-4, atol=1e-5)
    model.eval()
    from gnn_b200.trainers.util import get_args_tuple
    pred = model(get_args_tuple(data.test)).detach().cpu().numpy()
    # the golden predictions are those of the last epoch's PRE-step weights; one more step moves them very little
    assert restate.pearson(pred, g["plain_pred_test"]) >= 0.999
    assert res["epoch"] == E and res["model"] is model


@pytest.mark.parametrize("graphs", [False, True])
def test_train_batches_against_the_reference_trajectory(wheat_csv, golden, graphs, capsys):
    """gnn/trainers/batches.py:7-81 run by the reference on three partitioned WHEAT subgraphs (golden): every step's train
    and test loss (the model is called without edge weights, batches.py:37) and the per-epoch averages."""
    g = golden("wheat_trainers.npz")
    E, H, lr = int(g["epochs"]), int(g["hidden"]), float(g["lr"])
    ds = load_data(wheat_csv, num_neighbours=5, smoothing="laplacian", mode="distance", use_validation=False, split=0.8,
                   use_weights=True, algorithm="euclidean", batches=3, split_algorithm=split_dataset_graph,
                   split_algorithm_params={"partition": naive_partition})
    data = [d for d in ds]
    assert len(data) == 3
    model = _golden_model(g, "batches", H)
    losses, epoch = train_batches(model, torch.optim.Adam(model.parameters(), lr=lr), data, epochs=E, save_loss=False,
                                  cuda_graphs=graphs)
    assert epoch == E - 1
    np.testing.assert_allclose(losses, g["batches_losses"], rtol=1e-4)
    rows = [line for line in capsys.readouterr().out.splitlines() if line.startswith("Epoch:")]
    got = np.array([[float(r.split("Loss: ")[1].split(",")[0]), float(r.split("Test Loss: ")[1])] for r in rows])
    np.testing.assert_allclose(got, g["batches_steps"], rtol=1e-4, atol=2e-5)
    # early stopping: the rule of batches.py:69-76 on the same trajectory
    model = _golden_model(g, "batches", H)
    ref_losses = list(g["batches_losses"])
    stale, best, stop = 0, 1e8, E - 1
    for i, v in enumerate(ref_losses):
        stale, best = (0, v) if v < best else (stale + 1, best)
        if stale >= 1:
            stop = i
            break
    _, epoch = train_batches(model, torch.optim.Adam(model.parameters(), lr=lr), data, epochs=E, save_loss=False,
                             max_no_improvement=1, print_steps=False, cuda_graphs=graphs)
    assert epoch == stop


def test_distance_correlation_on_the_device(golden):
    """trainers.util.distance_correlation (plain.py:76-79) against the oracle's restatement of dcor's V-statistic."""
    from gnn_b200.trainers.util import distance_correlation
    g = golden("wheat_trainers.npz")
    for name in ("linear", "noise", "square", "const"):
        a = torch.from_numpy(g[f"dcor_{name}_a"]).cuda()
        b = torch.from_numpy(g[f"dcor_{name}_b"]).cuda()
        assert abs(float(distance_correlation(a, b)) - float(g[f"dcor_{name}"])) < 1e-9, name


def test_graphed_step_runs_each_call_once():
    """parallel.GraphedStep: eager warm-up, capture + replay, and every call does the work exactly once."""
    from gnn_b200.parallel import GraphedStep
    acc = torch.zeros(3, device="cuda")

    def body():
        acc.add_(1.0)
        return acc * 2.0

    step = GraphedStep(body, warmup=2)
    outs = [float(step()[0]) for _ in range(6)]
    assert outs == [2.0, 4.0, 6.0, 8.0, 10.0, 12.0] and step.captured and step.error is None
    off = GraphedStep(body, enabled=False)
    assert float(off()[0]) == 14.0 and not off.captured


def test_own_adam_matches_torch_adam_and_captures():
    """gnn_b200.optim.Adam (one launch per step, step count on the device) against torch.optim.Adam on the same gradients:
    parameters agree to rounding over several steps, through a CUDA-graph replay too."""
    from gnn_b200.optim import Adam
    from gnn_b200.parallel import GraphedStep, make_capturable
    torch.manual_seed(3)
    shapes = [(100, 5003), (100,), (100, 100), (1, 100), (1,)]
    mine = [torch.nn.Parameter(torch.randn(s, device="cuda") * 0.1) for s in shapes]
    ref = [torch.nn.Parameter(p.detach().clone()) for p in mine]
    o_mine, o_ref = Adam(mine, lr=2e-3), torch.optim.Adam(ref, lr=2e-3)
    assert make_capturable(o_mine)
    grads = [[torch.randn(s, device="cuda") * (10.0 ** -k) for s in shapes] for k in range(6)]
    for p in mine:
        p.grad = torch.zeros_like(p)
    cur = {"k": 0}

    def body():
        o_mine.step()
        return mine[0]

    step = GraphedStep(body, warmup=2)
    for k in range(6):
        for p, q, g in zip(mine, ref, grads[k]):
            p.grad.copy_(g)
            q.grad = g.clone()
        step()
        o_ref.step()
        for p, q in zip(mine, ref):
            assert float((p - q).abs().max()) <= 2e-6 * float(q.abs().max()) + 1e-9, k
    assert step.captured, step.error
    sd = o_mine.state_dict()
    assert float(sd["state"][0]["step"]) == 6.0
