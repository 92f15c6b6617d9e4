"""The "subsampling GNN" trainer -- gnn/trainers/ensemble2.py:14-256 of the reference.

Phase 1 (:39-124): one fresh model + optimiser per subgraph batch; every epoch is a train-mode forward on
the subgraph, an eval-mode forward on the test graph, MSE (+L1), backward, Adam step; the weights of the
epoch with the lowest test loss are kept.  Phase 2 (:125-212): the best sub-models, heads replaced by
Identity, feed a Linear aggregator trained on the last subgraph.  Final metrics (:214-256) on the
validation graph.  Same arguments and same returned dictionary.

What differs from the reference, deliberately: everything runs on the GPU; best weights are kept in
device memory instead of being written to model/ensemble/ on every improvement; per-epoch metrics are
reduced on the device and fetched with one small copy per epoch; `corrs` holds the plain Pearson r (the
reference's per-epoch value goes through a precedence slip at :78 and is not a metric of record).
"""
import copy
import time

import numpy as np
import torch
from torch.nn import Identity

from ..networks.networks import Ensemble
from ..parallel import GraphedStep, make_capturable
from .util import get_args_tuple, l1_regularize, pearson, save_csv, to_host


def _clone_state(model):
    return {k: v.detach().clone() for k, v in model.state_dict().items()}


def _keep_state(model, into):
    """Copy the current weights into the preallocated ``into`` (no allocation, no host work)."""
    with torch.no_grad():
        for k, v in model.state_dict().items():
            into[k].copy_(v)


class _Epoch:
    """The epoch body of phase 1 (ensemble2.py:51-88) as one fixed piece of device work; its three scalars (train loss,
    test loss, Pearson r of the test predictions) go to ``stats``."""

    def __init__(self, model, optimizer, batch, loss_func, l1):
        self.model, self.optimizer, self.loss_func, self.l1 = model, optimizer, loss_func, l1
        self.args, self.test_args = get_args_tuple(batch), get_args_tuple(batch.test)
        self.y, self.test_y = batch.y, batch.test.y
        self.stats = torch.zeros(3, dtype=torch.float64, device=batch.y.device)

    def __call__(self):
        model = self.model
        model.train()
        self.optimizer.zero_grad(set_to_none=False)
        out = model(*self.args)
        model.eval()
        with torch.no_grad():
            out_test = model(*self.test_args)
        loss = self.loss_func(out, self.y)
        if self.l1 is not None:
            loss = l1_regularize(model, loss, self.l1)
        loss.backward()
        self.optimizer.step()
        with torch.no_grad():
            self.stats[0] = loss.detach()
            self.stats[1] = self.loss_func(out_test, self.test_y)
            self.stats[2] = pearson(out_test, self.test_y)
        return self.stats


def train_ensemble2(model, optimizer, data, test_case="default", epochs=1000, aggregate_epochs=None,
                    max_no_improvement=-1, l1_lambda=0.01, use_l1_reg=False, save_loss=True,
                    early_stopper=lambda _, __, ___: None, print_epochs=True, cuda_graphs="auto", **_):
    if not torch.cuda.is_available():
        raise RuntimeError("gnn_b200 trainers need a CUDA device (no CPU fallback)")
    device = torch.device("cuda")
    test_name = f"{time.time()}-{test_case}"
    if aggregate_epochs is None:
        aggregate_epochs = epochs
    orig_model, orig_optimizer = model, optimizer
    data = [d.to(device) for d in data]
    loss_func = torch.nn.MSELoss()
    best_states, losses_dict = [], {}
    min_loss_so_far, best_model_so_far = 1000000000, None

    # ---------------------------------------------------------------- phase 1: one model per subgraph
    for i_batch, batch in enumerate(data[:-1]):
        model = orig_model().to(device)
        optimizer = orig_optimizer(model)
        no_improvement, least_loss = 0, 100000000
        losses, train_losses, corrs = [], [], []
        best_states.append(_clone_state(model))
        for p in model.parameters():
            if p.requires_grad and p.grad is None:
                p.grad = torch.zeros_like(p)
        graphs_on = (cuda_graphs is True or (cuda_graphs == "auto" and epochs > 3)) and make_capturable(optimizer)
        step = GraphedStep(_Epoch(model, optimizer, batch, loss_func, l1_lambda if use_l1_reg else None),
                           warmup=2, enabled=graphs_on)
        for epoch in range(epochs):
            tl, vl, r = step().cpu().tolist()                  # the epoch's only device -> host copy
            train_losses.append(tl)
            losses.append(vl)
            corrs.append(r)
            if vl < least_loss:                                # ensemble2.py:92-95 saves the post-step weights
                no_improvement, least_loss = 0, vl
                _keep_state(model, best_states[i_batch])
            else:
                no_improvement += 1
            if vl < min_loss_so_far:
                min_loss_so_far, best_model_so_far = vl, model
            if 0 < max_no_improvement <= no_improvement:
                break
            if print_epochs:
                print('Batch: {:02d}, Epoch: {:03d}, Loss: {:.5f}, Test Loss: {:.5f}'.format(i_batch, epoch, tl, vl))
        early_stopper(min(losses), i_batch, lambda: None)
        losses_dict[i_batch] = losses
        if save_loss:
            save_csv({"train_loss": train_losses, "loss": losses, "corr": corrs},
                     f"output/{test_name}-train-ensemble2-{i_batch}-data.csv")

    losses_and_epochs = {}
    for k, k_losses in losses_dict.items():
        losses_and_epochs[f"min_loss_{k}"] = min(k_losses)
        losses_and_epochs[f"min_epoch_{k}"] = k_losses.index(losses_and_epochs[f"min_loss_{k}"])

    # ---------------------------------------------------------------- phase 2: aggregator over frozen sub-models
    valid = data[-1]
    valid_tuple = get_args_tuple(valid)
    models, final_outs, size = [], [], 1
    for state in best_states:
        m = orig_model().to(device)
        m.load_state_dict(state)
        m.eval()
        seen = {}
        handle = m.get(-2).register_forward_hook(lambda _m, _i, output: seen.__setitem__("w", output.shape[1]))
        with torch.no_grad():
            final_outs.append(m(*valid_tuple).reshape(-1))
        handle.remove()
        size = seen.get("w", size)
        m.replace(-1, Identity())
        models.append(m)
    aggr_model = Ensemble(models, size).to(device)
    optimizer = orig_optimizer(aggr_model)
    aggr_losses, aggr_train_losses, corrs = [], [], []
    train_data = data[-2]
    least_loss, best_aggr = 1000000000000, copy.deepcopy(aggr_model.state_dict())
    tr_tuple, te_tuple = get_args_tuple(train_data), get_args_tuple(train_data.test)
    for aggr_epoch in range(aggregate_epochs):
        aggr_model.train()
        optimizer.zero_grad()
        out = aggr_model(*tr_tuple)
        aggr_model.eval()
        out_test = aggr_model(*te_tuple)
        loss = loss_func(out, train_data.y)
        plain_loss = loss
        if use_l1_reg:
            loss = l1_regularize(aggr_model, loss, l1_lambda)
        with torch.no_grad():
            tl, ll, vl, r = to_host(plain_loss, loss, loss_func(out_test, train_data.test.y),
                                    pearson(out_test, train_data.test.y))
        aggr_train_losses.append(tl)
        if vl < least_loss:                                    # saved BEFORE the step (ensemble2.py:190-195)
            least_loss = vl
            best_aggr = copy.deepcopy(aggr_model.state_dict())
        loss.backward()
        optimizer.step()
        aggr_losses.append(vl)
        corrs.append(r)
        if print_epochs:
            print('Batch: aggregation, Epoch: {:03d}, Loss: {:.5f}, Test Loss: {:.5f}'.format(aggr_epoch, ll, vl))
    aggr_model.load_state_dict(best_aggr)

    # ---------------------------------------------------------------- final metrics on the validation graph
    with torch.no_grad():
        yv = valid.y.flatten().double()
        pred_val = best_model_so_far(*valid_tuple)
        aggr_val = aggr_model(*valid_tuple)

        def mse_broadcast(p):
            # the reference compares [n, 1] predictions with [n] targets (ensemble2.py:221-222), which
            # broadcasts to all pairs: mean_ij (p_i - y_j)^2
            p = p.reshape(-1).double()
            return (p * p).mean() - 2.0 * p.mean() * yv.mean() + (yv * yv).mean()

        avg_val = torch.stack(final_outs).double().mean(dim=0)
        err = pred_val.double() - yv                           # [n, 1] - [n] -> [n, n], as in the reference (:234)
        n = yv.numel()
        var_phen = yv.var(unbiased=True)
        if n <= 8192:
            var_err = torch.from_numpy(np.atleast_2d(np.cov(err.cpu().numpy()))).to(device)
            her = ((var_phen - var_err) / var_phen).mean()
        else:
            # np.cov of the [n, n] error matrix: row i has variance var(y) and covariance var(y) with every
            # other row, so every entry equals var_phen and the "heritability" of :233-237 is 0
            her = torch.zeros((), dtype=torch.float64, device=device)
        best_loss, aggr_loss, loss, avg_corr, best_corr, her = to_host(
            mse_broadcast(pred_val), mse_broadcast(aggr_val), ((avg_val - yv) ** 2).mean(),
            pearson(avg_val, yv), pearson(pred_val, yv), her)
    return {
        "basics": {
            "losses": list(sorted(losses_dict.values(), key=lambda x: min(x))[-1]),
            "epoch": epochs,
            "best_loss": best_loss,
            "aggr_loss": aggr_loss,
            "min_loss_validation": loss,
            "model": best_model_so_far,
            "aggr_model": aggr_model,
            "cors": corrs,
            "cor_best": best_corr,
            "cor_valid": avg_corr,
            "heritability": her,
            **losses_and_epochs,
        }
    }
