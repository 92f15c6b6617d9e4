"""``train`` -- the full-batch trainer (gnn/trainers/plain.py:8-114 of the reference) on the device.

Same arguments and the same returned dictionary.  One epoch is one fixed piece of device work -- train-mode forward with
edge weights (plain.py:38-44), eval-mode forward on the test graph (:47-63) or the ``range`` slices of one graph
(:65-69), Pearson r (:71-74) and distance correlation (:76-79) of the test predictions, MSE (+L1), backward, optimiser
step -- whose four scalars land in a device vector; it is replayed as a CUDA graph (``parallel.GraphedStep``) and the
host reads the four scalars once per epoch for the log line and the early-stopping rule.
"""
import time

import torch

from ..parallel import GraphedStep, make_capturable
from .util import distance_correlation, get_args_tuple, pearson, save_csv


class _Epoch:
    def __init__(self, model, optimizer, data, loss_func, l1):
        self.model, self.optimizer, self.loss_func, self.l1 = model, optimizer, loss_func, l1
        self.args = get_args_tuple(data)
        self.stats = torch.zeros(4, dtype=torch.float64, device=data.x.device)     # train loss, test loss, r, dcor
        if hasattr(data, "range"):
            self.rows, self.test_rows, self.test_args = data.range, data.test.range, None
            self.y = data.y[self.rows.start:self.rows.stop]
            self.test_y = data.y[self.test_rows.start:self.test_rows.stop]
        else:
            self.rows = self.test_rows = None
            self.y = data.y
            if len(data.test.x) > 0:
                self.test_args, self.test_y = get_args_tuple(data.test), data.test.y
            else:
                self.test_args, self.test_y = self.args, data.y

    def __call__(self):
        model = self.model
        model.train()
        self.optimizer.zero_grad(set_to_none=False)
        out = model(self.args)
        model.eval()
        if self.test_args is None:
            out_test, out = out[self.test_rows.start:self.test_rows.stop], out[self.rows.start:self.rows.stop]
        else:
            with torch.no_grad():
                out_test = model(self.test_args)
        loss = self.loss_func(out, self.y)
        if self.l1 is not None:
            loss = model.l1_regularize(loss, self.l1)
        loss.backward()
        self.optimizer.step()
        with torch.no_grad():
            pred = out_test.detach()
            self.stats[0] = loss.detach()
            self.stats[1] = self.loss_func(pred, self.test_y)
            self.stats[2] = pearson(pred, self.test_y)
            self.stats[3] = distance_correlation(pred, self.test_y)
        return self.stats


def train(model, optimizer, data, test_case="default", epochs=1000, max_no_improvement=-1, l1_lambda=0.01,
          use_l1_reg=False, save_loss=True, cuda_graphs="auto", **_):
    if callable(model) and not isinstance(model, torch.nn.Module):
        model = model()
        optimizer = optimizer(model)
    if not torch.cuda.is_available():
        raise RuntimeError("gnn_b200 trainers need a CUDA device (no CPU fallback)")
    device = torch.device("cuda")
    data.to(device)                                            # plain.py:33-34
    model.to(device)
    for p in model.parameters():
        if p.requires_grad and p.grad is None:
            p.grad = torch.zeros_like(p)
    graphs_on = (cuda_graphs is True or (cuda_graphs == "auto" and epochs > 3)) and make_capturable(optimizer)
    body = _Epoch(model, optimizer, data, torch.nn.MSELoss(), l1_lambda if use_l1_reg else None)
    step = GraphedStep(body, warmup=2, enabled=graphs_on)
    history = {"train_loss": [], "loss": [], "corr": [], "dcor": []}
    stale, best = 0, 100000000
    test_name = f"{time.time()}-{test_case}"
    for epoch in range(epochs):
        tl, vl, r, dr = step().cpu().tolist()                  # the epoch's only device -> host copy
        for key, val in zip(("train_loss", "loss", "corr", "dcor"), (tl, vl, r, dr)):
            history[key].append(val)
        print('Epoch: {:03d}, Loss: {:.5f}, Test Loss: {:.5f}'.format(epoch, tl, vl))
        if vl < best:
            stale, best = 0, vl
        else:
            stale += 1
        if 0 < max_no_improvement <= stale:
            break
    if save_loss:
        save_csv({k: history[k] for k in ("train_loss", "loss", "corr")}, f"output/{test_name}-train-data.csv")
    return {"basics": {"model": model, "losses": history["loss"], "epoch": epochs, "cors": history["corr"],
                       "dcors": history["dcor"]}}
