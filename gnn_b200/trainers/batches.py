"""``train_batches`` -- the minibatch trainer over subsampled subgraphs (gnn/trainers/batches.py:7-81 of the reference:
one model, one optimiser step per subgraph, per-epoch averages of the train and the test loss; the model is called
WITHOUT edge weights, batches.py:37).

Same arguments and return value; the execution is organised for the GPU instead of following the reference's loop:

* every subgraph gets a prepared *step* -- its argument tuples, targets and a two-float slot of a device-resident ledger
  -- and the step body (train-mode forward, eval-mode forward on the test graph, MSE (+L1), backward, optimiser step,
  both losses written to the ledger) does no host work at all;
* steps are replayed as CUDA graphs once they have run a couple of times (``parallel.GraphedStep``): a WHEAT-sized
  subgraph step is ~100 kernel launches of a few microseconds each and would otherwise be bound by the host;
* the host reads the ledger ONCE per epoch (one small D2H copy), forms the averages, prints the per-step lines the
  reference prints (after the epoch instead of during it) and applies the early-stopping rule.
"""
import time

import torch

from ..parallel import GraphedStep, make_capturable
from .util import save_csv


class _Step:
    """The device work of one subgraph visit; results go to ``ledger[slot] = (train loss, test loss)``."""

    def __init__(self, model, optimizer, datum, ledger, slot, loss_func, l1):
        self.model, self.optimizer, self.loss_func, self.l1 = model, optimizer, loss_func, l1
        self.out_slot = ledger[slot]
        self.args = (datum.x, datum.edge_index)
        if hasattr(datum, "range"):                             # batches.py:48-52: train / test rows of ONE graph
            self.rows, self.test_rows = datum.range, datum.test.range
            self.y, self.test_y = datum.y[self.rows.start:self.rows.stop], datum.y[self.test_rows.start:self.test_rows.stop]
            self.test_args = None
        else:
            self.rows = self.test_rows = None
            self.y = datum.y
            if len(datum.test.x) > 0:                           # batches.py:41-43
                self.test_args, self.test_y = (datum.test.x, datum.test.edge_index), datum.test.y
            else:                                               # batches.py:44-46
                self.test_args, self.test_y = self.args, datum.y

    def __call__(self):
        model = self.model
        model.train()
        self.optimizer.zero_grad(set_to_none=False)
        out = model(self.args)
        model.eval()
        if self.test_args is None:
            out_test = out[self.test_rows.start:self.test_rows.stop]
            out = out[self.rows.start:self.rows.stop]
        else:
            with torch.no_grad():
                out_test = model(self.test_args)
        loss = self.loss_func(out, self.y)
        if self.l1 is not None:
            loss = model.l1_regularize(loss, self.l1)
        loss.backward()
        self.optimizer.step()
        with torch.no_grad():
            self.out_slot[0] = loss.detach()
            self.out_slot[1] = self.loss_func(out_test.detach(), self.test_y)
        return self.out_slot


def train_batches(model, optimizer, data, test_case="default", epochs=1000, max_no_improvement=-1, l1_lambda=0.01,
                  use_l1_reg=False, save_loss=True, print_steps=True, cuda_graphs="auto", **_):
    """Returns ``(losses, epoch)`` like the reference: the per-epoch mean test loss and the index of the last epoch run.
    ``cuda_graphs``: "auto" (replay captured steps when the optimiser allows it), True or False."""
    if callable(model) and not isinstance(model, torch.nn.Module):
        model = model()
        optimizer = optimizer(model)
    if not torch.cuda.is_available():
        raise RuntimeError("gnn_b200 trainers need a CUDA device (no CPU fallback)")
    device = torch.device("cuda")
    model.to(device)
    subgraphs = [d.to(device) for d in data]
    if not subgraphs:
        raise ValueError("train_batches needs at least one subgraph")
    for p in model.parameters():                                # gradients exist before the first step: zero_grad can keep them
        if p.requires_grad and p.grad is None:
            p.grad = torch.zeros_like(p)
    ledger = torch.zeros((len(subgraphs), 2), dtype=torch.float32, device=device)
    loss_func = torch.nn.MSELoss()
    graphs_on = cuda_graphs is True or (cuda_graphs == "auto" and epochs > 3)
    graphs_on = graphs_on and make_capturable(optimizer)
    pool = torch.cuda.graph_pool_handle() if graphs_on else None
    steps = [GraphedStep(_Step(model, optimizer, d, ledger, i, loss_func, l1_lambda if use_l1_reg else None),
                         warmup=2, pool=pool, enabled=graphs_on) for i, d in enumerate(subgraphs)]
    n = len(steps)
    test_name = f"{time.time()}-{test_case}"
    losses, train_losses = [], []
    stale, best, epoch = 0, 100000000, 0
    for epoch in range(epochs):
        for step in steps:
            step()
        host = ledger.cpu()                                     # the epoch's only device -> host copy
        train_losses.append(sum(float(host[i, 0]) / n for i in range(n)))
        losses.append(sum(float(host[i, 1]) / n for i in range(n)))
        if print_steps:
            for i in range(n):
                print('Epoch: {:03d}:{:03d}, Loss: {:.5f}, Test Loss: {:.5f}'.format(epoch, i, float(host[i, 0]), float(host[i, 1])))
        if losses[-1] < best:
            stale, best = 0, losses[-1]
        else:
            stale += 1
        if 0 < max_no_improvement <= stale:
            break
    if save_loss:
        save_csv({"train_loss": train_losses, "loss": losses}, f"output/{test_name}-train-data.csv")
    return losses, epoch
