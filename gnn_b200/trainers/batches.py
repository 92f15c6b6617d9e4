"""Minibatch trainer over subsampled subgraphs -- gnn/trainers/batches.py:7-81 of the reference: one
model, one optimiser step per subgraph, per-epoch averages of the train / test loss.  Like the
reference it calls the model without edge weights (batches.py:37)."""
import time

import torch

from .util import save_csv, to_host


def train_batches(model, optimizer, data, test_case="default", epochs=1000, max_no_improvement=-1, l1_lambda=0.01,
                  use_l1_reg=False, save_loss=True, print_steps=True, **_):
    if callable(model) and not isinstance(model, torch.nn.Module):
        model = model()
        optimizer = optimizer(model)
    if not torch.cuda.is_available():
        raise RuntimeError("gnn_b200 trainers need a CUDA device (no CPU fallback)")
    device = torch.device("cuda")
    model.to(device)
    for datum in data:
        datum.to(device)
    loss_func = torch.nn.MSELoss()
    losses, train_losses = [], []
    no_improvement, least_loss, epoch = 0, 100000000, 0
    test_name = f"{time.time()}-{test_case}"
    for epoch in range(epochs):
        losses.append(0)
        train_losses.append(0)
        for batch_no, datum in enumerate(data):
            model.train()
            optimizer.zero_grad()
            out = model((datum.x, datum.edge_index))
            model.eval()
            if not hasattr(datum, "range"):
                y = datum.y
                if len(datum.test.x) > 0:
                    out_test = model((datum.test.x, datum.test.edge_index))
                    test_y = datum.test.y
                else:
                    out_test = model((datum.x, datum.edge_index))
                    test_y = datum.y
            else:                                              # batches.py:48-52
                out_test = out[datum.test.range.start: datum.test.range.stop]
                out = out[datum.range.start: datum.range.stop]
                y = datum.y[datum.range.start: datum.range.stop]
                test_y = datum.y[datum.test.range.start: datum.test.range.stop]
            loss = loss_func(out, y)
            if use_l1_reg:
                loss = model.l1_regularize(loss, l1_lambda)
            loss.backward()
            optimizer.step()
            with torch.no_grad():
                tl, vl = to_host(loss, loss_func(out_test, test_y))
            losses[-1] += vl / len(data)
            train_losses[-1] += tl / len(data)
            if print_steps:
                print('Epoch: {:03d}:{:03d}, Loss: {:.5f}, Test Loss: {:.5f}'.format(epoch, batch_no, tl, vl))
        if losses[-1] < least_loss:
            no_improvement, least_loss = 0, losses[-1]
        else:
            no_improvement += 1
        if 0 < max_no_improvement <= no_improvement:
            break
    if save_loss:
        save_csv({"train_loss": train_losses, "loss": losses}, f"output/{test_name}-train-data.csv")
    return losses, epoch
