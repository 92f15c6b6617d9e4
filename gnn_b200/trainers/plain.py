"""Full-batch trainer -- gnn/trainers/plain.py:8-114 of the reference, on the device."""
import time

import torch

from .util import distance_correlation, get_args_tuple, pearson, save_csv, to_host


def train(model, optimizer, data, test_case="default", epochs=1000, max_no_improvement=-1, l1_lambda=0.01,
          use_l1_reg=False, save_loss=True, **_):
    if callable(model) and not isinstance(model, torch.nn.Module):
        model = model()
        optimizer = optimizer(model)
    if not torch.cuda.is_available():
        raise RuntimeError("gnn_b200 trainers need a CUDA device (no CPU fallback)")
    device = torch.device("cuda")
    loss_func = torch.nn.MSELoss()
    losses, train_losses, corrs, dcorrs = [], [], [], []
    no_improvement, least_loss = 0, 100000000
    test_name = f"{time.time()}-{test_case}"
    data.to(device)                                            # plain.py:33-34
    model.to(device)
    for epoch in range(epochs):
        model.train()
        optimizer.zero_grad()
        out = model(get_args_tuple(data))
        model.eval()
        if not hasattr(data, "range"):
            y = data.y
            if len(data.test.x) > 0:
                out_test = model(get_args_tuple(data.test))
                test_y = data.test.y
            else:
                out_test = model(get_args_tuple(data))
                test_y = data.y
        else:                                                  # plain.py:65-69
            out_test = out[data.test.range.start: data.test.range.stop]
            out = out[data.range.start: data.range.stop]
            y = data.y[data.range.start: data.range.stop]
            test_y = data.y[data.test.range.start: data.test.range.stop]
        loss = loss_func(out, y)
        if use_l1_reg:
            loss = model.l1_regularize(loss, l1_lambda)
        loss.backward()
        optimizer.step()
        with torch.no_grad():
            tl, vl, r, dr = to_host(loss, loss_func(out_test, test_y), pearson(out_test, test_y),
                                    distance_correlation(out_test, test_y))
        losses.append(vl)
        train_losses.append(tl)
        corrs.append(r)
        dcorrs.append(dr)
        print('Epoch: {:03d}, Loss: {:.5f}, Test Loss: {:.5f}'.format(epoch, tl, vl))
        if vl < least_loss:
            no_improvement, least_loss = 0, vl
        else:
            no_improvement += 1
        if 0 < max_no_improvement <= no_improvement:
            break
    if save_loss:
        save_csv({"train_loss": train_losses, "loss": losses, "corr": corrs}, f"output/{test_name}-train-data.csv")
    return {"basics": {"model": model, "losses": losses, "epoch": epochs, "cors": corrs, "dcors": dcorrs}}
