"""gnn/trainers/util.py of the reference: the model call tuple and the generic L1 penalty."""
import numpy as np
import torch


def l1_regularize(model, loss, l1_lambda):
    """trainers/util.py:4-13."""
    if hasattr(model, "l1_regularize"):
        return model.l1_regularize(loss, l1_lambda)
    flat = torch.cat([p.view(-1) for p in model.parameters()])
    return loss + l1_lambda * torch.abs(flat).sum().float()


def _features(x):
    # the reference calls x.float() (trainers/util.py:19); int8 genotype codes and 2-bit packed codes (uint8) are
    # consumed as they are
    return x if x.dtype in (torch.int8, torch.uint8) else x.float()


def _weights32(ew):
    """``data.edge_weight.float()`` (trainers/util.py:19), converted once per weight tensor and kept ON that tensor:
    the loaders hand out float64 weights, and a fresh float32 copy per call would defeat the gcn_norm / CSR cache
    (ops.gcn_graph is keyed on tensor identity) and rebuild both CSRs on every forward."""
    if ew.dtype == torch.float32:
        return ew
    memo = getattr(ew, "_gnnb_f32", None)
    if memo is None or memo[0] != ew._version:
        memo = (ew._version, ew.float())
        ew._gnnb_f32 = memo
    return memo[1]


def get_args_tuple(data):
    """trainers/util.py:16-24: (x, edge_index, edge_weight) / (x, edge_index) / x."""
    if getattr(data, "edge_index", None) is not None:
        if getattr(data, "edge_weight", None) is not None:
            return _features(data.x), data.edge_index, _weights32(data.edge_weight)
        return data.x, data.edge_index
    return data.x


def pearson(a, b):
    """Pearson r of two device vectors as a 0-d float64 tensor (np.corrcoef(...)[0][1] in the reference)."""
    a = a.detach().reshape(-1).double()
    b = b.detach().reshape(-1).double()
    a = a - a.mean()
    b = b - b.mean()
    return (a @ b) / torch.sqrt((a @ a) * (b @ b))


def distance_correlation(a, b, max_n=16384):
    """dcor.distance_correlation of plain.py:76-79 (biased V-statistic estimator) on the device."""
    a = a.detach().reshape(-1, 1).double()
    b = b.detach().reshape(-1, 1).double()
    if a.shape[0] > max_n:
        return torch.full((), float("nan"), device=a.device, dtype=torch.float64)      # no host copy: safe under graph capture

    def centred(v):
        d = (v - v.t()).abs()
        return d - d.mean(0, keepdim=True) - d.mean(1, keepdim=True) + d.mean()

    A, B = centred(a), centred(b)
    dcov2 = (A * B).mean()
    dvar = torch.sqrt((A * A).mean() * (B * B).mean())
    r2 = torch.where(dvar > 0, torch.clamp(dcov2, min=0.0) / dvar.clamp_min(1e-300), torch.zeros_like(dvar))
    return torch.sqrt(r2)                                       # 0 for a constant vector, as dcor returns


def to_host(*scalars):
    """One device->host copy for a handful of 0-d tensors."""
    return [float(v) for v in torch.stack([s.detach().double().reshape(()) for s in scalars]).cpu().numpy()]


def save_csv(columns, path):
    import os
    import pandas as pd
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    pd.DataFrame(columns).to_csv(path)
