"""The GCN model factories of the reference's gnn/networks/networks.py (the message-passing ones:
SURVEY.md 2.1 #3) and ``Ensemble``.  Same names, arguments and module order, so state_dict keys match.
The CNN / dense / k-hop factories of that file are outside the hot path and are not provided.
"""
import torch
from torch.nn import Dropout, ReLU

from .gcn_conv import GCNConv
from .head import Linear            # torch.nn.Linear whose few-output case runs on our kernels (same parameters and keys)
from .sequential import Sequential

_ARGS = 'x, edge_index, edge_weights?'
_CONV = 'x, edge_index, edge_weights -> x'


def _width(inp_size, internal_size):
    return inp_size if internal_size <= 0 else internal_size


class Ensemble(torch.nn.Module):
    """networks.py:100-127: concatenates the sub-models' outputs and feeds a Linear aggregator.  The
    sub-models sit in a plain list on purpose: like in the reference they are NOT registered, so only
    the aggregator is optimised and saved."""

    def __init__(self, models, in_size=1, out_size=1, aggregator=Linear):
        super().__init__()
        self.nets = models
        self.aggregator = aggregator(len(models) * in_size, out_size)

    def forward(self, *args):
        feats = []
        for net in self.nets:
            r = net(*args)
            feats.append(r.view(r.size(0), -1))
        return self.aggregator(torch.cat(tuple(feats), dim=1))

    def train(self, mode=True, **_):
        for net in self.nets:
            net.train()
        self.aggregator.train()
        return self

    def eval(self):
        for net in self.nets:
            net.eval()
        self.aggregator.eval()
        return self

    def to(self, *args, **kwargs):
        for net in self.nets:
            net.to(*args, **kwargs)
        return super().to(*args, **kwargs)

    def __repr__(self):
        return "Ensemble"

    __str__ = __repr__


def create_network_linear(inp_size, out_size, *_, **__):                                    # networks.py:130-134
    return Sequential(_ARGS, [(Linear(inp_size, out_size), "x -> x")])


def create_network_no_conv_relu(inp_size, out_size, internal_size=-1):                      # networks.py:151-160
    h = _width(inp_size, internal_size)
    return Sequential(_ARGS, [(GCNConv(inp_size, h), _CONV), ReLU(), Linear(h, out_size)])


def create_network_no_relu(inp_size, out_size, internal_size=-1, dropout=0.5, pool_size=2):  # networks.py:162-170
    h = _width(inp_size, internal_size)
    return Sequential(_ARGS, [(GCNConv(inp_size, h), _CONV), Linear(h, out_size)])


def create_network_no_conv_relu_dropout(inp_size, out_size, conv_kernel_size=30, pool_size=2, internal_size=-1,
                                        dropout=0.5):                                       # networks.py:173-184
    h = _width(inp_size, internal_size)
    return Sequential(_ARGS, [(GCNConv(inp_size, h), _CONV), ReLU(), Dropout(dropout), Linear(h, out_size)])


def create_network_two_no_conv_relu_dropout(inp_size, out_size, internal_size=-1, dropout=0.5, add_self_loops=True,
                                            **_):                                           # networks.py:187-205
    h = _width(inp_size, internal_size)
    return Sequential(_ARGS, [(GCNConv(inp_size, h, add_self_loops=add_self_loops), _CONV), (GCNConv(h, h), _CONV),
                              ReLU(), Dropout(dropout), Linear(h, out_size)])


def create_network_two_no_conv_relu(inp_size, out_size, internal_size=-1, add_self_loops=True, **_):   # :208-219
    h = _width(inp_size, internal_size)
    return Sequential(_ARGS, [(GCNConv(inp_size, h, add_self_loops=add_self_loops), _CONV), (GCNConv(h, h), _CONV),
                              ReLU(), Linear(h, out_size)])


def create_network_two_relu(inp_size, out_size, internal_size=-1, add_self_loops=True, **_):           # :221-233
    h = _width(inp_size, internal_size)
    return Sequential(_ARGS, [(GCNConv(inp_size, h, add_self_loops=add_self_loops), _CONV), ReLU(),
                              (GCNConv(h, h), _CONV), ReLU(), Linear(h, out_size)])


def create_network_no_conv_dropout(inp_size, out_size, internal_size=-1, dropout=0.5, add_self_loops=True, **_):  # :254-263
    h = _width(inp_size, internal_size)
    return Sequential(_ARGS, [(GCNConv(inp_size, h, add_self_loops=add_self_loops), _CONV), Dropout(dropout),
                              Linear(h, out_size)])


def create_network_two_no_conv_dropout(inp_size, out_size, internal_size=-1, dropout=0.5, add_self_loops=True, **_):
    """networks.py:266-276 -- the network of the shipped configuration (main.py:509)."""
    h = _width(inp_size, internal_size)
    return Sequential(_ARGS, [(GCNConv(inp_size, h, add_self_loops=add_self_loops), _CONV),
                              (GCNConv(h, h, add_self_loops=add_self_loops), _CONV), Dropout(dropout),
                              Linear(h, out_size)])


def create_ensemble_creator(*models, out_size=1, aggregator=Linear):                         # networks.py:279-292
    def creator(*args, **kwargs):
        nets = [m(*args, **kwargs) for m in models]
        return Sequential(_ARGS, [(Ensemble(nets, out_size=out_size, aggregator=aggregator), _CONV)])
    return creator


def create_network_test(inp_size, out_size, internal_size=-1, dropout=0.5, add_self_loops=True, **_):   # :295-306
    h = _width(inp_size, internal_size)
    return Sequential(_ARGS, [(Linear(inp_size, h), 'x -> x'), (GCNConv(h, h, add_self_loops=add_self_loops), _CONV),
                              (GCNConv(h, h, add_self_loops=add_self_loops), _CONV), Dropout(dropout),
                              Linear(h, out_size)])


def create_network_no_conv(inp_size, out_size, internal_size=-1, add_self_loops=True, **_):  # networks.py:337-345
    h = _width(inp_size, internal_size)
    return Sequential(_ARGS, [(GCNConv(inp_size, h, add_self_loops=add_self_loops), _CONV), Linear(h, out_size)])


def create_network_two_no_conv(inp_size, out_size, internal_size=-1, add_self_loops=True, **_):   # networks.py:348-357
    h = _width(inp_size, internal_size)
    return Sequential(_ARGS, [(GCNConv(inp_size, h, add_self_loops=add_self_loops), _CONV),
                              (GCNConv(h, h, add_self_loops=add_self_loops), _CONV), Linear(h, out_size)])
