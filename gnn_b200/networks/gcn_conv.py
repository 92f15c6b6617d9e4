"""GCNConv with the surface of torch_geometric 2.1.0's operator as the reference uses it
(gnn/networks/networks.py:5, 142-866): ``GCNConv(in_channels, out_channels, add_self_loops=True)``,
parameters ``lin.weight [out, in]`` (Glorot uniform) and ``bias [out]`` (zeros),
``forward(x, edge_index, edge_weight=None)``.  Semantics restated in SURVEY.md A.1.

Forward and backward run as CUDA kernels from libgnn_b200.so:
  gcn_norm + CSR (once per edge list, cached)  -> graph.cu
  H = x . W^T, dW = dH^T . x, dx = dH . W      -> linear_i8.cu / linear_tc.cu / linear.cu
                                                  (x: float32, int8 genotype codes, or uint8 = 2-bit packed codes,
                                                   four per byte as ops.pack_2bit writes them)
  out = A_hat H + b, dH = A_hat^T dOut         -> spmm.cu
"""
import math

import torch

from .. import ops


class _GCNConvFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, graph, layer):
        h = ops.linear_fwd(x, weight)
        out = ops.spmm(graph.rowptr, graph.col, graph.val, h, bias, n=graph.n, hub=graph.hub)
        ctx.graph = graph
        ctx.layer = layer
        ctx.has_bias = bias is not None
        ctx.x_needs_grad = x.is_floating_point() and x.requires_grad
        ctx.save_for_backward(x, weight)
        return out

    @staticmethod
    def backward(ctx, d_out):
        x, weight = ctx.saved_tensors
        g = ctx.graph
        d_out = d_out.contiguous()
        sink = getattr(ctx.layer, "grad_sink", None)
        # parallel.FlatGrads: this layer's dW / db are written straight into their slices of the flat gradient buffer.  The
        # first layer has the sink itself (it also starts the all-reduces); other layers are marked ``grad_inplace`` and
        # overwrite on their first backward of a step (a second use in the same step accumulates through autograd as usual)
        first_write = getattr(ctx.layer, "grad_inplace", False) and not ctx.layer._grad_written
        sunk = ((sink is not None or first_write) and ctx.has_bias and weight.grad is not None
                and ctx.layer.bias.grad is not None)
        if sunk and sink is None:
            ctx.layer._grad_written = True
        # the bias gradient (column sums of dOut) and the aggregation of dOut are independent and each fills only part of
        # the GPU: they run side by side, and so do the weight and the input gradient of an inner layer
        bias_job = None
        if ctx.has_bias:
            into = ctx.layer.bias.grad if sunk else None
            bias_job = ops.Fork(lambda: ops.colsum(d_out, out=into), reads=(d_out, into))
        d_h = ops.spmm(g.rowptr_t, g.col_t, g.val_t, d_out, None, n=g.n, hub=g.hub_t)     # transposed CSR
        if sunk and sink is not None and not x.is_floating_point():
            # data-parallel training (parallel.FlatGrads): dW and db go straight into the flat gradient buffer and their
            # all-reduces start here; autograd gets None for both and has nothing to accumulate
            bias_job.join()
            sink.write(ctx.layer, x, d_h)
            return None, None, None, None, None
        w_into = weight.grad if sunk else None
        if ctx.x_needs_grad:
            w_job = ops.Fork(lambda: ops.linear_bwd_w(x, d_h, p=weight.shape[1], out=w_into), reads=(x, d_h, w_into), slot=1)
            d_x = ops.linear_bwd_x(d_h, weight)
            d_w = w_job.join()
        else:
            d_w, d_x = ops.linear_bwd_w(x, d_h, p=weight.shape[1], out=w_into), None
        d_bias = bias_job.join() if bias_job is not None else None
        if sunk:        # float features / inner layers: same contract (the slices are overwritten, not accumulated into)
            return d_x, None, None, None, None
        return d_x, d_w, d_bias, None, None


class _Lin(torch.nn.Module):
    """Holds ``weight`` so that the parameter is named ``lin.weight`` like PyG's."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.weight = torch.nn.Parameter(torch.empty(out_channels, in_channels))
        self.reset_parameters()

    def reset_parameters(self):
        a = math.sqrt(6.0 / (self.in_channels + self.out_channels))
        with torch.no_grad():
            self.weight.uniform_(-a, a)

    def forward(self, x):
        return ops.linear_fwd(x, self.weight)


class GCNConv(torch.nn.Module):
    def __init__(self, in_channels, out_channels, improved=False, cached=False, add_self_loops=True,
                 normalize=True, bias=True, **_):
        super().__init__()
        if improved or not normalize:
            raise NotImplementedError("only the improved=False, normalize=True GCNConv of the reference is implemented")
        self.in_channels, self.out_channels = in_channels, out_channels
        self.add_self_loops = add_self_loops
        self.grad_sink = None           # set by parallel.FlatGrads: takes dW / db of this layer directly (see _GCNConvFn.backward)
        self.grad_inplace, self._grad_written = False, False    # ditto, for the layers behind the first one
        self.lin = _Lin(in_channels, out_channels)
        if bias:
            self.bias = torch.nn.Parameter(torch.zeros(out_channels))
        else:
            self.register_parameter("bias", None)

    def reset_parameters(self):
        self.lin.reset_parameters()
        if self.bias is not None:
            with torch.no_grad():
                self.bias.zero_()

    def forward(self, x, edge_index, edge_weight=None):
        if not x.is_cuda:
            raise RuntimeError("gnn_b200.GCNConv runs on CUDA tensors only (no CPU fallback)")
        graph = ops.gcn_graph(edge_index, edge_weight, x.shape[0], self.add_self_loops)
        return _GCNConvFn.apply(x, self.lin.weight, self.bias, graph, self)

    def __repr__(self):
        return f"GCNConv({self.in_channels}, {self.out_channels})"
