"""``Linear`` for the dense head of the GCN factories (gnn/networks/networks.py:273-274 ...): a ``torch.nn.Linear`` --
same parameters (``weight [out, in]``, ``bias [out]``), same initialisation, same state_dict keys -- whose forward and
backward run as one kernel each (csrc/head.cu) when the layer has a handful of outputs, which is every head the reference
builds (out_size = 1, gnn/main.py:438).  Anything else (many outputs, wide inputs, CPU tensors, doubles, 3-D inputs) goes
through torch's own implementation unchanged."""
import torch

from .. import ops


class _HeadFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, layer):
        ctx.save_for_backward(x, weight)
        ctx.has_bias = bias is not None
        ctx.layer = layer
        return ops.head_fwd(x, weight, bias)

    @staticmethod
    def backward(ctx, d_out):
        x, weight = ctx.saved_tensors
        layer = ctx.layer
        # parallel.FlatGrads(model=...): the first backward of a step writes dW / db straight into the flat gradient buffer
        inplace = (getattr(layer, "grad_inplace", False) and not layer._grad_written and ctx.has_bias
                   and weight.grad is not None and layer.bias.grad is not None)
        d_x, d_w, d_b = ops.head_bwd(d_out, x, weight, need_dA=ctx.needs_input_grad[0], need_db=ctx.has_bias,
                                     out_w=weight.grad if inplace else None, out_b=layer.bias.grad if inplace else None)
        if inplace:
            layer._grad_written = True
            return d_x, None, None, None
        return d_x, d_w, d_b, None


class Linear(torch.nn.Linear):
    grad_inplace, _grad_written = False, False      # set per instance by parallel.FlatGrads

    def forward(self, x):
        if (x.is_cuda and x.dim() == 2 and x.dtype == torch.float32 and self.weight.dtype == torch.float32
                and ops.head_supported(self.in_features, self.out_features)):
            return _HeadFn.apply(x, self.weight, self.bias, self)
        if self.grad_inplace and not self._grad_written and torch.is_grad_enabled():
            # marked for in-place gradients but this input goes through torch's Linear, whose autograd accumulates: the
            # slices are not zeroed by FlatGrads.zero(), so clear them here
            with torch.no_grad():
                self.weight.grad.zero_()
                self.bias.grad.zero_()
            self._grad_written = True
        return super().forward(x)
