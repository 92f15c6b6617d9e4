"""``Linear`` for the dense head of the GCN factories (gnn/networks/networks.py:273-274 ...): a ``torch.nn.Linear`` --
same parameters (``weight [out, in]``, ``bias [out]``), same initialisation, same state_dict keys -- whose forward and
backward run as one kernel each (csrc/head.cu) when the layer has a handful of outputs, which is every head the reference
builds (out_size = 1, gnn/main.py:438).  Anything else (many outputs, wide inputs, CPU tensors, doubles, 3-D inputs) goes
through torch's own implementation unchanged."""
import torch

from .. import ops


class _HeadFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, layer, drop_p):
        ctx.has_bias = bias is not None
        ctx.layer = layer
        ctx.drop_p = drop_p
        if drop_p > 0.0:
            if layer._drop_state is None or layer._drop_state.device != x.device:
                layer._drop_state = ops.dropout_state(x.device)
            out, used = ops.head_fwd(x, weight, bias, drop=(drop_p, layer._drop_state))
            ctx.save_for_backward(x, weight, used)
            return out
        ctx.save_for_backward(x, weight)
        return ops.head_fwd(x, weight, bias)

    @staticmethod
    def backward(ctx, d_out):
        x, weight = ctx.saved_tensors[:2]
        layer = ctx.layer
        drop = (ctx.drop_p, ctx.saved_tensors[2]) if ctx.drop_p > 0.0 else None
        # parallel.FlatGrads(model=...): the first backward of a step writes dW / db straight into the flat gradient buffer
        inplace = (getattr(layer, "grad_inplace", False) and not layer._grad_written and ctx.has_bias
                   and weight.grad is not None and layer.bias.grad is not None)
        d_x, d_w, d_b = ops.head_bwd(d_out, x, weight, need_dA=ctx.needs_input_grad[0], need_db=ctx.has_bias,
                                     out_w=weight.grad if inplace else None, out_b=layer.bias.grad if inplace else None, drop=drop)
        if inplace:
            layer._grad_written = True
            return d_x, None, None, None, None
        return d_x, d_w, d_b, None, None


class Linear(torch.nn.Linear):
    grad_inplace, _grad_written = False, False      # set per instance by parallel.FlatGrads
    _drop_state = None                              # generator state of the fused dropout, made on first use

    def runs_on_kernels(self, x):
        return (torch.is_tensor(x) and x.is_cuda and x.dim() == 2 and x.dtype == torch.float32
                and self.weight.dtype == torch.float32 and ops.head_supported(self.in_features, self.out_features))

    def forward(self, x, dropout_p=0.0):
        """``dropout_p`` > 0: a train-mode Dropout(dropout_p) on ``x`` applied inside the kernels (``Sequential`` passes it
        when a Dropout module sits directly in front of this layer; see ``ops.config["fused_dropout"]``)."""
        if self.runs_on_kernels(x):
            return _HeadFn.apply(x, self.weight, self.bias, self, float(dropout_p))
        if dropout_p > 0.0:
            x = torch.nn.functional.dropout(x, dropout_p, training=True)
        if self.grad_inplace and not self._grad_written and torch.is_grad_enabled():
            # marked for in-place gradients but this input goes through torch's Linear, whose autograd accumulates: the
            # slices are not zeroed by FlatGrads.zero(), so clear them here
            with torch.no_grad():
                self.weight.grad.zero_()
                self.bias.grad.zero_()
            self._grad_written = True
        return super().forward(x)
