"""``Linear`` for the dense head of the GCN factories (gnn/networks/networks.py:273-274 ...): a ``torch.nn.Linear`` --
same parameters (``weight [out, in]``, ``bias [out]``), same initialisation, same state_dict keys -- whose forward and
backward run as one kernel each (csrc/head.cu) when the layer has a handful of outputs, which is every head the reference
builds (out_size = 1, gnn/main.py:438).  Anything else (many outputs, wide inputs, CPU tensors, doubles, 3-D inputs) goes
through torch's own implementation unchanged."""
import torch

from .. import ops


class _HeadFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias):
        ctx.save_for_backward(x, weight)
        ctx.has_bias = bias is not None
        return ops.head_fwd(x, weight, bias)

    @staticmethod
    def backward(ctx, d_out):
        x, weight = ctx.saved_tensors
        d_x, d_w, d_b = ops.head_bwd(d_out, x, weight, need_dA=ctx.needs_input_grad[0], need_db=ctx.has_bias)
        return d_x, d_w, d_b


class Linear(torch.nn.Linear):
    def forward(self, x):
        if (x.is_cuda and x.dim() == 2 and x.dtype == torch.float32 and self.weight.dtype == torch.float32
                and ops.head_supported(self.in_features, self.out_features)):
            return _HeadFn.apply(x, self.weight, self.bias)
        return super().forward(x)
