"""``Adam`` -- the update rule of ``torch.optim.Adam(params, lr)`` as the reference uses it (gnn/main.py:318: default
betas and eps, no weight decay, no amsgrad), as a ``torch.optim.Optimizer`` whose ``step()`` is one kernel launch
(csrc/adam.cu) with the step counter on the device: drop it into the optimiser-creator closure the trainers take,
``lambda model: gnn_b200.optim.Adam(model.parameters(), lr=...)``.  It can be captured in a CUDA graph."""
import ctypes

import torch

from . import ops


class Adam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        if lr < 0 or eps < 0 or not 0 <= betas[0] < 1 or not 0 <= betas[1] < 1:
            raise ValueError("invalid Adam hyper-parameters")
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps, capturable=True))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            todo = [p for p in group["params"] if p.grad is not None]
            if not todo:
                continue
            for p in todo:
                if p.dtype != torch.float32 or not p.is_cuda or not p.is_contiguous() or not p.grad.is_contiguous():
                    raise RuntimeError("gnn_b200.optim.Adam updates contiguous float32 CUDA parameters (no CPU fallback)")
                st = self.state[p]
                if not st:
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
            holder = self.state[group["params"][0]]                 # the group's step count (device float32) lives with its first parameter
            step = holder.get("step")
            if step is None:
                step = holder["step"] = torch.zeros((), dtype=torch.float32, device=todo[0].device)
            step.add_(1.0)
            b1, b2 = group["betas"]
            with torch.cuda.device(todo[0].device):
                for i in range(0, len(todo), 8):
                    chunk = todo[i:i + 8]
                    k = len(chunk)
                    ptrs = lambda ts: (ctypes.c_void_p * k)(*[t.data_ptr() for t in ts])       # noqa: E731
                    ops._call("gnnb_adam_step", k, ptrs(chunk), ptrs([p.grad for p in chunk]),
                              ptrs([self.state[p]["exp_avg"] for p in chunk]), ptrs([self.state[p]["exp_avg_sq"] for p in chunk]),
                              (ctypes.c_int64 * k)(*[p.numel() for p in chunk]), ops._p(step),
                              float(group["lr"]), float(b1), float(b2), float(group["eps"]), ops._stream())
        return loss
