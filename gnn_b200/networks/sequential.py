"""``Sequential`` with the contract of the reference's gnn/networks/sequential.py (itself PyG's
Sequential plus optional trailing arguments): a model is declared by its global argument names and a
list of modules, each optionally followed by a header ``"in1, in2 -> out"``; a trailing ``?`` on a global
argument makes it optional.  ``nns`` / ``fns`` keep the same meaning, so ``state_dict`` keys
(``nns.{i}...``) match the reference's checkpoints.
"""
import torch


def parse_desc(desc):
    assert "->" in desc
    left, right = desc.split("->")
    names = lambda part: [t.strip() for t in part.split(",") if t.strip()]
    return names(left), names(right)


class Sequential(torch.nn.Module):
    def __init__(self, args, modules):
        super().__init__()
        self.args = [t.strip() for t in args.split(",") if t.strip()]
        assert len(modules) > 0
        assert isinstance(modules[0], (tuple, list))
        self.nns, self.fns, self.descs = torch.nn.ModuleList(), [], []
        for i, entry in enumerate(modules):
            header = None
            if isinstance(entry, (tuple, list)):
                assert len(entry) in (1, 2)
                header = entry[1] if len(entry) == 2 else None
                entry = entry[0]
            if header is not None:
                ins, outs = parse_desc(header)
            else:                       # no header: consume and produce what the previous module produced
                ins = outs = self.descs[i - 1][1]
            is_module = isinstance(entry, torch.nn.Module)
            self.nns.append(entry if is_module else None)
            self.fns.append(None if is_module else entry)
            self.descs.append((ins, outs))

    def reset_parameters(self):
        for m in self.nns:
            if hasattr(m, "reset_parameters"):
                m.reset_parameters()

    def forward(self, *args):
        if len(args) == 1 and len(self.args) > 1:      # model((x, edge_index, ...)) call form
            args = args[0]
        names = list(self.args)
        absent = []
        while len(args) < len(names) and names[-1].endswith("?"):
            absent.append(names.pop()[:-1])
        names = [t.strip("?") for t in names]
        assert len(args) == len(names)
        state = dict(zip(names, args))
        state["self"] = self
        out = None
        fused_p = 0.0
        for i, (m, f, (ins, outs)) in enumerate(zip(self.nns, self.fns, self.descs)):
            call = m if f is None else f
            vals = [state[k] for k in ins if k not in absent]
            if self._dropout_feeds_head(i, vals):
                # Dropout -> Linear head (networks.py:272-274): the head's kernels drop the activations as they read them
                # (no mask, no dropped copy in memory); this module passes its input on untouched
                fused_p = float(m.p)
                continue
            if fused_p > 0.0:
                out = (m(vals[0], dropout_p=fused_p),)
                fused_p = 0.0
                state.update(zip(outs, out))
                continue
            if isinstance(m, torch.nn.Linear) and vals and torch.is_tensor(vals[0]) and not vals[0].is_floating_point():
                # the loaders keep genotypes as int8 codes where the reference has float32 (load.py:50); GCNConv reads the
                # codes directly, a stock Linear in front of them (create_network_linear / _test) needs the float tensor
                if vals[0].dtype != torch.int8:
                    raise TypeError("2-bit packed genotypes feed GCNConv only; expand them (ops.unpack_2bit) for a Linear layer")
                vals[0] = vals[0].float()
            out = call(*vals)
            out = out if isinstance(out, tuple) else (out,)
            assert len(out) == len(outs)
            state.update(zip(outs, out))
        return out[0] if len(out) == 1 else out

    def _dropout_feeds_head(self, i, vals):
        """Module i is a train-mode Dropout whose one float32 CUDA matrix goes, under the same name, to a dense head that runs
        on our kernels (networks.head.Linear) as the next module."""
        from .. import ops
        from .head import Linear
        m = self.nns[i]
        if not (isinstance(m, torch.nn.Dropout) and m.training and 0.0 < m.p < 1.0 and not m.inplace
                and ops.config["fused_dropout"] and i + 1 < len(self.nns) and len(vals) == 1):
            return False
        nxt = self.nns[i + 1]
        (ins, outs), (nins, _) = self.descs[i], self.descs[i + 1]
        return (isinstance(nxt, Linear) and nxt.runs_on_kernels(vals[0]) and len(ins) == 1 and list(outs) == list(ins)
                and list(nins) == list(outs))

    def __repr__(self):
        rows = [f"  ({i}): {m if f is None else f.__name__}" for i, (m, f) in enumerate(zip(self.nns, self.fns))]
        return self.__class__.__name__ + "(\n" + "\n".join(rows) + "\n)"

    def l1_regularize(self, loss, l1_weight=0.1):
        flat = torch.cat([p.view(-1) for p in self.parameters()])
        return loss + l1_weight * torch.abs(flat).sum()

    def replace(self, index, module):
        is_module = isinstance(module, torch.nn.Module)
        self.nns[index] = module if is_module else None
        self.fns[index] = None if is_module else module

    def get(self, index):
        return self.nns[index] if self.nns[index] is not None else self.fns[index]
