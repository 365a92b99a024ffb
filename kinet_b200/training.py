"""Flat parameter / gradient / low-precision-mirror storage for mixed-precision data-parallel training.

    fp = FlatParams(model)                       # parameters and gradients become views of two flat fp32 buffers
    opt = torch.optim.AdamW([fp.flat_parameter], ..., fused=True, capturable=True)
    ...  backward()                              # every gradient accumulates into fp.grads
    torch.distributed.all_reduce(fp.grads, op=AVG)        # ONE collective per step (NCCL over NVLink)
    opt.step(); fp.refresh()                     # one optimizer kernel, ONE cast kernel for all bf16 copies

The reference trains under DistributedDataParallel (src/train.py:141-144), whose buckets do the same job
(flat gradient buffers, one all-reduce per bucket) from the autograd hooks; with the whole forward+backward in a
CUDA graph there are no hooks to run, so the buffers are laid out once, up front.  The bf16 mirror replaces the
~150 per-layer weight / bias casts autocast launches every step: kinet_b200's ``linear`` picks a layer's mirror
views up from ``layer._kinet_mirror``.
"""
import torch
from torch import nn

_ALIGN = 64     # elements: every parameter starts 128 / 256 bytes aligned in the bf16 / fp32 buffers


class FlatParams:
    def __init__(self, module, lp_dtype=torch.bfloat16):
        params = [p for p in module.parameters() if p.requires_grad]
        if not params:
            raise ValueError("module has no trainable parameters")
        dev = params[0].device
        if any(p.dtype != torch.float32 or p.device != dev for p in params):
            raise ValueError("FlatParams expects fp32 parameters on one device")
        offsets, total = [], 0
        for p in params:
            offsets.append(total)
            total += -(-p.numel() // _ALIGN) * _ALIGN
        self.master = torch.zeros(total, dtype=torch.float32, device=dev)
        self.grads = torch.zeros(total, dtype=torch.float32, device=dev)
        self.lp = torch.zeros(total, dtype=lp_dtype, device=dev)
        self.numel = sum(p.numel() for p in params)
        views = {}
        with torch.no_grad():
            for p, off in zip(params, offsets):
                n = p.numel()
                self.master[off:off + n].view_as(p).copy_(p)
                p.data = self.master[off:off + n].view_as(p)
                p.grad = self.grads[off:off + n].view_as(p)
                views[id(p)] = self.lp[off:off + n].view_as(p)
        self.flat_parameter = nn.Parameter(self.master)       # what the optimizer steps (same storage as the views)
        self.flat_parameter.grad = self.grads
        for m in module.modules():
            if isinstance(m, nn.Linear) and m.bias is not None and id(m.weight) in views and id(m.bias) in views:
                m.__dict__["_kinet_mirror"] = (views[id(m.weight)], views[id(m.bias)])
            elif isinstance(m, nn.MultiheadAttention) and m.in_proj_weight is not None and m.in_proj_bias is not None \
                    and id(m.in_proj_weight) in views and id(m.in_proj_bias) in views:
                m.__dict__["_kinet_mirror"] = (views[id(m.in_proj_weight)], views[id(m.in_proj_bias)])
        self.refresh()

    def refresh(self):
        """Bring the low-precision mirror up to date with the master parameters (after every optimizer step)."""
        self.lp.copy_(self.master)

    def zero_grad(self):
        self.grads.zero_()
