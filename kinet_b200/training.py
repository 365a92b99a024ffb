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

from .functions.layer_ops import advance_dropout_step
from torch import nn

_ALIGN = 64     # elements: every parameter starts 128 / 256 bytes aligned in the bf16 / fp32 buffers


def reference_param_groups(module, lr=2e-4, lr_backbone=2e-5, lr_backbone_names=("backbone.0",),
                           lr_linear_proj_names=("reference_points", "sampling_offsets"), lr_linear_proj_mult=0.1,
                           lr_track=1e-4, track_attention=False):
    """The reference's optimizer groups (src/train.py:95-117, defaults of cfgs/train.yaml:1-6): every trainable
    parameter whose name contains one of ``lr_backbone_names`` trains at ``lr_backbone``, the
    ``reference_points`` / ``sampling_offsets`` projections at ``lr * lr_linear_proj_mult``, the track-attention
    layers (if any) at ``lr_track``, the rest at ``lr``.  -> list of {"params": [...], "lr": ...} in that order,
    empty groups dropped; pass it to FlatParams(groups=...) / GraphedTrainStep(param_groups=...)."""
    def has(name, keys):
        return any(k in name for k in keys)
    special = tuple(lr_backbone_names) + tuple(lr_linear_proj_names) + ("layers_track_attention",)
    named = [(n, p) for n, p in module.named_parameters() if p.requires_grad]
    groups = [
        {"params": [p for n, p in named if not has(n, special)], "lr": lr},
        {"params": [p for n, p in named if has(n, lr_backbone_names)], "lr": lr_backbone},
        {"params": [p for n, p in named if has(n, lr_linear_proj_names)], "lr": lr * lr_linear_proj_mult},
    ]
    if track_attention:
        groups.append({"params": [p for n, p in named if has(n, ("layers_track_attention",))], "lr": lr_track})
    return [g for g in groups if g["params"]]


class FlatParams:
    """``groups``: optional list of {"params": [...], **optimizer options} (torch's param-group format).  Every
    group becomes ONE contiguous segment of the flat buffers and one flat ``nn.Parameter`` (``group_parameters``,
    ``optimizer_groups()``), so per-group learning rates / weight decay cost nothing extra in the fused optimizer."""

    def __init__(self, module, lp_dtype=torch.bfloat16, groups=None):
        trainable = [p for p in module.parameters() if p.requires_grad]
        if not trainable:
            raise ValueError("module has no trainable parameters")
        if groups is None:
            groups = [{"params": trainable}]
        groups = [dict(g, params=[p for p in g["params"] if p.requires_grad]) for g in groups]
        params = [p for g in groups for p in g["params"]]
        if len({id(p) for p in params}) != len(params):
            raise ValueError("a parameter appears in more than one group")
        if {id(p) for p in params} != {id(p) for p in trainable}:
            raise ValueError("the groups must cover exactly the module's trainable parameters")
        dev = params[0].device
        if any(p.dtype != torch.float32 or p.device != dev for p in params):
            raise ValueError("FlatParams expects fp32 parameters on one device")
        offsets, total, bounds = [], 0, []
        for g in groups:
            start = total
            for p in g["params"]:
                offsets.append(total)
                total += -(-p.numel() // _ALIGN) * _ALIGN
            bounds.append((start, total))
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
        self.params = params
        self.grad_views = [p.grad for p in params]
        self.flat_parameter = nn.Parameter(self.master)       # what the optimizer steps (same storage as the views)
        self.flat_parameter.grad = self.grads
        # one flat parameter per group (views of the same storage)
        self.group_parameters, self._group_options = [], []
        for g, (a, b) in zip(groups, bounds):
            gp = nn.Parameter(self.master[a:b])
            gp.grad = self.grads[a:b]
            self.group_parameters.append(gp)
            self._group_options.append({k: v for k, v in g.items() if k != "params"})
        self._module = module
        # parameters changed behind the mirror's back (checkpoint resume): bring the low-precision copies along
        self._hook = module.register_load_state_dict_post_hook(lambda m, incompatible: self.refresh())
        for m in module.modules():
            if isinstance(m, nn.Linear) and m.bias is not None and id(m.weight) in views and id(m.bias) in views:
                m.__dict__["_kinet_mirror"] = (views[id(m.weight)], views[id(m.bias)])
            elif isinstance(m, nn.MultiheadAttention) and m.in_proj_weight is not None and m.in_proj_bias is not None \
                    and id(m.in_proj_weight) in views and id(m.in_proj_bias) in views:
                m.__dict__["_kinet_mirror"] = (views[id(m.in_proj_weight)], views[id(m.in_proj_bias)])
        self.refresh()

    def optimizer_groups(self):
        """torch param groups over the flat per-group parameters: ``AdamW(fp.optimizer_groups(), ...)``."""
        return [dict(opts, params=[gp]) for gp, opts in zip(self.group_parameters, self._group_options)]

    def refresh(self):
        """Bring the low-precision mirror up to date with the master parameters: after every optimizer step, and
        after ANY other change of the parameters (``load_state_dict`` on the wrapped module calls it through a
        hook; manual edits of ``p`` / ``master`` must call it themselves) -- the autocast forward reads the mirror."""
        self.lp.copy_(self.master)

    def clip_grad_norm_(self, max_norm):
        """``torch.nn.utils.clip_grad_norm_(parameters, max_norm)`` (src/trackformer/engine.py:147-148) on the flat
        gradient buffer: two kernels, no host read, CUDA-graph capturable (the padding between parameters is zero
        and does not change the norm).  -> the total norm (device scalar)."""
        norm = torch.linalg.vector_norm(self.grads)
        self.grads.mul_(torch.clamp(max_norm / (norm + 1e-6), max=1.0))
        return norm

    def zero_grad(self):
        self.grads.zero_()

    def detach_grads(self):
        """Before a backward: let autograd hand every parameter its gradient tensor as is (``p.grad = None``)
        instead of adding it into the pre-set flat views -- one small add kernel per parameter and step
        (258 of them, 0.5 ms, in a 6+6-layer stack).  ``gather_grads`` moves them into the flat buffer."""
        for p in self.params:
            p.grad = None

    def gather_grads(self):
        """After the backward started with ``detach_grads``: copy all gradients into the flat buffer with one
        multi-tensor kernel (parameters that received none get zeros) and point ``p.grad`` at the views again."""
        have = [(v, p.grad) for v, p in zip(self.grad_views, self.params) if p.grad is not None]
        for v, p in zip(self.grad_views, self.params):
            if p.grad is None:
                v.zero_()
        if have:
            torch._foreach_copy_([v for v, _ in have], [g for _, g in have])
        for v, p in zip(self.grad_views, self.params):
            p.grad = v


class GraphedTrainStep:
    """One data-parallel training step as two CUDA graphs with a single all-reduce between them:

        graph 1: forward, loss, backward, ONE multi-tensor copy of all gradients into FlatParams.grads
        NCCL   : all_reduce(grads, AVG) when torch.distributed is initialised with more than one rank
        graph 2: optimizer step on the flat parameter, refresh of the low-precision weight mirror

    ``loss_fn(*inputs)`` runs the model (under whatever autocast it wants) and returns a scalar loss; ``inputs`` are
    example tensors whose shapes every later batch must share (they become the graph's static input buffers).
    Construction runs ``warmup`` REAL training steps on the example batch (eagerly, on a side stream) before
    recording.  ``make_optimizer(params)`` builds a capturable optimizer, e.g.
    ``lambda p: torch.optim.AdamW(p, lr=1e-4, fused=True, capturable=True)``.  ``param_groups``: the reference's
    per-group learning rates (``reference_param_groups(module)``; ``params`` is then a list of torch param groups
    over one flat parameter each); ``clip_max_norm`` > 0 adds the reference's gradient clipping (cfgs/train.yaml:14:
    0.1) at the start of graph 2, after the all-reduce.  Not supported: per-parameter weight-decay exclusions finer
    than the groups, and optimizers that are not capturable.
    Nothing in kinet_b200's kernels allocates, synchronises or reads device memory on the host, so the whole step
    replays without per-kernel host cost (a 6+6-layer stack launches ~1 100 kernels per step)."""

    def __init__(self, module, loss_fn, make_optimizer, example_inputs, lp_dtype=torch.bfloat16, warmup=3,
                 param_groups=None, clip_max_norm=0.0):
        import torch.distributed as dist
        self.flat = FlatParams(module, lp_dtype, groups=param_groups)
        self.clip_max_norm = float(clip_max_norm)
        self.optimizer = make_optimizer(self.flat.optimizer_groups() if param_groups is not None
                                        else [self.flat.flat_parameter])
        self.inputs = tuple(torch.empty_like(t) for t in example_inputs)
        for b, t in zip(self.inputs, example_inputs):
            b.copy_(t)
        self.loss = torch.zeros((), dtype=torch.float32, device=self.flat.master.device)
        self._world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        self._loss_fn = loss_fn
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):     # eager runs: fill the per-tensor caches, let the optimizer build its state
                self._forward_backward()
                self._exchange()
                self._update()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self._g_fb, self._g_opt = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._g_fb):
            self._forward_backward()
        with torch.cuda.graph(self._g_opt):
            self._update()

    def _forward_backward(self):
        self.flat.detach_grads()
        # the fused add + LayerNorm kernels draw their dropout masks from (seed, call site, THIS counter): advancing it
        # inside the graph gives every replay new masks although the call sites are baked in
        advance_dropout_step(self.flat.master.device)
        loss = self._loss_fn(*self.inputs)
        loss.backward()
        self.flat.gather_grads()
        self.loss.copy_(loss.detach())

    def _exchange(self):
        if self._world > 1:
            import torch.distributed as dist
            dist.all_reduce(self.flat.grads, op=dist.ReduceOp.AVG)

    def _update(self):
        if self.clip_max_norm > 0:      # the reference clips after the gradients are averaged (engine.py:146-149)
            self.flat.clip_grad_norm_(self.clip_max_norm)
        self.optimizer.step()
        self.flat.refresh()

    def __call__(self, *inputs):
        """Run one step on a new batch; returns the (device-resident, overwritten next step) loss tensor."""
        for b, t in zip(self.inputs, inputs):
            b.copy_(t, non_blocking=True)
        self._g_fb.replay()
        self._exchange()
        self._g_opt.replay()
        return self.loss
