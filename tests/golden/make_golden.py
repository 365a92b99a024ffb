#!/usr/bin/env python
"""Generate tests/golden/*.npz from the REAL reference function.

Run in the build container only (needs /root/reference; the GPU box has none):

    python tests/golden/make_golden.py

It imports the reference's ``ms_deform_attn_core_pytorch``
(src/trackformer/models/ops/functions/ms_deform_attn_func.py:34-54) unmodified.
That file does ``import MultiScaleDeformableAttention`` at :11 (the CUDA
extension, not buildable without a GPU), so an empty stub module is installed
first; the stub is never called by the function we use.

Every case stores the inputs, the reference forward output, and the three
reference gradients (autograd through grid_sample) for a stored upstream
gradient.  Cases 'recipe_f32' / 'recipe_f64' replay the reference's own test
recipes (ops/test.py:14-31,38-50 and ops/test_double_precision.py:14-27).
"""
import os
import sys
import types

import numpy as np
import torch

REF_OPS = "/root/reference/src/trackformer/models/ops"
HERE = os.path.dirname(os.path.abspath(__file__))


def load_reference():
    sys.modules.setdefault("MultiScaleDeformableAttention",
                           types.ModuleType("MultiScaleDeformableAttention"))
    sys.path.insert(0, REF_OPS)
    from functions.ms_deform_attn_func import ms_deform_attn_core_pytorch
    return ms_deform_attn_core_pytorch


def run_case(ref, value, shapes, loc, attn, grad_out=None):
    value = value.clone().requires_grad_(True)
    loc = loc.clone().requires_grad_(True)
    attn = attn.clone().requires_grad_(True)
    out = ref(value, shapes, loc, attn)
    if grad_out is None:
        # the reference recipe's loss is out.abs().sum() (ops/test.py:48-49)
        grad_out = torch.sign(out.detach())
    gv, gl, ga = torch.autograd.grad(out, (value, loc, attn), grad_out)
    return dict(value=value.detach().numpy(), shapes=shapes.numpy(), loc=loc.detach().numpy(),
                attn=attn.detach().numpy(), grad_out=grad_out.numpy(), out=out.detach().numpy(),
                grad_value=gv.numpy(), grad_loc=gl.numpy(), grad_attn=ga.numpy())


def recipe(dtype, level_shapes, N=2, M=2, D=4, Lq=3, P=2):
    """The reference test recipe, verbatim distributions."""
    shapes = torch.as_tensor(level_shapes, dtype=torch.long)
    L = len(level_shapes)
    S = int((shapes[:, 0] * shapes[:, 1]).sum())
    value = (torch.rand(N, S, M, D) * 0.01).to(dtype)
    loc = torch.rand(N, Lq, M, L, P, 2).to(dtype)
    attn = (torch.rand(N, Lq, M, L, P) + 1e-5).to(dtype)
    attn = attn / attn.sum(-1, keepdim=True).sum(-2, keepdim=True)
    return value, shapes, loc, attn


def main():
    ref = load_reference()
    cases = {}

    torch.manual_seed(3)
    cases["recipe_f32"] = run_case(ref, *recipe(torch.float32, [(8, 8), (4, 4), (2, 2)]))
    torch.manual_seed(3)
    cases["recipe_f64"] = run_case(ref, *recipe(torch.float64, [(12, 8), (6, 4), (3, 2)]))

    # Locations outside [0,1]: zero padding, partially valid corners, fully outside samples.
    g = torch.Generator().manual_seed(11)
    for name, dtype in (("edges_f32", torch.float32), ("edges_f64", torch.float64)):
        shapes = torch.as_tensor([(5, 7), (3, 4)], dtype=torch.long)
        N, M, D, Lq, L, P = 1, 2, 36, 5, 2, 3
        S = 5 * 7 + 3 * 4
        value = torch.randn(N, S, M, D, generator=g).to(dtype)
        loc = (torch.rand(N, Lq, M, L, P, 2, generator=g) * 1.6 - 0.3).to(dtype)
        attn = torch.softmax(torch.randn(N, Lq, M, L * P, generator=g), -1).view(N, Lq, M, L, P).to(dtype)
        go = torch.randn(N, Lq, M * D, generator=g).to(dtype)
        cases[name] = run_case(ref, value, shapes, loc, attn, go)

    # L = 8 (multi-frame decoder, deformable_transformer.py:439-441), M = 8, D = 32, P = 4.
    shapes = torch.as_tensor([(6, 9), (3, 5), (2, 3), (1, 2)] * 2, dtype=torch.long)
    N, M, D, Lq, L, P = 2, 8, 32, 7, 8, 4
    S = int((shapes[:, 0] * shapes[:, 1]).sum())
    value = torch.randn(N, S, M, D, generator=g)
    loc = torch.rand(N, Lq, M, L, P, 2, generator=g) * 1.1 - 0.05
    attn = torch.softmax(torch.randn(N, Lq, M, L * P, generator=g), -1).view(N, Lq, M, L, P)
    go = torch.randn(N, Lq, M * D, generator=g)
    cases["l8_f32"] = run_case(ref, value, shapes, loc, attn, go)

    # Exactly-integer pixel coordinates (loc*W - 0.5 integral) and the window borders
    # -0.5/W (pixel coordinate -1, excluded) and 1+0.5/W (pixel coordinate W, excluded).
    shapes = torch.as_tensor([(4, 6), (2, 3)], dtype=torch.long)
    N, M, D, Lq, L, P = 1, 1, 8, 6, 2, 4
    S = 4 * 6 + 2 * 3
    value = torch.randn(N, S, M, D, generator=g, dtype=torch.float64)
    loc = torch.empty(N, Lq, M, L, P, 2, dtype=torch.float64)
    for l, (h, w) in enumerate(shapes.tolist()):
        kx = torch.randint(-1, w + 1, (N, Lq, M, P), generator=g).double()
        ky = torch.randint(-1, h + 1, (N, Lq, M, P), generator=g).double()
        loc[:, :, :, l, :, 0] = (kx + 0.5) / w
        loc[:, :, :, l, :, 1] = (ky + 0.5) / h
    attn = torch.softmax(torch.randn(N, Lq, M, L * P, generator=g, dtype=torch.float64), -1).view(N, Lq, M, L, P)
    go = torch.randn(N, Lq, M * D, generator=g, dtype=torch.float64)
    # forward only is well defined here (the bilinear form is not differentiable at integer
    # coordinates; grid_sample and the kernels pick one-sided derivatives), grads still stored.
    cases["integer_f64"] = run_case(ref, value, shapes, loc, attn, go)

    for name, arrs in cases.items():
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **arrs)
        print(name, {k: v.shape for k, v in arrs.items()})


if __name__ == "__main__":
    main()
