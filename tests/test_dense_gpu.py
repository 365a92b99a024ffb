"""Parity of the dense-window (tcgen05 / TMEM) encoder-form backward, kinet_b200/csrc/msda_dense.cuh, called
through the C ABI like every other kernel.  ``-m gpu``.

The dense path runs for 16-bit values in the encoder form (Lq == S) with D in {32, 64}; it is compared with
  * the fp64 CPU oracle on the 16-bit-rounded inputs, at the library's stated 16-bit tolerance (1.2e-2 of max|ref|,
    tests/test_msda_gpu.py): the scatter product rounds the bilinear x attention weights to the operand type
    (bf16: 2^-9 per weight, fp32 accumulation), which stays below the final rounding of the 16-bit grad_value;
  * the library's own per-row kernels on the same inputs (MSDA_B200_DENSE=0): grad_sampling_loc / grad_attn_weight
    come from exact fp32 dot products with fp32 weights in both, so they agree to fp32 rounding; grad_value to two
    units in the last place of the 16-bit output.
The cases cover what the tiling and the windows can get wrong: levels of one row / one column, tiles that hang over
the level's border, P < 4, levels nobody samples, sampling locations outside the image, tiles with local and with
arbitrary locations in the same call (deferred to the per-row kernel), masked queries, 2-d and 4-d reference points.
"""
import ctypes
import os

import numpy as np
import pytest
import torch

from oracle import oracle_backward

pytestmark = pytest.mark.gpu


def _dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


def _rel(a, ref):
    a = np.asarray(a, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    return float(np.abs(a - ref).max() / max(np.abs(ref).max(), 1e-30))


def _encoder_case(levels, N, M, D, P, dtype, regime, seed=0, noise=1.0, mixed=False):
    """Encoder-form inputs: queries = the pyramid's pixels; 'local': reference grid + a few pixels of offset (what
    the model produces), 'uniform': anywhere in [-0.05, 1.05]^2; mixed: the second batch item / odd heads uniform."""
    dev = _dev()
    g = torch.Generator(device="cpu").manual_seed(seed)
    L = len(levels)
    S = sum(h * w for h, w in levels)
    value = torch.randn(N, S, M, D, generator=g).to(dtype)
    pts = []
    for h, w in levels:
        ys = (torch.arange(h, dtype=torch.float32) + 0.5) / h
        xs = (torch.arange(w, dtype=torch.float32) + 0.5) / w
        yy, xx = torch.meshgrid(ys, xs, indexing="ij")
        pts.append(torch.stack([xx.reshape(-1), yy.reshape(-1)], -1))
    ref = torch.cat(pts, 0)                                   # (S, 2)
    wh = torch.as_tensor([[w, h] for h, w in levels], dtype=torch.float32)
    off = torch.randn(N, S, M, L, P, 2, generator=g) * noise + torch.arange(1, P + 1).view(1, 1, 1, 1, P, 1) * 0.7
    loc = ref[None, :, None, None, None, :] + off / wh.view(1, 1, 1, L, 1, 2)
    uni = torch.rand(N, S, M, L, P, 2, generator=g) * 1.1 - 0.05
    if regime == "uniform":
        loc = uni
    elif mixed:
        pick = torch.zeros(N, S, M, 1, 1, 1, dtype=torch.bool)
        pick[N - 1:, S // 3:2 * S // 3, 1::2] = True           # a band of queries of the odd heads of the last batch item
        loc = torch.where(pick, uni, loc)
    attn = torch.softmax(torch.randn(N, S, M, L * P, generator=g), -1).view(N, S, M, L, P)
    go = torch.randn(N, S, M * D, generator=g).to(dtype)
    shapes = torch.as_tensor(levels, dtype=torch.int64)
    return [t.to(dev) for t in (value, shapes, loc.contiguous(), attn.contiguous(), go)]


def _backward(value, shapes, loc, attn, go, dense):
    from kinet_b200 import _lib
    from kinet_b200.functions import ms_deform_attn_backward
    old = os.environ.get("MSDA_B200_DENSE")
    os.environ["MSDA_B200_DENSE"] = "1" if dense else "0"
    try:
        grads = ms_deform_attn_backward(value, shapes, loc, attn, go, 64)
        launches = _lib.load().msda_b200_last_launch_count()
        torch.cuda.synchronize()
    finally:
        if old is None:
            os.environ.pop("MSDA_B200_DENSE", None)
        else:
            os.environ["MSDA_B200_DENSE"] = old
    return grads, launches


CASES = [
    # name, levels, N, M, D, P
    ("pyramid", [(20, 31), (10, 16), (5, 8), (3, 4)], 2, 2, 32, 4),
    ("thin_levels", [(37, 53), (19, 27), (1, 9), (7, 1)], 1, 3, 32, 4),     # H == 1 and W == 1 levels
    ("d64_p3", [(24, 40), (12, 20)], 1, 2, 64, 3),                          # D = 64, P = 3 (scalar sample access)
    ("one_level", [(33, 47)], 1, 2, 32, 4),                                 # L = 1: everything is the finest level
    ("tiny", [(3, 5), (2, 2)], 2, 1, 32, 2),                                # less than one tile
    ("finest_last", [(6, 9), (12, 18), (24, 36)], 1, 2, 32, 4),            # the finest level is not level 0
    ("d36", [(20, 31), (10, 16), (5, 8), (3, 4)], 2, 3, 36, 4),            # D = 36: K / N padded to 48, 8-byte row pieces, odd M
    ("d36_thin", [(37, 53), (19, 27), (1, 9), (7, 1)], 1, 8, 36, 4),
]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16], ids=["bf16", "f16"])
@pytest.mark.parametrize("regime", ["local", "uniform", "mixed"])
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_dense_backward_matches_oracle_and_per_row_kernels(case, regime, dtype):
    _, levels, N, M, D, P = case
    value, shapes, loc, attn, go = _encoder_case(levels, N, M, D, P, dtype, regime, mixed=regime == "mixed")
    (gv, gl, ga), launches = _backward(value, shapes, loc, attn, go, dense=True)
    (rv, rl, ra), base_launches = _backward(value, shapes, loc, attn, go, dense=False)
    # dense kernel + per-row kernel (split mode) + rounding pass, against per-row kernel + rounding pass
    assert base_launches == 2 and launches == 3, (launches, base_launches)
    # against the library's per-row kernels
    assert _rel(gl.float().cpu(), rl.float().cpu()) <= 2e-5
    assert _rel(ga.float().cpu(), ra.float().cpu()) <= 2e-5
    ulp = 2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -11
    assert _rel(gv.float().cpu(), rv.float().cpu()) <= 2.5 * ulp
    # against the fp64 oracle on the rounded inputs
    ref = oracle_backward(value.double().cpu().numpy(), shapes.cpu().numpy(), loc.double().cpu().numpy(),
                          attn.double().cpu().numpy(), go.double().cpu().numpy(), nthreads=8)
    assert _rel(gv.float().cpu().numpy(), ref[0]) <= 1.2e-2
    assert _rel(ga.float().cpu().numpy(), ref[2]) <= 1.2e-2
    # grad_sampling_loc: one-sided derivative at the bilinear kinks; fp32 vs fp64 may pick different sides there
    l = loc.double().cpu().numpy()
    hw = shapes.cpu().numpy().astype(np.float64)
    x = l[..., 0] * hw[None, None, None, :, None, 1] - 0.5
    y = l[..., 1] * hw[None, None, None, :, None, 0] - 0.5
    ok = (np.abs(x - np.round(x)) > 2e-4) & (np.abs(y - np.round(y)) > 2e-4)
    masked_fraction = 1.0 - ok.mean()
    assert masked_fraction < 0.005, masked_fraction
    err = (np.abs(gl.float().cpu().numpy().astype(np.float64) - ref[1]) * ok[..., None]).max() / max(np.abs(ref[1]).max(), 1e-30)
    assert err <= 1.2e-2, err


def test_dense_backward_weights_spanning_six_decades():
    """The scatter product rounds (bilinear x attention) weights to bf16 and accumulates in fp32: with |grad_out|
    spanning 1e-3 ... 1e3 across queries the result must still match the oracle at the 16-bit tolerance, row by
    row (relative to each row's own magnitude, not only to the global maximum)."""
    levels = [(20, 31), (10, 16), (5, 8), (3, 4)]
    value, shapes, loc, attn, go = _encoder_case(levels, 1, 2, 32, 4, torch.bfloat16, "local", seed=3)
    scale = torch.logspace(-3, 3, go.shape[1], device=go.device).view(1, -1, 1)
    go = (go.float() * scale).to(torch.bfloat16)
    (gv, gl, ga), launches = _backward(value, shapes, loc, attn, go, dense=True)
    assert launches == 3
    ref = oracle_backward(value.double().cpu().numpy(), shapes.cpu().numpy(), loc.double().cpu().numpy(),
                          attn.double().cpu().numpy(), go.double().cpu().numpy(), nthreads=8)
    gv = gv.float().cpu().numpy().astype(np.float64)
    assert _rel(gv, ref[0]) <= 1.2e-2
    # per pixel row: error against the sum of |contributions| bound, approximated by the row's own magnitude
    row_ref = np.abs(ref[0]).reshape(-1, 32).max(1)
    row_err = np.abs(gv - ref[0]).reshape(-1, 32).max(1)
    big = row_ref > 1e-2 * row_ref.max()
    assert (row_err[big] / row_ref[big]).max() <= 5e-2
    assert _rel(ga.float().cpu().numpy(), ref[2]) <= 1.2e-2


def test_dense_backward_run_to_run_and_zero_rows():
    """grad_value rows nobody samples stay exactly zero; repeated calls agree to summation-order noise."""
    levels = [(24, 36), (12, 18), (6, 9)]
    value, shapes, loc, attn, go = _encoder_case(levels, 1, 2, 32, 4, torch.bfloat16, "local", seed=5)
    loc = loc.clone()
    loc[..., 1, :, :] = 3.0                      # level 1: every sample outside the image -> nothing lands there
    (gv1, gl1, ga1), _ = _backward(value, shapes, loc, attn, go, dense=True)
    (gv2, gl2, ga2), _ = _backward(value, shapes, loc, attn, go, dense=True)
    s0, s1 = 24 * 36, 24 * 36 + 12 * 18
    assert float(gv1[:, s0:s1].abs().max()) == 0.0
    assert float(gl1[:, :, :, 1].abs().max()) == 0.0 and float(ga1[:, :, :, 1].abs().max()) == 0.0
    assert torch.equal(gl1, gl2) and torch.equal(ga1, ga2)          # single writer, fixed order
    assert _rel(gv1.float().cpu(), gv2.float().cpu()) <= 2.0 ** -7


@pytest.mark.parametrize("ref_dim,masked", [(2, False), (2, True), (4, True)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16], ids=["bf16", "f16"])
@pytest.mark.parametrize("D", [32, 36])
def test_dense_fused_prologue_matches_per_row_fused(dtype, ref_dim, masked, D):
    """The module-prologue entry point (raw offsets, logits, reference points): the dense kernel against the fused
    per-row kernels, and against the unfused PyTorch prologue + autograd in fp32 for the scalar gradients."""
    from kinet_b200.functions import MSDeformAttnFusedFunction
    dev = _dev()
    levels = [(20, 31), (10, 16), (5, 8), (3, 4)]
    N, M, P, L = 2, 2, 4, 4
    S = sum(h * w for h, w in levels)
    g = torch.Generator(device="cpu").manual_seed(7)
    value = torch.randn(N, S, M, D, generator=g).to(dtype).to(dev)
    shapes = torch.as_tensor(levels, dtype=torch.int64, device=dev)
    pts = []
    for h, w in levels:
        ys = (torch.arange(h, dtype=torch.float32) + 0.5) / h
        xs = (torch.arange(w, dtype=torch.float32) + 0.5) / w
        yy, xx = torch.meshgrid(ys, xs, indexing="ij")
        pts.append(torch.stack([xx.reshape(-1), yy.reshape(-1)], -1))
    ref = torch.cat(pts, 0)[None, :, None, :].expand(N, -1, L, -1)
    if ref_dim == 4:
        ref = torch.cat([ref, torch.rand(N, S, L, 2, generator=g) * 0.1 + 0.02], -1)
    ref = ref.contiguous().to(dev)
    off = (torch.randn(N, S, M, L, P, 2, generator=g) * 2.0).to(dev)
    logit = torch.randn(N, S, M, L * P, generator=g).to(dev)
    mask = (torch.rand(N, S, generator=g) < 0.1).to(dev) if masked else None
    go = torch.randn(N, S, M * D, generator=g).to(dtype).to(dev)
    got = []
    for dense in ("1", "0"):
        os.environ["MSDA_B200_DENSE"] = dense
        try:
            leaves = [t.clone().requires_grad_(True) for t in (value, ref, off, logit)]
            out = MSDeformAttnFusedFunction.apply(leaves[0], shapes, leaves[1], leaves[2], leaves[3], mask, shapes, 64)
            out.backward(go)
            torch.cuda.synchronize()
            got.append([t.grad for t in leaves])
        finally:
            os.environ.pop("MSDA_B200_DENSE", None)
    a, b = got
    ulp = 2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -11
    assert _rel(a[0].float().cpu(), b[0].float().cpu()) <= 2.5 * ulp
    for i in (1, 2, 3):
        assert _rel(a[i].float().cpu(), b[i].float().cpu()) <= 2e-5, i
    # unfused PyTorch prologue (reference ops/modules/ms_deform_attn.py:69-82) + the library's per-row kernels
    from kinet_b200.functions import MSDeformAttnFunction
    os.environ["MSDA_B200_DENSE"] = "0"
    try:
        leaves = [t.clone().requires_grad_(True) for t in (value, ref, off, logit)]
        attn = torch.softmax(leaves[3], -1).view(N, S, M, L, P)
        if mask is not None:
            attn = attn.masked_fill(mask[..., None, None, None], 0.0)
        if ref_dim == 2:
            loc = leaves[1][:, :, None, :, None, :] + leaves[2] / shapes[None, None, None, :, None, :].float()
        else:
            loc = leaves[1][:, :, None, :, None, :2] + leaves[2] / P * leaves[1][:, :, None, :, None, 2:] * 0.5
        out = MSDeformAttnFunction.apply(leaves[0], shapes, loc, attn, 64)
        out.backward(go)
        torch.cuda.synchronize()
    finally:
        os.environ.pop("MSDA_B200_DENSE", None)
    for i, tol in ((1, 2e-4), (2, 2e-4), (3, 2e-4)):
        assert _rel(a[i].float().cpu(), leaves[i].grad.float().cpu()) <= tol, i


def test_dense_path_not_taken_where_it_does_not_apply():
    """fp32 values, decoder form (Lq != S), D = 36: the per-row kernels run alone (launch counts tell)."""
    levels = [(20, 31), (10, 16)]
    value, shapes, loc, attn, go = _encoder_case(levels, 1, 2, 32, 4, torch.float32, "local")
    _, launches = _backward(value, shapes, loc, attn, go, dense=True)
    assert launches == 1                                     # fp32: one kernel, no rounding pass
    v16 = value.to(torch.bfloat16)
    _, launches = _backward(v16, shapes, loc[:, :100].contiguous(), attn[:, :100].contiguous(),
                            go[:, :100].to(torch.bfloat16).contiguous(), dense=True)
    assert launches == 2                                     # decoder form
