"""Parity of the sm_100a kernels (called through the C ABI, libmsda_b200.so) against the CPU
oracle and the reference's golden vectors.  All tests need a B200: ``-m gpu``.

Tolerances (north_star: 1e-5 relative in fp32; bf16 tolerance stated here):
  fp32: max|err| <= 1e-5 * max|ref| against the fp64 oracle on the same fp32 inputs
        (grad_value: atomic summation order varies run to run; same bound holds).
  fp64: 1e-11 relative.
  bf16/f16 value: inputs rounded to the 16-bit type first; compared with the fp64 oracle on the
        rounded inputs: max|err| <= 1.2e-2 * max|ref| (output rounding 2^-9 plus, for
        grad_sampling_loc, cancellation in corner differences).
"""
import os

import numpy as np
import pytest
import torch

from oracle import oracle_backward, oracle_forward

pytestmark = pytest.mark.gpu

GOLDEN = ["recipe_f32", "recipe_f64", "edges_f32", "edges_f64", "l8_f32", "integer_f64"]


def _dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


def _relerr(a, ref):
    a = np.asarray(a, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    scale = max(np.abs(ref).max(), 1e-30) if ref.size else 1.0
    return (np.abs(a - ref).max() / scale) if ref.size else 0.0


def _smooth_mask(shapes, loc, eps=2e-4):
    """True where the sample's pixel coordinates are at least eps away from every integer (and
    from the window borders -1, W, H, which are integers too).  grad_sampling_loc is a one-sided
    derivative at those kinks and which side is taken depends on fp32-vs-fp64 rounding of
    loc*W-0.5, so kink samples are excluded when comparing grad_sampling_loc (only)."""
    l = loc.detach().double().cpu().numpy()
    hw = shapes.cpu().numpy().astype(np.float64)
    x = l[..., 0] * hw[None, None, None, :, None, 1] - 0.5
    y = l[..., 1] * hw[None, None, None, :, None, 0] - 0.5
    ok = (np.abs(x - np.round(x)) > eps) & (np.abs(y - np.round(y)) > eps)
    return np.repeat(ok[..., None], 2, axis=-1)


def _check_grads(grads, ref_grads, shapes, loc, tol):
    m = _smooth_mask(shapes, loc)
    masked = 1.0 - float(m.mean()) if m.size else 0.0
    # the kink mask must stay a rounding-noise filter: a regression cannot hide behind it
    # (16-bit locations have 8 / 11 significant bits: loc * W - 0.5 is an integer far more often than for fp32 ones)
    limit = 0.05 if loc.dtype in (torch.bfloat16, torch.float16) else 0.005
    assert masked < limit or m.size < 4000, ("kink-masked fraction of grad_sampling_loc", masked)
    names = ("grad_value", "grad_sampling_loc", "grad_attn_weight")
    for i, (got, ref) in enumerate(zip(grads, ref_grads)):
        got = got.float().cpu().numpy() if got.dtype in (torch.bfloat16, torch.float16) else got.cpu().numpy()
        assert got.shape == ref.shape, names[i]
        if i == 1:
            scale = max(np.abs(ref).max(), 1e-30) if ref.size else 1.0
            err = (np.abs(got.astype(np.float64) - ref) * m).max() / scale if ref.size else 0.0
            assert np.isfinite(got).all()
        else:
            err = _relerr(got, ref)
        assert err <= tol, (names[i], err)


def _run_cuda(value, shapes, loc, attn, grad_out=None, level_start=None):
    from kinet_b200.functions import ms_deform_attn_backward, ms_deform_attn_forward
    out = ms_deform_attn_forward(value, shapes, loc, attn, 64, level_start)
    grads = None
    if grad_out is not None:
        grads = ms_deform_attn_backward(value, shapes, loc, attn, grad_out, 64, level_start)
    torch.cuda.synchronize()
    return out, grads


def _oracle64(value, shapes, loc, attn, grad_out=None):
    v = value.detach().double().cpu().numpy()
    l = loc.detach().double().cpu().numpy()
    a = attn.detach().double().cpu().numpy()
    s = shapes.cpu().numpy()
    out = oracle_forward(v, s, l, a, nthreads=8)
    grads = None
    if grad_out is not None:
        grads = oracle_backward(v, s, l, a, grad_out.detach().double().cpu().numpy(), nthreads=8)
    return out, grads


@pytest.mark.parametrize("name", GOLDEN)
def test_golden_forward_backward(golden, name):
    g = golden(name)
    dev = _dev()
    t = {k: torch.from_numpy(v).to(dev) for k, v in g.items()}
    out, grads = _run_cuda(t["value"], t["shapes"], t["loc"], t["attn"], t["grad_out"])
    tol = 1e-11 if name.endswith("f64") else 1e-5
    assert out.dtype == t["value"].dtype
    assert _relerr(out.cpu().numpy(), g["out"]) <= tol
    assert _relerr(grads[0].cpu().numpy(), g["grad_value"]) <= tol
    assert _relerr(grads[2].cpu().numpy(), g["grad_attn"]) <= tol
    if name != "integer_f64":   # grad_loc is one-sided at integer coordinates (see test_oracle.py)
        assert _relerr(grads[1].cpu().numpy(), g["grad_loc"]) <= tol


def _random_case(dev, N, M, D, Lq, levels, P, dtype=torch.float32, aux=None, spread=1.2, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    L = len(levels)
    S = sum(h * w for h, w in levels)
    aux = aux or (torch.float32 if dtype != torch.float64 else torch.float64)
    value = torch.randn(N, S, M, D, generator=g).to(dtype).to(dev)
    shapes = torch.as_tensor(levels, dtype=torch.int64, device=dev)
    loc = (torch.rand(N, Lq, M, L, P, 2, generator=g) * spread - (spread - 1) / 2).to(aux).to(dev)
    attn = torch.softmax(torch.randn(N, Lq, M, L * P, generator=g), -1).view(N, Lq, M, L, P).to(aux).to(dev)
    go = torch.randn(N, Lq, M * D, generator=g).to(dtype).to(dev)
    return value, shapes, loc, attn, go


VEC_CASES = [
    # (N, M, D, Lq, levels, P)
    (2, 8, 32, 37, [(9, 13), (5, 7), (3, 4), (2, 2)], 4),            # D=32 fast path, ragged Lq
    (1, 8, 36, 29, [(7, 11), (4, 6)] * 4, 4),                         # D=36 (GS=9), L=8
    (2, 4, 64, 19, [(6, 5), (3, 3)], 3),                              # D=64, P=3 (LP=6 < GS)
    (1, 2, 16, 70, [(10, 10)], 5),                                    # D=16, L=1, P=5
    (1, 3, 20, 11, [(4, 9), (2, 5)], 2),                              # D=20 -> generic path
    (3, 8, 32, 1, [(3, 3)], 1),                                       # single query, single sample
]


@pytest.mark.parametrize("case", VEC_CASES)
def test_fp32_matches_oracle(case):
    dev = _dev()
    N, M, D, Lq, levels, P = case
    value, shapes, loc, attn, go = _random_case(dev, N, M, D, Lq, levels, P)
    out, grads = _run_cuda(value, shapes, loc, attn, go)
    ref_out, ref_grads = _oracle64(value, shapes, loc, attn, go)
    assert _relerr(out.cpu().numpy(), ref_out) <= 1e-5
    _check_grads(grads, ref_grads, shapes, loc, 1e-5)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("aux_same", [False, True])
@pytest.mark.parametrize("case", VEC_CASES[:5])
def test_16bit_matches_oracle(case, dtype, aux_same):
    dev = _dev()
    N, M, D, Lq, levels, P = case
    aux = dtype if aux_same else torch.float32
    value, shapes, loc, attn, go = _random_case(dev, N, M, D, Lq, levels, P, dtype=dtype, aux=aux)
    out, grads = _run_cuda(value, shapes, loc, attn, go)
    assert out.dtype == dtype and grads[0].dtype == dtype
    assert grads[1].dtype == aux and grads[2].dtype == aux
    ref_out, ref_grads = _oracle64(value, shapes, loc, attn, go)
    tol = 1.2e-2
    assert _relerr(out.float().cpu().numpy(), ref_out) <= tol
    _check_grads(grads, ref_grads, shapes, loc, tol)


D36_WIDE_CASES = [
    # (N, M, Lq or None = encoder form, levels, P): the 16-bit D = 36 forward with 5 lanes of 8 channels on 16-byte
    # windows (msda_fwd.cuh "overhang"): even and odd heads read 4 channels of the neighbouring head's row, which
    # must never reach the output; forced here for small shapes (the launcher picks it for >= 57 k rows)
    (2, 8, None, [(13, 21), (7, 11), (4, 6), (2, 3)], 4),
    (1, 2, 53, [(9, 1), (1, 7), (3, 3), (1, 1)], 4),                  # W == 1 / H == 1 levels, ragged passes
    (3, 4, 7, [(5, 6), (3, 3)], 3),                                   # fewer samples than a chunk row
]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("case", D36_WIDE_CASES)
def test_d36_wide_forward_matches_oracle_and_narrow(case, dtype, monkeypatch):
    from kinet_b200.functions import ms_deform_attn_forward
    dev = _dev()
    N, M, Lq, levels, P = case
    S = sum(h * w for h, w in levels)
    value, shapes, loc, attn, go = _random_case(dev, N, M, 36, Lq or S, levels, P, dtype=dtype, aux=torch.float32,
                                                spread=1.1)
    # neighbouring heads carry large values: a leak of their channels into a row's output would be visible
    value = value.clone()
    value[:, :, 1::2] *= 50.0
    monkeypatch.setenv("MSDA_B200_D36_NARROW", "1")
    narrow = ms_deform_attn_forward(value, shapes, loc, attn, 64)
    monkeypatch.delenv("MSDA_B200_D36_NARROW")
    monkeypatch.setenv("MSDA_B200_D36_WIDE", "1")
    wide = ms_deform_attn_forward(value, shapes, loc, attn, 64)
    ref_out, _ = _oracle64(value, shapes, loc, attn, go)
    assert wide.dtype == dtype
    assert _relerr(wide.float().cpu().numpy(), ref_out) <= 1.2e-2
    # same fp32 arithmetic per channel, same rounding: the two lane mappings agree to the last bit
    assert torch.equal(wide, narrow)


def test_fp64_generic_matches_oracle():
    dev = _dev()
    value, shapes, loc, attn, go = _random_case(dev, 2, 2, 4, 5, [(12, 8), (6, 4), (3, 2)], 2,
                                                dtype=torch.float64)
    out, grads = _run_cuda(value, shapes, loc, attn, go)
    ref_out, ref_grads = _oracle64(value, shapes, loc, attn, go)
    assert _relerr(out.cpu().numpy(), ref_out) <= 1e-12
    for got, ref in zip(grads, ref_grads):
        assert _relerr(got.cpu().numpy(), ref) <= 1e-12


def test_encoder_form_2d_tiling_matches_oracle():
    """Lq == S switches the kernels to the band-ordered query walk; the result must not depend on it."""
    dev = _dev()
    levels = [(37, 45), (19, 23), (10, 12), (5, 6)]
    S = sum(h * w for h, w in levels)
    value, shapes, loc, attn, go = _random_case(dev, 2, 8, 32, S, levels, 4, spread=1.1)
    out, grads = _run_cuda(value, shapes, loc, attn, go)
    ref_out, ref_grads = _oracle64(value, shapes, loc, attn, go)
    assert _relerr(out.cpu().numpy(), ref_out) <= 1e-5
    _check_grads(grads, ref_grads, shapes, loc, 1e-5)


WALK_CASES = [
    # (levels, D, {launcher switches}): the encoder form walks each level band by band; every query must be
    # visited exactly once whatever the band height, the item length and the level shapes
    ([(37, 45), (19, 23), (10, 12), (5, 6)], 32, {}),
    ([(9, 1), (1, 7), (3, 3), (1, 1)], 32, {}),                                  # W == 1, H == 1, 1x1 levels
    ([(13, 21), (7, 11)], 36, {}),                                               # 24-row passes (D = 36)
    ([(13, 21), (7, 11)], 32, {"MSDA_B200_TILE_H": "3", "MSDA_B200_TILE_Q": "40"}),
    ([(13, 21), (7, 11)], 32, {"MSDA_B200_TILE_H": "16", "MSDA_B200_TILE_Q": "1000"}),
    ([(13, 21), (7, 11)], 64, {"MSDA_B200_TILE_H": "1"}),
    ([(13, 21), (7, 11)], 16, {"MSDA_B200_TILE_1D": "1"}),
]


@pytest.mark.parametrize("case", WALK_CASES)
def test_encoder_walk_covers_every_query(case, monkeypatch):
    levels, D, env = case
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    dev = _dev()
    S = sum(h * w for h, w in levels)
    value, shapes, loc, attn, go = _random_case(dev, 2, 8, D, S, levels, 4, spread=1.1, seed=5)
    out, grads = _run_cuda(value, shapes, loc, attn, go)
    ref_out, ref_grads = _oracle64(value, shapes, loc, attn, go)
    assert _relerr(out.cpu().numpy(), ref_out) <= 1e-5
    _check_grads(grads, ref_grads, shapes, loc, 1e-5)


def test_level_start_index_argument_and_6arg_form():
    from kinet_b200.functions import MSDeformAttnFunction
    dev = _dev()
    levels = [(9, 13), (5, 7), (3, 4)]
    value, shapes, loc, attn, go = _random_case(dev, 2, 8, 32, 21, levels, 4)
    lsi = torch.cat([shapes.new_zeros(1), (shapes[:, 0] * shapes[:, 1]).cumsum(0)[:-1]])
    o5 = MSDeformAttnFunction.apply(value, shapes, loc, attn, 64)
    o6 = MSDeformAttnFunction.apply(value, shapes, lsi, loc, attn, 64)
    assert torch.equal(o5, o6)


def test_autograd_matches_oracle_and_reference_recipe():
    """The reference's own check (ops/test.py:38-95): loss = |out|.sum(), three grads."""
    from kinet_b200.functions import MSDeformAttnFunction
    dev = _dev()
    torch.manual_seed(3)
    N, M, D, Lq, L, P = 2, 2, 4, 3, 3, 2
    shapes = torch.as_tensor([(8, 8), (4, 4), (2, 2)], dtype=torch.long, device=dev)
    S = 84
    value = (torch.rand(N, S, M, D) * 0.01).to(dev).requires_grad_(True)
    loc = torch.rand(N, Lq, M, L, P, 2).to(dev).requires_grad_(True)
    attn = torch.rand(N, Lq, M, L, P) + 1e-5
    attn = (attn / attn.sum(-1, keepdim=True).sum(-2, keepdim=True)).to(dev).requires_grad_(True)
    out = MSDeformAttnFunction.apply(value, shapes, loc, attn, 2)
    out.abs().sum().backward()
    ref_out, ref_grads = _oracle64(value, shapes, loc, attn, torch.sign(out.detach()))
    assert _relerr(out.detach().cpu().numpy(), ref_out) <= 1e-5
    for got, ref in zip((value.grad, loc.grad, attn.grad), ref_grads):
        assert _relerr(got.cpu().numpy(), ref) <= 1e-5


def test_gradcheck_double():
    """test_double_precision.py:117-119: gradcheck of the op in fp64."""
    from kinet_b200.functions import MSDeformAttnFunction
    dev = _dev()
    torch.manual_seed(3)
    shapes = torch.as_tensor([(12, 8), (6, 4), (3, 2)], dtype=torch.long, device=dev)
    N, M, D, Lq, L, P, S = 2, 2, 4, 3, 3, 2, 126
    value = (torch.rand(N, S, M, D, dtype=torch.float64) * 0.01).to(dev).requires_grad_(True)
    loc = torch.rand(N, Lq, M, L, P, 2, dtype=torch.float64).to(dev).requires_grad_(True)
    attn = torch.rand(N, Lq, M, L, P, dtype=torch.float64) + 1e-5
    attn = (attn / attn.sum(-1, keepdim=True).sum(-2, keepdim=True)).to(dev).requires_grad_(True)
    assert torch.autograd.gradcheck(MSDeformAttnFunction.apply, (value, shapes, loc, attn, 2))


def test_empty_queries_and_error_behaviour():
    from kinet_b200.functions import MSDeformAttnFunction, ms_deform_attn_backward
    dev = _dev()
    value, shapes, loc, attn, go = _random_case(dev, 2, 8, 32, 0, [(4, 4)], 4)
    out = MSDeformAttnFunction.apply(value, shapes, loc, attn, 64)
    assert out.shape == (2, 0, 256)
    gv, gl, ga = ms_deform_attn_backward(value, shapes, loc, attn, go, 64)
    assert not gv.any() and gl.numel() == 0 and ga.numel() == 0
    value, shapes, loc, attn, go = _random_case(dev, 3, 8, 32, 5, [(4, 4)], 4)
    with pytest.raises(RuntimeError):          # batch(3) % im2col_step(2) != 0, cu:46-48
        MSDeformAttnFunction.apply(value, shapes, loc, attn, 2)
    with pytest.raises(RuntimeError):          # value must be contiguous, cu:29
        MSDeformAttnFunction.apply(value.transpose(2, 3).contiguous().transpose(2, 3), shapes, loc, attn, 64)
    with pytest.raises(RuntimeError):          # CPU tensors: "Not implemented on the CPU", ms_deform_attn.h:27
        MSDeformAttnFunction.apply(value.cpu(), shapes.cpu(), loc.cpu(), attn.cpu(), 64)


def test_noncontiguous_loc_attn_are_handled():
    from kinet_b200.functions import ms_deform_attn_forward
    dev = _dev()
    value, shapes, loc, attn, go = _random_case(dev, 2, 8, 32, 9, [(6, 7), (3, 4)], 4)
    loc_nc = loc.permute(0, 2, 1, 3, 4, 5).contiguous().permute(0, 2, 1, 3, 4, 5)
    assert not loc_nc.is_contiguous()
    assert torch.equal(ms_deform_attn_forward(value, shapes, loc_nc, attn, 64),
                       ms_deform_attn_forward(value, shapes, loc, attn, 64))


def test_backward_run_to_run_spread():
    """grad_value uses floating-point reductions in global memory: order varies, spread is bounded."""
    dev = _dev()
    levels = [(20, 30), (10, 15)]
    value, shapes, loc, attn, go = _random_case(dev, 1, 8, 32, 600, levels, 4, spread=1.0)
    _, g1 = _run_cuda(value, shapes, loc, attn, go)
    _, g2 = _run_cuda(value, shapes, loc, attn, go)
    assert torch.equal(g1[1], g2[1]) and torch.equal(g1[2], g2[2])     # single-writer outputs
    assert _relerr(g1[0].cpu().numpy(), g2[0].cpu().numpy()) <= 2e-6


def test_module_matches_reference_math():
    """MSDeformAttn (module) on the GPU vs the same module with the sampling core replaced by the
    oracle's grid_sample port -- checks the host-side arithmetic (F4 quirk, query_attn_mask)."""
    from kinet_b200.modules import MSDeformAttn
    from oracle import torch_port_forward
    dev = _dev()
    torch.manual_seed(0)
    levels = [(12, 17), (6, 9), (3, 5), (2, 3)]
    S = sum(h * w for h, w in levels)
    m = MSDeformAttn(256, 4, 8, 4).to(dev)
    with torch.no_grad():
        m.sampling_offsets.weight.normal_(0, 0.02)
        m.attention_weights.weight.normal_(0, 0.5)
    N, Lq = 2, 23
    query = torch.randn(N, Lq, 256, device=dev)
    src = torch.randn(N, S, 256, device=dev)
    shapes = torch.as_tensor(levels, dtype=torch.long, device=dev)
    pad = torch.zeros(N, S, dtype=torch.bool, device=dev)
    pad[1, -40:] = True
    qmask = torch.zeros(N, Lq, dtype=torch.bool, device=dev)
    qmask[0, 3] = True
    for ref_dim in (2, 4):
        ref = torch.rand(N, Lq, 4, ref_dim, device=dev) * 0.6 + 0.2
        got = m(query, ref, src, shapes, pad, qmask)
        # replay the module's host arithmetic with the oracle core
        value = m.value_proj(src).masked_fill(pad[..., None], 0.).view(N, S, 8, 32)
        off = m.sampling_offsets(query).view(N, Lq, 8, 4, 4, 2)
        aw = torch.softmax(m.attention_weights(query).view(N, Lq, 8, 16), -1).view(N, Lq, 8, 4, 4)
        aw = aw.masked_fill(qmask[..., None, None, None], 0.)
        if ref_dim == 2:
            loc = ref[:, :, None, :, None, :] + off / shapes[None, None, None, :, None, :]
        else:
            loc = ref[:, :, None, :, None, :2] + off / 4 * ref[:, :, None, :, None, 2:] * 0.5
        want = m.output_proj(torch_port_forward(value, shapes, loc, aw))
        assert _relerr(got.detach().cpu().numpy(), want.detach().cpu().numpy()) <= 2e-5


def test_c_abi_calls_are_cuda_graph_capturable():
    """The library never allocates, synchronises or touches the host inside a call
    (include/msda_b200.h), so forward + backward can be captured into a CUDA graph and replayed --
    what SURVEY.md 7.3 asks for the per-frame tracker graph.  The replay must reproduce the eager result."""
    import ctypes
    from kinet_b200 import _lib
    dev = _dev()
    lib = _lib.load()
    N, M, D, Lq, levels, P = 2, 8, 32, 300, [(20, 27), (10, 14), (5, 7), (3, 4)], 4
    for dtype in (torch.float32, torch.bfloat16):
        value, shapes, loc, attn, go = _random_case(dev, N, M, D, Lq, levels, P, dtype=dtype, seed=5)
        S = value.shape[1]
        dims = _lib.MsdaDims(N, S, M, D, len(levels), Lq, P)
        vcode, acode = _lib.dtype_code(dtype), _lib.dtype_code(loc.dtype)
        out = torch.zeros(N, Lq, M * D, dtype=dtype, device=dev)
        gv, gl, ga = torch.zeros_like(value), torch.zeros_like(loc), torch.zeros_like(attn)
        wsb = int(lib.msda_b200_backward_workspace_bytes(vcode, ctypes.byref(dims)))
        ws = torch.empty(max(wsb // 4, 4), dtype=torch.float32, device=dev)
        p = lambda t: ctypes.c_void_p(t.data_ptr())

        def run(stream):
            st = ctypes.c_void_p(stream.cuda_stream)
            rc = lib.msda_b200_forward(p(value), vcode, p(shapes), None, p(loc), p(attn), acode, p(out),
                                       ctypes.byref(dims), st)
            assert rc == 0, rc
            rc = lib.msda_b200_backward(p(value), vcode, p(shapes), None, p(loc), p(attn), acode, p(go), p(gv),
                                        p(gl), p(ga), p(ws), ctypes.c_size_t(wsb), ctypes.byref(dims), st)
            assert rc == 0, rc

        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            run(side)                       # eager reference (also warms up lazy module loading)
        side.synchronize()
        want = [t.clone() for t in (out, gv, gl, ga)]
        for t in (out, gv, gl, ga):
            t.zero_()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            run(torch.cuda.current_stream())
        assert not out.any()                # capture records, it does not execute
        graph.replay()
        torch.cuda.synchronize()
        tol = 2e-6 if dtype == torch.float32 else 1e-2      # grad_value: atomic order differs run to run
        for got, ref in zip((out, gv, gl, ga), want):
            assert _relerr(got.float().cpu().numpy(), ref.float().cpu().numpy()) <= tol


def test_extension_module_shim_on_gpu():
    """`import MultiScaleDeformableAttention as MSDA` (reference ms_deform_attn_func.py:11) used exactly like
    the reference's autograd function uses it (:18-19, :26-29), 5/6 positional arguments."""
    import MultiScaleDeformableAttention as MSDA
    dev = _dev()
    value, shapes, loc, attn, go = _random_case(dev, 2, 8, 32, 41, [(9, 13), (5, 7), (3, 4)], 4, seed=8)
    out = MSDA.ms_deform_attn_forward(value, shapes, loc, attn, 64)
    gv, gl, ga = MSDA.ms_deform_attn_backward(value, shapes, loc, attn, go, 64)
    ref_out, ref_grads = _oracle64(value, shapes, loc, attn, go)
    assert out.shape == (2, 41, 256)
    assert _relerr(out.cpu().numpy(), ref_out) <= 1e-5
    _check_grads((gv, gl, ga), ref_grads, shapes, loc, 1e-5)


# ---- fused module prologue (softmax + query mask + location arithmetic inside the kernels) ----------
def _unfused_reference(value, shapes, ref, off, logit, qmask, norm):
    """The reference module's own op sequence (ms_deform_attn.py:69-82) + the plain autograd function."""
    from kinet_b200.functions import MSDeformAttnFunction
    N, Lq, M, L, P, _ = off.shape
    aw = torch.softmax(logit.float(), -1).view(N, Lq, M, L, P)
    if qmask is not None:
        aw = aw.masked_fill(qmask[..., None, None, None], 0.)
    if ref.shape[-1] == 2:
        loc = ref[:, :, None, :, None, :] + off.float() / norm[None, None, None, :, None, :]
    else:
        loc = ref[:, :, None, :, None, :2] + off.float() / P * ref[:, :, None, :, None, 2:] * 0.5
    return MSDeformAttnFunction.apply(value, shapes, loc, aw, 64)


FUSED_CASES = [
    # (N, M, D, Lq, levels, ref_dim, masked, dtype)
    (2, 8, 32, 57, [(12, 17), (6, 9), (3, 5), (2, 3)], 2, True, torch.float32),
    (2, 8, 32, 57, [(12, 17), (6, 9), (3, 5), (2, 3)], 4, False, torch.float32),
    (1, 8, 36, 31, [(7, 11), (4, 6)] * 4, 2, True, torch.float32),          # D = 36 (GS = 9), L*P = 32
    (1, 2, 16, 40, [(10, 10), (5, 5), (3, 3), (1, 1)], 4, True, torch.float32),   # D = 16: L*P = 16 = 4 chunks
    (2, 4, 64, 19, [(6, 5), (3, 3)], 2, False, torch.float32),
    (2, 8, 32, 57, [(12, 17), (6, 9), (3, 5), (2, 3)], 2, True, torch.bfloat16),
    (2, 8, 32, 57, [(12, 17), (6, 9), (3, 5), (2, 3)], 4, True, torch.float16),
    # 16-bit D = 36 with the 5-lane "overhang" forward forced (msda_fwd.cuh): softmax over 5-lane groups, 4 chunks of 4
    (2, 8, 36, 57, [(12, 17), (6, 9), (3, 5), (2, 3)], 2, True, torch.bfloat16, {"MSDA_B200_D36_WIDE": "1"}),
    (1, 4, 36, 31, [(12, 17), (6, 9), (3, 5), (2, 3)], 4, False, torch.float16, {"MSDA_B200_D36_WIDE": "1"}),
]


@pytest.mark.parametrize("case", FUSED_CASES)
def test_fused_prologue_matches_unfused(case, monkeypatch):
    from kinet_b200.functions import MSDeformAttnFusedFunction, fused_supported
    dev = _dev()
    if len(case) == 9:
        for k_, v_ in case[8].items():
            monkeypatch.setenv(k_, v_)
        case = case[:8]
    N, M, D, Lq, levels, ref_dim, masked, dtype = case
    L, P = len(levels), 4
    S = sum(h * w for h, w in levels)
    g = torch.Generator().manual_seed(31)
    value = torch.randn(N, S, M, D, generator=g).to(dtype).to(dev)
    off = (torch.randn(N, Lq, M, L, P, 2, generator=g) * 2.5).to(dtype).to(dev)
    logit = (torch.randn(N, Lq, M, L * P, generator=g) * 2).to(dtype).to(dev)
    ref = (torch.rand(N, Lq, L, ref_dim, generator=g) * (0.6 if ref_dim == 2 else 0.3) + 0.2).to(dev)
    shapes = torch.as_tensor(levels, dtype=torch.int64, device=dev)
    qmask = None
    if masked:
        qmask = torch.zeros(N, Lq, dtype=torch.bool, device=dev)
        qmask[0, 3] = True
        qmask[-1, -1] = True
    go = torch.randn(N, Lq, M * D, generator=g).to(dtype).to(dev)
    assert fused_supported(value, L, P)

    def run(fn):
        leaves = [t.clone().requires_grad_(True) for t in (value, ref, off, logit)]
        out = fn(*leaves)
        out.backward(go)
        return out.detach(), [t.grad for t in leaves]

    out_f, grads_f = run(lambda v, r, o, lg: MSDeformAttnFusedFunction.apply(v, shapes, r, o, lg, qmask, shapes, 64))
    out_u, grads_u = run(lambda v, r, o, lg: _unfused_reference(v, shapes, r, o, lg, qmask, shapes))
    tol = 1e-5 if dtype == torch.float32 else 1.2e-2
    assert out_f.dtype == value.dtype
    assert _relerr(out_f.float().cpu().numpy(), out_u.float().cpu().numpy()) <= tol
    names = ("grad_value", "grad_reference_points", "grad_offsets", "grad_logits")
    for name, a, b in zip(names, grads_f, grads_u):
        assert a is not None and a.dtype == b.dtype and a.shape == b.shape, name
        a, b = a.float(), b.float()
        scale = float(b.abs().max().clamp_min(1e-30))
        err = (a - b).abs() / scale
        if name in ("grad_offsets", "grad_reference_points"):
            # a location that lands within rounding distance of a bilinear kink may take the other
            # one-sided derivative in the two paths: allow a vanishing fraction of outliers
            assert float((err > tol).float().mean()) <= 2e-3, name
        else:
            assert float(err.max()) <= tol, (name, float(err.max()))
    if masked:
        assert not grads_f[3][0, 3].any() and not out_f[0, 3].any()


def test_module_fused_and_unfused_agree_under_autocast():
    """MSDeformAttn with fused_prologue on/off, fp32 and bf16 autocast, forward and parameter gradients."""
    from kinet_b200.modules import MSDeformAttn
    dev = _dev()
    torch.manual_seed(1)
    levels = [(12, 17), (6, 9), (3, 5), (2, 3)]
    S = sum(h * w for h, w in levels)
    m = MSDeformAttn(256, 4, 8, 4).to(dev)
    with torch.no_grad():
        m.sampling_offsets.weight.normal_(0, 0.02)
        m.attention_weights.weight.normal_(0, 0.5)
    N, Lq = 2, 37
    query = torch.randn(N, Lq, 256, device=dev)
    src = torch.randn(N, S, 256, device=dev)
    ref = torch.rand(N, Lq, 4, 2, device=dev) * 0.6 + 0.2
    shapes = torch.as_tensor(levels, dtype=torch.long, device=dev)
    for amp in (False, True):
        res = []
        for fused in (True, False):
            m.fused_prologue = fused
            m.zero_grad()
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
                out = m(query, ref, src, shapes)
            out.float().square().sum().backward()
            res.append((out.detach().float(), [p.grad.detach().clone() for p in m.parameters()]))
        tol = 2e-2 if amp else 2e-5
        assert _relerr(res[0][0].cpu().numpy(), res[1][0].cpu().numpy()) <= tol
        for (name, _), a, b in zip(m.named_parameters(), res[0][1], res[1][1]):
            if amp:
                # bf16 autocast: both paths round activations and gradients to bf16 at different places
                # (the unfused path even divides the offsets in bf16), so compare in the Frobenius norm
                err = float((a - b).norm() / b.norm().clamp_min(1e-30))
                # the PyTorch prologue under autocast divides the bf16 offsets in bf16 (bf16 / int64 ->
                # bf16), which moves ~1 % of the sampling locations of this white-noise value field across
                # a pixel boundary and flips their one-sided location derivative: the offset branch can
                # only be compared loosely here (test_fused_prologue_matches_unfused is the exact check)
                assert err <= (0.15 if name.startswith("sampling_offsets") else 3e-2), (name, err)
            else:
                assert _relerr(a.cpu().numpy(), b.cpu().numpy()) <= 1e-4, name
    m.fused_prologue = True


def test_concurrent_threads_and_streams():
    """Boundary row "streams / threading" (SURVEY.md 8(b)): the library keeps no mutable state shared between
    callers, ctypes releases the GIL around the calls, and every launch goes to the caller's stream -- two
    Python threads driving two streams at once get the same results as running alone."""
    import threading
    dev = _dev()
    cases = [_random_case(dev, 2, 8, 32, 700, [(20, 30), (10, 10)], 4, seed=s) for s in (1, 2)]
    want = [_run_cuda(*c) for c in cases]
    got, errs = [None, None], []

    def worker(i):
        try:
            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                for _ in range(20):
                    res = _run_cuda(*cases[i])
            s.synchronize()
            got[i] = res
        except Exception as e:          # surfaced below: an exception in a thread does not fail the test by itself
            errs.append(e)

    threads = [threading.Thread(target=worker, args=(i,)) for i in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errs, errs
    for (out, grads), (wout, wgrads) in zip(got, want):
        assert torch.equal(out, wout)
        assert torch.equal(grads[1], wgrads[1]) and torch.equal(grads[2], wgrads[2])
        assert _relerr(grads[0].cpu().numpy(), wgrads[0].cpu().numpy()) <= 1e-5     # atomics: order varies


def test_c_abi_from_a_cpp_host_program(tmp_path):
    """The drop-in boundary is a C ABI: tests/abi_c/abi_smoke.cpp includes include/msda_b200.h, links
    libmsda_b200.so and the CUDA runtime -- no torch, no Python in the process -- runs forward + backward in
    fp32 and bf16 on an encoder-form problem and checks them against the C oracle linked into the test."""
    import shutil
    import subprocess
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available on this box")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    libdir = os.path.join(root, "kinet_b200")
    exe, obj = str(tmp_path / "abi_smoke"), str(tmp_path / "msda_oracle.o")
    env = {k: v for k, v in os.environ.items() if k not in ("CC", "CXX")}
    subprocess.check_call(["gcc", "-O2", "-c", os.path.join(root, "oracle", "msda_oracle.c"), "-o", obj], env=env)
    subprocess.check_call([nvcc, "-O2", "-std=c++17", "-Wno-deprecated-gpu-targets", "-I", os.path.join(root, "include"),
                           "-o", exe, os.path.join(root, "tests", "abi_c", "abi_smoke.cpp"), obj,
                           "-L", libdir, "-lmsda_b200", "-Xlinker", "-rpath", "-Xlinker", libdir, "-lm"], env=env)
    res = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    print(res.stdout)
    assert res.returncode == 0, res.stdout + res.stderr


@pytest.mark.parametrize("case", [
    # (N, M, D, levels, P, Lq or None = encoder form, reference dim, masked)
    (2, 8, 32, [(9, 13), (5, 7), (3, 4), (2, 2)], 4, 37, 2, True),
    (1, 8, 36, [(7, 11), (4, 6)] * 4, 4, 29, 4, False),
    (1, 2, 32, [(12, 17), (6, 9), (3, 5), (2, 3)], 4, None, 2, True),        # encoder form: dense-window backward in bf16
], ids=["D32_L4_ref2_masked", "D36_L8_ref4", "encoder_D32_ref2_masked"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["f32", "bf16"])
def test_fused_forward_backward_against_torch_port_fp64(case, dtype):
    """The product default -- MSDeformAttnFusedFunction (module prologue inside the kernels), forward AND backward --
    directly against the oracle: the reference module's prologue arithmetic (ops/modules/ms_deform_attn.py:69-82)
    + torch_port_forward (the restated ms_deform_attn_core_pytorch) + autograd, all in fp64."""
    from kinet_b200.functions import MSDeformAttnFusedFunction
    from oracle import torch_port_forward
    N, M, D, levels, P, Lq, ref_dim, masked = case
    dev = _dev()
    L = len(levels)
    S = sum(h * w for h, w in levels)
    Lq = S if Lq is None else Lq
    g = torch.Generator(device="cpu").manual_seed(21)
    value = torch.randn(N, S, M, D, generator=g).to(dtype)
    shapes = torch.as_tensor(levels, dtype=torch.int64)
    if Lq == S:
        pts = []
        for h, w in levels:
            ys = (torch.arange(h, dtype=torch.float32) + 0.5) / h
            xs = (torch.arange(w, dtype=torch.float32) + 0.5) / w
            yy, xx = torch.meshgrid(ys, xs, indexing="ij")
            pts.append(torch.stack([xx.reshape(-1), yy.reshape(-1)], -1))
        ref = torch.cat(pts, 0)[None, :, None, :].expand(N, -1, L, -1)
    else:
        ref = torch.rand(N, Lq, 1, 2, generator=g).expand(-1, -1, L, -1) * 0.9 + 0.05
    if ref_dim == 4:
        ref = torch.cat([ref, torch.rand(N, Lq, L, 2, generator=g) * 0.2 + 0.05], -1)
    ref = ref.contiguous()
    off = torch.randn(N, Lq, M, L, P, 2, generator=g) * 1.5
    logit = torch.randn(N, Lq, M, L * P, generator=g)
    mask = (torch.rand(N, Lq, generator=g) < 0.15) if masked else None
    go = torch.randn(N, Lq, M * D, generator=g).to(dtype)

    leaves = [t.to(dev).clone().requires_grad_(True) for t in (value, ref, off, logit)]
    out = MSDeformAttnFusedFunction.apply(leaves[0], shapes.to(dev), leaves[1], leaves[2], leaves[3],
                                          mask.to(dev) if mask is not None else None, shapes.to(dev), 64)
    out.backward(go.to(dev))
    torch.cuda.synchronize()

    r = [t.double().clone().requires_grad_(True) for t in (value, ref, off, logit)]      # rounded inputs, fp64 arithmetic
    attn = torch.softmax(r[3], -1).view(N, Lq, M, L, P)
    if mask is not None:
        attn = attn.masked_fill(mask[..., None, None, None], 0.0)
    if ref_dim == 2:        # the reference divides by spatial_shapes in (H, W) order (:77-79); the module passes them as is
        loc = r[1][:, :, None, :, None, :] + r[2] / shapes.double()[None, None, None, :, None, :]
    else:
        loc = r[1][:, :, None, :, None, :2] + r[2] / P * r[1][:, :, None, :, None, 2:] * 0.5
    ref_out = torch_port_forward(r[0], shapes, loc, attn)
    ref_out.backward(go.double())
    tol = 2e-5 if dtype == torch.float32 else 1.2e-2
    assert _relerr(out.detach().float().cpu().numpy(), ref_out.detach().numpy()) <= tol
    assert _relerr(leaves[0].grad.float().cpu().numpy(), r[0].grad.numpy()) <= tol          # grad_value
    assert _relerr(leaves[3].grad.float().cpu().numpy(), r[3].grad.numpy()) <= tol          # grad of the logits
    # gradients through the sampling locations: kinks masked (one-sided derivatives), fraction bounded
    hw = shapes.double()
    x = loc.detach()[..., 0] * hw[None, None, None, :, None, 1] - 0.5
    y = loc.detach()[..., 1] * hw[None, None, None, :, None, 0] - 0.5
    ok = (((x - x.round()).abs() > 2e-4) & ((y - y.round()).abs() > 2e-4))
    assert 1.0 - float(ok.double().mean()) < 0.005
    goff, roff = leaves[2].grad.double().cpu(), r[2].grad
    err = ((goff - roff).abs() * ok[..., None]).max() / roff.abs().max()
    assert float(err) <= tol * (1 if dtype == torch.float32 else 2), float(err)
    if bool(ok.all()):      # the reference-point gradient sums over samples: compare only when no sample sits on a kink
        assert _relerr(leaves[1].grad.float().cpu().numpy(), r[1].grad.numpy()) <= tol * 4
