"""Harness around oracle/_ref/libmsda_ref_cuda.so -- the REFERENCE's own CUDA kernels compiled for
sm_100a (oracle/ref_cuda/).  TEST INFRASTRUCTURE ONLY: a second oracle at full sizes on the GPU box
and the same-box GPU baseline.  The host logic below restates ms_deform_attn_cuda_forward/_backward
(src/trackformer/models/ops/src/cuda/ms_deform_attn_cuda.cu:50-83, :119-163) with im2col_step = N:
level starts from the shapes, a `columns` scratch of L*P x the output reduced with sum(0), zero-filled
gradients, coord kernel then col2im kernel."""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(_HERE, "_ref", "libmsda_ref_cuda.so")
_lib = None


def available():
    return os.path.exists(SO)


def _load():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(SO)
    return _lib


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _sfx(t):
    return {torch.float32: "f32", torch.float64: "f64"}[t.dtype]


def _level_start(shapes):
    return torch.cat([shapes.new_zeros(1), (shapes[:, 0] * shapes[:, 1]).cumsum(0)[:-1]]).contiguous()


def ref_cuda_forward(value, shapes, loc, attn, columns=None):
    lib = _load()
    N, S, M, D = value.shape
    _, Lq, _, L, P, _ = loc.shape
    ls = _level_start(shapes)
    if columns is None:
        columns = torch.empty(L * P, N, Lq, M, D, dtype=value.dtype, device=value.device)
    st = torch.cuda.current_stream().cuda_stream
    rc = getattr(lib, "msda_ref_im2col_" + _sfx(value))(
        ctypes.c_void_p(st), _p(value), _p(shapes), _p(ls), _p(loc), _p(attn),
        N, S, M, D, L, Lq, P, _p(columns))
    assert rc == 0, rc
    return columns.sum(0).view(N, Lq, M * D)


def ref_cuda_backward(value, shapes, loc, attn, grad_out):
    lib = _load()
    N, S, M, D = value.shape
    _, Lq, _, L, P, _ = loc.shape
    ls = _level_start(shapes)
    gv, gl, ga = torch.zeros_like(value), torch.zeros_like(loc), torch.zeros_like(attn)
    go = grad_out.contiguous()
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    sfx = _sfx(value)
    rc = getattr(lib, "msda_ref_col2im_coord_" + sfx)(
        st, _p(go), _p(value), _p(shapes), _p(ls), _p(loc), _p(attn), N, S, M, D, L, Lq, P, _p(gl), _p(ga))
    assert rc == 0, rc
    rc = getattr(lib, "msda_ref_col2im_" + sfx)(
        st, _p(go), _p(shapes), _p(ls), _p(loc), _p(attn), N, S, M, D, L, Lq, P, _p(gv))
    assert rc == 0, rc
    return gv, gl, ga
