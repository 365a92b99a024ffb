"""ctypes front-end of oracle/msda_oracle.c plus a torch restatement of the
reference's CPU-capable path.  TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Two independent restatements live here:

* ``oracle_forward`` / ``oracle_backward`` call the plain-C scalar restatement
  of the reference CUDA kernels' arithmetic
  (src/trackformer/models/ops/src/cuda/ms_deform_im2col_cuda.cuh:24-163,
  218-235, 281-304, 352-376).  numpy in, numpy out, fp32 or fp64.
* ``torch_port_forward`` restates the reference's own CPU-runnable function
  ``ms_deform_attn_core_pytorch`` (src/trackformer/models/ops/functions/
  ms_deform_attn_func.py:34-54): one ``F.grid_sample`` per level, then the
  attention-weighted sum.  It is the timed "port" CPU baseline in bench.py
  because the reference file itself cannot travel to the GPU box.

Parity pin: both are checked against golden vectors produced by the real
reference function (tests/golden/, generator tests/golden/make_golden.py).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libmsda_oracle.so")
_lib = None


def build_oracle(force=False):
    """Compile oracle/msda_oracle.c with gcc (seconds)."""
    src = os.path.join(_HERE, "msda_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "libmsda_oracle.so"])
    return _SO


def _load():
    global _lib
    if _lib is None:
        build_oracle()
        lib = ctypes.CDLL(_SO)
        i = ctypes.c_int
        vp = ctypes.c_void_p
        for sfx in ("f32", "f64"):
            f = getattr(lib, "msda_oracle_forward_" + sfx)
            f.argtypes = [vp, vp, vp, vp, vp, i, i, i, i, i, i, i, i]
            f.restype = i
            g = getattr(lib, "msda_oracle_backward_" + sfx)
            g.argtypes = [vp] * 8 + [i] * 8
            g.restype = i
        lib.msda_oracle_max_threads.restype = i
        _lib = lib
    return _lib


def oracle_max_threads():
    return int(_load().msda_oracle_max_threads())


def _prep(value, shapes, loc, attn):
    dt = np.float64 if value.dtype == np.float64 else np.float32
    value = np.ascontiguousarray(value, dtype=dt)
    loc = np.ascontiguousarray(loc, dtype=dt)
    attn = np.ascontiguousarray(attn, dtype=dt)
    shapes = np.ascontiguousarray(shapes, dtype=np.int64)
    N, S, M, D = value.shape
    _, Lq, _, L, P, _ = loc.shape
    assert shapes.shape == (L, 2) and int((shapes[:, 0] * shapes[:, 1]).sum()) == S
    assert attn.shape == (N, Lq, M, L, P)
    return dt, value, shapes, loc, attn, (N, S, M, D, L, Lq, P)


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def oracle_forward(value, shapes, loc, attn, nthreads=1):
    """-> out (N, Lq, M*D), same dtype family as ``value`` (fp32 or fp64)."""
    dt, value, shapes, loc, attn, dims = _prep(value, shapes, loc, attn)
    N, S, M, D, L, Lq, P = dims
    out = np.empty((N, Lq, M * D), dtype=dt)
    fn = getattr(_load(), "msda_oracle_forward_" + ("f64" if dt == np.float64 else "f32"))
    rc = fn(_p(value), _p(shapes), _p(loc), _p(attn), _p(out), *dims, int(nthreads))
    if rc != 0:
        raise RuntimeError("msda oracle forward failed rc=%d" % rc)
    return out


def oracle_backward(value, shapes, loc, attn, grad_out, nthreads=1):
    """-> (grad_value, grad_sampling_loc, grad_attn_weight)."""
    dt, value, shapes, loc, attn, dims = _prep(value, shapes, loc, attn)
    N, S, M, D, L, Lq, P = dims
    grad_out = np.ascontiguousarray(grad_out, dtype=dt).reshape(N, Lq, M * D)
    gv = np.empty_like(value)
    gl = np.empty_like(loc)
    ga = np.empty_like(attn)
    fn = getattr(_load(), "msda_oracle_backward_" + ("f64" if dt == np.float64 else "f32"))
    rc = fn(_p(value), _p(shapes), _p(loc), _p(attn), _p(grad_out), _p(gv), _p(gl), _p(ga),
            *dims, int(nthreads))
    if rc != 0:
        raise RuntimeError("msda oracle backward failed rc=%d" % rc)
    return gv, gl, ga


def torch_port_forward(value, shapes, loc, attn):
    """Per-level grid_sample restatement of the reference's CPU path.

    value (N,S,M,D), shapes (L,2) [H,W], loc (N,Lq,M,L,P,2) in [0,1] (x,y),
    attn (N,Lq,M,L,P) -> (N,Lq,M*D).  torch tensors on any device.
    """
    import torch
    import torch.nn.functional as F

    N, S, M, D = value.shape
    Lq, L, P = loc.shape[1], loc.shape[3], loc.shape[4]
    hw = [(int(h), int(w)) for h, w in shapes.tolist()]
    # head-major image stack per level: (N*M, D, H, W)
    per_head = value.permute(0, 2, 3, 1).reshape(N * M, D, S)
    grid = (loc * 2 - 1).permute(0, 2, 1, 3, 4, 5).reshape(N * M, Lq, L, P, 2)
    taps = []
    off = 0
    for lvl, (h, w) in enumerate(hw):
        img = per_head[:, :, off:off + h * w].reshape(N * M, D, h, w)
        off += h * w
        taps.append(F.grid_sample(img, grid[:, :, lvl], mode="bilinear",
                                  padding_mode="zeros", align_corners=False))
    taps = torch.cat(taps, dim=-1)                       # (N*M, D, Lq, L*P)
    wts = attn.permute(0, 2, 1, 3, 4).reshape(N * M, 1, Lq, L * P)
    out = (taps * wts).sum(-1)                           # (N*M, D, Lq)
    return out.reshape(N, M * D, Lq).transpose(1, 2).contiguous()
