"""Autograd wrapper over the short-sequence attention kernels (include/kinet_layers_b200.h,
csrc/small_mha.cu): the attention core of ``nn.MultiheadAttention`` as the reference's KinematicTransformer
layers call it (src/trackformer/models/transformer.py:331-333, :393-401).

    out = small_mha(q, k, v, key_padding_mask=None, scale=None)

q (N, Lq, H, D), k / v (N, Lk, H, D): CUDA tensors of one floating dtype (fp32 / bf16 / fp16), any batch / row
strides (slices of a packed projection output and sequence-first ``permute`` views are fine) with the head and
channel dimensions packed (``stride(2) == D``, ``stride(3) == 1``).  key_padding_mask (N, Lk) bool, True =
ignore the key.  Returns (N, Lq, H * D), contiguous.  fp32 arithmetic inside for every dtype; attention-weight
dropout (``nn.MultiheadAttention(dropout=p)`` in training mode) is not implemented here -- the module applies it
to the output instead (see modules/kinematic_transformer.py).  CPU tensors raise, like the op itself.
"""
import ctypes

import torch
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from .. import _lib

_DTYPES = (torch.float32, torch.bfloat16, torch.float16)


def _strides(t, what):
    """(batch_stride, row_stride) in elements of a (N, L, H, D) view with packed heads."""
    N, L, H, D = t.shape
    if (D > 1 and t.stride(3) != 1) or (H > 1 and t.stride(2) != D):
        raise RuntimeError("kinet_b200.small_mha: %s must have packed (head, channel) dimensions" % what)
    return t.stride(0), t.stride(1)      # (the stride of a size-1 dimension is never multiplied by a non-zero index)


def _args(q, k, v, mask_u8, out, lse, scale):
    N, Lq, H, D = q.shape
    Lk = k.shape[1]
    qb, qr = _strides(q, "q")
    kb, kr = _strides(k, "k")
    vb, vr = _strides(v, "v")
    return _lib.KinetMhaArgs(q.data_ptr(), k.data_ptr(), v.data_ptr(), qb, qr, kb, kr, vb, vr,
                             None if mask_u8 is None else mask_u8.data_ptr(),
                             out.data_ptr(), Lq * H * D, H * D,
                             None if lse is None else lse.data_ptr(),
                             _lib.dtype_code(q.dtype), N, H, D, Lq, Lk, float(scale))


def supported(q, k, v):
    return (q.is_cuda and q.dtype in _DTYPES and k.dtype == q.dtype and v.dtype == q.dtype
            and q.dim() == 4 and 0 < q.shape[3] <= _lib.MHA_MAX_D)


class SmallMHAFunction(Function):
    @staticmethod
    def forward(ctx, q, k, v, key_padding_mask, scale):
        if not (q.is_cuda and k.is_cuda and v.is_cuda):
            raise RuntimeError("Not implemented on the CPU")
        if not supported(q, k, v):
            raise RuntimeError("kinet_b200.small_mha: q / k / v must share one of fp32 / bf16 / fp16 and D <= %d"
                               % _lib.MHA_MAX_D)
        N, Lq, H, D = q.shape
        if k.shape != v.shape or k.shape[0] != N or k.shape[2:] != q.shape[2:]:
            raise RuntimeError("kinet_b200.small_mha: shapes q %s k %s v %s" % (tuple(q.shape), tuple(k.shape), tuple(v.shape)))
        mask_u8 = None
        if key_padding_mask is not None:
            if tuple(key_padding_mask.shape) != (N, k.shape[1]):
                raise RuntimeError("kinet_b200.small_mha: key_padding_mask must be (N, Lk)")
            mask_u8 = key_padding_mask.to(torch.uint8).contiguous()
        need = any(ctx.needs_input_grad[:3])
        out = torch.empty(N, Lq, H * D, dtype=q.dtype, device=q.device)
        lse = torch.empty(N, H, Lq, dtype=torch.float32, device=q.device) if need else None
        a = _args(q, k, v, mask_u8, out, lse, scale)
        lib = _lib.load()
        with _lib.device_guard(q.device):
            rc = lib.kinet_b200_small_mha_forward(ctypes.byref(a), _lib.raw_stream(q.device))
        _lib.check(rc, "kinet_b200_small_mha_forward")
        if need:
            ctx.save_for_backward(q, k, v, mask_u8, out, lse)
            ctx.scale = float(scale)
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, dout):
        q, k, v, mask_u8, out, lse = ctx.saved_tensors
        N, Lq, H, D = q.shape
        Lk = k.shape[1]
        dout = dout.contiguous()
        if dout.dtype != q.dtype:
            dout = dout.to(q.dtype)
        dq = torch.empty(N, Lq, H, D, dtype=q.dtype, device=q.device)
        dk = torch.empty(N, Lk, H, D, dtype=q.dtype, device=q.device)
        dv = torch.empty(N, Lk, H, D, dtype=q.dtype, device=q.device)
        b = _lib.KinetMhaBwdArgs(_args(q, k, v, mask_u8, out, lse, ctx.scale), dout.data_ptr(), Lq * H * D, H * D,
                                 dq.data_ptr(), dk.data_ptr(), dv.data_ptr())
        lib = _lib.load()
        with _lib.device_guard(q.device):
            rc = lib.kinet_b200_small_mha_backward(ctypes.byref(b), _lib.raw_stream(q.device))
        _lib.check(rc, "kinet_b200_small_mha_backward")
        return dq, dk, dv, None, None


def small_mha(q, k, v, key_padding_mask=None, scale=None):
    """softmax(q k^T * scale + key_padding_mask) v per head; see the module docstring."""
    if scale is None:
        scale = q.shape[-1] ** -0.5
    return SmallMHAFunction.apply(q, k, v, key_padding_mask, scale)
