"""MSDeformAttnFunction -- autograd wrapper over the sm_100a kernels.

Mirrors the reference's src/trackformer/models/ops/functions/ms_deform_attn_func.py:
``MSDeformAttnFunction.apply(value, value_spatial_shapes, sampling_locations,
attention_weights, im2col_step)`` (:14-31, the 5-argument form the reference's
MSDeformAttn module uses) and additionally accepts the upstream Deformable-DETR
6-argument form with ``value_level_start_index`` in position 2.

The two module-level functions ``ms_deform_attn_forward`` / ``ms_deform_attn_backward``
have the signatures of the reference's native extension (src/vision.cpp:4-7,
src/ms_deform_attn.h:10-49) and are what MultiScaleDeformableAttention.py (the
drop-in shim for ``import MultiScaleDeformableAttention as MSDA``) re-exports.

There is no CPU path: CPU tensors raise, like the reference (ms_deform_attn.h:27,48).
"""
import ctypes
import os

import torch
import torch.nn.functional as F
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from .. import _lib
from .._cache import PerTensorCache


def _dims(value, loc):
    N, S, M, D = value.shape
    _, Lq, M2, L, P, two = loc.shape
    if M2 != M or two != 2 or loc.shape[0] != N:
        raise RuntimeError("sampling_loc must be (N, Lq, M, L, P, 2) matching value (N, S, M, D)")
    return _lib.MsdaDims(N, S, M, D, L, Lq, P)


def _check_inputs(value, spatial_shapes, sampling_loc, attn_weight, im2col_step):
    # same preconditions and order as ms_deform_attn_cuda.cu:29-34,46-48
    if not value.is_cuda:
        raise RuntimeError("Not implemented on the CPU")
    if not value.is_contiguous():
        raise RuntimeError("value tensor has to be contiguous")
    if not spatial_shapes.is_cuda:
        raise RuntimeError("spatial_shapes must be a CUDA tensor")
    if not sampling_loc.is_cuda:
        raise RuntimeError("sampling_loc must be a CUDA tensor")
    if not attn_weight.is_cuda:
        raise RuntimeError("attn_weight must be a CUDA tensor")
    batch = value.shape[0]
    step = min(batch, int(im2col_step))
    if step <= 0 or batch % step != 0:
        raise RuntimeError("batch(%d) must divide im2col_step(%d)" % (batch, step))
    if value.dim() != 4 or sampling_loc.dim() != 6 or attn_weight.dim() != 5:
        raise RuntimeError("expected value (N,S,M,D), sampling_loc (N,Lq,M,L,P,2), attn_weight (N,Lq,M,L,P)")
    if spatial_shapes.dim() != 2 or spatial_shapes.shape[0] != sampling_loc.shape[3]:
        raise RuntimeError("spatial_shapes must be (L, 2)")
    if tuple(attn_weight.shape) != tuple(sampling_loc.shape[:5]):
        raise RuntimeError("attn_weight shape must equal sampling_loc.shape[:5]")


def _aux_dtype(value, sampling_loc, attn_weight):
    """dtype the kernels read loc/attn in: the value dtype when both already have it, else fp32
    (fp64 stays fp64).  Sampling locations in bf16 would cost ~0.3 px at W=167, so mixed
    inputs are promoted, never demoted."""
    if value.dtype == torch.float64:
        return torch.float64
    if value.dtype == torch.float32:
        return torch.float32
    if sampling_loc.dtype == value.dtype and attn_weight.dtype == value.dtype:
        return value.dtype
    return torch.float32


def _prep(value, spatial_shapes, level_start_index, sampling_loc, attn_weight):
    aux = _aux_dtype(value, sampling_loc, attn_weight)
    loc = sampling_loc.to(aux).contiguous()
    attn = attn_weight.to(aux).contiguous()
    shapes = spatial_shapes.to(torch.int64).contiguous()
    lsi = None
    if level_start_index is not None:
        lsi = level_start_index.to(device=value.device, dtype=torch.int64).contiguous()
    return aux, loc, attn, shapes, lsi


def _ptr(t):
    # plain ints: the argtypes declared in _lib.load() convert them (None -> NULL)
    return t.data_ptr() if t is not None and t.numel() > 0 else None


def ms_deform_attn_forward(value, spatial_shapes, sampling_loc, attn_weight, im2col_step,
                           level_start_index=None):
    """-> output (N, Lq, M*D).  Reference: ms_deform_attn_cuda_forward (ms_deform_attn_cuda.cu:19-86)."""
    _check_inputs(value, spatial_shapes, sampling_loc, attn_weight, im2col_step)
    lib = _lib.load()
    aux, loc, attn, shapes, lsi = _prep(value, spatial_shapes, level_start_index, sampling_loc, attn_weight)
    dims = _dims(value, loc)
    out = torch.empty((dims.N, dims.Lq, dims.M * dims.D), dtype=value.dtype, device=value.device)
    with _lib.device_guard(value.device):
        rc = lib.msda_b200_forward(_ptr(value), _lib.dtype_code(value.dtype), _ptr(shapes), _ptr(lsi),
                                   _ptr(loc), _ptr(attn), _lib.dtype_code(aux), _ptr(out),
                                   ctypes.byref(dims), _lib.raw_stream(value.device))
    _lib.check(rc, "ms_deform_attn_forward")
    return out


def ms_deform_attn_backward(value, spatial_shapes, sampling_loc, attn_weight, grad_output, im2col_step,
                            level_start_index=None):
    """-> [grad_value, grad_sampling_loc, grad_attn_weight].
    Reference: ms_deform_attn_cuda_backward (ms_deform_attn_cuda.cu:89-168)."""
    _check_inputs(value, spatial_shapes, sampling_loc, attn_weight, im2col_step)
    if not grad_output.is_cuda:
        raise RuntimeError("grad_output must be a CUDA tensor")
    lib = _lib.load()
    aux, loc, attn, shapes, lsi = _prep(value, spatial_shapes, level_start_index, sampling_loc, attn_weight)
    dims = _dims(value, loc)
    go = grad_output.to(value.dtype).contiguous()
    grad_value = torch.empty_like(value)
    grad_loc = torch.empty_like(loc)
    grad_attn = torch.empty_like(attn)
    vcode = _lib.dtype_code(value.dtype)
    ws_bytes = int(lib.msda_b200_backward_workspace_bytes(vcode, ctypes.byref(dims)))
    ws = torch.empty(ws_bytes // 4, dtype=torch.float32, device=value.device) if ws_bytes else None
    with _lib.device_guard(value.device):
        rc = lib.msda_b200_backward(_ptr(value), vcode, _ptr(shapes), _ptr(lsi), _ptr(loc), _ptr(attn),
                                    _lib.dtype_code(aux), _ptr(go), _ptr(grad_value), _ptr(grad_loc),
                                    _ptr(grad_attn), _ptr(ws), ws_bytes,
                                    ctypes.byref(dims), _lib.raw_stream(value.device))
    _lib.check(rc, "ms_deform_attn_backward")
    if grad_loc.dtype != sampling_loc.dtype:
        grad_loc = grad_loc.to(sampling_loc.dtype)
    if grad_attn.dtype != attn_weight.dtype:
        grad_attn = grad_attn.to(attn_weight.dtype)
    return [grad_value, grad_loc, grad_attn]


class MSDeformAttnFunction(Function):
    """apply(value, spatial_shapes, sampling_locations, attention_weights, im2col_step)   [reference]
    apply(value, spatial_shapes, level_start_index, sampling_locations, attention_weights,
          im2col_step)                                                                   [upstream]"""

    @staticmethod
    def forward(ctx, *args):
        if len(args) == 5:
            value, shapes, loc, attn, im2col_step = args
            lsi = None
        elif len(args) == 6:
            value, shapes, lsi, loc, attn, im2col_step = args
        else:
            raise TypeError("MSDeformAttnFunction.apply takes 5 (reference) or 6 (upstream) arguments, got %d"
                            % len(args))
        ctx.im2col_step = im2col_step
        ctx.nargs = len(args)
        output = ms_deform_attn_forward(value, shapes, loc, attn, im2col_step, lsi)
        if lsi is None:
            ctx.save_for_backward(value, shapes, loc, attn)
        else:
            ctx.save_for_backward(value, shapes, loc, attn, lsi)
        return output

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_output):
        saved = ctx.saved_tensors
        value, shapes, loc, attn = saved[:4]
        lsi = saved[4] if len(saved) == 5 else None
        grad_value, grad_loc, grad_attn = ms_deform_attn_backward(
            value, shapes, loc, attn, grad_output, ctx.im2col_step, lsi)
        if ctx.nargs == 5:
            return grad_value, None, grad_loc, grad_attn, None
        return grad_value, None, None, grad_loc, grad_attn, None


def ms_deform_attn_core_pytorch(value, value_spatial_shapes, sampling_locations, attention_weights):
    """Debug/test helper with the reference's name (ms_deform_attn_func.py:34-54): per-level
    ``F.grid_sample`` + weighted sum in plain PyTorch.  It is NOT used by MSDeformAttn or
    MSDeformAttnFunction -- those always run the CUDA kernels -- and exists so code that imports
    this name from the ops package keeps working."""
    N, S, M, D = value.shape
    Lq, L, P = sampling_locations.shape[1], sampling_locations.shape[3], sampling_locations.shape[4]
    sizes = [(int(h), int(w)) for h, w in value_spatial_shapes.tolist()]
    per_head = value.permute(0, 2, 3, 1).reshape(N * M, D, S)
    grid = (sampling_locations * 2 - 1).permute(0, 2, 1, 3, 4, 5).reshape(N * M, Lq, L, P, 2)
    taps, off = [], 0
    for lvl, (h, w) in enumerate(sizes):
        img = per_head[:, :, off:off + h * w].reshape(N * M, D, h, w)
        off += h * w
        taps.append(F.grid_sample(img, grid[:, :, lvl], mode="bilinear", padding_mode="zeros",
                                  align_corners=False))
    taps = torch.cat(taps, dim=-1)
    wts = attention_weights.permute(0, 2, 1, 3, 4).reshape(N * M, 1, Lq, L * P)
    return (taps * wts).sum(-1).reshape(N, M * D, Lq).transpose(1, 2).contiguous()


# ---------------------------------------------------------------------------------------------------
# Fused module prologue: softmax over the logits, query mask and the sampling-
# location arithmetic of MSDeformAttn.forward (reference ops/modules/ms_deform_attn.py:69-82) run inside
# the kernels; the (N,Lq,M,L,P,2) location tensor and the softmax output are never materialised.
# ---------------------------------------------------------------------------------------------------
def fused_supported(value, n_levels, n_points):
    """True when libmsda_b200 has a fused kernel for this call (else: run the prologue in PyTorch)."""
    D = value.shape[-1]
    if not (value.is_cuda and value.dtype in (torch.float32, torch.bfloat16, torch.float16) and D in (16, 32, 36, 64)):
        return False
    # the vector kernels need 16-byte aligned rows (an offset view of a larger buffer may not be); the library
    # answers MSDA_E_SHAPE for those, and for MSDA_B200_FORCE_GENERIC, instead of running a fused kernel
    if value.data_ptr() % 16 or not value.is_contiguous() or os.environ.get("MSDA_B200_FORCE_GENERIC", "0") not in ("", "0"):
        return False
    # a lane keeps the softmax numerators of <= 4 samples of a row (kFusedMaxChunks, msda_common.cuh); samples
    # are dealt to the first 2^k <= D/VEC lanes of the row's group, VEC = 8 only in the 16-bit forward
    lanes = D // (4 if value.dtype == torch.float32 or D == 36 else 8)
    owners = 1 << (lanes.bit_length() - 1)
    return n_levels * n_points <= 4 * owners


_norm_f32 = PerTensorCache()      # offset_norm tensor object -> its fp32 copy, keyed additionally by device


def _offset_norm_f32(offset_norm, device):
    """fp32 copy of the (L, 2) offset divisors.  Every layer of a frame passes the same spatial_shapes tensor,
    so the int64 -> fp32 conversion (a 2 us kernel, twice per layer) is remembered per tensor object."""
    if offset_norm.dtype == torch.float32 and offset_norm.device == device and offset_norm.is_contiguous():
        return offset_norm
    hit = _norm_f32.get(offset_norm, str(device))
    if hit is not None:
        return hit
    return _norm_f32.put(offset_norm, offset_norm.to(device=device, dtype=torch.float32).contiguous(), str(device))


def _aligned16(t):
    """The kernels read these with 16-byte vector loads; an offset view can start anywhere."""
    return t if t.data_ptr() % 16 == 0 else t.clone()


def _fused_inputs(value, offsets, logits, reference_points, offset_norm, query_mask):
    if value.dtype == torch.float32:
        aux = torch.float32
    elif offsets.dtype == value.dtype and logits.dtype == value.dtype:
        aux = value.dtype
    else:
        aux = torch.float32
    off = offsets.to(aux).contiguous()
    lg = logits.to(aux).contiguous()
    ref = _aligned16(reference_points.to(torch.float32).contiguous())
    norm = _offset_norm_f32(offset_norm, value.device) if offset_norm is not None else None
    mask = query_mask.to(torch.uint8).contiguous() if query_mask is not None else None
    args = _lib.MsdaFusedInputs(off.data_ptr(), lg.data_ptr(), ref.data_ptr(),
                                norm.data_ptr() if norm is not None else None,
                                mask.data_ptr() if mask is not None else None, ref.shape[-1])
    return aux, args, (off, lg, ref, norm, mask)      # keep the tensors alive for the duration of the call


class MSDeformAttnFusedFunction(Function):
    """apply(value, spatial_shapes, reference_points, sampling_offsets, attention_logits, query_attn_mask,
             offset_norm, im2col_step) -> (N, Lq, M*D)

    value (N,S,M,D); reference_points (N,Lq,L,2|4); sampling_offsets (N,Lq,M,L,P,2) and attention_logits
    (N,Lq,M,L*P) straight from the module's Linear layers; query_attn_mask (N,Lq) bool or None;
    offset_norm (L,2): what the (x, y) offsets are divided by for 2-d reference points (the reference
    module divides by input_spatial_shapes, ms_deform_attn.py:77-79), ignored for 4-d boxes."""

    @staticmethod
    def forward(ctx, value, spatial_shapes, reference_points, sampling_offsets, attention_logits, query_attn_mask,
                offset_norm, im2col_step):
        N, S, M, D = value.shape
        _, Lq, M2, L, P, two = sampling_offsets.shape
        if M2 != M or two != 2 or tuple(attention_logits.shape) != (N, Lq, M, L * P):
            raise RuntimeError("sampling_offsets must be (N,Lq,M,L,P,2) and attention_logits (N,Lq,M,L*P)")
        if tuple(reference_points.shape[:3]) != (N, Lq, L) or reference_points.shape[-1] not in (2, 4):
            raise RuntimeError("reference_points must be (N, Lq, L, 2) or (N, Lq, L, 4)")
        _check_inputs(value, spatial_shapes, sampling_offsets, attention_logits.view(N, Lq, M, L, P), im2col_step)
        lib = _lib.load()
        shapes = spatial_shapes.to(torch.int64).contiguous()
        aux, args, keep = _fused_inputs(value, sampling_offsets, attention_logits, reference_points, offset_norm,
                                        query_attn_mask)
        dims = _lib.MsdaDims(N, S, M, D, L, Lq, P)
        out = torch.empty((N, Lq, M * D), dtype=value.dtype, device=value.device)
        with _lib.device_guard(value.device):
            rc = lib.msda_b200_forward_fused(_ptr(value), _lib.dtype_code(value.dtype), _ptr(shapes), None,
                                             ctypes.byref(args), _lib.dtype_code(aux), _ptr(out), ctypes.byref(dims),
                                             _lib.raw_stream(value.device))
        _lib.check(rc, "msda_b200_forward_fused")
        ctx.save_for_backward(value, shapes, reference_points, sampling_offsets, attention_logits,
                              query_attn_mask if query_attn_mask is not None else torch.empty(0),
                              offset_norm if offset_norm is not None else torch.empty(0))
        ctx.im2col_step = im2col_step
        del keep
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_output):
        value, shapes, ref, offsets, logits, mask, norm = ctx.saved_tensors
        mask = mask if mask.numel() else None
        norm = norm if norm.numel() else None
        N, S, M, D = value.shape
        _, Lq, _, L, P, _ = offsets.shape
        lib = _lib.load()
        aux, args, keep = _fused_inputs(value, offsets, logits, ref, norm, mask)
        off_c = keep[0]
        dims = _lib.MsdaDims(N, S, M, D, L, Lq, P)
        go = _aligned16(grad_output.to(value.dtype).contiguous())
        grad_value = torch.empty_like(value)
        grad_off = torch.empty((N, Lq, M, L, P, 2), dtype=aux, device=value.device)
        grad_logit = torch.empty((N, Lq, M, L * P), dtype=aux, device=value.device)
        need_ref = ctx.needs_input_grad[2]
        grad_loc = torch.empty((N, Lq, M, L, P, 2), dtype=torch.float32, device=value.device) if need_ref else None
        vcode = _lib.dtype_code(value.dtype)
        ws_bytes = int(lib.msda_b200_backward_workspace_bytes(vcode, ctypes.byref(dims)))
        ws = torch.empty(ws_bytes // 4, dtype=torch.float32, device=value.device) if ws_bytes else None
        with _lib.device_guard(value.device):
            rc = lib.msda_b200_backward_fused(_ptr(value), vcode, _ptr(shapes), None, ctypes.byref(args),
                                              _lib.dtype_code(aux), _ptr(go), _ptr(grad_value), _ptr(grad_off),
                                              _ptr(grad_logit), _ptr(grad_loc), _ptr(ws), ws_bytes,
                                              ctypes.byref(dims), _lib.raw_stream(value.device))
        _lib.check(rc, "msda_b200_backward_fused")
        grad_ref = None
        if need_ref:
            # location = ref_xy + ...: every (head, point) sample of a (query, level) shares the reference point
            if ref.shape[-1] == 2:
                grad_ref = grad_loc.sum(dim=(2, 4))
            else:
                g_xy = grad_loc.sum(dim=(2, 4))
                g_wh = (grad_loc * off_c.float()).sum(dim=(2, 4)) * (0.5 / P)
                grad_ref = torch.cat([g_xy, g_wh], -1)
            grad_ref = grad_ref.to(ref.dtype)
        return (grad_value, None, grad_ref, grad_off.to(offsets.dtype), grad_logit.to(logits.dtype), None, None, None)
