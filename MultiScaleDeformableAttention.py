"""Drop-in for the reference's native extension module ``MultiScaleDeformableAttention``
(LuisCossioUOH/Kinet, src/trackformer/models/ops/setup.py:47,57; bound in src/vision.cpp:4-7 and
imported as ``MSDA`` by ops/functions/ms_deform_attn_func.py:11).

With this file on ``sys.path`` the reference's own ``functions/ms_deform_attn_func.py`` and
``modules/ms_deform_attn.py`` run UNMODIFIED on the sm_100a kernels of libmsda_b200.so:

    MSDA.ms_deform_attn_forward(value, spatial_shapes, sampling_loc, attn_weight, im2col_step)
        -> output (N, Lq, M*D)
    MSDA.ms_deform_attn_backward(value, spatial_shapes, sampling_loc, attn_weight, grad_output,
                                 im2col_step) -> [grad_value, grad_sampling_loc, grad_attn_weight]

Positional arguments only, like the pybind functions.  The upstream Deformable-DETR arity
(``level_start_index`` after ``spatial_shapes``: 6 / 7 arguments) is accepted too.
There is no CPU implementation (the reference raises "Not implemented on the CPU" as well,
ops/src/ms_deform_attn.h:27,48).
"""
from kinet_b200.functions import ms_deform_attn_backward as _bwd
from kinet_b200.functions import ms_deform_attn_forward as _fwd

__all__ = ["ms_deform_attn_forward", "ms_deform_attn_backward"]


def ms_deform_attn_forward(*args):
    if len(args) == 5:
        value, shapes, loc, attn, step = args
        return _fwd(value, shapes, loc, attn, step)
    if len(args) == 6:
        value, shapes, lsi, loc, attn, step = args
        return _fwd(value, shapes, loc, attn, step, lsi)
    raise TypeError("ms_deform_attn_forward(): incompatible function arguments (expected 5 or 6, got %d)" % len(args))


def ms_deform_attn_backward(*args):
    if len(args) == 6:
        value, shapes, loc, attn, grad_output, step = args
        return _bwd(value, shapes, loc, attn, grad_output, step)
    if len(args) == 7:
        value, shapes, lsi, loc, attn, grad_output, step = args
        return _bwd(value, shapes, loc, attn, grad_output, step, lsi)
    raise TypeError("ms_deform_attn_backward(): incompatible function arguments (expected 6 or 7, got %d)" % len(args))
