from .ms_deform_attn_func import (  # noqa: F401
    MSDeformAttnFunction,
    MSDeformAttnFusedFunction,
    fused_supported,
    ms_deform_attn_backward,
    ms_deform_attn_core_pytorch,
    ms_deform_attn_forward,
)
