"""kinet_b200 -- B200-native (sm_100a) multi-scale deformable attention, a drop-in for
LuisCossioUOH/Kinet's src/trackformer/models/ops (MSDeformAttnFunction / MSDeformAttn)."""
from .functions import MSDeformAttnFunction, ms_deform_attn_core_pytorch  # noqa: F401
from .modules import MSDeformAttn  # noqa: F401

__all__ = ["MSDeformAttnFunction", "MSDeformAttn", "ms_deform_attn_core_pytorch"]
