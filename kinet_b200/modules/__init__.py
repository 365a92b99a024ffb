from .ms_deform_attn import MSDeformAttn  # noqa: F401
from .deformable_transformer import (  # noqa: F401
    DeformableTransformer,
    DeformableTransformerDecoder,
    DeformableTransformerDecoderLayer,
    DeformableTransformerEncoder,
    DeformableTransformerEncoderLayer,
)
