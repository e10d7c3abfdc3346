from .quant_primitives import (  # noqa: F401
    MappingType, ZeroPointDomain, choose_qparams_affine, dequantize_affine, quantize_affine,
)
from .utils import compute_error  # noqa: F401
from .quant_api import (  # noqa: F401
    Float8DynamicActivationInt4WeightConfig, Int4WeightOnlyConfig, Int8WeightOnlyConfig, PackingFormat, int4_weight_only, int8_weight_only, quantize_,
)
from .transform_module import register_quantize_module_handler  # noqa: F401
from .int4_tensor import Int4Tensor  # noqa: F401
