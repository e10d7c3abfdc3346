from .utils import AQTTensorImpl, Layout, PlainLayout  # noqa: F401
from .affine_quantized_tensor import (  # noqa: F401
    AffineQuantizedTensor, QuantizedLinearNotImplementedError, deregister_aqt_quantized_linear_dispatch,
    get_tensor_impl_constructor, register_aqt_quantized_linear_dispatch, register_layout, to_affine_quantized_intx,
)
from .layouts import PlainAQTTensorImpl, TensorCoreTiledAQTTensorImpl, TensorCoreTiledLayout  # noqa: F401
