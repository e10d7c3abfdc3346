"""AffineQuantizedTensor: the weight subclass `quantize_` installs on nn.Linear.

Drop-in for torchao/dtypes/affine_quantized_tensor.py:57-614 and
affine_quantized_tensor_ops.py:113-598 restricted to the weight-only int4 / int8 path:
same constructor fields, `from_hp_to_intx`, `dequantize`, layout registry
(`register_layout` / `get_tensor_impl_constructor`), linear dispatch table
(`register_aqt_quantized_linear_dispatch`), and the aten ops a weight needs
(detach / clone / _to_copy / copy_ / t / slice / view / select / index / linear / mm / addmm).
"""
import logging
from typing import Callable, Dict, Optional, Tuple

import torch
from torch.utils._python_dispatch import return_and_correct_aliasing

from ..quantization.quant_primitives import (
    MappingType,
    ZeroPointDomain,
    _dequantize_affine_tinygemm,
    _get_and_check_qmin_qmax,
    dequantize_affine,
)
from ..utils import fill_defaults
from .utils import AQTTensorImpl, Layout, PlainLayout

logger = logging.getLogger(__name__)
aten = torch.ops.aten

__all__ = [
    "AffineQuantizedTensor", "to_affine_quantized_intx", "register_layout", "get_tensor_impl_constructor",
    "register_aqt_quantized_linear_dispatch", "deregister_aqt_quantized_linear_dispatch",
    "QuantizedLinearNotImplementedError",
]


class QuantizedLinearNotImplementedError(NotImplementedError):
    """reference: affine_quantized_tensor_ops.py:163-166"""


_LAYOUT_CONSTRUCTOR_TABLE: Dict[type, Callable] = {}
_AQT_QLINEAR_DISPATCH_TABLE: Dict[Callable, Callable] = {}


def register_layout(layout_class):
    """Class decorator: `@register_layout(MyLayout) class MyImpl(AQTTensorImpl)` makes
    `MyImpl.from_plain` the constructor for `MyLayout` (reference: torchao/utils.py:618-644)."""

    def decorator(tensor_impl_class):
        _LAYOUT_CONSTRUCTOR_TABLE[layout_class] = tensor_impl_class.from_plain
        torch.serialization.add_safe_globals([layout_class, tensor_impl_class])
        return tensor_impl_class

    return decorator


def get_tensor_impl_constructor(layout_class):
    """reference: torchao/utils.py:646-664"""
    if layout_class not in _LAYOUT_CONSTRUCTOR_TABLE:
        raise ValueError(f"layout_name: {layout_class} is not supported yet for {AffineQuantizedTensor}")
    return _LAYOUT_CONSTRUCTOR_TABLE[layout_class]


def register_aqt_quantized_linear_dispatch(dispatch_condition, impl):
    """(check, impl) pair, both `(input[..., K], weight[N, K], bias[N] | None)`; first match wins.
    reference: affine_quantized_tensor_ops.py:116-130"""
    _AQT_QLINEAR_DISPATCH_TABLE[dispatch_condition] = impl


def deregister_aqt_quantized_linear_dispatch(dispatch_condition):
    if dispatch_condition in _AQT_QLINEAR_DISPATCH_TABLE:
        del _AQT_QLINEAR_DISPATCH_TABLE[dispatch_condition]
    else:
        logger.warning(f"Attempting to remove non-existent dispatch condition {dispatch_condition}")


class AffineQuantizedTensor(torch.Tensor):
    """Quantized weight: outer shape/dtype of the original tensor, data in `tensor_impl`.

    fields (reference: affine_quantized_tensor.py:57-132): tensor_impl, block_size, shape,
    quant_min, quant_max, zero_point_domain, dtype."""

    @staticmethod
    def __new__(cls, tensor_impl, block_size, shape, quant_min=None, quant_max=None,
                zero_point_domain=ZeroPointDomain.INT, dtype=None, strides=None):
        kwargs = {"device": tensor_impl.device, "layout": tensor_impl.layout, "requires_grad": False}
        if dtype is None:
            dtype = tensor_impl.dtype
        kwargs["dtype"] = dtype
        if strides is not None:
            kwargs["strides"] = strides
        return torch.Tensor._make_wrapper_subclass(cls, shape, **kwargs)

    def __init__(self, tensor_impl, block_size, shape, quant_min=None, quant_max=None,
                 zero_point_domain=ZeroPointDomain.INT, dtype=None, strides=None):
        self.tensor_impl = tensor_impl
        self.block_size = tuple(block_size)
        self.quant_min = quant_min
        self.quant_max = quant_max
        self.zero_point_domain = zero_point_domain

    # ------------------------------------------------------------------ construction
    @classmethod
    def from_hp_to_intx(cls, input_float, mapping_type, block_size, target_dtype, quant_min=None, quant_max=None,
                        eps=None, scale_dtype=None, zero_point_dtype=None, preserve_zero=True,
                        zero_point_domain=ZeroPointDomain.INT, _layout=PlainLayout(), use_hqq=False, *,
                        custom_scale=None, custom_zero_point=None):
        """reference: affine_quantized_tensor.py:231-373.  The reference runs pre_process ->
        choose_qparams -> quantize -> post_process -> from_plain as ~15 eager kernels; the two
        supported recipes are each ONE fused B200 kernel here (results bit-identical)."""
        from .layouts import PlainAQTTensorImpl, TensorCoreTiledAQTTensorImpl, TensorCoreTiledLayout

        if zero_point_domain is None:
            raise ValueError("please use ZeroPointDomain.NONE instead of None")
        if custom_scale is not None or custom_zero_point is not None:
            raise NotImplementedError("custom scale/zero_point are outside the B200 hot path")
        if not input_float.is_cuda:
            raise NotImplementedError("the B200 path quantizes CUDA tensors only (no CPU fallback)")
        if input_float.dim() == 3:
            # MoE expert weights [E, N, K] (reference: tensor_core_tiled_layout.py:283-294, block_size (1, 1, G)):
            # every expert is quantized by the 2-D path, the storage tensors get a leading expert dimension
            if not (isinstance(_layout, TensorCoreTiledLayout) and len(block_size) == 3 and block_size[0] == 1):
                raise NotImplementedError("3-D weights: TensorCoreTiledLayout with block_size (1, 1, G)")
            per_expert = [cls.from_hp_to_intx(input_float[e], mapping_type, tuple(block_size[1:]), target_dtype, quant_min,
                                              quant_max, eps, scale_dtype, zero_point_dtype, preserve_zero,
                                              zero_point_domain, _layout, use_hqq) for e in range(input_float.shape[0])]
            impl0 = per_expert[0].tensor_impl
            tensor_impl = impl0.__class__(torch.stack([w.tensor_impl.packed_weight for w in per_expert]),
                                          torch.stack([w.tensor_impl.scale_and_zero for w in per_expert]), False, _layout)
            return cls(tensor_impl, block_size, input_float.shape, per_expert[0].quant_min, per_expert[0].quant_max,
                       zero_point_domain, dtype=input_float.dtype)
        if input_float.dim() != 2:
            raise NotImplementedError("the B200 path quantizes 2-D linear weights (and 3-D MoE expert stacks)")
        original_shape = input_float.shape
        input_float = input_float.detach()
        qmin, qmax = _get_and_check_qmin_qmax(target_dtype, quant_min, quant_max)

        if isinstance(_layout, TensorCoreTiledLayout):
            tinygemm = (mapping_type is MappingType.ASYMMETRIC and (qmin, qmax) == (0, 15)
                        and zero_point_domain is ZeroPointDomain.FLOAT and not preserve_zero
                        and target_dtype == torch.int32 and len(block_size) == 2 and block_size[0] == 1)
            if not tinygemm:
                raise NotImplementedError(
                    "TensorCoreTiledLayout on B200 implements the tinygemm recipe: ASYMMETRIC uint4 in int32, "
                    "FLOAT zero-point domain, preserve_zero=False, block_size (1, G)")
            if zero_point_dtype not in (None, torch.bfloat16) or scale_dtype not in (None, torch.bfloat16):
                raise NotImplementedError("tinygemm qparams are bfloat16")
            if use_hqq:
                # reference: affine_quantized_tensor.py:252-288 -> _choose_qparams_and_quantize_affine_hqq with
                # nbits = log2(quant_max + 1), axis = 1, group_size = max(block_size), raw_output = False
                q, scale, zero = torch.ops.torchao.b200_int4_hqq_quantize(
                    input_float if input_float.dtype is torch.bfloat16 else input_float.to(torch.bfloat16), int(block_size[1]))
                tensor_impl = TensorCoreTiledAQTTensorImpl.from_plain(q, scale, zero, _layout)
                return cls(tensor_impl, block_size, original_shape, quant_min, quant_max, zero_point_domain,
                           dtype=input_float.dtype)
            tensor_impl = TensorCoreTiledAQTTensorImpl.from_hp(
                input_float, int(block_size[1]), 1e-6 if eps is None else float(eps), _layout)
            return cls(tensor_impl, block_size, original_shape, quant_min, quant_max, zero_point_domain,
                       dtype=input_float.dtype)

        if use_hqq:
            raise NotImplementedError("use_hqq=True on B200 is implemented for TensorCoreTiledLayout (Int4WeightOnlyConfig)")
        if isinstance(_layout, PlainLayout):
            int8 = (mapping_type is MappingType.SYMMETRIC and target_dtype == torch.int8 and (qmin, qmax) == (-128, 127)
                    and zero_point_domain is ZeroPointDomain.INT and len(block_size) == 2 and block_size[0] == 1)
            if not int8:
                raise NotImplementedError("PlainLayout on B200 implements SYMMETRIC int8 with block_size (1, G)")
            if block_size[1] != original_shape[1]:
                raise NotImplementedError("B200 int8 weight-only is per-channel (group_size=None)")
            tensor_impl = PlainAQTTensorImpl.from_hp(
                input_float, torch.finfo(input_float.dtype).eps if eps is None else float(eps),
                zero_point_dtype or torch.int64, _layout)
            return cls(tensor_impl, block_size, original_shape, quant_min, quant_max, zero_point_domain,
                       dtype=input_float.dtype)

        raise NotImplementedError(f"layout {_layout} is outside the B200 hot path (SURVEY.md §8)")

    # ------------------------------------------------------------------ views of the data
    @property
    def _layout(self) -> Layout:
        return self.tensor_impl._layout

    def dequantize(self, output_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
        """reference: affine_quantized_tensor.py:134-191"""
        if output_dtype is None:
            output_dtype = self.dtype
        if self.dim() == 3:
            return torch.stack([aten.select.int(self, 0, e).dequantize(output_dtype) for e in range(self.shape[0])])
        data, scale, zero_point = self.tensor_impl.get_plain()
        n, k = self.shape[-2], self.shape[-1]
        if self.zero_point_domain is ZeroPointDomain.FLOAT:
            dq = _dequantize_affine_tinygemm(data, self.block_size_for(data), scale, zero_point, data.dtype,
                                             self.quant_min, self.quant_max, output_dtype=output_dtype)
        else:
            dq = dequantize_affine(data, self.block_size_for(data), scale, zero_point, data.dtype, self.quant_min,
                                   self.quant_max, output_dtype=output_dtype)
        dq = dq[..., :n, :k] if not getattr(self.tensor_impl, "transposed", False) else dq[..., :k, :n].t()
        return dq

    def block_size_for(self, data):
        bs = self.block_size
        if getattr(self.tensor_impl, "transposed", False):
            bs = bs[::-1]
        return tuple(min(b, s) if b > 1 else 1 for b, s in zip(bs, data.shape)) if len(bs) == data.dim() else bs

    def _quantization_type(self):
        return (f"shape={self.shape}, block_size={self.block_size}, device={self.device}, _layout={self._layout}, "
                f"tensor_impl_dtype={self.tensor_impl.dtype}, quant_min={self.quant_min}, quant_max={self.quant_max}")

    def __repr__(self):
        return f"{self.__class__.__name__}({self._quantization_type()})"

    # ------------------------------------------------------------------ subclass plumbing
    def __tensor_flatten__(self):
        return ["tensor_impl"], [self.block_size, self.shape, self.quant_min, self.quant_max, self.zero_point_domain,
                                 self.dtype]

    @classmethod
    def __tensor_unflatten__(cls, tensor_data_dict, tensor_attributes, outer_size, outer_stride):
        block_size, shape, quant_min, quant_max, zero_point_domain, dtype = tensor_attributes
        return cls(tensor_data_dict["tensor_impl"], block_size, shape if outer_size is None else outer_size,
                   quant_min, quant_max, zero_point_domain, dtype=dtype, strides=outer_stride)

    def _apply_fn_to_data(self, fn):
        return self.__class__(fn(self.tensor_impl), self.block_size, self.shape, self.quant_min, self.quant_max,
                              self.zero_point_domain, dtype=self.dtype, strides=self.stride())

    def to(self, *args, **kwargs):
        device, dtype, _, _ = torch._C._nn._parse_to(*args, **kwargs)
        device = self.device if device is None else device
        if dtype is not None and dtype != self.dtype:
            raise NotImplementedError("changing the dtype of a quantized weight is not supported")
        return self.__class__(self.tensor_impl.to(device), self.block_size, self.shape, self.quant_min,
                              self.quant_max, self.zero_point_domain, dtype=self.dtype, strides=self.stride())

    @staticmethod
    def _quantized_linear_op(input_tensor, weight_tensor, bias):
        for check, impl in _AQT_QLINEAR_DISPATCH_TABLE.items():
            if check(input_tensor, weight_tensor, bias):
                return impl(input_tensor, weight_tensor, bias)
        raise QuantizedLinearNotImplementedError("No specialized dispatch found for quantized linear op")

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        kwargs = kwargs or {}
        if func is torch.nn.functional.linear:
            return _linear(func, args, kwargs)
        with torch._C.DisableTorchFunctionSubclass():
            return func(*args, **kwargs)

    @classmethod
    def __torch_dispatch__(cls, func, types, args, kwargs):
        kwargs = kwargs or {}
        handler = _ATEN_TABLE.get(func)
        if handler is None:
            raise NotImplementedError(f"{cls.__name__} dispatch: attempting to run unimplemented operator/function: "
                                      f"{func}")
        return handler(func, args, kwargs)


to_affine_quantized_intx = AffineQuantizedTensor.from_hp_to_intx
torch.serialization.add_safe_globals([AffineQuantizedTensor])


# ---------------------------------------------------------------------- op implementations
def _fallback_ok(weight_tensor):
    impl = getattr(weight_tensor._layout, "quantized_linear_impl", None) if isinstance(
        weight_tensor, AffineQuantizedTensor) else None
    return impl is None


def _linear(func, args, kwargs):
    """reference: affine_quantized_tensor_ops.py:267-296"""
    input_tensor, weight_tensor = args[0], args[1]
    bias = args[2] if len(args) > 2 else kwargs.get("bias")
    if not input_tensor.is_floating_point():
        raise NotImplementedError(f"{func} is not implemented for non floating point input")
    try:
        return AffineQuantizedTensor._quantized_linear_op(input_tensor, weight_tensor, bias)
    except QuantizedLinearNotImplementedError:
        if not _fallback_ok(weight_tensor):
            raise
        if isinstance(input_tensor, AffineQuantizedTensor):
            input_tensor = input_tensor.dequantize()
        if isinstance(weight_tensor, AffineQuantizedTensor):
            weight_tensor = weight_tensor.dequantize()
        return torch.nn.functional.linear(input_tensor, weight_tensor, bias)


def _mm_like(func, args, kwargs):
    """aten.mm(x, w_t) / aten.addmm(bias, x, w_t): reference affine_quantized_tensor_ops.py:347-419"""
    if func is aten.addmm.default:
        bias, input_tensor, weight_tensor = args[0], args[1], args[2]
    else:
        input_tensor, weight_tensor, bias = args[0], args[1], None
    if not input_tensor.is_floating_point():
        raise NotImplementedError(f"{func} is not implemented for non floating point input")
    assert input_tensor.shape[-1] == weight_tensor.shape[0], (
        f"need mat1 shape: {input_tensor.shape} final dim to match mat2 shape: {weight_tensor.shape} first dim")
    try:
        return AffineQuantizedTensor._quantized_linear_op(input_tensor, weight_tensor.t(), bias)
    except QuantizedLinearNotImplementedError:
        if not _fallback_ok(weight_tensor):
            raise
        if isinstance(input_tensor, AffineQuantizedTensor):
            input_tensor = input_tensor.dequantize()
        if isinstance(weight_tensor, AffineQuantizedTensor):
            weight_tensor = weight_tensor.dequantize()
        return func(bias, input_tensor, weight_tensor) if bias is not None else func(input_tensor, weight_tensor)


def _same_metadata(a, b):
    return (isinstance(a, AffineQuantizedTensor) and isinstance(b, AffineQuantizedTensor)
            and all(getattr(a, f) == getattr(b, f) for f in
                    ("block_size", "shape", "quant_min", "quant_max", "zero_point_domain", "dtype"))
            and isinstance(a.tensor_impl, type(b.tensor_impl)))


def _copy_(func, args, kwargs):
    dst, src = args[0], args[1]
    if _same_metadata(dst, src):
        dst.tensor_impl.copy_(src.tensor_impl)
        return dst
    raise ValueError(f"Not supported args for copy_ due to metadata mistach: {dst, src}")


def _t(func, args, kwargs):
    w = args[0]
    assert len(w.block_size) == 2
    new = w.__class__(w.tensor_impl.t(), w.block_size[::-1], w.shape[::-1], w.quant_min, w.quant_max,
                      w.zero_point_domain, dtype=w.dtype, strides=w.stride())
    return return_and_correct_aliasing(func, args, kwargs, new)


def _slice(func, args, kwargs):
    """reference: affine_quantized_tensor_ops.py:467-498"""
    w, dim, start, end, step = fill_defaults(args, 5, [0, None, None, 1])
    assert step == 1
    assert dim in (0, 1), f"Only dim==0 or 1 are supported, got: {dim}"
    start = 0 if start is None else start
    end = w.shape[dim] if end is None or end >= w.shape[dim] else end
    shape = list(w.shape)
    shape[dim] = end - start
    block_size = (min(shape[0], w.block_size[0]), min(shape[1], w.block_size[1]))
    new = w.__class__(aten.slice.Tensor(w.tensor_impl, dim, start, end, step), block_size, shape, w.quant_min,
                      w.quant_max, w.zero_point_domain, dtype=w.dtype, strides=w.stride())
    return return_and_correct_aliasing(func, args, kwargs, new)


def _view(func, args, kwargs):
    w, shape = args
    if tuple(w.shape) == tuple(shape):
        return w.__class__(w.tensor_impl, w.block_size, w.shape, w.quant_min, w.quant_max, w.zero_point_domain,
                           dtype=w.dtype, strides=w.stride())
    raise ValueError(f"{w.__class__.__name__} only supports .view() with same shape")


def _index(func, args, kwargs):
    """reference: affine_quantized_tensor_ops.py:519-538 (a single index tensor, on dim 0)"""
    w, indices = args
    assert len(indices) == 1, f"op {func} currently only implemented for single dimensional indexing but got indices: {indices}"
    new_impl = aten.index.Tensor(w.tensor_impl, indices)
    shape = (indices[0].numel(), *w.shape[1:])
    new = w.__class__(new_impl, w.block_size, shape, w.quant_min, w.quant_max, w.zero_point_domain, dtype=w.dtype)
    return return_and_correct_aliasing(func, args, kwargs, new)


def _select(func, args, kwargs):
    """reference: affine_quantized_tensor_ops.py:541-562 (dim 0 of a 3-D weight: one expert)"""
    w, dim, index = fill_defaults(args, 3, [0, 0])
    assert dim == 0, f"op {func} currently only implemented for dim=0 but got dim={dim}"
    assert w.dim() == 3, f"op {func} currently only implemented for 3 dimensional tensors but got shape={w.shape}"
    new_impl = aten.select.int(w.tensor_impl, dim, index)
    new = w.__class__(new_impl, w.block_size[1:], w.shape[1:], w.quant_min, w.quant_max, w.zero_point_domain, dtype=w.dtype)
    return return_and_correct_aliasing(func, args, kwargs, new)


_ATEN_TABLE = {
    aten.linear.default: _linear,
    aten.mm.default: _mm_like,
    aten.addmm.default: _mm_like,
    aten.detach.default: lambda f, a, k: return_and_correct_aliasing(f, a, k, a[0]._apply_fn_to_data(torch.detach)),
    aten.alias.default: lambda f, a, k: return_and_correct_aliasing(f, a, k, a[0]._apply_fn_to_data(torch.detach)),
    aten.clone.default: lambda f, a, k: return_and_correct_aliasing(f, a, k, a[0]._apply_fn_to_data(torch.clone)),
    aten._to_copy.default: lambda f, a, k: return_and_correct_aliasing(
        f, a, k, a[0].to(*a[1:], **{n: v for n, v in k.items() if n in ("device", "dtype")})._apply_fn_to_data(torch.clone)),
    aten.copy_.default: _copy_,
    aten.t.default: _t,
    aten.slice.Tensor: _slice,
    aten.view.default: _view,
    aten.index.Tensor: _index,
    aten.select.int: _select,
}
