"""Layouts + storage tensors + linear dispatch of the weight-only path.

* TensorCoreTiledLayout / TensorCoreTiledAQTTensorImpl — int4, drop-in for
  torchao/dtypes/uintx/tensor_core_tiled_layout.py, but the packed weight lives in the B200
  strip-major layout (include/ao_b200.h) instead of the mma.sync m16n8k16 tinygemm layout, and the
  linear calls torch.ops.torchao.b200_int4_weight_only_mm instead of aten._weight_int4pack_mm.
* PlainAQTTensorImpl — int8, drop-in for torchao/dtypes/uintx/plain_layout.py (weight-only half).
"""
from dataclasses import dataclass
from typing import Optional, Tuple

import torch
from torch.utils._python_dispatch import is_traceable_wrapper_subclass, return_and_correct_aliasing

from .. import ops as _ops
from ..quantization.quant_primitives import CUDA_DIV_SEMANTICS, ZeroPointDomain
from ..quantization import quant_primitives as _qp
from ..quantization.utils import pack_tinygemm_scales_and_zeros, unpack_tinygemm_scales_and_zeros
from ..utils import fill_defaults, find_multiple
from .affine_quantized_tensor import (
    AffineQuantizedTensor,
    register_aqt_quantized_linear_dispatch,
    register_layout,
)
from .utils import AQTTensorImpl, Layout, PlainLayout

aten = torch.ops.aten

__all__ = ["TensorCoreTiledLayout", "TensorCoreTiledAQTTensorImpl", "PlainLayout", "PlainAQTTensorImpl"]


@dataclass(frozen=True)
class TensorCoreTiledLayout(Layout):
    """reference: tensor_core_tiled_layout.py:117-188.  `inner_k_tiles` is accepted for API
    compatibility (Int4WeightOnlyConfig's default is TensorCoreTiledLayout(inner_k_tiles=8)); the
    strip-major layout has no such knob.  Padding: N -> x128, K -> x max(G, 64) (reference: x8 / x1024)."""

    inner_k_tiles: int = 8

    def pre_process(self, input: torch.Tensor) -> torch.Tensor:
        n, k = input.shape[-2:]
        return torch.nn.functional.pad(input, (0, find_multiple(k, 64) - k, 0, find_multiple(n, 128) - n))

    def post_process(self, input, scale, zero_point, block_size):
        n, k = input.shape[-2:]
        np_, kp = find_multiple(n, 128), find_multiple(k, max(64, block_size[-1]))
        input = torch.nn.functional.pad(input, (0, kp - k, 0, np_ - n))
        pad_g = (kp - k) // block_size[-1]
        scale = torch.nn.functional.pad(scale, (0, pad_g, 0, np_ - n))
        zero_point = torch.nn.functional.pad(zero_point, (0, pad_g, 0, np_ - n))
        return input, scale, zero_point

    def extra_repr(self):
        return f"inner_k_tiles={self.inner_k_tiles}"


def _wrapper(cls, like: torch.Tensor, shape):
    return torch.Tensor._make_wrapper_subclass(cls, shape, device=like.device, dtype=like.dtype,
                                               layout=like.layout, requires_grad=False)


@register_layout(TensorCoreTiledLayout)
class TensorCoreTiledAQTTensorImpl(AQTTensorImpl):
    """int4 storage, both strip-major (include/ao_b200.h): packed_weight int32 (Np/16, Kp/64, 32, 4) and
    scale_and_zero bf16 (Np/16, Kp/G, 16, 2) — the numbers of the reference's (Kp/G, Np, 2) tensor with the 16
    rows of a strip made contiguous (ops.sz_b200_to_tinygemm converts).  reference: tensor_core_tiled_layout.py:191-520."""

    def __new__(cls, packed_weight, scale_and_zero, transposed, _layout):
        return _wrapper(cls, packed_weight, packed_weight.shape)

    def __init__(self, packed_weight, scale_and_zero, transposed, _layout):
        self.packed_weight = packed_weight
        self.scale_and_zero = scale_and_zero
        self.transposed = transposed
        self._layout = _layout

    def __tensor_flatten__(self):
        return ["packed_weight", "scale_and_zero"], [self.transposed, self._layout]

    @classmethod
    def __tensor_unflatten__(cls, tensor_data_dict, tensor_attributes, outer_size, outer_stride):
        transposed, _layout = tensor_attributes
        return cls(tensor_data_dict["packed_weight"], tensor_data_dict["scale_and_zero"], transposed, _layout)

    # -- construction -------------------------------------------------------------------
    @classmethod
    def from_hp(cls, weight: torch.Tensor, group_size: int, eps: float, _layout: Layout):
        """bf16 (N, K) -> storage in ONE kernel (pad + qparams + quantize + pack)."""
        packed, sz = torch.ops.torchao.b200_int4_quantize_pack(weight, group_size, eps, _qp.CUDA_DIV_SEMANTICS)
        return cls(packed, sz, False, _layout)

    @classmethod
    def from_plain(cls, int_data, scale, zero_point, _layout):
        """reference: tensor_core_tiled_layout.py:262-307.  int_data int32 (N, K) in 0..15,
        scale / zero_point (N, K/G) bf16."""
        assert isinstance(_layout, TensorCoreTiledLayout)
        if int_data.dim() == 3:  # for moe quant (reference :283-294): one packed weight per expert, stacked
            e_n = int_data.shape[0]
            scale = scale.reshape(e_n, int_data.shape[-2], -1)
            zero_point = zero_point.reshape(e_n, int_data.shape[-2], -1)
            per = [cls.from_plain(int_data[e], scale[e], zero_point[e], _layout) for e in range(e_n)]
            return cls(torch.stack([p.packed_weight for p in per]), torch.stack([p.scale_and_zero for p in per]), False, _layout)
        assert int_data.dtype == torch.int32 and int_data.dim() == 2, "int_data must be 2-D int32"
        n, k = int_data.shape
        scale = scale.reshape(n, -1)
        zero_point = zero_point.reshape(n, -1)
        group_size = k // scale.shape[1]
        int_data, scale, zero_point = _layout.post_process(int_data, scale, zero_point, (1, group_size))
        packed = torch.ops.torchao.b200_int4_pack(int_data)
        sz = _ops.pack_scales_and_zeros_b200(scale, zero_point)
        return cls(packed, sz, False, _layout)

    @classmethod
    def from_tinygemm(cls, packed_weight, scale_and_zero, inner_k_tiles, _layout=None):
        """Adopt a weight packed by the reference (aten._convert_weight_to_int4pack) — checkpoint interchange."""
        q = _ops.unpack_tensor_core_tiled_layout(packed_weight, inner_k_tiles)
        s, z = unpack_tinygemm_scales_and_zeros(scale_and_zero)
        n = q.shape[0]
        return cls.from_plain(q, s.reshape(n, -1).contiguous(), z.reshape(n, -1).contiguous(),
                              _layout or TensorCoreTiledLayout(inner_k_tiles))

    # -- views ----------------------------------------------------------------------------
    @property
    def padded_shape(self) -> Tuple[int, int]:
        return self.packed_weight.shape[-4] * 16, self.packed_weight.shape[-3] * 64

    @property
    def group_size(self) -> int:
        return self.padded_shape[1] // self.scale_and_zero.shape[-3]

    @property
    def block_size(self):
        return (1, self.group_size)

    def get_plain(self):
        """-> int_data int32 (Np, Kp), scale (Np, Kp/G), zero (Np, Kp/G).  The reference recovers
        int_data with an identity matmul + re-quantize (tensor_core_tiled_layout.py:465-517); here it
        is a plain unpack kernel (exact by construction)."""
        if self.packed_weight.dim() == 5:  # expert stack
            per = [self.__class__(self.packed_weight[e], self.scale_and_zero[e], self.transposed, self._layout).get_plain()
                   for e in range(self.packed_weight.shape[0])]
            return tuple(torch.stack([p[i] for p in per]) for i in range(3))
        int_data = torch.ops.torchao.b200_int4_unpack(self.packed_weight)
        s, z = _ops.unpack_scales_and_zeros_b200(self.scale_and_zero)
        return int_data, s, z

    def get_layout(self):
        return self._layout

    def _apply_fn_to_data(self, fn):
        return self.__class__(fn(self.packed_weight), fn(self.scale_and_zero), self.transposed, self._layout)

    def to(self, *args, **kwargs):
        device, _, _, _ = torch._C._nn._parse_to(*args, **kwargs)
        device = self.device if device is None else device
        return self.__class__(self.packed_weight.to(device), self.scale_and_zero.to(device), self.transposed,
                              self._layout)

    def t(self):
        return self.__class__(self.packed_weight, self.scale_and_zero, not self.transposed, self._layout)

    def copy_(self, src):
        if not (isinstance(src, TensorCoreTiledAQTTensorImpl) and self.packed_weight.shape == src.packed_weight.shape
                and self.scale_and_zero.shape == src.scale_and_zero.shape and self.transposed == src.transposed):
            raise ValueError(f"Not supported args for copy_ due to metadata mistach: {self, src}")
        self.packed_weight.copy_(src.packed_weight)
        self.scale_and_zero.copy_(src.scale_and_zero)
        return self

    @classmethod
    def __torch_dispatch__(cls, func, types, args, kwargs):
        kwargs = kwargs or {}
        self = args[0]
        if func is aten.detach.default:
            return return_and_correct_aliasing(func, args, kwargs, self._apply_fn_to_data(torch.detach))
        if func is aten.clone.default:
            return return_and_correct_aliasing(func, args, kwargs, self._apply_fn_to_data(torch.clone))
        if func is aten.copy_.default:
            return self.copy_(args[1])
        if func is aten.t.default:
            return return_and_correct_aliasing(func, args, kwargs, self.t())
        if func in (aten.select.int, aten.index.Tensor):
            # reference :363-372: experts of a 3-D weight.  Both storage tensors carry the expert dimension in front.
            assert not (func is aten.select.int and args[1] != 0), "aten.select.int currently only has support for dim=0"
            assert self.packed_weight.dim() == 5, f"{func}: only on a stack of expert weights"
            return return_and_correct_aliasing(
                func, args, kwargs, self._apply_fn_to_data(lambda x: func(x, *args[1:], **kwargs).contiguous()))
        if func is aten.slice.Tensor:
            assert self.packed_weight.dim() == 4, "slice: 2-D weights only"
            # logical (N, K) ranges -> 16-row strips / 64-wide k-tiles / groups (reference: :386-440 does the same on
            # its tiles).  Both tensors are strip-major: dim 0 = strip, dim 1 = k-tile (weight) or group (qparams).
            _, dim, start, end, step = fill_defaults(args, 5, [0, None, None, 1])
            assert step == 1 and dim in (0, 1), f"slice: dim={dim}, step={step} not supported"
            np_, kp = self.padded_shape
            full = np_ if dim == 0 else kp
            start = 0 if start is None else start
            end = full if end is None or end > full else end
            pw_unit = 16 if dim == 0 else 64
            sz_unit = 16 if dim == 0 else self.group_size
            if start % pw_unit or start % sz_unit:
                raise NotImplementedError(f"int4 slice start must be a multiple of {max(pw_unit, sz_unit)} on dim {dim}")
            if dim == 1 and (-(-end // 64) * 64) % self.group_size:
                raise NotImplementedError(f"int4 slice end on dim 1 must cover whole groups of {self.group_size}")
            pw = aten.slice.Tensor(self.packed_weight, dim, start // pw_unit, -(-end // pw_unit), 1)
            sz_end = -(-end // 16) if dim == 0 else (-(-end // 64) * 64) // self.group_size
            sz = aten.slice.Tensor(self.scale_and_zero, dim, start // sz_unit, sz_end, 1)
            new = self.__class__(pw.contiguous(), sz.contiguous(), self.transposed, self._layout)
            return return_and_correct_aliasing(func, args, kwargs, new)
        raise NotImplementedError(f"TensorCoreTiledAQTTensorImpl dispatch: attempting to run {func}, this is not supported")


# ---------------------------------------------------------------------------------------- int4 linear
def _aqt_is_tensor_core_tile_uint4(aqt):
    return aqt.tensor_impl.dtype == torch.int32 and aqt.quant_min == 0 and aqt.quant_max == 15


def _linear_bf16_act_uint4_weight_check(input_tensor, weight_tensor, bias):
    """reference: tensor_core_tiled_layout.py:58-71"""
    return (
        not is_traceable_wrapper_subclass(input_tensor)
        and input_tensor.dtype == torch.bfloat16
        and isinstance(weight_tensor, AffineQuantizedTensor)
        and _aqt_is_tensor_core_tile_uint4(weight_tensor)
        and weight_tensor.dtype == torch.bfloat16
        and len(weight_tensor.shape) == 2
        and weight_tensor.zero_point_domain == ZeroPointDomain.FLOAT
        and isinstance(weight_tensor._layout, TensorCoreTiledLayout)
    )


def _linear_bf16_act_uint4_weight_impl(input_tensor, weight_tensor, bias):
    """reference: tensor_core_tiled_layout.py:74-114.  No activation padding (the kernel treats k beyond
    the activation's width as zero) and no output slice (the kernel writes only the first N columns)."""
    assert weight_tensor.block_size[0] == 1, f"Requires groupwise quantization, got block_size: {weight_tensor.block_size}"
    assert input_tensor.shape[-1] == weight_tensor.shape[1], (
        f"need input_tensor shape: {input_tensor.shape} final dim to match weight_tensor shape: "
        f"{weight_tensor.shape} second dim")
    impl = weight_tensor.tensor_impl
    orig_act_size = input_tensor.size()
    act_mat = input_tensor.reshape(-1, input_tensor.shape[-1])
    out_features = weight_tensor.shape[-2]
    y = torch.ops.torchao.b200_int4_weight_only_mm(act_mat, impl.packed_weight, weight_tensor.block_size[-1],
                                                   impl.scale_and_zero, out_features)
    y = y.reshape(*orig_act_size[:-1], out_features)
    if bias is not None:
        y += bias
    return y


register_aqt_quantized_linear_dispatch(_linear_bf16_act_uint4_weight_check, _linear_bf16_act_uint4_weight_impl)


# ---------------------------------------------------------------------------------------- int8
@register_layout(PlainLayout)
class PlainAQTTensorImpl(AQTTensorImpl):
    """int8 storage.  The reference keeps int_data int8 (N, K) row-major (plain_layout.py:47-229); here the bytes live in
    the strip-major biased layout (include/ao_b200.h: `packed` uint8 (Np/16, Kp/64, 1024)) that the TMA-ring GEMV and
    the tcgen05 GEMM stream, and `int_data` is recovered exactly (unpack kernel) whenever it is asked for.
    scale bf16 (N,), zero_point int64 (N,) == 0."""

    def __new__(cls, packed, scale, zero_point, _layout, shape, transposed=False):
        n, k = shape
        return torch.Tensor._make_wrapper_subclass(cls, (k, n) if transposed else (n, k), device=packed.device,
                                                   dtype=torch.int8, layout=packed.layout, requires_grad=False)

    def __init__(self, packed, scale, zero_point, _layout, shape, transposed=False):
        self.packed = packed
        self.scale = scale
        self.zero_point = zero_point
        self._layout = _layout
        self._nk = (int(shape[0]), int(shape[1]))
        self.transposed = transposed

    def __tensor_flatten__(self):
        names = ["packed", "scale"] + (["zero_point"] if self.zero_point is not None else [])
        return names, [self._layout, self._nk, self.transposed]

    @classmethod
    def __tensor_unflatten__(cls, tensor_data_dict, tensor_attributes, outer_size, outer_stride):
        _layout, nk, transposed = tensor_attributes
        return cls(tensor_data_dict["packed"], tensor_data_dict["scale"], tensor_data_dict.get("zero_point"),
                   _layout, nk, transposed)

    @classmethod
    def from_hp(cls, weight, eps, zero_point_dtype, _layout):
        """bf16 (N, K) -> int8 data + scale in ONE kernel, then the strip-major pack."""
        q, scale = torch.ops.torchao.b200_int8_quantize(weight, eps, _qp.CUDA_DIV_SEMANTICS)
        zp = torch.zeros(scale.shape, dtype=zero_point_dtype, device=weight.device)
        return cls(torch.ops.torchao.b200_int8_pack(q), scale, zp, _layout, q.shape)

    @classmethod
    def from_plain(cls, int_data, scale, zero_point, _layout):
        assert isinstance(_layout, PlainLayout)
        assert int_data.dtype == torch.int8 and int_data.dim() == 2, "int_data must be 2-D int8"
        return cls(torch.ops.torchao.b200_int8_pack(int_data), scale, zero_point, _layout, int_data.shape)

    @property
    def int_data(self):
        """the reference's field: int8 (N, K) row-major (materialised by the unpack kernel, bit-exact)."""
        return torch.ops.torchao.b200_int8_unpack(self.packed, self._nk[0], self._nk[1])

    def get_plain(self):
        return self.int_data, self.scale, self.zero_point

    def get_layout(self):
        return self._layout

    def _apply_fn_to_data(self, fn):
        return self.__class__(fn(self.packed), fn(self.scale), fn(self.zero_point) if self.zero_point is not None else None,
                              self._layout, self._nk, self.transposed)

    def to(self, *args, **kwargs):
        device, _, _, _ = torch._C._nn._parse_to(*args, **kwargs)
        device = self.device if device is None else device
        return self._apply_fn_to_data(lambda t: t.to(device))

    def t(self):
        return self.__class__(self.packed, self.scale, self.zero_point, self._layout, self._nk, not self.transposed)

    def copy_(self, src):
        if not (isinstance(src, PlainAQTTensorImpl) and self._nk == src._nk and self.transposed == src.transposed):
            raise ValueError(f"Not supported args for copy_ due to metadata mistach: {self, src}")
        self.packed.copy_(src.packed)
        self.scale.copy_(src.scale)
        if self.zero_point is not None and src.zero_point is not None:
            self.zero_point.copy_(src.zero_point)
        return self

    @classmethod
    def __torch_dispatch__(cls, func, types, args, kwargs):
        kwargs = kwargs or {}
        self = args[0]
        if func is aten.detach.default:
            return return_and_correct_aliasing(func, args, kwargs, self._apply_fn_to_data(torch.detach))
        if func is aten.clone.default:
            return return_and_correct_aliasing(func, args, kwargs, self._apply_fn_to_data(torch.clone))
        if func is aten.copy_.default:
            return self.copy_(args[1])
        if func is aten.t.default:
            return return_and_correct_aliasing(func, args, kwargs, self.t())
        if func is aten.slice.Tensor:
            # reference: plain_layout.py:160-200 — rows slice data + qparams, columns slice data only.  Done on the plain
            # view (unpack -> slice -> pack): slicing happens at load / shard time, not on the hot path.
            _, dim, start, end, step = fill_defaults(args, 5, [0, None, None, 1])
            assert step == 1 and dim in (0, 1) and not self.transposed
            data = aten.slice.Tensor(self.int_data, dim, start, end, 1).contiguous()
            if dim == 0:
                scale = aten.slice.Tensor(self.scale, 0, start, end, 1)
                zp = aten.slice.Tensor(self.zero_point, 0, start, end, 1) if self.zero_point is not None else None
            else:
                scale, zp = self.scale, self.zero_point
            new = self.__class__(torch.ops.torchao.b200_int8_pack(data), scale, zp, self._layout, data.shape)
            return return_and_correct_aliasing(func, args, kwargs, new)
        raise NotImplementedError(f"PlainAQTTensorImpl dispatch: attempting to run {func}, this is not supported")


def _aqt_is_int8(aqt):
    return (aqt.tensor_impl.dtype == torch.int8 and (aqt.quant_min is None or aqt.quant_min == -128)
            and (aqt.quant_max is None or aqt.quant_max == 127))


def _linear_fp_act_int8_weight_check(input_tensor, weight_tensor, bias):
    """reference: plain_layout.py:232-247"""
    return (
        not is_traceable_wrapper_subclass(input_tensor)
        and input_tensor.is_floating_point()
        and isinstance(weight_tensor, AffineQuantizedTensor)
        and _aqt_is_int8(weight_tensor)
        and len(weight_tensor.shape) == 2
        and len(weight_tensor.block_size) == 2
        and weight_tensor.block_size[0] == 1
        and weight_tensor.block_size[1] == weight_tensor.shape[1]
        and weight_tensor.zero_point_domain == ZeroPointDomain.INT
        and isinstance(weight_tensor._layout, PlainLayout)
    )


def _linear_fp_act_int8_weight_impl(input_tensor, weight_tensor, bias):
    """reference: plain_layout.py:250-266:  y = mm(x, W_int8^T.to(x.dtype)) * scale (+ bias)."""
    if input_tensor.dtype != torch.bfloat16:
        raise NotImplementedError("B200 int8 weight-only linear takes bfloat16 activations")
    impl = weight_tensor.tensor_impl
    x2d = input_tensor.reshape(-1, input_tensor.shape[-1])
    y = torch.ops.torchao.b200_int8_weight_only_mm(x2d, impl.packed, impl.scale)
    y = y.reshape(*input_tensor.shape[:-1], y.shape[-1])
    if bias is not None:
        y += bias.to(y.dtype)
    return y


register_aqt_quantized_linear_dispatch(_linear_fp_act_int8_weight_check, _linear_fp_act_int8_weight_impl)
