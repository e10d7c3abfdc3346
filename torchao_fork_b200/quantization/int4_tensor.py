"""Int4Tensor — `Int4WeightOnlyConfig(version=2, packing_format="plain")`.

Mirror of torchao/quantization/quantize_/workflows/int4/int4_tensor.py:29-490: same tensor attributes
(`qdata` int8 (N, K/2) with two signed nibbles per byte, `scale` / `zero_point` (K/G, N) in the weight's dtype;
3-D weights carry a leading batch dimension), same non-tensor attributes (`block_size`, `shape`), same ops
(linear, bmm, slice with aliasing preserved, view, squeeze, detach / clone / to / copy_).

The reference quantizes and multiplies through a third-party dependency that is NOT in /root/reference:
fbgemm-gpu-genai >= 1.2.0 (`fbgemm_gpu.experimental.gen_ai.quantize.int4_row_quantize_zp` / `pack_int4`,
`torch.ops.fbgemm.bf16i4bf16_rowwise[_batched]`).  Their published algorithm is restated here:

    x = w.reshape(-1, G).float();  lo = min(x), hi = max(x)                       (per group)
    scale = clamp(hi - lo, min=1e-6) / 15;   zero = lo + 8 * scale
    q = clamp(round((x - lo) / scale), 0, 15) - 8                                   (int8 in [-8, 7])
    qdata = (q[:, 0::2] & 0xF) | (q[:, 1::2] << 4);   scale, zero -> (K/G, N), cast to w.dtype
    linear:  y = x @ (q * scale + zero).T                                           (bf16i4bf16_rowwise)

Parity: unpinned against fbgemm itself (absent); pinned to this restatement by tests/test_int4_tensor.py, which
re-states the reference's own test file (test/quantization/quantize_/workflows/int4/test_int4_tensor.py).

The GEMM runs on this repo's kernels: q + 8 is the unsigned nibble of the strip-major int4 storage
(include/ao_b200.h) and (scale, zero) are exactly its group parameters — w = (u - 8) * scale + zero — so the
B200 storage is built once per tensor (lazily, cached beside the reference's fields) and `linear` / `bmm` call
`torchao::b200_int4_weight_only_mm`.  The cache doubles the int4 bytes of a v2 weight (the price of keeping the
reference's fields addressable for slicing / copy_ / state_dict); `copy_` into a tensor drops the cache of every
Int4Tensor that aliases the written storage (writing `t.qdata` by hand does not: call `t.copy_` or re-wrap).
"""
import weakref
from typing import List

import torch
from torch.utils._python_dispatch import return_and_correct_aliasing

from ..utils import fill_defaults

__all__ = ["Int4Tensor", "int4_row_quantize_zp", "pack_int4", "unpack_int4"]

aten = torch.ops.aten
# Int4Tensor -> [(packed_weight, scale_and_zero)] per batch entry: the strip-major storage the kernels read.  Kept outside
# the tensor objects so that it is neither pickled with a state_dict nor seen by __tensor_flatten__.
# Keyed by id() (tensor `==` is elementwise) -> (data_ptr of qdata's storage, storage list); dropped by a finalizer.
_B200_CACHE = {}


def int4_row_quantize_zp(x: torch.Tensor, group_size: int = 128):
    """fbgemm_gpu.experimental.gen_ai.quantize.int4_row_quantize_zp, restated (see the module docstring).
    x (N, K) -> q int8 (N, K) in [-8, 7], scales fp32 (K/G, N), zeros fp32 (K/G, N)."""
    n_bit = 4
    to_quant = x.reshape(-1, group_size).to(torch.float)
    max_val = to_quant.amax(dim=1, keepdim=True)
    min_val = to_quant.amin(dim=1, keepdim=True)
    max_int = 2 ** n_bit - 1
    scales = (max_val - min_val).clamp(min=1e-6) / max_int
    zeros = min_val + scales * (2 ** (n_bit - 1))
    out = to_quant.sub(min_val).div(scales).round().clamp_(0, max_int)
    out = (out - 2 ** (n_bit - 1)).to(dtype=torch.int8).reshape(x.shape)
    scales = scales.view(x.shape[0], -1).t().contiguous()
    zeros = zeros.view(x.shape[0], -1).t().contiguous()
    return out, scales, zeros


def pack_int4(x: torch.Tensor) -> torch.Tensor:
    """fbgemm pack_int4: int8 (N, K) in [-8, 7] -> int8 (N, K/2), even k in the low nibble."""
    low_x = torch.bitwise_and(x[:, ::2], 0xF)
    high_x = torch.bitwise_left_shift(x[:, 1::2], 4)
    return torch.bitwise_or(low_x, high_x).contiguous()


def unpack_int4(qdata: torch.Tensor) -> torch.Tensor:
    """inverse of pack_int4 -> int8 (..., K) in [-8, 7]"""
    low = torch.bitwise_left_shift(qdata, 4) >> 4   # arithmetic shifts on int8: sign-extend the low nibble
    high = qdata >> 4
    return torch.stack((low, high), dim=-1).reshape(*qdata.shape[:-1], qdata.shape[-1] * 2)


class Int4Tensor(torch.Tensor):
    """reference: int4_tensor.py:29-106"""

    tensor_data_names = ["qdata", "scale", "zero_point"]
    tensor_attribute_names = ["block_size", "shape"]

    @staticmethod
    def __new__(cls, qdata, scale, zero_point, block_size, shape):
        return torch.Tensor._make_wrapper_subclass(cls, shape, device=qdata.device, dtype=scale.dtype, requires_grad=False)

    def __init__(self, qdata, scale, zero_point, block_size, shape):
        self.qdata = qdata
        self.scale = scale
        self.zero_point = zero_point
        self.block_size = list(block_size)

    def _quantization_type(self):
        return f"shape={self.shape}, block_size={self.block_size}, device={self.device}"

    def __repr__(self):
        return f"{self.__class__.__name__}({self._quantization_type()})"

    def __tensor_flatten__(self):
        return list(self.tensor_data_names), [self.block_size, self.shape]

    @classmethod
    def __tensor_unflatten__(cls, tensor_data_dict, tensor_attributes, outer_size, outer_stride):
        block_size, shape = tensor_attributes
        return cls(tensor_data_dict["qdata"], tensor_data_dict["scale"], tensor_data_dict["zero_point"], block_size,
                   shape if outer_size is None else outer_size)

    # ------------------------------------------------------------------ construction
    @classmethod
    def from_hp(cls, w: torch.Tensor, block_size: List[int]):
        """reference: int4_tensor.py:63-103"""
        assert len(block_size) == w.ndim, (
            f"Expecting the length of block_size to be equal to the dimension of the weight, got {block_size=} and {w.ndim=}")
        assert all(x == 1 for x in block_size[:-1]) and block_size[-1] != 1, "Only groupwise quant is supported right now"
        if not w.is_cuda:
            raise NotImplementedError("the B200 path quantizes CUDA tensors only (no CPU fallback)")
        group_size = block_size[-1]
        original_shape = w.shape
        if w.ndim >= 3:
            wq, scale, zero_point = zip(*[int4_row_quantize_zp(i, group_size) for i in w])
            wq = torch.stack([pack_int4(i) for i in wq], dim=0)
            scale = torch.stack(scale, dim=0)
            zero_point = torch.stack(zero_point, dim=0)
        else:
            wq, scale, zero_point = int4_row_quantize_zp(w, group_size)
            wq = pack_int4(wq)
        return cls(wq, scale.to(w.dtype), zero_point.to(w.dtype), list(block_size), original_shape)

    # ------------------------------------------------------------------ the B200 storage
    def _b200_storage(self):
        """-> list (one entry per batch element; one entry for a 2-D weight) of (packed_weight, scale_and_zero)."""
        hit = _B200_CACHE.get(id(self))
        if hit is None:
            from ..dtypes.layouts import TensorCoreTiledAQTTensorImpl, TensorCoreTiledLayout
            if self.scale.dtype != torch.bfloat16:
                raise NotImplementedError("Int4Tensor on B200: bfloat16 weights (tinygemm-style bf16 group parameters)")
            q = self.qdata if self.qdata.dim() == 3 else self.qdata.unsqueeze(0)
            s = self.scale if self.scale.dim() == 3 else self.scale.unsqueeze(0)
            z = self.zero_point if self.zero_point.dim() == 3 else self.zero_point.unsqueeze(0)
            out = []
            for e in range(q.shape[0]):
                u = (unpack_int4(q[e]).to(torch.int32) + 8).contiguous()          # (N, K) in 0..15
                impl = TensorCoreTiledAQTTensorImpl.from_plain(u, s[e].t().contiguous(), z[e].t().contiguous(),
                                                               TensorCoreTiledLayout(8))
                out.append((impl.packed_weight, impl.scale_and_zero))
            weakref.finalize(self, _B200_CACHE.pop, id(self), None)
            hit = _B200_CACHE[id(self)] = (self.qdata.untyped_storage().data_ptr(), out)
        return hit[1]

    def dequantize(self, output_dtype=None):
        q = unpack_int4(self.qdata).to(self.scale.dtype)                              # (..., N, K)
        g = self.block_size[-1]
        s = self.scale.transpose(-1, -2).repeat_interleave(g, dim=-1)
        z = self.zero_point.transpose(-1, -2).repeat_interleave(g, dim=-1)
        return (q * s + z).to(output_dtype or self.dtype)

    def _apply_fn_to_data(self, fn):
        return self.__class__(fn(self.qdata), fn(self.scale), fn(self.zero_point), self.block_size, self.shape)

    # ------------------------------------------------------------------ dispatch
    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        kwargs = kwargs or {}
        if func is torch.nn.functional.linear:
            return _linear(args, kwargs)
        if func is torch.bmm:
            return _bmm(args, kwargs)
        with torch._C.DisableTorchFunctionSubclass():
            return func(*args, **kwargs)

    @classmethod
    def __torch_dispatch__(cls, func, types, args, kwargs):
        kwargs = kwargs or {}
        handler = _ATEN_TABLE.get(func)
        if handler is None:
            raise NotImplementedError(f"{cls.__name__} dispatch: attempting to run unimplemented operator/function: {func}")
        return handler(func, args, kwargs)


torch.serialization.add_safe_globals([Int4Tensor])


def _mm(x2d, packed, group_size, sz, n_out):
    return torch.ops.torchao.b200_int4_weight_only_mm(x2d, packed, group_size, sz, n_out)


def _linear(args, kwargs):
    """reference: int4_tensor.py:109-135 (fbgemm.bf16i4bf16_rowwise -> torchao::b200_int4_weight_only_mm)"""
    input_tensor, weight_tensor = args[0], args[1]
    bias = args[2] if len(args) > 2 else kwargs.get("bias")
    assert weight_tensor.dim() == 2, "linear needs a 2-D Int4Tensor"
    orig_act_size = input_tensor.size()
    orig_out_features = weight_tensor.shape[-2]
    x2d = input_tensor.reshape(-1, input_tensor.shape[-1])
    if x2d.dtype != torch.bfloat16:
        raise NotImplementedError("Int4Tensor on B200: bfloat16 activations")
    packed, sz = weight_tensor._b200_storage()[0]
    res = _mm(x2d, packed, weight_tensor.block_size[-1], sz, orig_out_features)
    res = res.reshape(*orig_act_size[:-1], orig_out_features)
    if bias is not None:
        res = res + bias
    return res


def _bmm(args, kwargs):
    """reference: int4_tensor.py:138-159.  input (B, M, K) x weight Int4Tensor (B, N, K) -> (B, M, N)"""
    input_tensor, weight_tensor = args[0], args[1]
    assert weight_tensor.dim() == 3 and input_tensor.dim() == 3 and input_tensor.shape[0] == weight_tensor.shape[0]
    orig_out_features = weight_tensor.shape[-2]
    stor = weight_tensor._b200_storage()
    outs = [_mm(input_tensor[b].contiguous(), stor[b][0], weight_tensor.block_size[-1], stor[b][1], orig_out_features)
            for b in range(input_tensor.shape[0])]
    return torch.stack(outs, dim=0)


def _slice(func, args, kwargs):
    """reference: int4_tensor.py:162-235 — dim 0 / 1 of the (N, K) weight mapped onto qdata (N, K/2) and the
    transposed scale / zero_point (K/G, N); the three results are views (aliasing preserved)."""
    self, dim, start, end, step = fill_defaults(args, 5, [0, None, None, 1])
    assert step == 1
    assert dim == 0 or dim == 1, f"Only dim==0 or 1 are supported, got: {dim}"
    start = 0 if start is None else start
    if end is None or end >= self.shape[dim]:
        end = self.shape[dim]
    assert self.qdata.ndim == 2, f"Expected packed weight to have dim 2, got {self.qdata.dim}"
    n, k_by_2 = self.qdata.shape
    sz_dim0, sz_dim1 = self.scale.shape
    data_len = self.shape[dim]
    pw_len, sz_len = (n, sz_dim1) if dim == 0 else (k_by_2, sz_dim0)
    sz_dim = 1 - dim
    if pw_len == 0 or sz_len == 0:
        return return_and_correct_aliasing(func, args, kwargs, self.__class__(
            self.qdata, self.scale, self.zero_point, self.block_size, self.shape))
    pw_ratio = data_len / pw_len
    sz_ratio = data_len / sz_len
    qdata = aten.slice.Tensor(self.qdata, dim, int(start / pw_ratio), int(end / pw_ratio), step)
    scale = aten.slice.Tensor(self.scale, sz_dim, int(start / sz_ratio), int(end / sz_ratio), step)
    zero_point = aten.slice.Tensor(self.zero_point, sz_dim, int(start / sz_ratio), int(end / sz_ratio), step)
    new = self.__class__(qdata, scale, zero_point, self.block_size, (qdata.shape[0], qdata.shape[1] * 2))
    return return_and_correct_aliasing(func, args, kwargs, new)


def _view(func, args, kwargs):
    self, size = args
    if tuple(self.shape) == tuple(size):
        return return_and_correct_aliasing(func, args, kwargs, self.__class__(
            self.qdata, self.scale, self.zero_point, self.block_size, self.shape))
    raise NotImplementedError(f"Int4Tensor.view to a different shape ({tuple(self.shape)} -> {tuple(size)}) is not supported")


def _squeeze(func, args, kwargs):
    """reference: int4_tensor.py:446-490 (dim 0 of a (1, N, K) weight)"""
    self, dim = args
    dims = [dim] if isinstance(dim, int) else list(dim)
    assert all(d == 0 or d == -self.dim() for d in dims) and self.shape[0] == 1, "squeeze: only a leading dimension of 1"
    new = self.__class__(self.qdata.squeeze(0), self.scale.squeeze(0), self.zero_point.squeeze(0), self.block_size[1:],
                         self.shape[1:])
    return return_and_correct_aliasing(func, args, kwargs, new)


def _select(func, args, kwargs):
    self, dim, index = fill_defaults(args, 3, [0, 0])
    assert dim == 0 and self.dim() == 3, "select: dim 0 of a 3-D Int4Tensor"
    new = self.__class__(self.qdata[index], self.scale[index], self.zero_point[index], self.block_size[1:], self.shape[1:])
    return return_and_correct_aliasing(func, args, kwargs, new)


def _copy_(func, args, kwargs):
    dst, src = args[0], args[1]
    if (isinstance(src, Int4Tensor) and dst.shape == src.shape and list(dst.block_size) == list(src.block_size)
            and dst.qdata.shape == src.qdata.shape):
        dst.qdata.copy_(src.qdata)
        dst.scale.copy_(src.scale)
        dst.zero_point.copy_(src.zero_point)
        # every tensor that aliases the written storage (the parent of a slice, its .data, ...) loses its kernel storage
        base = dst.qdata.untyped_storage().data_ptr()
        for key in [k for k, v in _B200_CACHE.items() if v[0] == base]:
            _B200_CACHE.pop(key, None)
        return dst
    raise ValueError(f"Not supported args for copy_ due to metadata mismatch: {args[0], args[1]}")


def _to_copy(func, args, kwargs):
    self = args[0]
    device = kwargs.get("device", None)
    dtype = kwargs.get("dtype", None)
    if dtype is not None and dtype != self.dtype:
        raise NotImplementedError("changing the dtype of an Int4Tensor is not supported")
    return return_and_correct_aliasing(func, args, kwargs, self._apply_fn_to_data(
        lambda x: x.to(device=device).clone() if device is not None else x.clone()))


_ATEN_TABLE = {
    aten.linear.default: lambda f, a, k: _linear(a, k),
    aten.bmm.default: lambda f, a, k: _bmm(a, k),
    aten.slice.Tensor: _slice,
    aten.view.default: _view,
    aten.squeeze.dim: _squeeze,
    aten.squeeze.dims: _squeeze,
    aten.select.int: _select,
    aten.detach.default: lambda f, a, k: return_and_correct_aliasing(f, a, k, a[0]._apply_fn_to_data(torch.detach)),
    aten.alias.default: lambda f, a, k: return_and_correct_aliasing(f, a, k, a[0]._apply_fn_to_data(torch.detach)),
    aten.clone.default: lambda f, a, k: return_and_correct_aliasing(f, a, k, a[0]._apply_fn_to_data(torch.clone)),
    aten.contiguous.default: lambda f, a, k: a[0]._apply_fn_to_data(lambda x: x.contiguous()),
    aten._to_copy.default: _to_copy,
    aten.copy_.default: _copy_,
}
