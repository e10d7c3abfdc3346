"""Affine quantization primitives of the weight-only path, backed by the B200 kernels.

Same names, argument meaning and error behaviour as torchao/quantization/quant_primitives.py for
the two branches the weight-only path uses (tinygemm uint4 group-wise; symmetric int8 per-row).
Everything else in the reference file (HQQ, float8, floatx, GGUF, QQQ ...) is outside the hot
path and raises NotImplementedError rather than silently taking a slow path.

The arithmetic contract is the reference's eager bf16 arithmetic: every op rounds to bf16.
`CUDA_DIV_SEMANTICS` selects how `tensor / python_scalar` is evaluated: torch CUDA eager
multiplies by the fp32 reciprocal (the reference quantizing a model on the GPU does that, so it
is the default); torch CPU eager divides (what the committed golden vectors hold).
"""
from enum import Enum, auto
from typing import List, Optional, Tuple, Union

import torch

from .. import ops as _ops  # noqa: F401  (registers torch.ops.torchao.*)

__all__ = [
    "MappingType", "ZeroPointDomain", "TorchAODType",
    "choose_qparams_affine", "quantize_affine", "dequantize_affine",
    "_choose_qparams_affine_tinygemm", "_quantize_affine_tinygemm", "_dequantize_affine_tinygemm",
]

CUDA_DIV_SEMANTICS = True


class MappingType(Enum):
    """reference: quant_primitives.py:54-75"""
    SYMMETRIC = auto()
    SYMMETRIC_NO_CLIPPING_ERR = auto()
    ASYMMETRIC = auto()


class ZeroPointDomain(Enum):
    """reference: quant_primitives.py:78-90"""
    INT = auto()
    FLOAT = auto()
    NONE = auto()


class TorchAODType(Enum):
    INT1 = auto()
    INT2 = auto()
    INT3 = auto()
    INT4 = auto()
    INT5 = auto()
    INT6 = auto()
    INT7 = auto()


torch.serialization.add_safe_globals([MappingType, ZeroPointDomain])

_DTYPE_BOUNDS = {
    torch.uint8: (0, 255), torch.int8: (-128, 127), torch.int16: (-(2**15), 2**15 - 1),
    torch.int32: (-(2**31), 2**31 - 1), torch.uint4: (0, 15) if hasattr(torch, "uint4") else (0, 15),
}


def _get_and_check_qmin_qmax(dtype, quant_min, quant_max):
    """reference: quant_primitives.py:230-261"""
    if dtype not in _DTYPE_BOUNDS:
        raise ValueError(f"Unsupported dtype: {dtype}")
    lo, hi = _DTYPE_BOUNDS[dtype]
    quant_min = lo if quant_min is None else quant_min
    quant_max = hi if quant_max is None else quant_max
    assert quant_min >= lo, f"quant_min out of bound for dtype, quant_min_lower_bound: {lo} quant_min: {quant_min}"
    assert quant_max <= hi, f"quant_max out of bound for dtype, quant_max_upper_bound: {hi} quant_max: {quant_max}"
    return quant_min, quant_max


def _group_size_of(block_size, shape):
    """block_size (1, ..., 1, G) over a 2-D weight -> G.  Anything else is outside the hot path."""
    assert len(block_size) == len(shape), f"Got input dim:{len(shape)}, block_size: {block_size}"
    if len(shape) != 2 or block_size[0] != 1:
        raise NotImplementedError(f"B200 path supports 2-D weights with block_size (1, G); got {block_size}")
    g = block_size[-1]
    assert shape[-1] % g == 0, (
        f"Expecting input size at 1 dimension: {shape[-1]} to be divisible by block_size at 1 dimension: {g}")
    return g


def _require_cuda_bf16(t, what):
    if not t.is_cuda:
        raise NotImplementedError(f"{what}: the B200 path has no CPU implementation (got device {t.device})")
    if t.dtype != torch.bfloat16:
        raise NotImplementedError(f"{what}: the B200 path quantizes bfloat16 weights (got {t.dtype})")


@torch.no_grad()
def _choose_qparams_affine_tinygemm(input, mapping_type, block_size, target_dtype, quant_min=None, quant_max=None,
                                    eps=None, scale_dtype=None, zero_point_dtype=None):
    """reference: quant_primitives.py:1238-1306.  Returns (scale, zero_point), each (N, K/G) bf16."""
    quant_min, quant_max = _get_and_check_qmin_qmax(target_dtype, quant_min, quant_max)
    assert mapping_type is MappingType.ASYMMETRIC, f"Unsupported mapping type: {mapping_type}"
    _require_cuda_bf16(input, "_choose_qparams_affine_tinygemm")
    if (quant_min, quant_max) != (0, 15):
        raise NotImplementedError("tinygemm qparams are implemented for uint4 (quant_min=0, quant_max=15)")
    if scale_dtype not in (None, torch.bfloat16) or zero_point_dtype not in (None, torch.bfloat16):
        raise NotImplementedError("tinygemm qparams are bf16 on the B200 path")
    if eps is None:
        eps = torch.finfo(input.dtype).eps
    g = _group_size_of(block_size, input.shape)
    return torch.ops.torchao.b200_int4_choose_qparams(input, g, float(eps), CUDA_DIV_SEMANTICS)


def _quantize_affine_tinygemm(input, block_size, scale, zero_point, output_dtype, quant_min=None, quant_max=None):
    """reference: quant_primitives.py:462-573.  int32 (N, K) in 0..15."""
    quant_min, quant_max = _get_and_check_qmin_qmax(output_dtype, quant_min, quant_max)
    _require_cuda_bf16(input, "_quantize_affine_tinygemm")
    if (quant_min, quant_max) != (0, 15):
        raise NotImplementedError("tinygemm quantize is implemented for uint4 (quant_min=0, quant_max=15)")
    g = _group_size_of(block_size, input.shape)
    q = torch.ops.torchao.b200_int4_quantize(input, scale.to(torch.bfloat16), zero_point.to(torch.bfloat16), g)
    return q.to(output_dtype)


def _dequantize_affine_tinygemm(input, block_size, scale, zero_point, input_dtype, quant_min=None, quant_max=None, *,
                                output_dtype=torch.float32):
    """reference: quant_primitives.py:939-1031: (q - mid) -> output_dtype, *= scale, += zero_point
    (one rounding per op).  Plain device-side tensor ops; not on the hot path (the linear kernels
    dequantize in registers)."""
    quant_min, quant_max = _get_and_check_qmin_qmax(input_dtype, quant_min, quant_max)
    g = _group_size_of(block_size, input.shape)
    n, k = input.shape
    mid_point = (quant_max + quant_min + 1) / 2
    dq = (input.view(n, k // g, g) - mid_point).to(output_dtype)
    dq = dq * scale.view(n, k // g, 1)
    if zero_point is not None:
        dq = dq + zero_point.view(n, k // g, 1)
    return dq.view(n, k).to(output_dtype)


@torch.no_grad()
def choose_qparams_affine(input, mapping_type, block_size, target_dtype, quant_min=None, quant_max=None, eps=None,
                          scale_dtype=None, zero_point_dtype=None):
    """reference: quant_primitives.py:1193-1234, 1497-1577.  Implemented branch: SYMMETRIC int8 with one
    block per row (what Int8WeightOnlyConfig uses).  Returns (scale (N,), zero_point (N,) zeros)."""
    quant_min, quant_max = _get_and_check_qmin_qmax(target_dtype, quant_min, quant_max)
    _require_cuda_bf16(input, "choose_qparams_affine")
    if mapping_type is not MappingType.SYMMETRIC or target_dtype != torch.int8 or (quant_min, quant_max) != (-128, 127):
        raise NotImplementedError("B200 path implements SYMMETRIC int8 (-128..127) choose_qparams_affine")
    g = _group_size_of(block_size, input.shape)
    if g != input.shape[-1]:
        raise NotImplementedError("B200 int8 weight-only path is per-channel (group_size == in_features)")
    if eps is None:
        eps = torch.finfo(input.dtype).eps
    _, scale = torch.ops.torchao.b200_int8_quantize(input, float(eps), CUDA_DIV_SEMANTICS)
    zero_point = torch.zeros(scale.shape, dtype=zero_point_dtype or torch.int32, device=input.device)
    return scale.to(scale_dtype or input.dtype), zero_point


def quantize_affine(input, block_size, scale, zero_point, output_dtype, quant_min=None, quant_max=None):
    """reference: quant_primitives.py:305-459 (int8, zero_point == 0 branch)."""
    quant_min, quant_max = _get_and_check_qmin_qmax(output_dtype, quant_min, quant_max)
    _require_cuda_bf16(input, "quantize_affine")
    if output_dtype != torch.int8 or (quant_min, quant_max) != (-128, 127):
        raise NotImplementedError("B200 path implements int8 (-128..127) quantize_affine")
    g = _group_size_of(block_size, input.shape)
    if g != input.shape[-1]:
        raise NotImplementedError("B200 int8 weight-only path is per-channel (group_size == in_features)")
    if zero_point is not None and zero_point.numel() and bool((zero_point != 0).any()):
        raise NotImplementedError("B200 int8 weight-only path is symmetric (zero_point == 0)")
    # the kernel derives q from the given scale exactly as quantize_affine does: round(w * (1/scale))
    inv = (1.0 / scale.to(torch.bfloat16)).view(-1, 1)
    return torch.clamp(torch.round(input * inv), quant_min, quant_max).to(output_dtype)


def dequantize_affine(input, block_size, scale, zero_point, input_dtype, quant_min=None, quant_max=None, *,
                      output_dtype=torch.float32):
    """reference: quant_primitives.py:726-833 (INT zero-point domain): (q - zp) * scale."""
    assert input.dtype == input_dtype, f"Expected: {input_dtype}, got: {input.dtype}"
    g = _group_size_of(block_size, input.shape)
    n, k = input.shape
    dq = input.view(n, k // g, g).to(output_dtype)
    if zero_point is not None:
        dq = dq - zero_point.view(n, k // g, 1).to(output_dtype)
    dq = dq * scale.view(n, k // g, 1)
    return dq.view(n, k).to(output_dtype)
