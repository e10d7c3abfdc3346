"""`quantize_` and the weight-only configs (reference: torchao/quantization/quant_api.py).

Kept: `quantize_` (:482-546), `_replace_with_custom_fn_if_matches_filter` (:173-218), `_is_linear`
(:271-283), `Int4WeightOnlyConfig` (:997-1157, version 1 / TensorCoreTiledLayout),
`Int8WeightOnlyConfig` (:1200-1255).  The other ~18 workflow configs are outside the hot path.
"""
import logging
import types
from dataclasses import dataclass
from enum import Enum
from typing import Callable, Optional, Union

import torch
import torch.nn as nn

from ..core.config import AOBaseConfig
from ..dtypes.affine_quantized_tensor import AffineQuantizedTensor, to_affine_quantized_intx
from ..dtypes.layouts import PlainLayout, TensorCoreTiledLayout
from .quant_primitives import MappingType, ZeroPointDomain
from .transform_module import _QUANTIZE_CONFIG_HANDLER, register_quantize_module_handler

logger = logging.getLogger(__name__)

__all__ = ["quantize_", "Int4WeightOnlyConfig", "Int8WeightOnlyConfig", "int4_weight_only", "int8_weight_only",
           "PackingFormat"]


class PackingFormat(str, Enum):
    """reference: torchao/quantization/quantize_/common/packing_format.py (only PLAIN is on this path)."""
    PLAIN = "plain"
    PRESHUFFLED = "preshuffled"
    MARLIN_SPARSE = "marlin_sparse"


# reference: quant_api.py:158-170
LAYOUT_TO_ZERO_POINT_DOMAIN = {TensorCoreTiledLayout: [ZeroPointDomain.FLOAT]}
LAYOUT_TO_PRESERVE_ZEROS = {TensorCoreTiledLayout: False}


def _is_linear(mod, *args):
    """reference: quant_api.py:271-283 — an nn.Linear whose weight is still a plain tensor."""
    return (isinstance(mod, torch.nn.Linear) and hasattr(mod, "weight")
            and not isinstance(mod.weight, AffineQuantizedTensor))


def _replace_with_custom_fn_if_matches_filter(model, replacement_fn, filter_fn, cur_fqn="", device=None,
                                              extra_args=()):
    """reference: quant_api.py:173-218 (recursive walk; optional streaming move to `device`)."""
    if filter_fn(model, cur_fqn[:-1]):
        if device is not None:
            model.to(device=device)
        return replacement_fn(model, *extra_args)
    for name, child in list(model.named_children()):
        new_child = _replace_with_custom_fn_if_matches_filter(child, replacement_fn, filter_fn,
                                                              f"{cur_fqn}{name}.", device, extra_args)
        if new_child is not child and new_child is not None:
            setattr(model, name, new_child)
    if device is not None:
        model.to(device=device)
    return model


def quantize_(model: torch.nn.Module, config: AOBaseConfig,
              filter_fn: Optional[Callable[[torch.nn.Module, str], bool]] = None,
              device: Optional[Union[str, torch.device]] = None):
    """Quantize `model` in place.  reference: quant_api.py:482-546."""
    torch._C._log_api_usage_once("torchao.quantization.quantize_")
    filter_fn = _is_linear if filter_fn is None else filter_fn
    if not isinstance(config, AOBaseConfig):
        raise AssertionError("Passing a generic Callable to `quantize_` is no longer recommended; pass an AOBaseConfig")
    if type(config) not in _QUANTIZE_CONFIG_HANDLER:
        raise NotImplementedError(f"{type(config).__name__} is outside the B200 weight-only path")
    handler = _QUANTIZE_CONFIG_HANDLER[type(config)]
    _replace_with_custom_fn_if_matches_filter(model, handler, filter_fn, device=device, extra_args=(config,))


def _linear_extra_repr(self):
    w = self.weight
    desc = w._quantization_type() if isinstance(w, AffineQuantizedTensor) else str(tuple(w.shape))
    return f"in_features={w.shape[1]}, out_features={w.shape[0]}, weight={type(w).__name__}({desc})"


@dataclass
class Int4WeightOnlyConfig(AOBaseConfig):
    """uint4 weight-only asymmetric per-group quantization for the tinygemm recipe.
    reference: quant_api.py:997-1031 (same fields and defaults)."""

    group_size: int = 128
    layout: Optional[TensorCoreTiledLayout] = TensorCoreTiledLayout(inner_k_tiles=8)
    use_hqq: bool = False
    zero_point_domain: Optional[ZeroPointDomain] = ZeroPointDomain.NONE
    set_inductor_config: bool = True
    preserve_zero: Optional[bool] = None
    packing_format: PackingFormat = PackingFormat.PLAIN
    version: int = 1

    def __post_init__(self):
        torch._C._log_api_usage_once("torchao.quantization.Int4WeightOnlyConfig")


int4_weight_only = Int4WeightOnlyConfig


def _int4_weight_only_quantize_tensor(weight, config):
    """reference: quant_api.py:1042-1140"""
    group_size = config.group_size
    layout = config.layout
    if weight.shape[-1] % group_size != 0:
        logger.info(f"Skipping quantizing weight with int4 weight only quantization because the shape of weight "
                    f"{weight.shape} is not compatible with group_size {group_size}")
        return weight
    if config.version == 2:
        # reference: quant_api.py:1062-1084
        block_size = [1 for _ in range(weight.ndim - 1)] + [group_size]
        fmt = config.packing_format
        fmt = fmt.value if isinstance(fmt, PackingFormat) else str(fmt)
        if fmt == PackingFormat.PLAIN.value:
            from .int4_tensor import Int4Tensor
            return Int4Tensor.from_hp(weight, block_size)
        if fmt in ("preshuffled", "marlin_sparse"):
            raise NotImplementedError(
                f"packing_format={fmt!r}: a storage order of fbgemm's / marlin's own kernels; on B200 every int4 weight is "
                "served from the strip-major storage — use packing_format='plain'")
        raise ValueError(f"Unsupported packing format: {config.packing_format}")
    assert config.version == 1
    assert type(layout) in LAYOUT_TO_ZERO_POINT_DOMAIN, f"Only support layout: {LAYOUT_TO_ZERO_POINT_DOMAIN.keys()}"
    zero_point_domain = config.zero_point_domain
    if zero_point_domain == ZeroPointDomain.NONE:
        zero_point_domain = LAYOUT_TO_ZERO_POINT_DOMAIN[type(layout)][0]
    else:
        assert zero_point_domain in LAYOUT_TO_ZERO_POINT_DOMAIN[type(layout)], (
            f"Layout only support {LAYOUT_TO_ZERO_POINT_DOMAIN[type(layout)]}")
    preserve_zero = config.preserve_zero if config.preserve_zero is not None else LAYOUT_TO_PRESERVE_ZEROS[type(layout)]
    block_size = tuple([1 for _ in range(weight.ndim - 1)] + [group_size])
    return to_affine_quantized_intx(
        weight, MappingType.ASYMMETRIC, block_size, torch.int32, 0, 15, 1e-6,
        zero_point_dtype=torch.bfloat16, preserve_zero=preserve_zero, zero_point_domain=zero_point_domain,
        _layout=layout, use_hqq=config.use_hqq)


@register_quantize_module_handler(Int4WeightOnlyConfig)
def _int4_weight_only_transform(module, config):
    """reference: quant_api.py:1143-1157.  `set_inductor_config` is accepted and ignored: this path
    does not go through inductor (CUDA graphs, not a tracing compiler)."""
    assert hasattr(module, "weight"), "applying int4 weight only quant requires module to have weight attribute"
    new_weight = _int4_weight_only_quantize_tensor(module.weight, config)
    module.weight = torch.nn.Parameter(new_weight, requires_grad=False)
    module.extra_repr = types.MethodType(_linear_extra_repr, module)
    return module


@dataclass
class Float8DynamicActivationInt4WeightConfig(AOBaseConfig):
    """reference: quant_api.py:1159-1197 — float8 per-row DYNAMIC ACTIVATION quantization on top of preshuffled int4
    weights (fbgemm's Int4PreshuffledTensor and its fp8 x int4 kernel).  The config exists so that code naming it imports;
    applying it raises: it is not a weight-only path (the activations are quantized too), outside SURVEY.md section 8."""

    packing_format: PackingFormat = PackingFormat.PRESHUFFLED


@register_quantize_module_handler(Float8DynamicActivationInt4WeightConfig)
def _float8_dynamic_activation_int4_weight_transform(module, config):
    raise NotImplementedError(
        "Float8DynamicActivationInt4WeightConfig quantizes the activations to float8 and needs fbgemm's preshuffled int4 "
        "kernel; the B200 path is weight-only (bf16 activations): use Int4WeightOnlyConfig(group_size=128[, version=2])")


@dataclass
class Int8WeightOnlyConfig(AOBaseConfig):
    """int8 weight-only symmetric per-channel quantization.  reference: quant_api.py:1200-1216"""

    group_size: Optional[int] = None
    set_inductor_config: bool = True

    def __post_init__(self):
        torch._C._log_api_usage_once("torchao.quantization.Int8WeightOnlyConfig")


int8_weight_only = Int8WeightOnlyConfig


def _int8_weight_only_quantize_tensor(weight, config):
    """reference: quant_api.py:1223-1240"""
    group_size = config.group_size if config.group_size is not None else weight.shape[-1]
    block_size = tuple([1 for _ in range(weight.dim() - 1)] + [group_size])
    return to_affine_quantized_intx(weight, MappingType.SYMMETRIC, block_size, torch.int8,
                                    eps=torch.finfo(torch.float32).eps, zero_point_dtype=torch.int64)


@register_quantize_module_handler(Int8WeightOnlyConfig)
def _int8_weight_only_transform(module, config):
    """reference: quant_api.py:1243-1255"""
    assert hasattr(module, "weight"), "applying int8 weight only quant requires module to have weight attribute"
    new_weight = _int8_weight_only_quantize_tensor(module.weight, config)
    module.weight = torch.nn.Parameter(new_weight, requires_grad=False)
    module.extra_repr = types.MethodType(_linear_extra_repr, module)
    return module
