"""config type -> module transform registry (reference: torchao/quantization/transform_module.py:13-52)."""
import functools
from typing import Callable, Dict, Type

import torch

from ..core.config import AOBaseConfig

_QUANTIZE_CONFIG_HANDLER: Dict[Type[AOBaseConfig], Callable[[torch.nn.Module, AOBaseConfig], torch.nn.Module]] = {}


def register_quantize_module_handler(config_type):
    """`@register_quantize_module_handler(MyConfig) def handler(module, config) -> module`"""

    @functools.wraps(config_type)
    def decorator(func):
        _QUANTIZE_CONFIG_HANDLER[config_type] = func
        return func

    return decorator
