"""Layout / tensor-impl base types (reference: torchao/dtypes/utils.py:33-135)."""
from dataclasses import dataclass
from typing import Tuple

import torch


@dataclass(frozen=True)
class Layout:
    """reference: dtypes/utils.py:33-69"""

    def pre_process(self, input: torch.Tensor) -> torch.Tensor:
        return input

    def post_process(self, input, scale, zero_point, block_size):
        return input, scale, zero_point

    def pre_process_static(self, input, scale, zero_point, block_size):
        return self.pre_process(input), scale, zero_point

    def __repr__(self):
        return f"{self.__class__.__name__}({self.extra_repr()})"

    def extra_repr(self) -> str:
        return ""


@dataclass(frozen=True)
class PlainLayout(Layout):
    """reference: dtypes/utils.py:72-82"""


def is_device(target_device_str: str, device) -> bool:
    return torch.device(device).type == target_device_str


class AQTTensorImpl(torch.Tensor):
    """Base of the per-layout storage tensors (reference: dtypes/utils.py:102-135)."""

    def get_plain(self) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        raise NotImplementedError

    def get_layout(self) -> Layout:
        raise NotImplementedError

    @classmethod
    def from_plain(cls, data, scale, zero_point, _layout):
        raise NotImplementedError

    def __repr__(self):
        return f"{self.__class__.__name__}(layout={self.get_layout()}, shape={tuple(self.shape)})"

    __torch_function__ = torch._C._disabled_torch_function_impl
