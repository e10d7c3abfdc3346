"""torchao_fork_b200 — B200-native weight-only quantized linear behind the torchao API.

    from torchao_fork_b200 import quantize_, Int4WeightOnlyConfig
    quantize_(model, Int4WeightOnlyConfig(group_size=128))

The CUDA library (lib/libao_b200.so, built by `python -m torchao_fork_b200.build`) is loaded lazily by the
first op call; a missing library raises — there is no fallback path.
"""
from . import ops  # noqa: F401  registers torch.ops.torchao.*
from .quantization import (  # noqa: F401
    Int4WeightOnlyConfig, Int8WeightOnlyConfig, int4_weight_only, int8_weight_only, quantize_,
)
from . import dtypes, quantization  # noqa: F401

__version__ = "0.1.0"
