from .model import ModelArgs, Transformer, model_weight_bytes, transformer_configs  # noqa: F401
from .decode_engine import DecodeEngine  # noqa: F401
from .generate import Generator, generate  # noqa: F401
