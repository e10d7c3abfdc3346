"""AOBaseConfig + JSON (de)serialisation (reference: torchao/core/config.py:27-303)."""
import abc
import dataclasses
import enum
import importlib
import json
from typing import Any, Dict

import torch

__all__ = ["AOBaseConfig", "config_to_dict", "config_from_dict", "ConfigJSONEncoder"]


class AOBaseConfig(abc.ABC):
    """Base of every workflow config; `quantize_(model, config)` looks the handler up by type(config).
    reference: core/config.py:27-60"""

    VERSION = 1


def _encode(o):
    if isinstance(o, AOBaseConfig) or dataclasses.is_dataclass(o) and not isinstance(o, type):
        data = {f.name: _encode(getattr(o, f.name)) for f in dataclasses.fields(o)} if dataclasses.is_dataclass(o) else {}
        return {"_type": type(o).__name__, "_version": getattr(o, "VERSION", 1), "_module": type(o).__module__,
                "_data": data}
    if isinstance(o, enum.Enum):
        return {"_type": type(o).__name__, "_module": type(o).__module__, "_data": o.name}
    if isinstance(o, torch.dtype):
        return {"_type": "torch.dtype", "_data": str(o).split(".")[-1]}
    if isinstance(o, (list, tuple)):
        return [_encode(v) for v in o]
    if isinstance(o, dict):
        return {k: _encode(v) for k, v in o.items()}
    return o


class ConfigJSONEncoder(json.JSONEncoder):
    def default(self, o):
        return _encode(o)


def config_to_dict(config: AOBaseConfig) -> Dict[str, Any]:
    """reference: core/config.py:174-199"""
    if not isinstance(config, AOBaseConfig):
        raise TypeError(f"Expected AOBaseConfig instance, got {type(config)}")
    return _encode(config)


def _decode(d):
    if isinstance(d, list):
        return [_decode(v) for v in d]
    if not isinstance(d, dict):
        return d
    if "_type" not in d:
        return {k: _decode(v) for k, v in d.items()}
    if d["_type"] == "torch.dtype":
        return getattr(torch, d["_data"])
    mod = importlib.import_module(d["_module"])
    cls = getattr(mod, d["_type"])
    if isinstance(cls, type) and issubclass(cls, enum.Enum):
        return cls[d["_data"]]
    expected = getattr(cls, "VERSION", 1)
    if d.get("_version", 1) != expected:
        raise ValueError(f"Version mismatch: stored {d.get('_version')}, current {expected}")
    return cls(**{k: _decode(v) for k, v in d["_data"].items()})


def config_from_dict(data: Dict[str, Any]) -> AOBaseConfig:
    """reference: core/config.py:202-303"""
    if not isinstance(data, dict) or "_type" not in data or "_data" not in data:
        raise ValueError("Input dictionary missing required '_type' or '_data' fields")
    return _decode(data)
