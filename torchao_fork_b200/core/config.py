"""AOBaseConfig + JSON (de)serialisation — host-side mirror of torchao/core/config.py:27-303.

Wire format (the reference's, so dicts travel both ways): {"_type": <class NAME>, "_version": <instance version>,
"_data": {...}}; enums {"_type": name, "_data": member}; torch dtypes {"_type": "torch.dtype", "_data": "bfloat16"};
layouts {"_type": name, "_data": {...}}.  There is NO module path in the format: `config_from_dict` resolves `_type`
only inside an allowlist of this package's modules (reference: ALLOWED_AO_MODULES, config.py:202-210) and only to
config / enum / layout classes — a dict can never name an arbitrary importable callable."""
import abc
import dataclasses
import enum
import importlib
import json
import warnings
from typing import Any, Dict

import torch

__all__ = ["AOBaseConfig", "config_to_dict", "config_from_dict", "ConfigJSONEncoder", "ALLOWED_AO_MODULES"]

_DEFAULT_VERSION = 1


class AOBaseConfig(abc.ABC):
    """Base of every workflow config; `quantize_(model, config)` looks the handler up by type(config).
    `version` is an INSTANCE attribute (subclasses declare `version: int = N`) so that several versions of one config
    can co-exist (reference: core/config.py:27-66)."""

    version: int = _DEFAULT_VERSION


class ConfigJSONEncoder(json.JSONEncoder):
    """reference: core/config.py:69-171"""

    def default(self, o):
        if isinstance(o, AOBaseConfig):
            fields = ({f.name: getattr(o, f.name) for f in dataclasses.fields(o)} if dataclasses.is_dataclass(o)
                      else {k: v for k, v in o.__dict__.items() if not k.startswith("_")})
            return {"_type": type(o).__name__, "_version": getattr(o, "version", _DEFAULT_VERSION),
                    "_data": {k: self.encode_value(v) for k, v in fields.items() if k != "version"}}
        if dataclasses.is_dataclass(o) and not isinstance(o, type):
            if "Layout" in type(o).__name__:  # layouts carry no version (reference :129-137)
                return {"_type": type(o).__name__,
                        "_data": {f.name: self.encode_value(getattr(o, f.name)) for f in dataclasses.fields(o)}}
            return {"_type": type(o).__name__, "_version": getattr(o, "version", _DEFAULT_VERSION),
                    "_data": {f.name: self.encode_value(getattr(o, f.name)) for f in dataclasses.fields(o) if f.name != "version"}}
        if "Layout" in type(o).__name__ and hasattr(o, "__dict__"):
            return {"_type": type(o).__name__,
                    "_data": {k: self.encode_value(v) for k, v in o.__dict__.items() if not k.startswith("_")}}
        if isinstance(o, enum.Enum):
            return {"_type": type(o).__name__, "_data": o.name}
        if isinstance(o, torch.dtype):
            return {"_type": "torch.dtype", "_data": str(o).split(".")[-1]}
        if isinstance(o, list):
            return [self.encode_value(v) for v in o]
        if isinstance(o, tuple):
            raise NotImplementedError(f"Tuples are serialized as lists in JSON; use a list instead. got: {o}")
        if isinstance(o, dict):
            return {k: self.encode_value(v) for k, v in o.items()}
        return super().default(o)

    def encode_value(self, value):
        try:
            return self.default(value)
        except TypeError:
            return value


def config_to_dict(config: AOBaseConfig) -> Dict[str, Any]:
    """reference: core/config.py:174-199"""
    if not isinstance(config, AOBaseConfig):
        raise TypeError(f"Expected AOBaseConfig instance, got {type(config)}")
    return json.loads(json.dumps(config, cls=ConfigJSONEncoder))


# the only places a `_type` may resolve to (reference: ALLOWED_AO_MODULES)
ALLOWED_AO_MODULES = (
    "torchao_fork_b200.quantization",
    "torchao_fork_b200.dtypes",
    "torchao_fork_b200.quantization.quant_primitives",
)


def _resolve(type_name: str):
    if not isinstance(type_name, str) or not type_name.isidentifier():
        raise ValueError(f"Invalid config type name {type_name!r}")
    from ..dtypes.utils import Layout

    for module_path in ALLOWED_AO_MODULES:
        module = importlib.import_module(module_path)
        cls = getattr(module, type_name, None)
        if isinstance(cls, type) and issubclass(cls, (AOBaseConfig, enum.Enum, Layout)):
            return cls
    raise ValueError(f"Failed to find class {type_name} in any of the allowed modules: {', '.join(ALLOWED_AO_MODULES)}")


def _is_encoded(v):
    return isinstance(v, dict) and "_type" in v and "_data" in v


def config_from_dict(data: Dict[str, Any]):
    """reference: core/config.py:213-303"""
    if not isinstance(data, dict):
        raise TypeError(f"Expected dictionary, got {type(data)}")
    if "_type" not in data or "_data" not in data:
        raise ValueError("Input dictionary missing required '_type' or '_data' fields")
    type_name, obj_data = data["_type"], data["_data"]
    stored_version = data.get("_version", _DEFAULT_VERSION)
    if type_name == "torch.dtype":
        dt = getattr(torch, obj_data, None) if isinstance(obj_data, str) else None
        if not isinstance(dt, torch.dtype):
            raise ValueError(f"Unknown torch dtype {obj_data!r}")
        return dt
    cls = _resolve(type_name)
    if issubclass(cls, enum.Enum):
        if not isinstance(obj_data, str) or obj_data not in cls.__members__:
            raise ValueError(f"{obj_data!r} is not a member of {cls.__name__}")
        return cls[obj_data]
    if not isinstance(obj_data, dict):
        raise ValueError(f"'_data' of {type_name} must be a dictionary")
    current = getattr(cls, "version", _DEFAULT_VERSION)
    processed = {}
    if stored_version != current:
        warnings.warn(f"Stored version is not the same as current default version of the config: "
                      f"{stored_version=}, current_default_version={current}, please check the deprecation warning")
        processed["version"] = stored_version
    for key, value in obj_data.items():
        if _is_encoded(value):
            processed[key] = config_from_dict(value)
        elif isinstance(value, list):
            processed[key] = [config_from_dict(v) if _is_encoded(v) else v for v in value]
        elif isinstance(value, dict):
            processed[key] = {k: config_from_dict(v) if _is_encoded(v) else v for k, v in value.items()}
        else:
            processed[key] = value
    try:
        return cls(**processed)
    except Exception as e:
        raise ValueError(f"Failed to create instance of {cls.__name__}: {e}")
