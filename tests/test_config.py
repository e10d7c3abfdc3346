"""Config (de)serialisation: the reference's wire format (torchao/core/config.py:69-303) both ways, and no way to name
an arbitrary importable class from a dictionary."""
import json
import subprocess
import sys
import warnings

import pytest
import torch

import torchao_fork_b200 as ao
from torchao_fork_b200.core.config import config_from_dict, config_to_dict
from torchao_fork_b200.dtypes import TensorCoreTiledLayout
from torchao_fork_b200.quantization.quant_primitives import ZeroPointDomain


def test_round_trip_int4_and_int8_configs():
    c = ao.Int4WeightOnlyConfig(group_size=64, layout=TensorCoreTiledLayout(inner_k_tiles=4), use_hqq=True,
                                zero_point_domain=ZeroPointDomain.FLOAT)
    d = config_to_dict(c)
    assert d["_type"] == "Int4WeightOnlyConfig" and d["_version"] == 1 and "_module" not in d
    assert "version" not in d["_data"]
    assert d["_data"]["layout"] == {"_type": "TensorCoreTiledLayout", "_data": {"inner_k_tiles": 4}}
    assert d["_data"]["zero_point_domain"] == {"_type": "ZeroPointDomain", "_data": "FLOAT"}
    json.dumps(d)  # plain JSON
    assert config_from_dict(d) == c
    c8 = ao.Int8WeightOnlyConfig(group_size=None)
    assert config_from_dict(config_to_dict(c8)) == c8


def test_decodes_a_dictionary_produced_by_the_reference():
    # literally what the reference's config_to_dict(Int4WeightOnlyConfig(group_size=32)) emits (core/config.py:80-96):
    # class NAME only, instance version, enum / layout / dtype sub-objects
    ref = {"_type": "Int4WeightOnlyConfig", "_version": 1,
           "_data": {"group_size": 32, "layout": {"_type": "TensorCoreTiledLayout", "_data": {"inner_k_tiles": 8}},
                     "use_hqq": False, "zero_point_domain": {"_type": "ZeroPointDomain", "_data": "NONE"},
                     "set_inductor_config": True, "preserve_zero": None,
                     "int4_packing_format": {"_type": "Int4PackingFormat", "_data": "PLAIN"},
                     "int4_choose_qparams_algorithm": {"_type": "Int4ChooseQParamsAlgorithm", "_data": "TINYGEMM"}}}
    known = {f for f in ao.Int4WeightOnlyConfig.__dataclass_fields__}
    ref["_data"] = {k: v for k, v in ref["_data"].items() if k in known}   # fields this build has
    c = config_from_dict(ref)
    assert isinstance(c, ao.Int4WeightOnlyConfig) and c.group_size == 32 and c.layout == TensorCoreTiledLayout(8)
    assert c.zero_point_domain == ZeroPointDomain.NONE


def test_dtype_and_version_mismatch():
    assert config_from_dict({"_type": "torch.dtype", "_data": "bfloat16"}) is torch.bfloat16
    with pytest.raises(ValueError):
        config_from_dict({"_type": "torch.dtype", "_data": "manual_seed"})
    d = config_to_dict(ao.Int8WeightOnlyConfig())
    d["_version"] = 7
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        try:
            config_from_dict(d)
        except ValueError:
            pass
        assert any("version" in str(x.message) for x in w)


@pytest.mark.parametrize("evil", [
    {"_type": "Popen", "_module": "subprocess", "_data": {"args": ["true"]}},
    {"_type": "Popen", "_data": {"args": ["true"]}},
    {"_type": "subprocess.Popen", "_data": {"args": ["true"]}},
    {"_type": "quantize_", "_data": {}},                 # a function of an allowed module, not a config class
    {"_type": "AffineQuantizedTensor", "_data": {}},     # a class of an allowed module, not a config / enum / layout
    {"_type": "__import__", "_data": {"name": "os"}},
])
def test_foreign_types_are_rejected(evil):
    with pytest.raises(ValueError):
        config_from_dict(evil)


def test_rejecting_does_not_run_anything():
    # the advisor's proof of concept: this dictionary used to spawn a process
    code = ("import sys; sys.path.insert(0, %r)\n"
            "from torchao_fork_b200.core.config import config_from_dict\n"
            "try:\n"
            "    config_from_dict({'_type': 'Popen', '_module': 'subprocess', '_data': {'args': ['sh', '-c', 'echo PWNED']}})\n"
            "except ValueError as e:\n"
            "    print('rejected')\n") % (str(__import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__)))),)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert "rejected" in out.stdout and "PWNED" not in out.stdout + out.stderr


def test_fp8_activation_int4_config_exists_and_says_why_it_is_not_applied():
    """quant_api.py:1159-1197 is not a weight-only path: the name imports, applying it raises with the reason."""
    import pytest
    import torch
    from torchao_fork_b200.quantization import Float8DynamicActivationInt4WeightConfig, quantize_
    cfg = Float8DynamicActivationInt4WeightConfig()
    assert str(getattr(cfg.packing_format, "value", cfg.packing_format)) == "preshuffled"
    with pytest.raises(NotImplementedError, match="weight-only"):
        quantize_(torch.nn.Sequential(torch.nn.Linear(128, 128, bias=False)), cfg)
