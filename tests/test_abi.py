"""The drop-in boundary on a machine without a GPU: libao_b200.so loads, exports every symbol include/ao_b200.h
declares, the ctypes table in torchao_fork_b200/_lib.py covers exactly those symbols, and argument validation answers
with error codes (no CUDA call is made)."""
import ctypes
import os
import re

import pytest

from torchao_fork_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "ao_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"AO_B200_API\s+[\w\s\*]+?\b(ao_b200_\w+)\s*\(", src)))


def test_header_declares_the_expected_surface():
    names = declared_symbols()
    assert len(names) >= 35
    for must in ("ao_b200_int4_quantize_pack", "ao_b200_int4_mm", "ao_b200_int4_gemv_fused", "ao_b200_int8_quantize",
                 "ao_b200_int8_pack", "ao_b200_int8_unpack", "ao_b200_int8_mm", "ao_b200_int8_gemv_fused",
                 "ao_b200_tinygemm_unpack", "ao_b200_tinygemm_dequant", "ao_b200_decode_attention_fused", "ao_b200_sample",
                 "ao_b200_tp_allreduce_add"):
        assert must in names


def test_library_exports_every_declared_symbol():
    if not os.path.exists(_lib.LIB_PATH):
        pytest.skip("libao_b200.so not built (python -m torchao_fork_b200.build)")
    handle = ctypes.CDLL(_lib.LIB_PATH)
    missing = [n for n in declared_symbols() if not hasattr(handle, n)]
    assert not missing, f"declared in include/ao_b200.h but not exported: {missing}"


def test_ctypes_table_matches_the_header():
    assert sorted(_lib.SIGNATURES) == declared_symbols()


def test_version_and_argument_validation_without_a_gpu():
    if not os.path.exists(_lib.LIB_PATH):
        pytest.skip("libao_b200.so not built")
    L = _lib.lib()
    assert L.ao_b200_abi_version() == 5
    assert L.ao_b200_error_string(0) == b"ok"
    assert b"invalid" in L.ao_b200_error_string(1)
    # null pointers / bad shapes are rejected before any launch
    assert L.ao_b200_int4_mm(None, 1, 64, 64, None, None, 16, 64, 128, None, 16, 16, None, 0, None) == 1
    assert L.ao_b200_int8_mm(None, 1, 64, 64, None, None, 16, 64, None, 16, 16, None, 0, None) == 1
    assert L.ao_b200_int8_pack(None, 16, 64, 16, 64, None, None) == 1
    assert L.ao_b200_int4_quantize_pack(None, 16, 64, 64, 16, 64, 128, 1e-6, 1, None, None, None) == 1
    with pytest.raises(RuntimeError, match="invalid argument"):
        _lib.check(1, "probe")
