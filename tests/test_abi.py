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


def test_persistent_kernel_size_queries_without_a_gpu():
    """ao_b200_decode_mega_{plan,xchg,state}_bytes are host arithmetic: the exchange region grows by one slot per rank plus
    the rank-local sum slot under tensor parallelism, and shapes the kernel does not cover are refused (0 bytes)."""
    if not os.path.exists(_lib.LIB_PATH):
        pytest.skip("libao_b200.so not built")
    L = _lib.lib()
    keep = []

    def model(world=1, **over):
        m = _lib.MegaModel()
        layers = (_lib.MegaLayer * 2)()
        keep.append(layers)
        m.n_layer, m.dim, m.n_head, m.n_kv_head, m.head_dim, m.inter, m.vocab = 2, 1024, 8, 2, 128, 2048, 1024
        m.group_size, m.max_seq, m.world, m.rank, m.n_cta = 128, 64, world, 0, 148
        m.layers = ctypes.cast(layers, ctypes.POINTER(_lib.MegaLayer))
        one = ctypes.c_void_p(16)                                  # any non-null pointer: nothing is dereferenced here
        for f in ("out_packed", "out_sz", "out_norm", "emb", "freqs", "tok", "pos", "logits", "state"):
            setattr(m, f, one)
        for r in range(world):
            m.xchg[r] = 16
        for k, v in over.items():
            setattr(m, k, v)
        return m

    assert L.ao_b200_decode_mega_state_bytes() == 64
    m1 = model()
    assert L.ao_b200_decode_mega_plan_bytes(ctypes.byref(m1)) == 2 * 112
    x1 = L.ao_b200_decode_mega_xchg_bytes(ctypes.byref(m1))
    assert x1 > 0 and x1 % 128 == 0
    x2 = L.ao_b200_decode_mega_xchg_bytes(ctypes.byref(model(world=2)))
    x8 = L.ao_b200_decode_mega_xchg_bytes(ctypes.byref(model(world=8)))
    slot = 1024 // 2 * 8                                           # one slot of dim / 2 packets of 8 bytes
    assert x2 - x1 == 2 * 2 * slot                                 # wo and w2 partial sums: 1 -> 2 + 1 slots each
    assert x8 - x1 == 2 * 8 * slot
    assert L.ao_b200_decode_mega_plan_bytes(ctypes.byref(model(world=3))) == 0      # world must be a power of two
    assert L.ao_b200_decode_mega_plan_bytes(ctypes.byref(model(head_dim=64))) == 0
    assert L.ao_b200_decode_mega_plan_bytes(ctypes.byref(model(group_size=48))) == 0
