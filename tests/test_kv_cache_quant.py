"""AffineQuantizedKVCache / _quantize_activation_per_token_absmax on the CPU against golden vectors produced by the
reference (tests/golden/make_golden_kvq.py: torchao/_models/llama/model.py:198-240, torchao/quantization/utils.py:152-182).
The eager module is the prefill side of the int8 KV cache; its CUDA evaluation and the decode kernel are pinned by
tests/test_reference_on_device.py against the reference running on the GPU."""
import glob
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from torchao_fork_b200._models.llama.model import AffineQuantizedKVCache, KVCache
from torchao_fork_b200.quantization.utils import _quantize_activation_per_token_absmax

KVQ = sorted(glob.glob(os.path.join(GOLDEN, "kvq_*.npz")))


def bf16(b):
    return torch.from_numpy(b.astype(np.int16)).view(torch.bfloat16)


def bits(t):
    return t.contiguous().view(torch.int16).numpy().view(np.uint16)


@pytest.mark.parametrize("path", KVQ, ids=os.path.basename)
def test_int8_kv_cache_matches_reference_golden(path):
    g = np.load(path)
    k, v = bf16(g["k"]), bf16(g["v"])
    B, H, T, D = k.shape
    cache = AffineQuantizedKVCache.from_float(KVCache(B, T, H, D, torch.bfloat16))
    assert cache.k_cache.dtype == torch.int8 and cache.k_cache_scale.shape == (B, H, T, 1)
    with torch.no_grad():
        k_out, v_out = cache.update(torch.arange(5), k[:, :, :5], v[:, :, :5])
        for t in range(5, T):
            k_out, v_out = cache.update(torch.tensor([t]), k[:, :, t : t + 1], v[:, :, t : t + 1])
    assert np.array_equal(cache.k_cache.numpy(), g["k_cache"]) and np.array_equal(cache.v_cache.numpy(), g["v_cache"])
    assert np.array_equal(bits(cache.k_cache_scale), g["k_cache_scale"]) and np.array_equal(bits(cache.v_cache_scale), g["v_cache_scale"])
    assert np.array_equal(bits(k_out), g["k_out"]) and np.array_equal(bits(v_out), g["v_out"])
    q, s = _quantize_activation_per_token_absmax(k)
    assert q.dtype == torch.int8 and s.dtype == torch.bfloat16 and s.shape == (B, H, T)
    assert np.array_equal(q.numpy(), g["q_all"]) and np.array_equal(bits(s), g["s_all"])
    assert int(q.min()) >= -127          # -128 is never produced (quant_min = -127)


def test_setup_caches_switches_cache_type():
    from torchao_fork_b200._models.llama import ModelArgs, Transformer
    from torchao_fork_b200._models.llama.model import transformer_configs
    m = Transformer(ModelArgs(**{**transformer_configs["tiny-llama3"], "n_layer": 1}))
    m.setup_caches(1, 16, device="cpu")
    assert isinstance(m.layers[0].attention.kv_cache, KVCache)
    m.setup_caches(1, 16, device="cpu", kv_cache_quantization=True)     # same sizes, other cache type: rebuilt
    assert isinstance(m.layers[0].attention.kv_cache, AffineQuantizedKVCache)
    m.setup_caches(1, 8, device="cpu", kv_cache_quantization=True)      # smaller request: kept
    assert m.layers[0].attention.kv_cache.k_cache.shape[2] == 16


@pytest.mark.parametrize("path", KVQ, ids=os.path.basename)
def test_oracle_kv_quantizer_matches_reference_golden(path):
    """The numpy oracle's restatement (oracle/ao_oracle.py, ieee division = the CPU reference) against the same vectors."""
    import ao_oracle as o
    g = np.load(path)
    k, v = o.from_bits(g["k"]), o.from_bits(g["v"])
    B, H, T, D = k.shape
    q, s = o.quantize_activation_per_token_absmax(k)
    assert np.array_equal(q, g["q_all"]) and np.array_equal(o.to_bits(s), g["s_all"])
    kc, vc = np.zeros((B, H, T, D), np.int8), np.zeros((B, H, T, D), np.int8)
    ks, vs = np.ones((B, H, T, 1), np.float32), np.ones((B, H, T, 1), np.float32)
    k_out = o.kv_cache_update(kc, ks, np.arange(5), k[:, :, :5])
    v_out = o.kv_cache_update(vc, vs, np.arange(5), v[:, :, :5])
    for t in range(5, T):
        k_out = o.kv_cache_update(kc, ks, np.array([t]), k[:, :, t : t + 1])
        v_out = o.kv_cache_update(vc, vs, np.array([t]), v[:, :, t : t + 1])
    assert np.array_equal(kc, g["k_cache"]) and np.array_equal(vc, g["v_cache"])
    assert np.array_equal(o.to_bits(ks), g["k_cache_scale"]) and np.array_equal(o.to_bits(vs), g["v_cache_scale"])
    assert np.array_equal(o.to_bits(k_out), g["k_out"]) and np.array_equal(o.to_bits(v_out), g["v_out"])
