"""Generate tests/golden/kvq_*.npz by running the REFERENCE's AffineQuantizedKVCache (torchao/_models/llama/model.py:198-240)
and _quantize_activation_per_token_absmax (torchao/quantization/utils.py:152-182) on CPU.

Run in the build container only (needs /root/reference):

    PYTHONPATH=/root/reference python tests/golden/make_golden_kvq.py

bf16 tensors are stored as their uint16 bit patterns.  (On CUDA the same expressions round differently — tensor / scalar is
a multiplication by the reciprocal there —; that mode is pinned by tests/test_reference_on_device.py against the
reference running on the GPU.)
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from torchao._models.llama.model import AffineQuantizedKVCache  # noqa: E402
from torchao.quantization.utils import _quantize_activation_per_token_absmax  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def bits(t):
    assert t.dtype == torch.bfloat16
    return t.contiguous().view(torch.int16).numpy().view(np.uint16).copy()


def rows(B, H, T, D, seed):
    g = torch.Generator().manual_seed(seed)
    k = torch.randn(B, H, T, D, generator=g)
    v = torch.randn(B, H, T, D, generator=g) * 3.0
    k[0, 0, 1] = 0.0
    k[0, 0, 2] *= 1e4
    v[0, 0, 3] *= 1e-6
    k[0, 1 % H, 4] = 0.5
    return k.to(torch.bfloat16), v.to(torch.bfloat16)


for name, (B, H, T, D, seed) in {"small_1x2x12x128": (1, 2, 12, 128, 5), "batch_2x1x9x64": (2, 1, 9, 64, 6)}.items():
    k, v = rows(B, H, T, D, seed)
    cache = AffineQuantizedKVCache(B, T, H, D, scale_dtype=torch.bfloat16)
    with torch.no_grad():
        # prefill of the first 5 positions at once, then one position at a time
        k_out, v_out = cache.update(torch.arange(5), k[:, :, :5], v[:, :, :5])
        for t in range(5, T):
            k_out, v_out = cache.update(torch.tensor([t]), k[:, :, t : t + 1], v[:, :, t : t + 1])
        q, s = _quantize_activation_per_token_absmax(k)
    np.savez(os.path.join(OUT, f"kvq_{name}.npz"), k=bits(k), v=bits(v), k_cache=cache.k_cache.numpy(), v_cache=cache.v_cache.numpy(),
             k_cache_scale=bits(cache.k_cache_scale), v_cache_scale=bits(cache.v_cache_scale), k_out=bits(k_out), v_out=bits(v_out),
             q_all=q.numpy(), s_all=bits(s))
    print(name, "ok")
