"""Generate tests/golden/hqq_*.npz by running the REFERENCE's HQQ quantizer on CPU (fp32 optimizer path):
torchao/quantization/quant_primitives.py:1998-2111 (_choose_qparams_and_quantize_affine_hqq) as called by
AffineQuantizedTensor.from_hp_to_intx(use_hqq=True) (affine_quantized_tensor.py:252-288) for Int4WeightOnlyConfig.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden_hqq.py

Nothing under tests/ imports this at test time; the committed .npz files are the fixtures.  bf16 tensors are stored as
their uint16 bit patterns."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from torchao.quantization.quant_primitives import _choose_qparams_and_quantize_affine_hqq  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def bits(t):
    assert t.dtype == torch.bfloat16
    return t.contiguous().view(torch.int16).numpy().view(np.uint16).copy()


CASES = [("randn", 32, 256, 64, 1), ("small", 48, 1024, 128, 2), ("randn", 16, 64, 32, 3), ("heavy", 24, 512, 256, 4)]

for kind, n, k, gs, seed in CASES:
    g = torch.Generator().manual_seed(seed)
    if kind == "randn":
        w = torch.randn(n, k, generator=g)
    elif kind == "small":
        w = (torch.rand(n, k, generator=g) * 2 - 1) / (k ** 0.5)
    else:  # heavy tails: what the lp < 1 proximal step is for
        w = torch.randn(n, k, generator=g) * torch.exp(torch.randn(n, k, generator=g))
    w = w.to(torch.bfloat16)
    q, scale, zero, shape = _choose_qparams_and_quantize_affine_hqq(
        w, nbits=4, group_size=gs, axis=1, compute_dtype=torch.bfloat16, device="cpu", raw_output=False)
    assert tuple(shape) == (n, k) and q.dtype == torch.uint8
    np.savez_compressed(os.path.join(OUT, f"hqq_{kind}_{n}x{k}_g{gs}.npz"), w=bits(w), q=q.reshape(n, k).numpy(),
                        scale=bits(scale.reshape(n, k // gs)), zero=bits(zero.reshape(n, k // gs)), group_size=gs)
    print(kind, n, k, gs, "ok")
