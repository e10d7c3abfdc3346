"""Generate tests/golden/*.npz by running the REFERENCE's own Python functions on CPU.

Run in the build container only (needs /root/reference, which does not exist on
the GPU box):

    PYTHONPATH=/root/reference python tests/golden/make_golden.py

Nothing under tests/ imports this at test time; the committed .npz files are the
fixtures.  bf16 tensors are stored as their uint16 bit patterns.
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
import torchao  # noqa: E402
from torchao.dtypes import Int4CPULayout  # noqa: E402
from torchao.quantization import Int4WeightOnlyConfig, Int8WeightOnlyConfig, quantize_  # noqa: E402
from torchao.quantization.quant_primitives import (  # noqa: E402
    MappingType,
    _choose_qparams_affine_tinygemm,
    _dequantize_affine_tinygemm,
    _quantize_affine_tinygemm,
    choose_qparams_affine,
    dequantize_affine,
    quantize_affine,
)
from torchao.quantization.utils import pack_tinygemm_scales_and_zeros  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def bits(t):
    assert t.dtype == torch.bfloat16
    return t.contiguous().view(torch.int16).numpy().view(np.uint16).copy()


def make_weight(kind, n, k, seed):
    g = torch.Generator().manual_seed(seed)
    if kind == "randn":
        w = torch.randn(n, k, generator=g)
    elif kind == "small":          # nn.Linear-like magnitudes
        w = (torch.rand(n, k, generator=g) * 2 - 1) / (k ** 0.5)
    elif kind == "edge":
        # constant groups (scale clamps to eps), zeros, huge/small magnitudes, ties
        w = torch.randn(n, k, generator=g)
        w[0, :] = 0.0
        w[1, :] = 3.0
        w[2, : k // 2] = -1e-8
        w[3, :] = torch.linspace(-1, 1, k)
        w[4, :] *= 1e4
        w[5, :] *= 1e-5
        w[6, ::2] = 0.5
        w[6, 1::2] = -0.5
        w[7, :] = torch.arange(k) % 16
    else:
        raise ValueError(kind)
    return w.to(torch.bfloat16)


def int4_case(name, kind, n, k, gs, seed, m=3):
    w = make_weight(kind, n, k, seed)
    s, z = _choose_qparams_affine_tinygemm(
        w, MappingType.ASYMMETRIC, (1, gs), torch.int32, 0, 15, 1e-6, torch.bfloat16, torch.bfloat16
    )
    q = _quantize_affine_tinygemm(w, (1, gs), s, z, torch.int32, 0, 15)
    sz = pack_tinygemm_scales_and_zeros(s.reshape(n, -1), z.reshape(n, -1), torch.bfloat16)
    dq2 = _dequantize_affine_tinygemm(q, (1, gs), s, z, torch.int32, 0, 15, output_dtype=torch.bfloat16)

    # reference CPU linear: quantize_ with Int4CPULayout -> aten._weight_int4pack_mm_for_cpu
    lin = torch.nn.Linear(k, n, bias=False).to(torch.bfloat16)
    lin.weight.data.copy_(w)
    quantize_(lin, Int4WeightOnlyConfig(group_size=gs, layout=Int4CPULayout(), set_inductor_config=False))
    g = torch.Generator().manual_seed(seed + 1000)
    x = torch.randn(m, k, generator=g).to(torch.bfloat16)
    y_cpu = lin(x)
    # the AQT built by the CPU path carries the same qparams (zero_point_dtype = weight dtype)
    sz_cpu = lin.weight.tensor_impl.scale_and_zero
    assert torch.equal(sz_cpu, sz)
    eye = torch.eye(k, dtype=torch.bfloat16)
    dq_cpu_mm = lin(eye).t().contiguous()       # fp32-dequant of the CPU kernel, rounded once to bf16

    np.savez_compressed(
        os.path.join(OUT, f"int4_{name}.npz"),
        group_size=np.int32(gs),
        w=bits(w), scale=bits(s.reshape(n, -1)), zero=bits(z.reshape(n, -1)),
        q=q.numpy().astype(np.uint8), scale_and_zero=bits(sz), dequant_2round=bits(dq2),
        x=bits(x), y_cpu=bits(y_cpu), dequant_cpu_mm=bits(dq_cpu_mm),
    )
    print("int4", name, w.shape, gs)


def int8_case(name, kind, n, k, seed, m=3):
    w = make_weight(kind, n, k, seed)
    eps = torch.finfo(torch.float32).eps
    s, zp = choose_qparams_affine(
        w, MappingType.SYMMETRIC, (1, k), torch.int8, eps=eps, zero_point_dtype=torch.int64
    )
    q = quantize_affine(w, (1, k), s, zp, torch.int8)
    dq = dequantize_affine(q, (1, k), s, zp, torch.int8, output_dtype=torch.bfloat16)
    lin = torch.nn.Linear(k, n, bias=False).to(torch.bfloat16)
    lin.weight.data.copy_(w)
    quantize_(lin, Int8WeightOnlyConfig(set_inductor_config=False))
    assert torch.equal(lin.weight.tensor_impl.int_data, q)
    assert torch.equal(lin.weight.tensor_impl.scale.reshape(-1), s.reshape(-1))
    assert int(lin.weight.tensor_impl.zero_point.abs().max()) == 0
    g = torch.Generator().manual_seed(seed + 1000)
    x = torch.randn(m, k, generator=g).to(torch.bfloat16)
    y_cpu = lin(x)
    np.savez_compressed(
        os.path.join(OUT, f"int8_{name}.npz"),
        w=bits(w), scale=bits(s.reshape(-1)), q=q.numpy(), dequant=bits(dq),
        x=bits(x), y_cpu=bits(y_cpu),
    )
    print("int8", name, w.shape)


def padded_case():
    """Linear(256 -> 321): TensorCoreTiledLayout.pre_process pads K->1024, N->328 with zeros
    BEFORE qparams (tensor_core_tiled_layout.py:127-135; test_affine_quantized.py test_slice_int4wo)."""
    from torchao.dtypes import TensorCoreTiledLayout
    w = make_weight("randn", 321, 256, 7)
    wp = TensorCoreTiledLayout().pre_process(w)
    gs = 64
    n, k = wp.shape
    s, z = _choose_qparams_affine_tinygemm(
        wp, MappingType.ASYMMETRIC, (1, gs), torch.int32, 0, 15, 1e-6, torch.bfloat16, torch.bfloat16
    )
    q = _quantize_affine_tinygemm(wp, (1, gs), s, z, torch.int32, 0, 15)
    np.savez_compressed(
        os.path.join(OUT, "int4_padded_321x256.npz"),
        group_size=np.int32(gs), w=bits(w), padded_shape=np.array(wp.shape),
        scale=bits(s.reshape(n, -1)), zero=bits(z.reshape(n, -1)), q=q.numpy().astype(np.uint8),
    )
    print("padded", wp.shape)


if __name__ == "__main__":
    torch.set_num_threads(4)
    int4_case("randn_16x64_g32", "randn", 16, 64, 32, 1)
    int4_case("randn_32x256_g64", "randn", 32, 256, 64, 2)
    int4_case("small_48x1024_g128", "small", 48, 1024, 128, 3)
    int4_case("randn_64x512_g256", "randn", 64, 512, 256, 4)
    int4_case("edge_16x256_g128", "edge", 16, 256, 128, 5)
    int4_case("edge_16x256_g32", "edge", 16, 256, 32, 6)
    int8_case("randn_32x256", "randn", 32, 256, 11)
    int8_case("small_48x1024", "small", 48, 1024, 12)
    int8_case("edge_16x256", "edge", 16, 256, 13)
    padded_case()
    print("torch", torch.__version__, "torchao", torchao.__version__)
