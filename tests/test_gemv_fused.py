"""Fused decode GEMV (csrc/gemv_int4_fused.cu): plain / RMSNorm prologue / residual / SiLU*mul epilogues against a
float64 restatement on the bit-exact dequantised weight, and against aten._weight_int4pack_mm."""
import numpy as np
import pytest
import torch

import ao_oracle as o
import torchao_fork_b200 as ao  # noqa: F401
from torchao_fork_b200 import ops as aops

pytestmark = pytest.mark.gpu
T = torch.ops.torchao
DEV = "cuda"


def sqnr(ref, x):
    ref, x = ref.double(), x.double()
    return (10 * torch.log10(ref.pow(2).mean() / (ref - x).pow(2).mean().clamp_min(1e-300))).item()


def _quant(n, k, gs, seed):
    torch.manual_seed(seed)
    w = torch.randn(n, k, dtype=torch.bfloat16, device=DEV)
    packed, sz = T.b200_int4_quantize_pack(w, gs, 1e-6, True)
    # exact affine weight in float64: (q - 8) * s + z  (no bf16 rounding of W': the fp32-dequant semantics)
    q = T.b200_int4_unpack(packed)[:n, :k].double()
    s_, z_ = aops.unpack_scales_and_zeros_b200(sz)
    s = s_[:n].double().repeat_interleave(gs, dim=1)[:, :k]
    z = z_[:n].double().repeat_interleave(gs, dim=1)[:, :k]
    return packed, sz, (q - 8.0) * s + z


# the last two are the K of the row-parallel wo / w2 shards of Llama-3-8B at TP = 8 (one partial stage, idle consumer warps)
SHAPES = [(4096, 4096), (1280, 8192), (512, 3584), (4096, 14336), (384, 1024), (4096, 512), (4096, 1792)]


@pytest.mark.parametrize("shape", SHAPES, ids=str)
@pytest.mark.parametrize("gs", [32, 64, 128, 256])
@pytest.mark.parametrize("m", [1, 2, 5, 8, 13, 16])
def test_plain_matches_float64(shape, gs, m):
    n, k = shape
    if m * (2 * k + 16) > 190 * 1024:
        pytest.skip("activation does not fit the fused kernel's staging (ao_b200_int4_mm falls back)")
    packed, sz, wd = _quant(n, k, gs, n + k + gs + m)
    x = torch.randn(m, k, dtype=torch.bfloat16, device=DEV)
    try:
        y = T.b200_int4_gemv_fused(x, packed, gs, sz, n, None, 0.0, None, aops.EPILOGUE_PLAIN)
    except RuntimeError as e:
        if "workspace too small" in str(e):   # staging + >= 3 ring stages must fit; ao_b200_int4_mm falls back then
            assert m * (2 * k + 16) > 64 * 1024
            y = T.b200_int4_weight_only_mm(x, packed, gs, sz, n)
        else:
            raise
    ref = x.double() @ wd.t()
    assert y.shape == (m, n)
    assert sqnr(ref, y) >= 45.0  # bf16 output rounding alone is ~ 50 dB


@pytest.mark.parametrize("m", [1, 4, 16])
def test_dispatch_through_int4_mm_matches_aten(m):
    n, k, gs = 6144, 4096, 128
    packed, sz, _ = _quant(n, k, gs, 11)
    x = torch.randn(m, k, dtype=torch.bfloat16, device=DEV)
    y = T.b200_int4_weight_only_mm(x, packed, gs, sz, n)
    q = T.b200_int4_unpack(packed)
    u8 = (q[:, ::2] << 4 | q[:, 1::2]).to(torch.uint8)
    y_ref = torch.ops.aten._weight_int4pack_mm(x, torch.ops.aten._convert_weight_to_int4pack(u8, 8), gs,
                                               aops.sz_b200_to_tinygemm(sz))
    assert sqnr(y_ref, y) >= 40.0
    assert ((y_ref.float() - y.float()).abs().mean() / y_ref.float().abs().mean()).item() < 1e-2


@pytest.mark.parametrize("m", [1, 3, 8])
@pytest.mark.parametrize("shape,gs", [((6144, 4096), 128), ((1024, 8192), 64), ((2048, 1024), 32)])
def test_rmsnorm_prologue(shape, gs, m):
    n, k = shape
    packed, sz, wd = _quant(n, k, gs, 3 * n + m)
    x = (torch.randn(m, k, device=DEV) * 3).to(torch.bfloat16)
    nw = (1 + 0.1 * torch.randn(k, device=DEV)).to(torch.bfloat16)
    eps = 1e-5
    y = T.b200_int4_gemv_fused(x, packed, gs, sz, n, nw, eps, None, aops.EPILOGUE_PLAIN)
    xd = x.double()
    a = xd * torch.rsqrt(xd.pow(2).mean(-1, keepdim=True) + eps) * nw.double()
    assert sqnr(a @ wd.t(), y) >= 40.0


@pytest.mark.parametrize("m", [1, 6, 16])
def test_residual_epilogue_in_place(m):
    n, k, gs = 4096, 4096, 128
    packed, sz, wd = _quant(n, k, gs, 77 + m)
    x = torch.randn(m, k, dtype=torch.bfloat16, device=DEV)
    h = torch.randn(m, n, dtype=torch.bfloat16, device=DEV)
    y = T.b200_int4_gemv_fused(x, packed, gs, sz, n, None, 0.0, h, aops.EPILOGUE_RESIDUAL)
    ref = h.double() + (x.double() @ wd.t())
    assert sqnr(ref, y) >= 45.0
    # the C ABI allows y to alias the residual (the decode engine updates the residual stream in place)
    from torchao_fork_b200 import _lib
    h2 = h.clone()
    L = _lib.lib()
    ws = torch.zeros(max(16, int(L.ao_b200_int4_mm_workspace_bytes(m, n, k))), dtype=torch.uint8, device=DEV)
    _lib.check(L.ao_b200_int4_gemv_fused(x.data_ptr(), m, k, k, packed.data_ptr(), sz.data_ptr(), n, k, gs,
                                         h2.data_ptr(), n, n, None, 0.0, h2.data_ptr(), n, 1, ws.data_ptr(), ws.numel(),
                                         torch.cuda.current_stream().cuda_stream), "fused")
    assert torch.equal(h2, y)


@pytest.mark.parametrize("m", [1, 7])
@pytest.mark.parametrize("inter,k,gs", [(14336, 4096, 128), (1792, 1024, 64)])
def test_silu_mul_epilogue_with_interleaved_gate_up(inter, k, gs, m):
    p1, sz1, w1 = _quant(inter, k, gs, 5)
    p3, sz3, w3 = _quant(inter, k, gs, 6)
    pf, szf = aops.interleave_gate_up(p1, sz1, p3, sz3)
    # the interleave is a pure permutation
    qf = T.b200_int4_unpack(pf)
    q1, q3 = T.b200_int4_unpack(p1), T.b200_int4_unpack(p3)
    exp = torch.stack([q1.reshape(-1, 8, k), q3.reshape(-1, 8, k)], 1).reshape(-1, k)
    assert torch.equal(qf, exp)
    x = torch.randn(m, k, dtype=torch.bfloat16, device=DEV)
    nw = (1 + 0.1 * torch.randn(k, device=DEV)).to(torch.bfloat16)
    y = T.b200_int4_gemv_fused(x, pf, gs, szf, 2 * inter, nw, 1e-5, None, aops.EPILOGUE_SILU_MUL)
    assert y.shape == (m, inter)
    xd = x.double()
    a = xd * torch.rsqrt(xd.pow(2).mean(-1, keepdim=True) + 1e-5) * nw.double()
    gate, up = a @ w1.t(), a @ w3.t()
    ref = torch.nn.functional.silu(gate) * up
    assert sqnr(ref, y) >= 38.0


def test_identity_rows_reproduce_the_bf16_dequant():
    # x = e_k: one non-zero term, so the fp32 group affine equals the single-rounded bf16 fma of the CUDA reference op
    n, k, gs = 1024, 512, 128
    packed, sz, _ = _quant(n, k, gs, 9)
    dq = T.b200_int4_dequantize(packed, sz, gs)
    eye = torch.eye(k, dtype=torch.bfloat16, device=DEV)
    for r0 in (0, 16 * 7, k - 16):
        y = T.b200_int4_gemv_fused(eye[r0:r0 + 16].contiguous(), packed, gs, sz, n, None, 0.0, None, 0)
        assert torch.equal(y, dq[:, r0:r0 + 16].t())


def test_shape_errors_are_loud():
    packed, sz, _ = _quant(256, 256, 128, 1)
    x = torch.randn(17, 256, dtype=torch.bfloat16, device=DEV)
    with pytest.raises(RuntimeError):
        T.b200_int4_gemv_fused(x, packed, 128, sz, 256, None, 0.0, None, 0)
    with pytest.raises(RuntimeError):
        T.b200_int4_gemv_fused(x[:2], packed, 128, sz, 256, None, 0.0, None, 1)  # residual epilogue without residual
