"""int8 weight-only on the strip-major layout: pack / unpack against the oracle's restatement, the TMA-ring GEMV
(csrc/gemv_int8_fused.cu) with its prologue / epilogues, the fallback GEMV and the tcgen05 GEMM (M > 16), all against
the reference formula  y = bf16(bf16(x @ q^T) * scale)  (plain_layout.py:250-266)."""
import numpy as np
import pytest
import torch

import ao_oracle as o
import torchao_fork_b200 as ao
from torchao_fork_b200 import ops as aops

pytestmark = pytest.mark.gpu
T = torch.ops.torchao
DEV = "cuda"
EPS = torch.finfo(torch.float32).eps


def sqnr(ref, x):
    ref, x = ref.double(), x.double()
    return (10 * torch.log10(ref.pow(2).mean() / (ref - x).pow(2).mean().clamp_min(1e-300))).item()


def _quant(n, k, seed):
    torch.manual_seed(seed)
    w = torch.randn(n, k, dtype=torch.bfloat16, device=DEV)
    q, s = T.b200_int8_quantize(w, EPS, True)
    return q, s, T.b200_int8_pack(q)


def _ref_mm(x, q, s):
    return torch.mm(x, q.t().to(torch.bfloat16)) * s


@pytest.mark.parametrize("shape", [(16, 64), (48, 192), (1000, 4096 + 16), (130, 200)], ids=str)
def test_pack_matches_oracle_and_round_trips(shape):
    n, k = shape
    torch.manual_seed(n)
    q = torch.randint(-128, 128, (n, k), dtype=torch.int8, device=DEV)
    packed = T.b200_int8_pack(q)
    np_, kp = aops.padded_int8_shape(n, k)
    assert tuple(packed.shape) == (np_ // 16, kp // 64, 1024)
    qpad = np.zeros((np_, kp), dtype=np.int8)
    qpad[:n, :k] = q.cpu().numpy()
    assert np.array_equal(packed.cpu().numpy(), o.pack_int8_b200(qpad))
    assert torch.equal(T.b200_int8_unpack(packed, n, k), q)


SHAPES = [(4096, 4096), (1280, 8192), (512, 3584), (4096, 14336), (384, 1024), (1000, 4096 + 16), (4096, 512), (4096, 1792)]


@pytest.mark.parametrize("shape", SHAPES, ids=str)
@pytest.mark.parametrize("m", [1, 2, 5, 8, 13, 16])
def test_mm_small_m(shape, m):
    """M <= 16 through ao_b200_int8_mm: the fused kernel where the activations fit, the fallback GEMV otherwise."""
    n, k = shape
    q, s, packed = _quant(n, k, n + k + m)
    x = torch.randn(m, k, dtype=torch.bfloat16, device=DEV)
    y = T.b200_int8_weight_only_mm(x, packed, s)
    ref = _ref_mm(x, q, s)
    assert y.shape == (m, n)
    assert sqnr(ref, y) >= 45.0
    assert ((ref.float() - y.float()).abs().mean() / ref.float().abs().mean()).item() < 1e-2
    # exact-integer activations make the fp32 accumulation order-independent -> bit-exact
    xi = torch.randint(-4, 5, (m, k), device=DEV).to(torch.bfloat16)
    assert torch.equal(T.b200_int8_weight_only_mm(xi, packed, s), _ref_mm(xi, q, s))


@pytest.mark.parametrize("shape", [(4096, 4096), (6144, 4096), (512, 3584), (1000, 4096 + 16), (4096, 14336)], ids=str)
@pytest.mark.parametrize("m", [17, 64, 200, 1024])
def test_mm_prefill_gemm(shape, m):
    n, k = shape
    q, s, packed = _quant(n, k, n + m)
    x = torch.randn(m, k, dtype=torch.bfloat16, device=DEV)
    y = T.b200_int8_weight_only_mm(x, packed, s)
    ref = _ref_mm(x, q, s)
    assert sqnr(ref, y) >= 45.0
    assert ((ref.float() - y.float()).abs().mean() / ref.float().abs().mean()).item() < 1e-2
    xi = torch.randint(-4, 5, (m, k), device=DEV).to(torch.bfloat16)
    assert torch.equal(T.b200_int8_weight_only_mm(xi, packed, s), _ref_mm(xi, q, s))


@pytest.mark.parametrize("m", [1, 3, 8, 16])
@pytest.mark.parametrize("shape", [(6144, 4096), (1024, 8192), (2048, 1024)], ids=str)
def test_rmsnorm_prologue_and_residual(shape, m):
    n, k = shape
    if m * (2 * k + 16) > 150 * 1024:
        pytest.skip("activations do not fit the fused kernel's staging (ao_b200_int8_mm falls back to the plain GEMV)")
    q, s, packed = _quant(n, k, 3 * n + m)
    x = (torch.randn(m, k, device=DEV) * 3).to(torch.bfloat16)
    nw = (1 + 0.1 * torch.randn(k, device=DEV)).to(torch.bfloat16)
    eps = 1e-5
    y = T.b200_int8_gemv_fused(x, packed, s, nw, eps, None, aops.EPILOGUE_PLAIN)
    xf = x.float()
    xn = (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).to(torch.bfloat16) * nw   # RMSNorm.forward, model.py:489-503
    ref = _ref_mm(xn, q, s)
    assert sqnr(ref, y) >= 38.0   # the fused norm applies rsqrt after the GEMV (one rounding of x * w instead of two)
    res = torch.randn(m, n, dtype=torch.bfloat16, device=DEV)
    y2 = T.b200_int8_gemv_fused(x, packed, s, None, 0.0, res, aops.EPILOGUE_RESIDUAL)
    ref2 = res + _ref_mm(x, q, s)
    assert sqnr(ref2, y2) >= 45.0


@pytest.mark.parametrize("m", [1, 4, 9])
def test_silu_mul_epilogue(m):
    i, k = 1024, 2048
    q1, s1, _ = _quant(i, k, 5)
    q3, s3, _ = _quant(i, k, 6)
    qf, sf = aops.interleave_gate_up_int8(q1, s1, q3, s3)
    packed = T.b200_int8_pack(qf)
    x = torch.randn(m, k, dtype=torch.bfloat16, device=DEV)
    y = T.b200_int8_gemv_fused(x, packed, sf, None, 0.0, None, aops.EPILOGUE_SILU_MUL)
    ref = torch.nn.functional.silu(_ref_mm(x, q1, s1)) * _ref_mm(x, q3, s3)   # model.py:486
    assert y.shape == (m, i)
    assert sqnr(ref, y) >= 40.0


def test_quantize_api_int8_impl_fields_and_linear():
    torch.manual_seed(0)
    lin = torch.nn.Linear(1024, 520, bias=True, device=DEV, dtype=torch.bfloat16)
    w = lin.weight.detach().clone()
    ao.quantize_(lin, ao.Int8WeightOnlyConfig())
    impl = lin.weight.tensor_impl
    q, s = T.b200_int8_quantize(w, EPS, True)
    assert impl.dtype == torch.int8 and tuple(impl.shape) == (520, 1024)
    assert torch.equal(impl.int_data, q) and torch.equal(impl.scale, s)
    int_data, scale, zp = impl.get_plain()
    assert int_data.dtype == torch.int8 and zp.dtype == torch.int64 and int(zp.abs().sum()) == 0
    for m in (1, 40):
        x = torch.randn(m, 1024, dtype=torch.bfloat16, device=DEV)
        ref = _ref_mm(x, q, s) + lin.bias
        assert sqnr(ref, lin(x)) >= 45.0
    sl = lin.weight[:256]                      # rows slice data + qparams (plain_layout.py:160-200)
    assert torch.equal(sl.tensor_impl.int_data, q[:256]) and torch.equal(sl.tensor_impl.scale, s[:256])
    assert torch.equal(lin.weight.dequantize(), (q.to(torch.bfloat16) * s.unsqueeze(-1)))
