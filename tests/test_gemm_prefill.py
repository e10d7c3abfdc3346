"""Prefill path (M > 16): the tcgen05 / tensor-memory dequant-GEMM behind torch.ops.torchao.b200_int4_weight_only_mm,
against aten._weight_int4pack_mm — the kernel the reference dispatches to for the same call
(torchao/dtypes/uintx/tensor_core_tiled_layout.py:_linear_bf16_act_uint4_weight_impl).

Bars: identity activations bit-exact (the dequantised weights and the k order of both operands are exactly the
reference's); random activations SQNR >= 40 dB and mean-abs-rel < 1e-2 (test/test_ops.py:113-117's bf16 figure — the two
kernels sum the same bf16 x bf16 products in fp32, in a different order); an exact-arithmetic case (small integers:
every product and partial sum is representable) bit-exact against an fp64 matmul."""
import pytest
import torch

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available():
    pytest.skip("needs a CUDA device", allow_module_level=True)

import torchao_fork_b200 as ao  # noqa: E402
from torchao_fork_b200 import _lib  # noqa: E402
from torchao_fork_b200 import ops as aops  # noqa: E402
from torchao_fork_b200.quantization.utils import compute_error  # noqa: E402

T = torch.ops.torchao
DEV = "cuda"


def _aten_pack(q, ikt=8):
    u8 = (q[:, ::2] << 4 | q[:, 1::2]).to(torch.uint8)
    return torch.ops.aten._convert_weight_to_int4pack(u8.contiguous(), ikt)


def _aten_mm(x, q, sz_b200, gs, n):
    k = q.shape[1]
    kp = (k + 1023) // 1024 * 1024
    qpad = torch.nn.functional.pad(q, (0, kp - k))
    sz_ref = aops.sz_b200_to_tinygemm(sz_b200)
    szpad = torch.nn.functional.pad(sz_ref, (0, 0, 0, 0, 0, kp // gs - sz_ref.shape[0]))
    xpad = torch.nn.functional.pad(x, (0, kp - x.shape[1]))
    return torch.ops.aten._weight_int4pack_mm(xpad, _aten_pack(qpad), gs, szpad)[:, :n]


@pytest.mark.parametrize("m", [17, 64, 65, 100, 128, 129, 200, 256, 300, 777, 1024])
@pytest.mark.parametrize("shape,gs", [((4096, 4096), 128), ((6144, 4096), 32), ((1024, 14336), 64), ((512, 2048), 256)])
def test_prefill_matches_aten(m, shape, gs):
    n, k = shape
    torch.manual_seed(m * 31 + n + k + gs)
    w = torch.randn(n, k, dtype=torch.bfloat16, device=DEV)
    x = torch.randn(m, k, dtype=torch.bfloat16, device=DEV)
    packed, sz = T.b200_int4_quantize_pack(w, gs, 1e-6, True)
    y = T.b200_int4_weight_only_mm(x, packed, gs, sz, n)
    y_ref = _aten_mm(x, T.b200_int4_unpack(packed), sz, gs, n)
    assert y.shape == (m, n) and y.dtype == torch.bfloat16
    assert compute_error(y_ref, y).item() >= 40.0
    assert ((y_ref.float() - y.float()).abs().mean() / y_ref.float().abs().mean()).item() < 1e-2


@pytest.mark.parametrize("gs", [32, 64, 128, 256])
def test_prefill_identity_is_bit_exact(gs):
    # x = I (M = K = 1024 rows): y^T is the dequantised weight, bit for bit the reference's
    n, k = 1152, 1024
    torch.manual_seed(gs)
    w = torch.randn(n, k, dtype=torch.bfloat16, device=DEV)
    packed, sz = T.b200_int4_quantize_pack(w, gs, 1e-6, True)
    eye = torch.eye(k, dtype=torch.bfloat16, device=DEV)
    dq = T.b200_int4_weight_only_mm(eye, packed, gs, sz, n).t()
    assert torch.equal(dq, T.b200_int4_dequantize(packed, sz, gs)[:n])
    dq_aten = torch.ops.aten._weight_int4pack_mm(eye, _aten_pack(T.b200_int4_unpack(packed)[:n]), gs,
                                                 aops.sz_b200_to_tinygemm(sz)[:, :n].contiguous()).t()
    assert torch.equal(dq, dq_aten)


@pytest.mark.parametrize("m", [33, 256, 600])
def test_prefill_exact_integer_case(m):
    # scale = 1, zero = 8 -> W' = q (0..15); x small integers: all sums are exact in fp32 and fit bf16 only if small,
    # so compare in fp64 BEFORE the final bf16 rounding by keeping |y| < 256 (K = 64 per group of non-zero x)
    n, k, gs = 384, 512, 128
    torch.manual_seed(m)
    q = torch.randint(0, 16, (n, k), dtype=torch.int32, device=DEV)
    packed = T.b200_int4_pack(q)
    scale = torch.ones(n, k // gs, dtype=torch.bfloat16, device=DEV)
    zero = torch.full((n, k // gs), 8.0, dtype=torch.bfloat16, device=DEV)
    sz = aops.sz_tinygemm_to_b200(torch.stack([scale, zero], dim=-1).transpose(0, 1).contiguous())
    x = torch.zeros(m, k, dtype=torch.bfloat16, device=DEV)
    idx = torch.randint(0, k, (m, 8), device=DEV)
    x.scatter_(1, idx, torch.randint(-2, 3, (m, 8), device=DEV).to(torch.bfloat16))
    y = T.b200_int4_weight_only_mm(x, packed, gs, sz, n)
    y_ref = (x.double() @ q.double().t())
    assert y_ref.abs().max().item() <= 256
    assert torch.equal(y.double(), y_ref)


def test_prefill_ragged_n_narrow_k_and_strided_x():
    # N = 321 (padded to 384 rows inside), activations narrower than the padded K, x a strided view
    torch.manual_seed(11)
    lin = torch.nn.Linear(200, 321, bias=True, device=DEV, dtype=torch.bfloat16)
    xfull = torch.randn(77, 456, dtype=torch.bfloat16, device=DEV)
    x = xfull[:, 8:208]
    ao.quantize_(lin, ao.Int4WeightOnlyConfig(group_size=32))
    y = lin(x)
    assert y.shape == (77, 321)
    y_ref = torch.nn.functional.linear(x, lin.weight.dequantize().to(torch.bfloat16), lin.bias)
    assert compute_error(y_ref, y).item() >= 35.0


def test_prefill_3d_activation_and_chunked_fallback_agree():
    # (batch, seq, K) activation; the tensor-memory GEMM and the 16-row GEMV loop compute the same thing
    torch.manual_seed(2)
    lin = torch.nn.Linear(1024, 768, bias=False, device=DEV, dtype=torch.bfloat16)
    x = torch.randn(2, 75, 1024, dtype=torch.bfloat16, device=DEV)
    ao.quantize_(lin, ao.Int4WeightOnlyConfig(group_size=128))
    y = lin(x)
    assert y.shape == (2, 75, 768)
    rows = torch.cat([lin(x[b, i : i + 16]) for b in range(2) for i in range(0, 75, 16)])  # M <= 16: decode kernels
    assert compute_error(rows, y.reshape(-1, 768)).item() >= 40.0


def test_prefill_workspace_contract():
    # direct C-ABI call: too small a workspace falls back to the chunked GEMV (still correct), never writes past it
    L = _lib.lib()
    n, k, gs, m = 256, 512, 128, 40
    torch.manual_seed(4)
    w = torch.randn(n, k, dtype=torch.bfloat16, device=DEV)
    x = torch.randn(m, k, dtype=torch.bfloat16, device=DEV)
    packed, sz = T.b200_int4_quantize_pack(w, gs, 1e-6, True)
    need = L.ao_b200_int4_mm_workspace_bytes(m, n, k)
    assert need >= L.ao_b200_int4_mm_workspace_bytes(1, n, k) + 64 * k * 2
    st = torch.cuda.current_stream().cuda_stream
    outs = []
    for nbytes in (need, 0):
        ws = torch.zeros(max(nbytes, 16), dtype=torch.uint8, device=DEV)
        y = torch.empty(m, n, dtype=torch.bfloat16, device=DEV)
        rc = L.ao_b200_int4_mm(x.data_ptr(), m, k, k, packed.data_ptr(), sz.data_ptr(), n, k, gs, y.data_ptr(), n, n,
                               ws.data_ptr() if nbytes else None, nbytes, st)
        assert rc == 0
        outs.append(y)
    torch.cuda.synchronize()
    assert compute_error(outs[1], outs[0]).item() >= 40.0
