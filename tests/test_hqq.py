"""Int4WeightOnlyConfig(use_hqq=True) on the GPU (csrc/hqq.cu) against the reference's golden
vectors (tests/golden/hqq_*.npz, produced by _choose_qparams_and_quantize_affine_hqq on CPU) and the numpy oracle.
The solver iterates in fp32 with a mean over each group, so agreement is stated as a tolerance: scale bit-exact (no
iteration in it), q identical on >= 99.9 % of the weights and never off by more than one step, zero within a bf16 ulp
(or 1/128 of the scale) on >= 99 % of the groups."""
import glob
import os

import numpy as np
import pytest
import torch

import ao_oracle as o
import torchao_fork_b200 as ao  # noqa: F401
from conftest import GOLDEN
from test_oracle_golden import hqq_agreement
from torchao_fork_b200.quantization import Int4WeightOnlyConfig, quantize_

pytestmark = pytest.mark.gpu
T = torch.ops.torchao
HQQ = sorted(glob.glob(os.path.join(GOLDEN, "hqq_*.npz")))


def bf16_from_bits(b):
    return torch.from_numpy(b.astype(np.int16)).view(torch.bfloat16).cuda()


def f32_of(t):
    return t.float().cpu().numpy()


@pytest.mark.parametrize("path", HQQ, ids=os.path.basename)
def test_hqq_matches_reference_golden(path):
    g = np.load(path)
    q, s, z = T.b200_int4_hqq_quantize(bf16_from_bits(g["w"]), int(g["group_size"]))
    same_q, max_dq, scale_exact, zero_ok = hqq_agreement(q.cpu().numpy().astype(np.uint8), f32_of(s), f32_of(z), g)
    assert scale_exact
    assert same_q >= 0.999 and max_dq <= 1
    assert zero_ok >= 0.99


@pytest.mark.parametrize("shape,gs", [((256, 1024), 128), ((96, 512), 32), ((64, 2048), 256), ((130, 448), 64)])
def test_hqq_matches_oracle_on_seeded_weights(shape, gs):
    torch.manual_seed(shape[0] + gs)
    w = (torch.randn(*shape) * torch.exp(0.5 * torch.randn(*shape))).to(torch.bfloat16)
    q, s, z = T.b200_int4_hqq_quantize(w.cuda(), gs)
    oq, os_, oz = o.hqq_quantize(w.float().numpy(), gs)
    ref = {"q": oq, "scale": o.to_bits(os_), "zero": o.to_bits(oz)}
    same_q, max_dq, scale_exact, zero_ok = hqq_agreement(q.cpu().numpy().astype(np.uint8), f32_of(s), f32_of(z), ref)
    assert scale_exact and same_q >= 0.999 and max_dq <= 1 and zero_ok >= 0.99


def test_quantize_with_use_hqq_runs_the_same_kernels_and_lowers_the_lp_error():
    torch.manual_seed(5)
    lin = torch.nn.Linear(1024, 384, bias=False, dtype=torch.bfloat16, device="cuda")
    with torch.no_grad():
        lin.weight.mul_(torch.exp(torch.randn_like(lin.weight.float())).to(torch.bfloat16))  # heavy tails
    w = lin.weight.detach().clone()
    import copy
    hqq, mm = copy.deepcopy(lin), copy.deepcopy(lin)
    quantize_(hqq, Int4WeightOnlyConfig(group_size=64, use_hqq=True))
    quantize_(mm, Int4WeightOnlyConfig(group_size=64))
    e_hqq = (w.float() - hqq.weight.dequantize().float()).abs().pow(0.7).mean().item()
    e_mm = (w.float() - mm.weight.dequantize().float()).abs().pow(0.7).mean().item()
    assert e_hqq < e_mm
    # the quantized tensor matches the op's plain outputs and goes through the same linear kernels
    q, s, z = T.b200_int4_hqq_quantize(w, 64)
    assert torch.equal(T.b200_int4_unpack(hqq.weight.tensor_impl.packed_weight)[:384, :1024], q)
    x = torch.randn(3, 1024, dtype=torch.bfloat16, device="cuda")
    y = hqq(x)
    ref = x.float() @ hqq.weight.dequantize().float().t()
    assert (10 * torch.log10(ref.pow(2).mean() / (ref - y.float()).pow(2).mean())).item() >= 40.0


def test_hqq_constant_and_zero_groups():
    w = torch.zeros(16, 256, dtype=torch.bfloat16, device="cuda")
    w[1] = 3.0
    w[2, :128] = -1.5
    q, s, z = T.b200_int4_hqq_quantize(w, 128)
    oq, os_, oz = o.hqq_quantize(w.float().cpu().numpy(), 128)
    assert np.array_equal(q.cpu().numpy().astype(np.uint8), oq)
    assert np.array_equal(o.to_bits(f32_of(s)), o.to_bits(os_)) and np.array_equal(o.to_bits(f32_of(z)), o.to_bits(oz))
    deq = (q.float() - 8) * s.float().repeat_interleave(128, 1) + z.float().repeat_interleave(128, 1)
    assert torch.allclose(deq, w.float(), atol=1e-2)


@pytest.mark.parametrize("shape,gs", [((256, 512), 64), ((512, 1024), 128), ((64, 256), 32)])
def test_hqq_device_side_early_stop_equals_the_host_driven_loop(shape, gs):
    """ao_b200_hqq_solve (the early stop of quant_primitives.py:1964-1972 decided by a one-block kernel, no host round
    trip) against the same iterations driven from the host with ao_b200_hqq_step: same number of iterations, bit-identical
    zero."""
    from torchao_fork_b200 import _lib
    from torchao_fork_b200.ops import HQQ_BETA, HQQ_ITERS, HQQ_KAPPA, HQQ_LP_NORM
    torch.manual_seed(3)
    w = (torch.randn(*shape, device="cuda") * torch.rand(shape[0], 1, device="cuda")).to(torch.bfloat16).contiguous()
    w[0, :gs] = 0.25                                   # a constant group
    L = _lib.lib()
    n_groups = shape[0] * shape[1] // gs
    st = torch.cuda.current_stream().cuda_stream
    err = torch.empty(int(L.ao_b200_hqq_blocks(n_groups)), dtype=torch.float32, device="cuda")

    def init():
        scale = torch.empty(n_groups, dtype=torch.float32, device="cuda")
        zero = torch.empty_like(scale)
        _lib.check(L.ao_b200_hqq_init(w.data_ptr(), n_groups, gs, scale.data_ptr(), zero.data_ptr(), st), "init")
        return scale, zero

    scale_h, zero_h = init()
    beta, best, ran = HQQ_BETA, 1e4, 0
    for _ in range(HQQ_ITERS):
        _lib.check(L.ao_b200_hqq_step(w.data_ptr(), n_groups, gs, scale_h.data_ptr(), zero_h.data_ptr(), beta, HQQ_LP_NORM,
                                      err.data_ptr(), st), "step")
        beta *= HQQ_KAPPA
        ran += 1
        current = float(err.sum(dtype=torch.float64)) / w.numel()
        if current < best:
            best = current
        else:
            break
    scale_d, zero_d = init()
    state = torch.empty(2, dtype=torch.float64, device="cuda")
    _lib.check(L.ao_b200_hqq_solve(w.data_ptr(), n_groups, gs, scale_d.data_ptr(), zero_d.data_ptr(), HQQ_BETA, HQQ_KAPPA,
                                   HQQ_LP_NORM, HQQ_ITERS, err.data_ptr(), state.data_ptr(), st), "solve")
    torch.cuda.synchronize()
    flags = state[1:].view(torch.int32).cpu()
    assert int(flags[1]) == ran, (int(flags[1]), ran)
    assert int(flags[0]) == (1 if ran < HQQ_ITERS else int(flags[0]))
    assert torch.equal(zero_d, zero_h) and torch.equal(scale_d, scale_h)
    assert abs(float(state[0]) - best) <= 1e-12 * max(best, 1.0)
