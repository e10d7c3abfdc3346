"""ao_b200_sample (csrc/sample.cu) against the reference's op sequence (torchao/_models/llama/generate.py:79-100:
logits / T -> topk pivot mask -> softmax -> argmax(probs / q)) on the SAME exponential noise."""
import pytest
import torch

import torchao_fork_b200 as ao  # noqa: F401
from torchao_fork_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda"


def reference_sample(logits, temperature, top_k, noise):
    if temperature == 0.0:
        return torch.argmax(logits, dim=-1)
    lg = logits.float() / max(temperature, 1e-5)
    if top_k is not None:
        v, _ = torch.topk(lg, min(top_k, lg.size(-1)))
        lg = torch.where(lg < v[..., -1:], -float("inf"), lg)
    probs = torch.softmax(lg, dim=-1)
    return torch.argmax(probs / noise, dim=-1), probs


def run_kernel(logits, temperature, top_k, noise):
    out = torch.full((logits.shape[0],), -1, dtype=torch.int64, device=DEV)
    _lib.check(_lib.lib().ao_b200_sample(
        logits.data_ptr(), logits.stride(0), logits.shape[0], logits.shape[1], float(temperature),
        0 if top_k is None else top_k, None if noise is None else noise.data_ptr(),
        0 if noise is None else noise.stride(0), out.data_ptr(), torch.cuda.current_stream().cuda_stream), "sample")
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("B,V", [(1, 128256), (4, 32000), (16, 1000), (2, 199), (3, 4099)])
@pytest.mark.parametrize("top_k", [200, 1, 50, None])
@pytest.mark.parametrize("temperature", [0.8, 1.0, 0.05])
def test_matches_reference_ops_on_same_noise(B, V, top_k, temperature):
    torch.manual_seed(B * 1000 + V + (top_k or 0))
    logits = (torch.randn(B, V, device=DEV) * 3).to(torch.bfloat16)
    agree = 0
    trials = 8
    for _ in range(trials):
        noise = torch.empty(B, V, device=DEV).exponential_(1)
        ref, probs = reference_sample(logits, temperature, top_k, noise)
        got = run_kernel(logits, temperature, top_k, noise)
        assert got.min() >= 0 and got.max() < V
        # the chosen token must be inside the kept set, and its score must equal the reference's best score up to the
        # fp32 rounding of the softmax sum (different reduction order): identical token unless two scores nearly tie
        score = probs / noise
        best = score.gather(1, ref[:, None])[:, 0]
        mine = score.gather(1, got[:, None])[:, 0]
        assert torch.all(mine >= best * (1 - 1e-5))
        agree += int((got == ref).sum())
    assert agree >= trials * B - 1


def test_bf16_ties_at_the_pivot_are_kept():
    # many equal logits around the k-th place: the reference keeps every entry >= pivot
    V = 4096
    logits = torch.zeros(1, V, device=DEV, dtype=torch.bfloat16)
    logits[0, :10] = 5.0
    logits[0, 10:500] = 2.0   # pivot value for top_k = 100 -> 490 ties kept
    logits[0, 500:] = -3.0
    noise = torch.empty(1, V, device=DEV).exponential_(1)
    ref, probs = reference_sample(logits, 1.0, 100, noise)
    got = run_kernel(logits, 1.0, 100, noise)
    assert int((probs > 0).sum()) == 500
    assert torch.equal(got, ref)


@pytest.mark.parametrize("B,V", [(1, 128256), (5, 777)])
def test_greedy_is_first_argmax(B, V):
    torch.manual_seed(V)
    logits = torch.randn(B, V, device=DEV).to(torch.bfloat16)
    logits[:, V // 3] = 20.0
    logits[:, V // 2] = 20.0   # tie: first index wins
    got = run_kernel(logits, 0.0, 200, None)
    assert torch.all(got == V // 3)


def test_negative_and_mixed_sign_logits():
    torch.manual_seed(5)
    logits = (torch.randn(2, 5000, device=DEV) * 2 - 30).to(torch.bfloat16)
    logits[1] = -logits[1] - 25
    noise = torch.empty(2, 5000, device=DEV).exponential_(1)
    ref, _ = reference_sample(logits, 0.7, 40, noise)
    got = run_kernel(logits, 0.7, 40, noise)
    assert torch.equal(got, ref)


def test_argument_errors():
    logits = torch.zeros(1, 16, device=DEV, dtype=torch.bfloat16)
    out = torch.zeros(1, dtype=torch.int64, device=DEV)
    rc = _lib.lib().ao_b200_sample(logits.data_ptr(), 16, 1, 16, 0.8, 4, None, 0, out.data_ptr(), None)
    assert rc != 0  # sampling with temperature needs the noise
