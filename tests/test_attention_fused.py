"""ao_b200_decode_attention_fused (RoPE + KV append + GQA attention + split merge, one launch) against a float64
restatement of Attention.forward for one token (torchao/_models/llama/model.py:405-470, 547-559) and against the
two-kernel path (ao_b200_rope_kv_append + ao_b200_decode_attention)."""
import math

import pytest
import torch

import torchao_fork_b200 as ao  # noqa: F401
from torchao_fork_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda"
HD = 128


def _sqnr(ref, x):
    ref, x = ref.double(), x.double()
    return (10 * torch.log10(ref.pow(2).mean() / (ref - x).pow(2).mean().clamp_min(1e-300))).item()


def _freqs(S):
    f = 1.0 / (500000.0 ** (torch.arange(0, HD, 2, device=DEV).float() / HD))
    ang = torch.outer(torch.arange(S, device=DEV).float(), f)
    return torch.stack([ang.cos(), ang.sin()], -1).to(torch.bfloat16).contiguous()  # [S][64][2]


def _rope_ref(x, cs):  # x [..., 128] bf16, cs [64][2] bf16 -> bf16 (fp32 math, model.py:547-559)
    xs = x.float().reshape(*x.shape[:-1], HD // 2, 2)
    c, s = cs[:, 0].float(), cs[:, 1].float()
    o = torch.stack([xs[..., 0] * c - xs[..., 1] * s, xs[..., 1] * c + xs[..., 0] * s], -1)
    return o.flatten(-2).to(torch.bfloat16)


def _run_fused(qkv, freqs, pos, kc, vc, B, H, Hkv, S, n_split):
    L = _lib.lib()
    wsb = int(L.ao_b200_decode_attention_fused_workspace_bytes(B, H, Hkv, n_split))
    ws = torch.zeros(wsb, dtype=torch.uint8, device=DEV)
    out = torch.zeros(B, H * HD, dtype=torch.bfloat16, device=DEV)
    p = torch.tensor([pos], dtype=torch.int32, device=DEV)
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(2):  # twice: the ticket counters must come back to zero
        _lib.check(L.ao_b200_decode_attention_fused(qkv.data_ptr(), freqs.data_ptr(), p.data_ptr(), kc.data_ptr(),
                                                    vc.data_ptr(), out.data_ptr(), B, H, Hkv, HD, S, 1 / math.sqrt(HD),
                                                    n_split, ws.data_ptr(), ws.numel(), st), "attn_fused")
    torch.cuda.synchronize()
    assert int(ws[-4 * B * Hkv:].view(torch.int32).abs().sum()) == 0
    return out


@pytest.mark.parametrize("B,H,Hkv", [(1, 32, 8), (3, 8, 8), (2, 16, 2), (1, 4, 4)])
@pytest.mark.parametrize("pos,S,n_split", [(0, 64, 1), (5, 64, 4), (100, 256, 16), (255, 256, 16), (700, 1024, 64), (33, 64, 7)])
def test_fused_attention_matches_float64(B, H, Hkv, pos, S, n_split):
    torch.manual_seed(pos + B + H)
    qkv = torch.randn(B, (H + 2 * Hkv) * HD, device=DEV).to(torch.bfloat16)
    freqs = _freqs(S)
    kc = torch.randn(B, Hkv, S, HD, device=DEV).to(torch.bfloat16)
    vc = torch.randn(B, Hkv, S, HD, device=DEV).to(torch.bfloat16)
    kc[:, :, pos:] = 7.0   # garbage beyond the current length must not be read
    vc[:, :, pos:] = 7.0
    kc2, vc2 = kc.clone(), vc.clone()
    out = _run_fused(qkv, freqs, pos, kc, vc, B, H, Hkv, S, n_split)

    q = _rope_ref(qkv[:, : H * HD].view(B, H, HD), freqs[pos])
    k_new = _rope_ref(qkv[:, H * HD:(H + Hkv) * HD].view(B, Hkv, HD), freqs[pos])
    v_new = qkv[:, (H + Hkv) * HD:].view(B, Hkv, HD)
    assert torch.equal(kc[:, :, pos], k_new) and torch.equal(vc[:, :, pos], v_new)          # append is bit-exact
    assert torch.equal(kc[:, :, :pos], kc2[:, :, :pos]) and torch.equal(kc[:, :, pos + 1:], kc2[:, :, pos + 1:])
    K = kc[:, :, : pos + 1].double().repeat_interleave(H // Hkv, dim=1)
    V = vc[:, :, : pos + 1].double().repeat_interleave(H // Hkv, dim=1)
    att = torch.softmax(torch.einsum("bhd,bhsd->bhs", q.double(), K) / math.sqrt(HD), -1)
    ref = torch.einsum("bhs,bhsd->bhd", att, V).reshape(B, H * HD)
    assert _sqnr(ref, out) > 40.0

    # the two-kernel path gives the same caches and (up to the split partition) the same output
    L = _lib.lib()
    st = torch.cuda.current_stream().cuda_stream
    qb = torch.zeros(B, H * HD, dtype=torch.bfloat16, device=DEV)
    p = torch.tensor([pos], dtype=torch.int32, device=DEV)
    _lib.check(L.ao_b200_rope_kv_append(qkv.data_ptr(), freqs.data_ptr(), p.data_ptr(), qb.data_ptr(), kc2.data_ptr(),
                                        vc2.data_ptr(), B, H, Hkv, HD, S, st), "rope")
    wsb = int(L.ao_b200_decode_attention_workspace_bytes(B, H, n_split))
    ws = torch.zeros(max(wsb, 16), dtype=torch.uint8, device=DEV)
    out2 = torch.zeros_like(out)
    _lib.check(L.ao_b200_decode_attention(qb.data_ptr(), kc2.data_ptr(), vc2.data_ptr(), p.data_ptr(), out2.data_ptr(), B, H,
                                          Hkv, HD, S, 1 / math.sqrt(HD), n_split, ws.data_ptr(), ws.numel(), st), "attn")
    torch.cuda.synchronize()
    assert torch.equal(kc, kc2) and torch.equal(vc, vc2)
    assert _sqnr(out2, out) > 45.0


def test_workspace_too_small_is_an_error():
    L = _lib.lib()
    t = torch.zeros(1024, dtype=torch.bfloat16, device=DEV)
    p = torch.zeros(1, dtype=torch.int32, device=DEV)
    rc = L.ao_b200_decode_attention_fused(t.data_ptr(), t.data_ptr(), p.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(), 1,
                                          8, 8, HD, 64, 0.1, 4, t.data_ptr(), 16, None)
    assert rc == 4
