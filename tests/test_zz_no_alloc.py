"""ao_b200_decode_attention_static — the allocation-free entry point (static __device__ scratch inside the library;
no counterpart in the reference) — against ao_b200_decode_attention with
a caller-owned workspace, through the C ABI.  Same kernels, same order of operations: the outputs must be IDENTICAL
(and within the 35 dB the attention tests use against an fp64 restatement).  Collected last on purpose (file name):
a failure here must not hide other tests under `pytest -x`.  Passes alone in a fresh process and after the others:
the shared array needs no particular starting contents."""
import math

import pytest
import torch

from torchao_fork_b200 import _lib

pytestmark = pytest.mark.gpu

STATIC_BYTES = 1 << 20  # AO_B200_STATIC_SCRATCH_BYTES (include/ao_b200.h)


def _case(seed, B, H, Hkv, S, pos):
    g = torch.Generator(device="cuda").manual_seed(seed)
    q = torch.randn(B, H, 128, device="cuda", dtype=torch.bfloat16, generator=g)
    k = torch.randn(B, Hkv, S, 128, device="cuda", dtype=torch.bfloat16, generator=g)
    v = torch.randn(B, Hkv, S, 128, device="cuda", dtype=torch.bfloat16, generator=g)
    p = torch.tensor([pos], device="cuda", dtype=torch.int32)
    return q, k, v, p


def _reference_fp64(q, k, v, pos, H, Hkv):
    grp = H // Hkv
    qq = q.double()
    kk = k[:, :, : pos + 1].double().repeat_interleave(grp, dim=1)
    vv = v[:, :, : pos + 1].double().repeat_interleave(grp, dim=1)
    s = torch.einsum("bhd,bhsd->bhs", qq, kk) / math.sqrt(128)
    return torch.einsum("bhs,bhsd->bhd", torch.softmax(s, dim=-1), vv)


def _sqnr(ref, x):
    ref, x = ref.double(), x.double()
    return (10 * torch.log10(ref.pow(2).mean() / (ref - x).pow(2).mean())).item()


def test_static_scratch_entry_point_matches_workspace_entry_point_twice_in_a_row():
    L = _lib.lib()
    st = torch.cuda.current_stream().cuda_stream
    scale = 1.0 / math.sqrt(128)
    # two calls in a row with DIFFERENT shapes and contents: the second call finds the first call's partials in the array
    for seed, (B, H, Hkv, S, pos, n_split) in enumerate([(2, 32, 8, 256, 200, 8), (1, 16, 2, 128, 77, 5)]):
        q, k, v, p = _case(seed, B, H, Hkv, S, pos)
        wb = int(L.ao_b200_decode_attention_workspace_bytes(B, H, n_split))
        assert 0 < wb <= STATIC_BYTES
        ws = torch.empty(wb, dtype=torch.uint8, device="cuda")
        out_ws = torch.empty(B, H, 128, device="cuda", dtype=torch.bfloat16)
        out_st = torch.empty_like(out_ws)
        _lib.check(L.ao_b200_decode_attention(q.data_ptr(), k.data_ptr(), v.data_ptr(), p.data_ptr(), out_ws.data_ptr(),
                                              B, H, Hkv, 128, S, scale, n_split, ws.data_ptr(), wb, st), "decode_attention")
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        _lib.check(L.ao_b200_decode_attention_static(q.data_ptr(), k.data_ptr(), v.data_ptr(), p.data_ptr(),
                                                     out_st.data_ptr(), B, H, Hkv, 128, S, scale, n_split, st),
                   "decode_attention_static")
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        assert torch.equal(out_st, out_ws), f"case {seed}: static-scratch output differs from the workspace entry point"
        assert _sqnr(_reference_fp64(q, k, v, pos, H, Hkv), out_st) > 35.0


def test_static_scratch_entry_point_rejects_sizes_beyond_the_array():
    L = _lib.lib()
    B, H, n_split = 4, 64, 8  # 4 * 64 * 8 * 130 * 4 B = 1,064,960 B > 1 MiB
    assert int(L.ao_b200_decode_attention_workspace_bytes(B, H, n_split)) > STATIC_BYTES
    q, k, v, p = _case(3, B, H, 8, 64, 10)
    out = torch.full((B, H, 128), 7.0, device="cuda", dtype=torch.bfloat16)
    rc = L.ao_b200_decode_attention_static(q.data_ptr(), k.data_ptr(), v.data_ptr(), p.data_ptr(), out.data_ptr(),
                                           B, H, 8, 128, 64, 0.1, n_split, torch.cuda.current_stream().cuda_stream)
    assert rc == 5  # AO_B200_ERR_STATIC_SCRATCH
    assert b"static scratch" in L.ao_b200_error_string(rc)
    torch.cuda.synchronize()
    assert bool((out == 7.0).all()), "a rejected call must not launch anything"
    with pytest.raises(RuntimeError, match="static scratch"):
        _lib.check(rc, "decode_attention_static")
