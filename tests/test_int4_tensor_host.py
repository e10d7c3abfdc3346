"""Host-side logic of Int4Tensor (quantization/int4_tensor.py) on the CPU: the restated fbgemm recipe against numpy,
nibble packing, and the slice / select / view / copy_ / state_dict mechanics of the reference's tensor
(torchao/quantization/quantize_/workflows/int4/int4_tensor.py:162-490).  The GEMM side is in tests/test_int4_tensor.py."""
import io

import numpy as np
import pytest
import torch

import torchao_fork_b200  # noqa: F401
from torchao_fork_b200.quantization.int4_tensor import Int4Tensor, int4_row_quantize_zp, pack_int4, unpack_int4


def _make(n=32, k=256, g=64, seed=0):
    torch.manual_seed(seed)
    w = (torch.randn(n, k) * 0.1).to(torch.bfloat16)
    q, s, z = int4_row_quantize_zp(w, g)
    return w, q, Int4Tensor(pack_int4(q), s.to(torch.bfloat16), z.to(torch.bfloat16), [1, g], w.shape)


def test_row_quantize_and_pack_follow_the_recipe():
    w, q, t = _make()
    import ao_oracle as o
    qn, scale_t, zero_t = o.int4_row_quantize_zp(w.float().numpy(), 64)      # CPU: tensor / scalar is a true division
    scale, zero = scale_t.T.reshape(-1, 1), zero_t.T.reshape(-1, 1)
    assert np.array_equal(q.numpy(), qn)
    assert np.array_equal(t.qdata.numpy(), o.pack_int4(qn))
    assert torch.equal(unpack_int4(t.qdata), q) and int(q.min()) >= -8 and int(q.max()) <= 7
    assert t.scale.shape == (4, 32) and t.zero_point.shape == (4, 32)
    assert torch.equal(t.scale, torch.from_numpy(scale.reshape(32, 4).T.copy()).to(torch.bfloat16))
    assert torch.equal(t.zero_point, torch.from_numpy(zero.reshape(32, 4).T.copy()).to(torch.bfloat16))
    # dequantize == q * scale + zero, and it is close to the weight (4-bit, groups of 64)
    dq = t.dequantize(torch.float32)
    ref = q.float() * t.scale.float().t().repeat_interleave(64, 1) + t.zero_point.float().t().repeat_interleave(64, 1)
    assert torch.allclose(dq, ref, rtol=2 ** -6, atol=2e-3)      # dequantize() works in the weight's dtype (bf16)
    assert (dq - w.float()).abs().max() < 0.05


def test_slice_select_view_copy_and_state_dict():
    w, q, t = _make()
    top = t.narrow(0, 0, 8)                                   # rows: qdata dim 0, scale / zero dim 1
    assert top.shape == (8, 256) and torch.equal(top.qdata, t.qdata[:8]) and torch.equal(top.scale, t.scale[:, :8])
    assert top.qdata.data_ptr() == t.qdata.data_ptr()        # aliasing preserved
    left = t.narrow(1, 0, 128)                                # columns: qdata dim 1 (two per byte), scale / zero dim 0
    assert left.shape == (32, 128) and torch.equal(left.qdata, t.qdata[:, :64]) and torch.equal(left.zero_point, t.zero_point[:2])
    assert t.view(32, 256).shape == (32, 256)
    with pytest.raises(NotImplementedError):
        t.view(64, 128)
    # copy_ into a slice lands in the parent's storage
    w2, q2, t2 = _make(seed=1)
    top.copy_(t2.narrow(0, 0, 8))
    assert torch.equal(t.qdata[:8], t2.qdata[:8]) and torch.equal(t.scale[:, :8], t2.scale[:, :8])
    with pytest.raises(ValueError):
        top.copy_(t2)
    # 3-D (expert / batch stacks): select dim 0, squeeze a leading 1
    st = Int4Tensor(torch.stack([t.qdata, t2.qdata]), torch.stack([t.scale, t2.scale]), torch.stack([t.zero_point, t2.zero_point]),
                    [1, 1, 64], (2, 32, 256))
    e1 = st[1]
    assert e1.shape == (32, 256) and e1.block_size == [1, 64] and torch.equal(e1.qdata, t2.qdata)
    one = Int4Tensor(t.qdata[None], t.scale[None], t.zero_point[None], [1, 1, 64], (1, 32, 256)).squeeze(0)
    assert one.shape == (32, 256) and torch.equal(one.qdata, t.qdata)
    # state_dict round trip with weights_only=True
    lin = torch.nn.Linear(256, 32, bias=False, dtype=torch.bfloat16)
    lin.weight = torch.nn.Parameter(t, requires_grad=False)
    buf = io.BytesIO()
    torch.save(lin.state_dict(), buf)
    buf.seek(0)
    sd = torch.load(buf, weights_only=True)
    assert isinstance(sd["weight"], Int4Tensor) and torch.equal(sd["weight"].qdata, t.qdata) and sd["weight"].block_size == [1, 64]


def test_no_cpu_product_path():
    w, _, t = _make()
    with pytest.raises(NotImplementedError):
        Int4Tensor.from_hp(w, [1, 64])                                       # quantization of CPU tensors is not offered
    with pytest.raises(Exception):
        torch.nn.functional.linear(torch.randn(1, 256, dtype=torch.bfloat16), t)   # no CPU GEMM either
