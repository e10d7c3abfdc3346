"""Decode engine (CUDA-graph single-token step) vs the eager gpt-fast Transformer on the same quantized weights."""
import pytest
import torch

import torchao_fork_b200 as ao
from torchao_fork_b200._models.llama import Generator, ModelArgs, Transformer, model_weight_bytes

cuda = pytest.mark.gpu


def _sqnr(ref, x):
    ref, x = ref.float(), x.float()
    return (10 * torch.log10(ref.pow(2).mean() / (ref - x).pow(2).mean())).item()


def _tiny(quant, seed=0):
    torch.manual_seed(seed)
    old = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    try:
        with torch.device("cuda"):
            m = Transformer(ModelArgs.from_name("tiny-llama3")).eval()
    finally:
        torch.set_default_dtype(old)
    if quant == "int4":
        ao.quantize_(m, ao.Int4WeightOnlyConfig(group_size=128))
    elif quant == "int8":
        ao.quantize_(m, ao.Int8WeightOnlyConfig())
    return m


def test_model_args_match_reference_table():
    # reference: torchao/_models/llama/model.py:104-113
    a = ModelArgs.from_name("Llama-3-8B")
    assert (a.dim, a.n_layer, a.n_head, a.n_local_heads, a.intermediate_size, a.vocab_size, a.head_dim) == (
        4096, 32, 32, 8, 14336, 128256, 128)
    b = ModelArgs.from_name("Llama-3-70B")
    assert (b.dim, b.n_layer, b.n_head, b.n_local_heads, b.intermediate_size, b.head_dim) == (8192, 80, 64, 8, 28672, 128)
    with torch.device("meta"):
        m = Transformer(a)
    n_lin = sum(p.numel() for n, p in m.named_parameters() if p.dim() == 2 and "tok_embeddings" not in n)
    assert n_lin == 7_504_658_432


@cuda
@pytest.mark.parametrize("quant", ["int4", "int8"])
@pytest.mark.parametrize("batch", [1, 4, 32])
def test_engine_step_matches_eager_forward(quant, batch):
    m = _tiny(quant)
    gen = Generator(m, batch, 64, temperature=0.0, top_k=None)
    eng = gen.engine
    torch.manual_seed(1)
    T = 5
    prompt = torch.randint(0, m.config.vocab_size, (batch, T), device="cuda")
    with torch.no_grad():
        m(prompt, torch.arange(T, device="cuda", dtype=torch.int32))  # fills the caches
    tok = torch.randint(0, m.config.vocab_size, (batch,), device="cuda")
    for step in range(3):
        pos = T + step
        # eager reference on a copy of the caches (the engine appends to the live ones)
        saved = [(b.attention.kv_cache.k_cache.clone(), b.attention.kv_cache.v_cache.clone()) for b in m.layers]
        with torch.no_grad():
            ref = m(tok.view(batch, 1), torch.tensor([pos], device="cuda", dtype=torch.int32))[:, -1]
        ref_k = [b.attention.kv_cache.k_cache.clone() for b in m.layers]
        for b, (k, v) in zip(m.layers, saved):
            b.attention.kv_cache.k_cache.copy_(k)
            b.attention.kv_cache.v_cache.copy_(v)
        eng.set_state(tok, pos)
        nxt = eng.step().clone()
        torch.cuda.synchronize()
        snr = _sqnr(ref, eng.logits)
        assert snr > 30.0, f"step {step}: logits SQNR {snr:.1f} dB"
        for b, k in zip(m.layers, ref_k):
            assert _sqnr(k[:, :, pos], b.attention.kv_cache.k_cache[:, :, pos]) > 35.0
        assert int(eng.pos.item()) == pos + 1
        assert torch.equal(eng.tok, nxt)
        tok = ref.argmax(-1)


@cuda
def test_generate_shapes_and_streaming_agree():
    m = _tiny("int4")
    gen = Generator(m, 1, 48, temperature=0.0, top_k=None)
    prompt = torch.randint(0, m.config.vocab_size, (6,))
    a = gen.generate(prompt, 10)
    b = gen.generate(prompt, 10, stream_to_host=True)
    assert a.shape == (1, 16) and a.device.type == "cpu"
    assert torch.equal(a[:, :6], prompt.view(1, -1).to(torch.int64))
    assert torch.equal(a, b)  # greedy: device loop and host-streamed loop produce the same tokens


@cuda
def test_model_weight_bytes_counts_quantized_storage():
    m = _tiny("int4")
    cfg = m.config
    n_w = sum(int(p.shape[0]) * int(p.shape[1]) for n, p in m.named_parameters() if p.dim() == 2 and "tok_emb" not in n)
    got = model_weight_bytes(m)
    expect_min = n_w // 2 + (n_w // 128) * 4
    assert expect_min <= got <= expect_min * 1.05 + 64 * cfg.dim
