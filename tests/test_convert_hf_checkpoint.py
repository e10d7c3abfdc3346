"""HF checkpoint -> gpt-fast model.pth (torchao_fork_b200/_models/llama/convert_hf_checkpoint.py; reference:
scripts/convert_hf_checkpoint.py:20-139), on the CPU: (1) the row re-ordering of q / k turns HuggingFace's half-split
rotary embedding into the model's pair-interleaved one, so attention scores are unchanged; (2) a checkpoint directory with
a safetensors shard, a torch shard and an index json converts into exactly the state_dict the Transformer loads."""
import json

import torch

from torchao_fork_b200._models.llama import ModelArgs, Transformer
from torchao_fork_b200._models.llama.convert_hf_checkpoint import convert_hf_checkpoint, convert_state_dict, halves_to_pairs
from torchao_fork_b200._models.llama.model import apply_rotary_emb, precompute_freqs_cis, transformer_configs


def _hf_rope(x, pos, base):
    """transformers' LlamaRotaryEmbedding + apply_rotary_pos_emb: x [H][hd]; cos / sin repeated over the two halves."""
    hd = x.shape[-1]
    inv_freq = 1.0 / (base ** (torch.arange(0, hd, 2).float() / hd))
    ang = pos * inv_freq
    cos, sin = torch.cat([ang.cos(), ang.cos()]), torch.cat([ang.sin(), ang.sin()])
    rot = torch.cat([-x[..., hd // 2:], x[..., : hd // 2]], dim=-1)
    return x * cos + rot * sin


def test_row_reordering_maps_half_split_rope_onto_pair_interleaved_rope():
    torch.manual_seed(0)
    H, hd, dim, base, pos = 4, 128, 64, 500000.0, 17
    wq, wk = torch.randn(H * hd, dim, dtype=torch.float64), torch.randn(H * hd, dim, dtype=torch.float64)
    x, y = torch.randn(dim, dtype=torch.float64), torch.randn(dim, dtype=torch.float64)
    q_hf = _hf_rope((wq @ x).view(H, hd).float(), pos, base)
    k_hf = _hf_rope((wk @ y).view(H, hd).float(), 5, base)
    fc = precompute_freqs_cis(32, hd, base, dtype=torch.float32)
    q_gf = apply_rotary_emb((halves_to_pairs(wq, H, hd) @ x).view(1, 1, H, hd).float(), fc[pos : pos + 1])[0, 0]
    k_gf = apply_rotary_emb((halves_to_pairs(wk, H, hd) @ y).view(1, 1, H, hd).float(), fc[5:6])[0, 0]
    # the same numbers in the other order: pair (2i, 2i + 1) of the model = (i, i + hd / 2) of HF
    assert torch.allclose(q_gf[:, 0::2], q_hf[:, : hd // 2], atol=1e-4) and torch.allclose(q_gf[:, 1::2], q_hf[:, hd // 2:], atol=1e-4)
    # ... hence identical attention scores
    assert torch.allclose((q_gf * k_gf).sum(-1), (q_hf * k_hf).sum(-1), atol=1e-3)


def _to_hf(sd, cfg):
    """inverse of the conversion, for building a synthetic HF checkpoint"""
    def pairs_to_halves(w, n_head):
        return w.view(n_head, cfg.head_dim // 2, 2, w.shape[1]).transpose(1, 2).reshape(w.shape)
    out = {"model.embed_tokens.weight": sd["tok_embeddings.weight"], "model.norm.weight": sd["norm.weight"], "lm_head.weight": sd["output.weight"]}
    nq, nk = cfg.n_head * cfg.head_dim, cfg.n_local_heads * cfg.head_dim
    for l in range(cfg.n_layer):
        a, h = f"layers.{l}.", f"model.layers.{l}."
        wqkv = sd[a + "attention.wqkv.weight"]
        out[h + "self_attn.q_proj.weight"] = pairs_to_halves(wqkv[:nq], cfg.n_head).contiguous()
        out[h + "self_attn.k_proj.weight"] = pairs_to_halves(wqkv[nq : nq + nk], cfg.n_local_heads).contiguous()
        out[h + "self_attn.v_proj.weight"] = wqkv[nq + nk:].contiguous()
        out[h + "self_attn.o_proj.weight"] = sd[a + "attention.wo.weight"]
        out[h + "self_attn.rotary_emb.inv_freq"] = torch.zeros(cfg.head_dim // 2)
        out[h + "mlp.gate_proj.weight"] = sd[a + "feed_forward.w1.weight"]
        out[h + "mlp.up_proj.weight"] = sd[a + "feed_forward.w3.weight"]
        out[h + "mlp.down_proj.weight"] = sd[a + "feed_forward.w2.weight"]
        out[h + "input_layernorm.weight"] = sd[a + "attention_norm.weight"]
        out[h + "post_attention_layernorm.weight"] = sd[a + "ffn_norm.weight"]
    return out


def test_checkpoint_directory_converts_into_the_models_state_dict(tmp_path):
    from safetensors.torch import save_file
    cfg = ModelArgs(**transformer_configs["tiny-llama3"])
    torch.manual_seed(1)
    model = Transformer(cfg)
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    hf = _to_hf(sd, cfg)
    assert halves_to_pairs(hf["model.layers.0.self_attn.q_proj.weight"], cfg.n_head, cfg.head_dim).equal(
        sd["layers.0.attention.wqkv.weight"][: cfg.n_head * cfg.head_dim])
    # two shards, one of each kind, plus the index
    ckpt = tmp_path / "tiny-llama3"
    ckpt.mkdir()
    names = sorted(hf)
    first, second = {k: hf[k].contiguous() for k in names[::2]}, {k: hf[k].contiguous() for k in names[1::2]}
    save_file(first, str(ckpt / "model-00001-of-00002.safetensors"))
    torch.save(second, ckpt / "pytorch_model-00002-of-00002.bin")
    index = {"weight_map": {**{k: "model-00001-of-00002.safetensors" for k in first}, **{k: "pytorch_model-00002-of-00002.bin" for k in second}}}
    (ckpt / "model.safetensors.index.json").write_text(json.dumps(index))
    target = convert_hf_checkpoint(ckpt)
    got = torch.load(target, weights_only=True)
    assert sorted(got) == sorted(sd)
    for k in sd:
        assert torch.equal(got[k], sd[k]), k
    fresh = Transformer(cfg)
    assert not fresh.load_state_dict(got, strict=True).missing_keys
    # the pure function refuses tensors it does not know
    bad = dict(hf, **{"model.layers.0.self_attn.q_norm.weight": torch.zeros(1)})
    try:
        convert_state_dict(bad, cfg)
        raise AssertionError("unknown tensor accepted")
    except KeyError:
        pass


def test_every_reference_model_name_resolves():
    """torchao/_models/llama/model.py:75-158: the names generate.py / eval.py / the converter look up from a checkpoint path."""
    want = {"CodeLlama-7b-Python-hf": (32, 4096, 32, 32, 128), "7B": (32, 4096, 32, 32, 128), "13B": (40, 5120, 40, 40, 128),
            "30B": (60, 6656, 52, 52, 128), "34B": (48, 8192, 64, 8, 128), "70B": (80, 8192, 64, 8, 128),
            "Mistral-7B": (32, 4096, 32, 8, 128), "stories15M": (6, 288, 6, 6, 48), "stories110M": (12, 768, 12, 12, 64),
            "Llama-3-8B": (32, 4096, 32, 8, 128), "Llama-3.1-8B": (32, 4096, 32, 8, 128), "Llama-3.1-70B": (80, 8192, 64, 8, 128),
            "Llama-3.1-405B": (126, 16384, 128, 8, 128), "Llama-3.2-3B": (28, 3072, 24, 8, 128)}
    for name, (nl, dim, h, hkv, hd) in want.items():
        c = ModelArgs.from_name(name)
        assert (c.n_layer, c.dim, c.n_head, c.n_local_heads, c.head_dim) == (nl, dim, h, hkv, hd), name
    # a checkpoint path resolves to the longest matching name (generate.py:247 `Transformer.from_name(checkpoint_path)`)
    assert ModelArgs.from_name("checkpoints/meta-llama/Meta-Llama-3.1-70B/model.pth").use_scaled_rope
    assert ModelArgs.from_name("checkpoints/meta-llama/Llama-3.2-3B/model.pth").tie_word_embeddings
    assert ModelArgs.from_name("Llama-3.1-405B").intermediate_size == 53248
    assert ModelArgs.from_name("7B").intermediate_size == 11008      # find_multiple(2 * 4 * dim / 3, 256)
