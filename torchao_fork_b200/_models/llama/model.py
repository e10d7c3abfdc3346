"""gpt-fast style Llama used by the decode benchmark — host-side mirror of the reference's
torchao/_models/llama/model.py (same module / parameter names, so `quantize_` walks it identically and
state dicts interchange).  This eager module is the *caller* of the quantized linears:
it runs prefill; single-token decode goes through DecodeEngine (decode_engine.py), which drives the same
weights through the CUDA decode kernels inside one CUDA graph.
"""
import math
from dataclasses import dataclass
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch import Tensor

from ...utils import find_multiple


@dataclass
class ModelArgs:
    """reference: model.py:29-72"""

    block_size: int = 2048
    vocab_size: int = 32000
    n_layer: int = 32
    n_head: int = 32
    dim: int = 4096
    intermediate_size: Optional[int] = None
    n_local_heads: int = -1
    head_dim: int = 64
    rope_base: float = 10000
    norm_eps: float = 1e-5
    use_scaled_rope: bool = False
    tie_word_embeddings: bool = False

    def __post_init__(self):
        if self.n_local_heads == -1:
            self.n_local_heads = self.n_head
        if self.intermediate_size is None:
            self.intermediate_size = find_multiple(int(2 * (4 * self.dim) / 3), 256)
        self.head_dim = self.dim // self.n_head

    @classmethod
    def from_name(cls, name: str):
        if name in transformer_configs:
            return cls(**transformer_configs[name])
        hits = sorted((c for c in transformer_configs if c in str(name).upper() or c in str(name)), key=len, reverse=True)
        if not hits:
            raise KeyError(f"unknown model name {name!r}")
        assert len(hits) == 1 or len(hits[0]) != len(hits[1]), name
        return cls(**transformer_configs[hits[0]])


_L3 = dict(n_local_heads=8, vocab_size=128256, rope_base=500000)
# reference: model.py:75-158.  "Llama-3-70B" is the harness's addition: the reference only
# has "70B" (Llama-2 vocab) and "Llama-3.1-70B" (scaled rope); BASELINE.json config 4 names Llama-3-70B.
transformer_configs = {
    "CodeLlama-7b-Python-hf": dict(block_size=16384, vocab_size=32000, n_layer=32, dim=4096, rope_base=1000000),
    "7B": dict(n_layer=32, n_head=32, dim=4096),
    "13B": dict(n_layer=40, n_head=40, dim=5120),
    "30B": dict(n_layer=60, n_head=52, dim=6656),
    "34B": dict(n_layer=48, n_head=64, dim=8192, vocab_size=32000, n_local_heads=8, intermediate_size=22016,
                rope_base=1000000),  # CodeLlama-34B-Python-hf
    "70B": dict(n_layer=80, n_head=64, dim=8192, n_local_heads=8, intermediate_size=28672),
    "Mistral-7B": dict(n_layer=32, n_head=32, n_local_heads=8, dim=4096, intermediate_size=14336, vocab_size=32000),
    "stories15M": dict(n_layer=6, n_head=6, dim=288),
    "stories110M": dict(n_layer=12, n_head=12, dim=768),
    "Llama-3-8B": dict(block_size=8192, n_layer=32, n_head=32, dim=4096, intermediate_size=14336, **_L3),
    "Llama-3.1-8B": dict(block_size=131072, n_layer=32, n_head=32, dim=4096, intermediate_size=14336,
                         use_scaled_rope=True, **_L3),
    "Llama-3-70B": dict(block_size=8192, n_layer=80, n_head=64, dim=8192, intermediate_size=28672, **_L3),
    "Llama-3.1-70B": dict(block_size=131072, n_layer=80, n_head=64, dim=8192, intermediate_size=28672,
                          use_scaled_rope=True, **_L3),
    "Llama-3.1-405B": dict(block_size=131072, n_layer=126, n_head=128, dim=16384, intermediate_size=53248,
                           use_scaled_rope=True, **_L3),
    "Llama-3.2-3B": dict(block_size=131072, n_layer=28, n_head=24, dim=3072, intermediate_size=8192, use_scaled_rope=True,
                         tie_word_embeddings=True, **_L3),
    # small shapes with head_dim 128 for tests (not in the reference)
    "tiny-llama3": dict(block_size=256, n_layer=2, n_head=8, n_local_heads=2, dim=1024, intermediate_size=2048,
                        vocab_size=1024, rope_base=500000),
}


class KVCache(nn.Module):
    """reference: model.py:166-192.  [B][n_local_heads][S][head_dim]"""

    def __init__(self, max_batch_size, max_seq_length, n_heads, head_dim, dtype=torch.bfloat16):
        super().__init__()
        shape = (max_batch_size, n_heads, max_seq_length, head_dim)
        self.register_buffer("k_cache", torch.zeros(shape, dtype=dtype))
        self.register_buffer("v_cache", torch.zeros(shape, dtype=dtype))

    def update(self, input_pos, k_val, v_val):
        assert input_pos.shape[0] == k_val.shape[2]
        self.k_cache[:, :, input_pos] = k_val
        self.v_cache[:, :, input_pos] = v_val
        return self.k_cache, self.v_cache


class AffineQuantizedKVCache(nn.Module):
    """reference: model.py:198-240.  int8 rows [B][n_local_heads][S][head_dim] + one scale per row
    [B][n_local_heads][S][1]; update() returns the dequantized caches with the positions just written replaced by the
    unquantized values.  (Decode steps do not call update(): csrc/decode_attn_q8.cu does the same thing for the new
    row inside the attention launch and never materialises the dequantized cache.)"""

    def __init__(self, max_batch_size, max_seq_length, n_heads, head_dim, scale_dtype=torch.bfloat16):
        super().__init__()
        cache_shape = (max_batch_size, n_heads, max_seq_length, head_dim)
        scale_shape = (max_batch_size, n_heads, max_seq_length, 1)
        self.register_buffer("k_cache", torch.zeros(cache_shape, dtype=torch.int8))
        self.register_buffer("v_cache", torch.zeros(cache_shape, dtype=torch.int8))
        self.register_buffer("k_cache_scale", torch.ones(scale_shape, dtype=scale_dtype))
        self.register_buffer("v_cache_scale", torch.ones(scale_shape, dtype=scale_dtype))

    def update(self, input_pos, k_val, v_val):
        from ...quantization.utils import _quantize_activation_per_token_absmax
        q_k_val, k_scale = _quantize_activation_per_token_absmax(k_val)
        self.k_cache[:, :, input_pos] = q_k_val
        self.k_cache_scale[:, :, input_pos] = k_scale.unsqueeze(-1)
        k_out = self.k_cache * self.k_cache_scale
        k_out[:, :, input_pos] = k_val
        q_v_val, v_scale = _quantize_activation_per_token_absmax(v_val)
        self.v_cache[:, :, input_pos] = q_v_val
        self.v_cache_scale[:, :, input_pos] = v_scale.unsqueeze(-1)
        v_out = self.v_cache * self.v_cache_scale
        v_out[:, :, input_pos] = v_val
        return k_out, v_out

    @classmethod
    def from_float(cls, kv_cache):
        max_batch_size, n_heads, max_seq_length, head_dim = kv_cache.k_cache.shape
        return cls(max_batch_size, max_seq_length, n_heads, head_dim, kv_cache.k_cache.dtype)


class RMSNorm(nn.Module):
    """reference: model.py:489-503 — normalise in fp32, cast back, multiply by the weight."""

    def __init__(self, dim: int, eps: float = 1e-5):
        super().__init__()
        self.eps = eps
        self.weight = nn.Parameter(torch.ones(dim))

    def forward(self, x: Tensor) -> Tensor:
        xf = x.float()
        return (xf * torch.rsqrt(xf.square().mean(-1, keepdim=True) + self.eps)).type_as(x) * self.weight


def apply_scaling(freqs: Tensor) -> Tensor:
    """Llama-3.1 rope scaling (reference: model.py:506-527)."""
    factor, lo, hi, old_len = 8, 1, 4, 8192
    out = []
    for f in freqs.tolist():
        wavelen = 2 * math.pi / f
        if wavelen < old_len / hi:
            out.append(f)
        elif wavelen > old_len / lo:
            out.append(f / factor)
        else:
            smooth = (old_len / wavelen - lo) / (hi - lo)
            out.append((1 - smooth) * f / factor + smooth * f)
    return torch.tensor(out, dtype=freqs.dtype, device=freqs.device)


def precompute_freqs_cis(seq_len, n_elem, base=10000, dtype=torch.bfloat16, use_scaled=False) -> Tensor:
    """-> [seq_len][n_elem/2][2] = (cos, sin)  (reference: model.py:530-544)"""
    freqs = 1.0 / (base ** (torch.arange(0, n_elem, 2)[: n_elem // 2].float() / n_elem))
    if use_scaled:
        freqs = apply_scaling(freqs)
    ang = torch.outer(torch.arange(seq_len, device=freqs.device), freqs)
    return torch.stack([torch.cos(ang), torch.sin(ang)], dim=-1).to(dtype=dtype)


def apply_rotary_emb(x: Tensor, freqs_cis: Tensor) -> Tensor:
    """x [B][T][H][hd], rotation of consecutive pairs in fp32 (reference: model.py:547-559)."""
    xs = x.float().reshape(*x.shape[:-1], -1, 2)
    fc = freqs_cis.view(1, xs.size(1), 1, xs.size(3), 2)
    out = torch.stack([xs[..., 0] * fc[..., 0] - xs[..., 1] * fc[..., 1],
                       xs[..., 1] * fc[..., 0] + xs[..., 0] * fc[..., 1]], -1)
    return out.flatten(3).type_as(x)


class Attention(nn.Module):
    """reference: model.py:411-473"""

    def __init__(self, config: ModelArgs):
        super().__init__()
        assert config.dim % config.n_head == 0
        total_head_dim = (config.n_head + 2 * config.n_local_heads) * config.head_dim
        self.wqkv = nn.Linear(config.dim, total_head_dim, bias=False)
        self.wo = nn.Linear(config.n_head * config.head_dim, config.dim, bias=False)
        self.kv_cache = None
        self.n_head, self.head_dim, self.n_local_heads, self.dim = (
            config.n_head, config.head_dim, config.n_local_heads, config.dim)
        self._register_load_state_dict_pre_hook(self._load_hook)

    @staticmethod
    def _load_hook(state_dict, prefix, *args):
        if prefix + "wq.weight" in state_dict:
            state_dict[prefix + "wqkv.weight"] = torch.cat(
                [state_dict.pop(prefix + n + ".weight") for n in ("wq", "wk", "wv")])

    def forward(self, x, freqs_cis, mask, input_pos=None):
        b, t, _ = x.shape
        q_size, kv_size = self.n_head * self.head_dim, self.n_local_heads * self.head_dim
        q, k, v = self.wqkv(x).split([q_size, kv_size, kv_size], dim=-1)
        q = apply_rotary_emb(q.view(b, t, self.n_head, self.head_dim), freqs_cis).transpose(1, 2)
        k = apply_rotary_emb(k.view(b, t, self.n_local_heads, self.head_dim), freqs_cis).transpose(1, 2)
        v = v.view(b, t, self.n_local_heads, self.head_dim).transpose(1, 2)
        if self.kv_cache is not None:
            k, v = self.kv_cache.update(input_pos, k, v)
        rep = self.n_head // self.n_local_heads
        k, v = k.repeat_interleave(rep, dim=1), v.repeat_interleave(rep, dim=1)
        if mask is not None:
            y = F.scaled_dot_product_attention(q, k, v, attn_mask=mask, dropout_p=0.0)
        else:
            y = F.scaled_dot_product_attention(q, k, v, dropout_p=0.0, is_causal=True)
        return self.wo(y.transpose(1, 2).contiguous().view(b, t, q_size))


class FeedForward(nn.Module):
    """reference: model.py:476-486"""

    def __init__(self, config: ModelArgs):
        super().__init__()
        self.w1 = nn.Linear(config.dim, config.intermediate_size, bias=False)
        self.w3 = nn.Linear(config.dim, config.intermediate_size, bias=False)
        self.w2 = nn.Linear(config.intermediate_size, config.dim, bias=False)

    def forward(self, x):
        return self.w2(F.silu(self.w1(x)) * self.w3(x))


class TransformerBlock(nn.Module):
    """reference: model.py:390-408"""

    def __init__(self, config: ModelArgs):
        super().__init__()
        self.attention = Attention(config)
        self.feed_forward = FeedForward(config)
        self.ffn_norm = RMSNorm(config.dim, config.norm_eps)
        self.attention_norm = RMSNorm(config.dim, config.norm_eps)

    def forward(self, x, input_pos, freqs_cis, mask):
        h = x + self.attention(self.attention_norm(x), freqs_cis, mask, input_pos)
        return h + self.feed_forward(self.ffn_norm(h))


class Transformer(nn.Module):
    """reference: model.py:243-387"""

    def __init__(self, config: ModelArgs):
        super().__init__()
        self.config = config
        self.tok_embeddings = nn.Embedding(config.vocab_size, config.dim)
        self.layers = nn.ModuleList(TransformerBlock(config) for _ in range(config.n_layer))
        self.norm = RMSNorm(config.dim, eps=config.norm_eps)
        self.output = nn.Linear(config.dim, config.vocab_size, bias=False)
        self.freqs_cis: Optional[Tensor] = None
        self.causal_mask: Optional[Tensor] = None
        self.max_batch_size = -1
        self.max_seq_length = -1

    @classmethod
    def from_name(cls, name: str):
        return cls(ModelArgs.from_name(name))

    def setup_caches(self, max_batch_size, max_seq_length, dtype=torch.bfloat16, device=None, kv_cache_quantization=False):
        """reference: model.py:259-330 (`training`, `linear_causal_mask` are outside this path)"""
        want = AffineQuantizedKVCache if kv_cache_quantization else KVCache
        have = type(self.layers[0].attention.kv_cache) if len(self.layers) else want
        if self.max_seq_length >= max_seq_length and self.max_batch_size >= max_batch_size and have is want:
            return
        cfg = self.config
        device = device or self.tok_embeddings.weight.device
        max_seq_length = find_multiple(max_seq_length, 8)
        self.max_seq_length, self.max_batch_size = max_seq_length, max_batch_size
        self.causal_mask = torch.tril(torch.ones(max_seq_length, max_seq_length, dtype=torch.bool, device=device))
        for blk in self.layers:
            if kv_cache_quantization:
                blk.attention.kv_cache = AffineQuantizedKVCache(max_batch_size, max_seq_length, blk.attention.n_local_heads,
                                                                cfg.head_dim, dtype).to(device)
            else:
                blk.attention.kv_cache = KVCache(max_batch_size, max_seq_length, blk.attention.n_local_heads,
                                                 cfg.head_dim, dtype).to(device)
        self.freqs_cis = precompute_freqs_cis(cfg.block_size, cfg.head_dim, cfg.rope_base, dtype,
                                              cfg.use_scaled_rope).to(device)

    def forward(self, idx: Tensor, input_pos: Optional[Tensor] = None) -> Tensor:
        assert self.freqs_cis is not None, "Caches must be initialized first"
        if input_pos is None:
            mask, freqs_cis = None, self.freqs_cis[: idx.shape[1]]
        else:
            mask, freqs_cis = self.causal_mask[None, None, input_pos], self.freqs_cis[input_pos]
        x = self.tok_embeddings(idx)
        for layer in self.layers:
            x = layer(x, input_pos, freqs_cis, mask)
        return self.output(self.norm(x))


def model_weight_bytes(model: nn.Module) -> int:
    """Bytes a decode step must stream: every parameter/buffer except the embedding table and caches
    (reference: torchao/_models/utils.py get_model_size_in_bytes(ignore_embeddings=True), generate.py:855)."""
    from ...dtypes.affine_quantized_tensor import AffineQuantizedTensor

    def tensor_bytes(t):
        if isinstance(t, AffineQuantizedTensor):
            names, _ = t.tensor_impl.__tensor_flatten__()
            return sum(getattr(t.tensor_impl, n).numel() * getattr(t.tensor_impl, n).element_size() for n in names)
        return t.numel() * t.element_size()

    total = 0
    for name, mod in model.named_modules():
        if isinstance(mod, (nn.Embedding, KVCache)):
            continue
        for p in mod.parameters(recurse=False):
            total += tensor_bytes(p)
    return total
