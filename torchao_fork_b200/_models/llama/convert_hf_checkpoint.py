"""HuggingFace Llama checkpoint -> the gpt-fast `model.pth` this package (and the reference) loads.

Counterpart of /root/reference/scripts/convert_hf_checkpoint.py:20-139 (real-checkpoint path).  Three
things happen to the tensors:

  * names: `model.layers.N.self_attn.q_proj.weight` -> `layers.N.attention.wq.weight` etc. (table below);
  * RoPE layout: HF keeps a head's rotary dimensions as two halves (x[: d/2], x[d/2 :] — `rotate_half`), gpt-fast's
    `apply_rotary_emb` (model.py:547-559) pairs neighbouring dimensions (x[2i], x[2i + 1]); the rows of q_proj / k_proj
    are re-ordered accordingly (per head: row r of the first half -> 2r, row r of the second half -> 2r + 1);
  * q / k / v are stacked into the fused `wqkv` the model's Attention uses (model.py:422).

`convert_state_dict` is the pure function (tested on CPU); `main` reads the shards named by the checkpoint's index json
(safetensors or torch), writes `model.pth`, and copies the tokenizer for Llama-3 checkpoints.

    python -m torchao_fork_b200._models.llama.convert_hf_checkpoint --checkpoint_dir checkpoints/meta-llama/Meta-Llama-3-8B
"""
import json
import re
import shutil
from pathlib import Path
from typing import Dict, Optional

import torch

from .model import ModelArgs

# HF name (layer number abstracted) -> gpt-fast name; None = dropped
_NAMES = {
    "model.embed_tokens.weight": "tok_embeddings.weight",
    "model.norm.weight": "norm.weight",
    "lm_head.weight": "output.weight",
    "self_attn.q_proj.weight": "attention.wq.weight",
    "self_attn.k_proj.weight": "attention.wk.weight",
    "self_attn.v_proj.weight": "attention.wv.weight",
    "self_attn.o_proj.weight": "attention.wo.weight",
    "self_attn.rotary_emb.inv_freq": None,
    "mlp.gate_proj.weight": "feed_forward.w1.weight",
    "mlp.up_proj.weight": "feed_forward.w3.weight",
    "mlp.down_proj.weight": "feed_forward.w2.weight",
    "input_layernorm.weight": "attention_norm.weight",
    "post_attention_layernorm.weight": "ffn_norm.weight",
}
_LAYER = re.compile(r"^model\.layers\.(\d+)\.(.+)$")


def halves_to_pairs(w: torch.Tensor, n_head: int, head_dim: int) -> torch.Tensor:
    """Rows of a q / k projection from HF's half-split rotary layout to gpt-fast's interleaved one."""
    out_features, in_features = w.shape
    assert out_features == n_head * head_dim
    return w.view(n_head, 2, head_dim // 2, in_features).transpose(1, 2).reshape(out_features, in_features)


@torch.inference_mode()
def convert_state_dict(hf: Dict[str, torch.Tensor], config: ModelArgs) -> Dict[str, torch.Tensor]:
    if config.tie_word_embeddings and "lm_head.weight" not in hf:
        hf = dict(hf, **{"lm_head.weight": hf["model.embed_tokens.weight"].clone()})
    out: Dict[str, torch.Tensor] = {}
    for key, value in hf.items():
        m = _LAYER.match(key)
        if m is None:
            if key not in _NAMES:
                raise KeyError(f"unexpected tensor in the checkpoint: {key}")
            new = _NAMES[key]
        else:
            if m.group(2) not in _NAMES:
                raise KeyError(f"unexpected tensor in the checkpoint: {key}")
            tail = _NAMES[m.group(2)]
            new = None if tail is None else f"layers.{m.group(1)}.{tail}"
        if new is not None:
            out[new] = value
    for layer in range(config.n_layer):
        pre = f"layers.{layer}.attention."
        q = halves_to_pairs(out.pop(pre + "wq.weight"), config.n_head, config.head_dim)
        k = halves_to_pairs(out.pop(pre + "wk.weight"), config.n_local_heads, config.head_dim)
        v = out.pop(pre + "wv.weight")
        out[pre + "wqkv.weight"] = torch.cat([q, k, v])
    return out


def _load_shards(checkpoint_dir: Path) -> Dict[str, torch.Tensor]:
    for index_name in ("model.safetensors.index.json", "pytorch_model.bin.index.json"):
        index = checkpoint_dir / index_name
        if index.is_file():
            break
    else:
        raise FileNotFoundError(f"no model.safetensors.index.json / pytorch_model.bin.index.json under {checkpoint_dir}")
    with open(index) as f:
        shards = sorted({checkpoint_dir / name for name in json.load(f)["weight_map"].values()})
    merged: Dict[str, torch.Tensor] = {}
    for shard in shards:
        if shard.suffix == ".safetensors":
            from safetensors.torch import load_file
            merged.update(load_file(str(shard), device="cpu"))
        else:
            merged.update(torch.load(str(shard), map_location="cpu", mmap=True, weights_only=True))
    return merged


def convert_hf_checkpoint(checkpoint_dir: Path, model_name: Optional[str] = None) -> Path:
    checkpoint_dir = Path(checkpoint_dir)
    model_name = model_name or checkpoint_dir.name
    config = ModelArgs.from_name(model_name)
    result = convert_state_dict(_load_shards(checkpoint_dir), config)
    target = checkpoint_dir / "model.pth"
    torch.save(result, target)
    if any(tag in model_name.lower() for tag in ("llama-3-", "llama-3.1-", "llama-3.2-")):
        original = checkpoint_dir / "original" / ("mp16" if "llama-3.1-405b" in model_name.lower() else "")
        if (original / "tokenizer.model").is_file():
            shutil.copy(original / "tokenizer.model", checkpoint_dir / "tokenizer.model")
    return target


def main():
    import argparse
    ap = argparse.ArgumentParser(description="Convert a HuggingFace Llama checkpoint to the gpt-fast model.pth")
    ap.add_argument("--checkpoint_dir", type=Path, required=True)
    ap.add_argument("--model_name", type=str, default=None)
    args = ap.parse_args()
    print(f"saved {convert_hf_checkpoint(args.checkpoint_dir, args.model_name)}")


if __name__ == "__main__":
    main()
