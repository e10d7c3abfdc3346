"""Perplexity evaluation of the gpt-fast Llama on the B200 weight-only path.

The reference evaluates through lm-evaluation-harness (torchao/_models/llama/eval.py:33-236 `run_evaluation`,
torchao/_models/_eval.py:29-126 `TransformerEvalWrapper`): third-party, absent from /root/reference and from this image,
and its task data (wikitext) needs the network.  What the harness does for a perplexity task is restated here so that a
token file is all it takes:

  * `TransformerEvalWrapper._model_call` (_eval.py:54-67): caches sized for the window, one forward over
    `prepare_inputs_for_model(inps)` = (tokens [1][T], input_pos = arange(T)) (model.py:19-26);
  * lm-eval's `loglikelihood_rolling`: the token stream is cut into disjoint windows of `max_length` predicted tokens, each
    conditioned on the token before it (the first window on `prefix_token`), the log-likelihoods of all predicted tokens
    are summed;
  * `word_perplexity` = exp(-sum logp / n_words), `byte_perplexity` likewise; without the text only the token-level
    perplexity exp(-sum logp / n_tokens) is defined, and that is what `rolling_perplexity` returns.

    python -m torchao_fork_b200._models.llama.eval --random_init tiny-llama3 -q int4wo-128 --tokens tokens.pt
"""
import math
from typing import Optional

import torch

from .model import Transformer


def prepare_inputs_for_model(inps: torch.Tensor):
    """reference: model.py:19-26"""
    if inps.dim() > 2:
        raise ValueError(f"Expected input to be of dim 1 or 2, but got {inps.dim()}")
    input_pos = torch.arange(0, inps.shape[-1], device=inps.device).to(torch.int32)
    return inps.view(-1, inps.shape[-1]), input_pos


@torch.no_grad()
def model_call(model: Transformer, inps: torch.Tensor, max_length: int) -> torch.Tensor:
    """reference: _eval.py:54-67 — logits [1][T][V] of one window."""
    max_seq_length = min(max(inps.size()), max_length)
    model.setup_caches(1, max_seq_length, device=inps.device)
    return model(*prepare_inputs_for_model(inps))


@torch.no_grad()
def rolling_perplexity(model: Transformer, tokens: torch.Tensor, max_length: Optional[int] = None,
                       prefix_token: int = 0, limit: Optional[int] = None) -> dict:
    """tokens: int64 [N] on the model's device.  Returns {"token_perplexity", "nll_sum", "n_tokens", "n_windows"}."""
    if max_length is None:
        max_length = model.config.block_size
    dev = model.tok_embeddings.weight.device
    tokens = tokens.to(dev).view(-1).to(torch.int64)
    n = tokens.numel() if limit is None else min(tokens.numel(), int(limit))
    nll, cnt, windows = 0.0, 0, 0
    for start in range(0, n, max_length):
        end = min(start + max_length, n)
        target = tokens[start:end]
        first = tokens[start - 1 : start] if start > 0 else torch.tensor([prefix_token], device=dev, dtype=torch.int64)
        inp = torch.cat([first, target[:-1]])
        logits = model_call(model, inp.view(1, -1), max_length)[0].float()
        logp = torch.log_softmax(logits, dim=-1)
        nll -= float(logp.gather(1, target.view(-1, 1)).sum())
        cnt += int(target.numel())
        windows += 1
    return {"token_perplexity": math.exp(nll / max(cnt, 1)), "nll_sum": nll, "n_tokens": cnt, "n_windows": windows}


def main():
    import argparse
    from pathlib import Path

    from .generate import _load_model, _random_init_model, apply_quantization

    ap = argparse.ArgumentParser(description="token-level rolling perplexity (see the module docstring)")
    ap.add_argument("--checkpoint_path", type=Path, default=None)
    ap.add_argument("--random_init", type=str, default=None, help="model name: seeded random weights (no checkpoints offline)")
    ap.add_argument("-q", "--quantization", type=str, default=None, help="int8wo, int4wo-<groupsize>[-hqq]")
    ap.add_argument("--tokens", type=Path, default=None, help="torch-saved int64 token tensor; default: seeded random tokens")
    ap.add_argument("--limit", type=int, default=None, help="number of tokens to evaluate")
    ap.add_argument("--max_length", type=int, default=None)
    ap.add_argument("--device", type=str, default="cuda")
    ap.add_argument("--precision", type=lambda x: getattr(torch, x.split(".")[-1]), default=torch.bfloat16)
    args = ap.parse_args()
    assert (args.checkpoint_path is None) != (args.random_init is None), "give --checkpoint_path or --random_init <model name>"
    model = (_load_model(args.checkpoint_path, args.device, args.precision) if args.checkpoint_path is not None
             else _random_init_model(args.random_init, args.device, args.precision))
    apply_quantization(model, args.quantization)
    if args.tokens is not None:
        tokens = torch.load(str(args.tokens), weights_only=True)
    else:
        tokens = torch.randint(0, model.config.vocab_size, (args.limit or 4096,), generator=torch.Generator().manual_seed(0))
    res = rolling_perplexity(model, tokens, args.max_length, limit=args.limit)
    print(res)


if __name__ == "__main__":
    main()
