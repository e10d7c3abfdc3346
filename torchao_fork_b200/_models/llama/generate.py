"""generate() for the gpt-fast Llama — host-side mirror of torchao/_models/llama/generate.py:79-231.

prefill runs the eager Transformer (prompt-sized M -> the prefill GEMM); every following token runs
DecodeEngine's CUDA graph.  `stream_to_host=True` is the serving loop: each step the token to feed is
copied from pinned host memory to the device and the sampled token is read back to pinned host memory
(what a server needs to detokenize / stop on EOS); otherwise tokens stay on the device until the end,
like the reference's decode_n_tokens (generate.py:120-143).
"""
from typing import Optional

import torch

from .decode_engine import DecodeEngine
from .model import Transformer


def multinomial_sample_one_no_sync(probs):
    """reference: generate.py:79-83"""
    q = torch.empty_like(probs).exponential_(1)
    return torch.argmax(probs / q, dim=-1, keepdim=True).to(dtype=torch.int)


def logits_to_probs(logits, temperature: float = 1.0, top_k: Optional[int] = None):
    """reference: generate.py:86-94"""
    logits = logits / max(temperature, 1e-5)
    if top_k is not None:
        v, _ = torch.topk(logits, min(top_k, logits.size(-1)))
        logits = torch.where(logits < v.select(-1, -1).unsqueeze(-1), -float("Inf"), logits)
    return torch.nn.functional.softmax(logits, dim=-1)


def sample(logits, temperature: float = 1.0, top_k: Optional[int] = None):
    """reference: generate.py:97-100"""
    if temperature == 0.0:
        return torch.argmax(logits[:, -1], dim=-1, keepdim=True).to(torch.int), None
    probs = logits_to_probs(logits[:, -1].float(), temperature, top_k)
    return multinomial_sample_one_no_sync(probs), probs


def prefill(model: Transformer, x: torch.Tensor, input_pos: torch.Tensor, **sampling_kwargs) -> torch.Tensor:
    """reference: generate.py:103-108"""
    return sample(model(x, input_pos), **sampling_kwargs)[0]


class Generator:
    """Owns the caches, the DecodeEngine and its CUDA graph for one (model, batch_size, max_seq_length)."""

    def __init__(self, model: Transformer, batch_size: int, max_seq_length: int, temperature: float = 0.8,
                 top_k: Optional[int] = 200, use_graph: bool = True, tp=None, mega: Optional[bool] = None):
        self.model = model
        self.batch_size = batch_size
        dev = model.tok_embeddings.weight.device
        model.setup_caches(batch_size, max_seq_length, device=dev)
        self.engine = DecodeEngine(model, batch_size, temperature, top_k, tp=tp, mega=mega)
        if use_graph:
            self.engine.capture()
        self.sampling = dict(temperature=temperature, top_k=top_k)
        self.h_tok = torch.zeros(batch_size, dtype=torch.int64).pin_memory()
        self.h_out = torch.zeros(batch_size, dtype=torch.int64).pin_memory()
        self._done = torch.cuda.Event()

    @torch.no_grad()
    def generate(self, prompt: torch.Tensor, max_new_tokens: int, stream_to_host: bool = False,
                 callback=lambda x: x) -> torch.Tensor:
        """prompt: int [T] or [B][T] (host or device).  Returns [B][T + new] on the prompt's device.
        reference: generate.py:149-231."""
        eng, model, B = self.engine, self.model, self.batch_size
        dev = eng.device
        out_device = prompt.device
        prompt = prompt.to(dev, non_blocking=True).view(-1, prompt.shape[-1]).to(torch.int64)
        if prompt.shape[0] == 1 and B > 1:
            prompt = prompt.repeat(B, 1)
        T = prompt.shape[1]
        max_seq = min(T + max_new_tokens, model.config.block_size, model.max_seq_length)
        new_tokens = max_seq - T
        input_pos = torch.arange(0, T, device=dev, dtype=torch.int32)
        next_token = prefill(model, prompt, input_pos, **self.sampling).view(B)
        seq = torch.empty(B, max_seq, dtype=torch.int64, device=dev)
        seq[:, :T] = prompt
        seq[:, T] = next_token
        eng.set_state(next_token, T)
        if new_tokens > 1:
            seq[:, T + 1:] = self.decode_tokens(new_tokens - 1, stream_to_host, callback).to(dev)
        return seq.to(out_device)


    @torch.no_grad()
    def decode_tokens(self, n: int, stream_to_host: bool = False, callback=lambda x: x) -> torch.Tensor:
        """Decode n tokens from the engine's current state (decode_n_tokens, generate.py:120-143) -> int64 [B][n].
        stream_to_host=True: every step copies the token to feed from pinned host memory to the device and
        reads the sampled token back into pinned host memory (result on the host); otherwise the tokens
        stay on the device (result on the device), as in the reference."""
        eng, B, dev = self.engine, self.batch_size, self.engine.device
        if eng._host_pos + n > eng.S:
            raise RuntimeError(f"decoding {n} tokens from position {eng._host_pos} would run past the KV cache "
                               f"(max_seq_length {eng.S})")
        if stream_to_host:
            out = torch.empty(B, n, dtype=torch.int64)
            self.h_tok.copy_(eng.tok)
            stream = torch.cuda.current_stream(dev)
            if eng.graph is not None and eng.graph_stream is None:
                eng.capture_streaming(self.h_tok, self.h_out)
            if eng.graph_stream is not None:
                # one replay = H2D of the fed token + the step + D2H of the sampled token (memcpy nodes in the graph)
                h_tok, h_out, o = self.h_tok.numpy(), self.h_out.numpy(), out.numpy()
                eng._host_pos += n
                done = self._done
                for i in range(n):
                    eng.graph_stream.replay()
                    # an event recorded behind the graph reaches the host sooner than the stream's own completion
                    # (measured: ~50 us per token less for the one-kernel graph, benchmarks/e2e_probe2.py)
                    done.record(stream)
                    done.synchronize()
                    callback(self.h_out)
                    o[:, i] = h_out
                    h_tok[:] = h_out
                eng.check_status()
                return out
            for i in range(n):
                eng.tok.copy_(self.h_tok, non_blocking=True)        # H2D: the token the host feeds
                tok = eng.step()
                self.h_out.copy_(tok, non_blocking=True)            # D2H: the sampled token
                stream.synchronize()
                callback(self.h_out)
                out[:, i] = self.h_out
                self.h_tok.copy_(self.h_out)
            return out
        out = torch.empty(B, n, dtype=torch.int64, device=dev)
        for i in range(n):
            out[:, i].copy_(eng.step())
            callback(out[:, i])
        eng.check_status()
        return out


@torch.no_grad()
def generate(model: Transformer, prompt: torch.Tensor, max_new_tokens: int, batch_size: int = 1, *,
             temperature: float = 0.8, top_k: Optional[int] = 200, callback=lambda x: x) -> torch.Tensor:
    """One-shot convenience with the reference's signature shape (generate.py:149-163)."""
    g = Generator(model, batch_size, prompt.shape[-1] + max_new_tokens, temperature, top_k)
    return g.generate(prompt, max_new_tokens, callback=callback)
