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
                 top_k: Optional[int] = 200, use_graph: bool = True, tp=None, mega: Optional[bool] = None,
                 kv_cache_quantization: bool = False):
        self.model = model
        self.batch_size = batch_size
        dev = model.tok_embeddings.weight.device
        model.setup_caches(batch_size, max_seq_length, device=dev, kv_cache_quantization=kv_cache_quantization)
        self.engine = DecodeEngine(model, batch_size, temperature, top_k, tp=tp, mega=mega)
        if use_graph:
            self.engine.capture()
        self.sampling = dict(temperature=temperature, top_k=top_k)
        self.h_tok = torch.zeros(batch_size, dtype=torch.int64).pin_memory()
        self.h_out = torch.zeros(batch_size, dtype=torch.int64).pin_memory()
        self._done = torch.cuda.Event()
        self.h_flag = torch.zeros(1, dtype=torch.int32).pin_memory()   # release counter of the pre-launching serving loop
        import os
        self._prelaunch = os.environ.get("AO_B200_E2E_PRELAUNCH", "1") != "0"

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


    def _decode_prelaunched(self, n: int, out: torch.Tensor, callback) -> torch.Tensor:
        """The serving loop with the NEXT step's launch hidden behind the current step: step i + 1 is enqueued while the
        host still waits for token i, behind a stream-ordered wait (ao_b200_stream_wait_value32) on a flag in pinned host
        memory that the host bumps once it has stored the token to feed; the step's own H2D copy node runs after that
        wait.  The host learns of a token by watching the pinned result word (a token id is never negative) instead of an
        event.  Same copies per step as the plain loop (8 B in, 8 B out).  Measured: profiles/r02_e2e_prelaunch.txt."""
        from ... import _lib
        eng = self.engine
        L = _lib.lib()
        stream = torch.cuda.current_stream(eng.device)
        sp = stream.cuda_stream
        h_tok, h_out, o, flag = self.h_tok.numpy(), self.h_out.numpy(), out.numpy(), self.h_flag.numpy()
        fptr = self.h_flag.data_ptr()
        seq = int(flag[0])
        h_out[:] = -1

        def enqueue(i):
            _lib.check(L.ao_b200_stream_wait_value32(sp, fptr, (seq + i + 1) & 0xFFFFFFFF), "stream_wait_value32")
            eng.graph_stream.replay()

        try:
            enqueue(0)
            flag[0] = seq + 1                      # h_tok already holds the first token to feed
            for i in range(n):
                if i + 1 < n:
                    enqueue(i + 1)
                spins = 0
                while h_out.min() < 0:
                    spins += 1
                    if spins % 500000 == 0 and stream.query():
                        if h_out.min() < 0:
                            raise RuntimeError("decode step finished without delivering a token")
                callback(self.h_out)
                o[:, i] = h_out
                h_tok[:] = h_out
                h_out[:] = -1
                flag[0] = seq + i + 2              # releases step i + 1 (if enqueued)
        except BaseException:
            flag[0] = seq + n + 2                  # never leave the stream waiting: releases every step that was enqueued
            stream.synchronize()
            raise
        stream.synchronize()
        if seq + n > 0x3FFFFFFF:                   # the wait compares cyclically ((int32)(*flag - value) >= 0): stay far from the wrap
            flag[0] = 0
        eng.check_status()
        return out

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
                if self._prelaunch:
                    return self._decode_prelaunched(n, out, callback)
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
             temperature: float = 0.8, top_k: Optional[int] = 200, callback=lambda x: x,
             kv_cache_quantization: bool = False, cache_size: Optional[int] = None) -> torch.Tensor:
    """One-shot convenience with the reference's signature shape (generate.py:149-163)."""
    g = Generator(model, batch_size, cache_size or (prompt.shape[-1] + max_new_tokens), temperature, top_k,
                  kv_cache_quantization=kv_cache_quantization)
    return g.generate(prompt, max_new_tokens, callback=callback)


# ------------------------------------------------------------------------------------------------ command line
def _load_model(checkpoint_path, device, precision):
    """reference: generate.py:241-253.  Also loads what `--save` wrote: a state_dict whose linear weights are already
    quantized tensor subclasses (AffineQuantizedTensor / Int4Tensor) — `assign=True` keeps them as they are."""
    checkpoint = torch.load(str(checkpoint_path), mmap=True, weights_only=True, map_location="cpu")
    if "model" in checkpoint and "stories" in str(checkpoint_path):
        checkpoint = checkpoint["model"]
    with torch.device("meta"):
        model = Transformer.from_name(str(checkpoint_path))
    model.load_state_dict(checkpoint, assign=True)
    return model.to(device=device, dtype=precision).eval()


def _random_init_model(name, device, precision):
    """No checkpoints exist offline: the architecture `name` with seeded random weights (what bench.py measures)."""
    torch.manual_seed(1234)
    old = torch.get_default_dtype()
    torch.set_default_dtype(precision)
    try:
        with torch.device(device):
            model = Transformer.from_name(name)
    finally:
        torch.set_default_dtype(old)
    return model.eval()


def apply_quantization(model, quantization: Optional[str]):
    """`-q` of the reference CLI (generate.py:395-428), the techniques on this path: int8wo, int4wo-<groupsize>,
    int4wo-<groupsize>-hqq.  Everything else the reference lists is outside SURVEY.md §8 and says so."""
    from ... import Int4WeightOnlyConfig, Int8WeightOnlyConfig, quantize_
    if not quantization:
        return model
    if "int8wo" in quantization:
        quantize_(model, Int8WeightOnlyConfig())
    elif "int4wo" in quantization:
        use_hqq = "hqq" in quantization
        group_size = int(quantization.split("-")[1])
        assert group_size in [32, 64, 128, 256], f"int4wo group_size needs to be one of [32,64,128,256] but got {group_size}"
        quantize_(model, Int4WeightOnlyConfig(group_size=group_size, use_hqq=use_hqq))
    else:
        raise NotImplementedError(f"-q {quantization}: this path implements int8wo, int4wo-<groupsize>[-hqq]")
    return model


def get_model_size_in_bytes(model, ignore_embeddings=False):
    """reference: torchao/utils.py:234-262 — bytes of every parameter / buffer, tensor subclasses counted by their inner
    tensors."""
    def nbytes(t):
        if hasattr(t, "__tensor_flatten__"):
            names, _ = t.__tensor_flatten__()
            return sum(nbytes(getattr(t, n)) for n in names)
        return t.numel() * t.element_size()
    size = 0
    for name, child in model.named_children():
        if not (isinstance(child, torch.nn.Embedding) and ignore_embeddings):
            for p in list(child.parameters(recurse=False)) + list(child.buffers(recurse=False)):
                size += nbytes(p)
            size += get_model_size_in_bytes(child, ignore_embeddings)
    return size


def main(prefill_size: Optional[int] = None, prompt="Hello, my name is", num_samples: int = 5, max_new_tokens: int = 200,
         batch_size: int = 1, top_k: int = 200, temperature: float = 0.8, checkpoint_path=None, random_init: Optional[str] = None,
         quantization: Optional[str] = None, kv_cache_quantization: bool = False, cache_size: Optional[int] = None,
         save: bool = False, device="cuda", precision=torch.bfloat16, write_result=None):
    """reference: generate.py:258-1115 (`main`), the arguments that exist on this path.  `--compile` has no counterpart:
    decode always runs the captured graph / persistent kernel.  The prompt is token ids (comma separated) unless a
    `tokenizer.model` (sentencepiece, or tiktoken BPE for Llama-3 names) sits next to the checkpoint."""
    import os
    import time
    from datetime import datetime
    from pathlib import Path
    assert (checkpoint_path is None) != (random_init is None), "give --checkpoint_path or --random_init <model name>"
    t0 = time.time()
    tokenizer = None
    if checkpoint_path is not None:
        checkpoint_path = Path(checkpoint_path)
        assert checkpoint_path.is_file(), checkpoint_path
        model = _load_model(checkpoint_path, device, precision)
        tok_path = checkpoint_path.parent / "tokenizer.model"
        if tok_path.is_file():
            from .tokenizer import get_tokenizer
            tokenizer = get_tokenizer(tok_path, checkpoint_path)  # generate.py:314 — the path carries the model name
    else:
        model = _random_init_model(random_init, device, precision)
    torch.cuda.synchronize()
    print(f"Time to load model: {time.time() - t0:.02f} seconds")
    vocab = model.config.vocab_size
    if prefill_size is not None and prefill_size > 0:
        g = torch.Generator().manual_seed(0)
        encoded = torch.randint(0, vocab, (int(prefill_size),), generator=g).to(device)
    elif tokenizer is not None:
        encoded = torch.tensor([tokenizer.bos_id()] + tokenizer.encode(prompt), dtype=torch.int64, device=device)
    else:
        try:
            encoded = torch.tensor([int(t) for t in str(prompt).split(",")], dtype=torch.int64, device=device)
        except ValueError:
            encoded = torch.tensor([1] + [ord(c) % vocab for c in str(prompt)], dtype=torch.int64, device=device)  # no tokenizer offline
    prompt_length = encoded.size(0)
    torch.manual_seed(1234)
    apply_quantization(model, quantization)
    model_size = get_model_size_in_bytes(model, ignore_embeddings=True) / 1e9
    if save:
        name = checkpoint_path.name.split(".")[0] if checkpoint_path is not None else random_init
        out_path = os.path.join(os.getcwd(), f"{name}-{quantization}.pt")
        torch.save(model.state_dict(), out_path)
        print(f"saved {out_path}")
    gen = Generator(model, batch_size, cache_size or (prompt_length + max_new_tokens), temperature, top_k,
                    kv_cache_quantization=kv_cache_quantization)
    agg = {"tokens_per_sec": [], "decode_tokens_per_sec": [], "prefill_time": []}
    start = -1  # one untimed sample first, like the reference's compile pass (generate.py:931-935)
    for i in range(start, num_samples):
        torch.cuda.synchronize()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        t0 = time.perf_counter()
        e[0].record()
        eng, B = gen.engine, batch_size
        p2 = encoded.view(1, -1).repeat(B, 1)
        input_pos = torch.arange(0, prompt_length, device=device, dtype=torch.int32)
        with torch.no_grad():
            next_token = prefill(model, p2, input_pos, **gen.sampling).view(B)
        e[1].record()
        eng.set_state(next_token, prompt_length)
        new_tokens = min(max_new_tokens, model.max_seq_length - prompt_length)
        toks = gen.decode_tokens(new_tokens - 1) if new_tokens > 1 else torch.empty(B, 0, dtype=torch.int64, device=device)
        e[2].record()
        torch.cuda.synchronize()
        t = time.perf_counter() - t0
        if i < 0:
            continue
        y = torch.cat([p2, next_token.view(B, 1), toks], dim=1)
        if tokenizer is not None and prefill_size is None:
            print(tokenizer.decode(y[0].tolist()))
        generated = y.size(-1) - prompt_length
        agg["tokens_per_sec"].append(generated / t)
        agg["decode_tokens_per_sec"].append(generated / (e[1].elapsed_time(e[2]) / 1000))
        agg["prefill_time"].append(e[0].elapsed_time(e[1]) / 1000)
        print(f"Sample {i + 1} | overall time {t:.04f} s {generated / t:.02f} tokens/sec",
              f"| prefill time {agg['prefill_time'][-1]:.04f} s decode {agg['decode_tokens_per_sec'][-1]:.02f} tokens/sec")
        print(f"Bandwidth achieved: {model_size * generated / t:.02f} GB/s")
    print("==========")
    tokpersec = sum(agg["tokens_per_sec"]) / len(agg["tokens_per_sec"])
    decode_tokpersec = sum(agg["decode_tokens_per_sec"]) / len(agg["decode_tokens_per_sec"])
    ttft = sum(agg["prefill_time"]) / len(agg["prefill_time"])
    bandwidth = model_size * tokpersec
    mem = torch.cuda.max_memory_reserved() / 1e9
    print(f"Average overall tokens/sec: {tokpersec:.2f}")
    print(f"Average decode tokens/sec: {decode_tokpersec:.04f} s")
    print(f"Average TTFT: {ttft:.04f} s")
    print(f"Average tokens/sec: {tokpersec:.2f}")
    if batch_size > 1:
        print(f"Average tokens/sec including batches {batch_size * tokpersec:.2f}")
    print(f"Average Bandwidth: {bandwidth:.02f} GB/s")
    print(f"Peak Memory Usage: {mem:.02f} GB")
    print(f"Model Size: {model_size:.02f} GB")
    if write_result:
        name = checkpoint_path.parent.name if checkpoint_path is not None else random_init
        txt = (f"\n{datetime.today().strftime('%Y%m%d%H%M%S')}, tok/s={tokpersec:6.2f}, tok/s_decode={decode_tokpersec:6.2f}, "
               f"ttft={ttft:5.4f}, mem/s={bandwidth:7.2f} GB/s, peak_mem={mem:5.2f} GB, model_size={model_size:5.2f} GB "
               f"quant: {quantization}, mod: {name}, kv_quant: {kv_cache_quantization}, dtype: {precision}, device: {device} ")
        with open(write_result, "a") as f:
            f.write(txt)
    return {"tokens_per_sec": tokpersec, "decode_tokens_per_sec": decode_tokpersec, "ttft": ttft, "model_size_gb": model_size}


if __name__ == "__main__":
    import argparse
    from pathlib import Path

    parser = argparse.ArgumentParser(description="gpt-fast Llama generation on the B200 weight-only path "
                                                 "(the reference's generate.py flags that exist on this path)")
    parser.add_argument("--prefill_size", type=int, default=None, help="Whether to run in ttft mode")
    parser.add_argument("--prompt", type=str, default="Hello, my name is", help="Input prompt.")
    parser.add_argument("--num_samples", type=int, default=5, help="Number of samples.")
    parser.add_argument("--max_new_tokens", type=int, default=200, help="Maximum number of new tokens.")
    parser.add_argument("--batch_size", type=int, default=1, help="Batch size to benchmark with")
    parser.add_argument("--top_k", type=int, default=200, help="Top-k for sampling.")
    parser.add_argument("--temperature", type=float, default=0.8, help="Temperature for sampling.")
    parser.add_argument("--checkpoint_path", type=Path, default=None, help="Model checkpoint path.")
    parser.add_argument("--random_init", type=str, default=None,
                        help="model name (e.g. Llama-3-8B): seeded random weights instead of a checkpoint (no checkpoints offline)")
    parser.add_argument("-q", "--quantization", type=str, help="int8wo, int4wo-<groupsize>, int4wo-<groupsize>-hqq")
    parser.add_argument("--kv_cache_quantization", action="store_true", help="Whether to quantize the KV cache")
    parser.add_argument("--cache_size", type=int, default=None,
                        help="Force size of cache to be a certain number of tokens, if not set, will use max_new_tokens+prompt_size")
    parser.add_argument("--save", action="store_true", help="Whether to save the quantized model.")
    parser.add_argument("--device", type=str, default="cuda", help="Device to use")
    parser.add_argument("--precision", type=lambda x: getattr(torch, x.split(".")[-1]), default=torch.bfloat16,
                        help="dtype precision to use")
    parser.add_argument("--write_result", type=Path, default=None, help="Path where to write the result")
    args = parser.parse_args()
    print(args)
    main(args.prefill_size, args.prompt, args.num_samples, args.max_new_tokens, args.batch_size, args.top_k, args.temperature,
         args.checkpoint_path, args.random_init, args.quantization, args.kv_cache_quantization, args.cache_size, args.save,
         args.device, args.precision, args.write_result)
