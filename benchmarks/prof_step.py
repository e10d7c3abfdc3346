"""Where does a decode step go?  Builds the Llama-3-8B int4 decode engine and times, each in its own CUDA graph
replayed back to back, (a) the full step, (b) only one kind of launch over all 32 layers (every layer has its own
weights, so a kind's graph streams what the real step streams), (c) the sampling tail.  Not a benchmark: the kinds
are timed out of their real order, so PDL overlap between DIFFERENT kinds is not reproduced.
Usage: python benchmarks/prof_step.py [--batch 1] [--seq 256]"""
import argparse
import json
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torchao_fork_b200 as ao  # noqa: E402
from torchao_fork_b200 import _lib  # noqa: E402
from torchao_fork_b200._models.llama import Generator, ModelArgs, Transformer  # noqa: E402


def time_graph(fn, iters=20):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters  # us per replay


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--seq", type=int, default=256)
    ap.add_argument("--pos", type=int, default=100)
    args = ap.parse_args()
    cfg = ModelArgs.from_name("Llama-3-8B")
    torch.manual_seed(0)
    torch.set_default_dtype(torch.bfloat16)
    with torch.device("cuda"):
        model = Transformer(cfg).eval()
    torch.set_default_dtype(torch.float32)
    ao.quantize_(model, ao.Int4WeightOnlyConfig(group_size=128))
    gen = Generator(model, args.batch, args.seq, use_graph=True)
    e = gen.engine
    e.set_state(torch.zeros(args.batch, dtype=torch.int64, device="cuda"), args.pos)
    L = _lib.lib()
    B = e.B

    def st():
        return torch.cuda.current_stream().cuda_stream

    kinds = {
        "wqkv": lambda: [l["wqkv"].launch_fused(e.h, e.qkv, st(), norm_w=l["attn_norm"], eps=e.eps, ws=e.ws) for l in e.layers],
        "rope_kv_append": lambda: [_lib.check(L.ao_b200_rope_kv_append(
            e.qkv.data_ptr(), e.freqs.data_ptr(), e.pos.data_ptr(), e.q.data_ptr(), l["k_cache"].data_ptr(),
            l["v_cache"].data_ptr(), B, e.H, e.Hkv, e.hd, e.S, st()), "rope") for l in e.layers],
        "attention": lambda: [_lib.check(L.ao_b200_decode_attention(
            e.q.data_ptr(), l["k_cache"].data_ptr(), l["v_cache"].data_ptr(), e.pos.data_ptr(), e.o.data_ptr(), B, e.H,
            e.Hkv, e.hd, e.S, 1.0 / math.sqrt(e.hd), e.n_split, e.attn_ws.data_ptr(), e.attn_ws.numel(), st()), "attn")
            for l in e.layers],
        "attention_fused": lambda: [_lib.check(L.ao_b200_decode_attention_fused(
            e.qkv.data_ptr(), e.freqs.data_ptr(), e.pos.data_ptr(), l["k_cache"].data_ptr(), l["v_cache"].data_ptr(),
            e.o.data_ptr(), B, e.H, e.Hkv, e.hd, e.S, 1.0 / math.sqrt(e.hd), e.n_split, e.attn_fused_ws.data_ptr(),
            e.attn_fused_ws.numel(), st()), "attn_fused") for l in e.layers],
        "wo": lambda: [l["wo"].launch_fused(e.o, e.h, st(), residual=e.h, epilogue=1, ws=e.ws) for l in e.layers],
        "w13": lambda: [l["w13"].launch_fused(e.h, e.s, st(), norm_w=l["ffn_norm"], eps=e.eps, epilogue=2, ws=e.ws) for l in e.layers],
        "w2": lambda: [l["w2"].launch_fused(e.s, e.h, st(), residual=e.h, epilogue=1, ws=e.ws) for l in e.layers],
        "output": lambda: e.output.launch_fused(e.h, e.logits, st(), norm_w=e.out_norm, eps=e.eps, ws=e.ws),
        "sample": lambda: e._sample(e.logits),
        "embed+bookkeeping": lambda: (torch.index_select(e.emb, 0, e.tok, out=e.h), e.tok.copy_(e.next_tok), e.pos.add_(1)),
    }
    out = {"full_step_us": time_graph(e._launch_step), "n_split": e.n_split, "pos": args.pos}
    tot = 0.0
    for name, fn in kinds.items():
        us = time_graph(fn)
        out[name + "_us"] = us
        tot += us
    out["sum_of_kinds_us"] = tot
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
