"""Profiling driver (NOT a benchmark): a Llama-3-8B-shaped model cut to --layers transformer blocks, int4
quantized, a few CUDA-graph decode steps.  Run under `ncu -k regex:ao` to list / capture this repo's kernels
with the shapes of the real decode step without paying for the full 32-layer build inside the profiler."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torchao_fork_b200 as ao  # noqa: E402
from torchao_fork_b200._models.llama import Generator, ModelArgs, Transformer  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=2)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--quant", default="int4")
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--prefill", type=int, default=0, help="also run one eager prefill of this many tokens (prefill GEMM)")
    args = ap.parse_args()
    cfg = ModelArgs.from_name("Llama-3-8B")
    cfg.n_layer = args.layers
    torch.manual_seed(1234)
    torch.set_default_dtype(torch.bfloat16)
    with torch.device("cuda"):
        model = Transformer(cfg).eval()
    torch.set_default_dtype(torch.float32)
    if args.quant == "int4":
        ao.quantize_(model, ao.Int4WeightOnlyConfig(group_size=128))
    elif args.quant == "int8":
        ao.quantize_(model, ao.Int8WeightOnlyConfig())
    gen = Generator(model, args.batch, 256, use_graph=not args.no_graph)
    gen.engine.set_state(torch.zeros(args.batch, dtype=torch.int64, device="cuda"), 200)
    for _ in range(args.steps):
        gen.engine.step()
    torch.cuda.synchronize()
    if args.prefill:
        model.setup_caches(args.batch, args.prefill + 8)
        idx = torch.randint(0, cfg.vocab_size, (args.batch, args.prefill), device="cuda")
        with torch.no_grad():
            model(idx, torch.arange(args.prefill, device="cuda", dtype=torch.int32))
        torch.cuda.synchronize()
    print("ok", gen.engine.launches_per_step, "launches/step")


if __name__ == "__main__":
    main()
