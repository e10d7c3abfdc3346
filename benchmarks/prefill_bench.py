"""BASELINE.json config 3 (second half): the 2048-token prefill of Llama-3-8B through the model, end to end
(reference: torchao/_models/llama/generate.py:103-108 `prefill` = one eager forward over the prompt + sampling).
Linears run the tcgen05 dequant-GEMM (csrc/gemm_int4_tc.cu, int4 and int8 instantiations); attention is torch SDPA,
norms / RoPE torch ops, as in the reference.  Prints one JSON line per workload:

    {"workload", "tokens", "ms", "tok_per_s", "linear_tflops", "frac_tensor_sustained", "frac_tensor_burst"}

linear_tflops counts 2 * tokens * (weights of every nn.Linear incl. the output projection) only.

    python benchmarks/prefill_bench.py --workloads llama3-8b-int8 llama3-8b-int4 --tokens 2048 --out profiles/r02_prefill_2048.jsonl
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workloads", nargs="+", default=["llama3-8b-int8", "llama3-8b-int4"])
    ap.add_argument("--tokens", type=int, default=2048)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    peaks = {}
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peaks = json.load(open(pk))
    burst = float(peaks.get("bf16_tflops", 1653.7))
    sustained = float(peaks.get("bf16_tflops_sustained", 1363.6))
    from torchao_fork_b200._models.llama.generate import prefill
    lines = []
    for wl in args.workloads:
        spec = bench.WORKLOADS[wl]
        model, cfg = bench.build_model(spec, torch.device("cuda"))
        T = args.tokens
        model.setup_caches(1, T, device="cuda")
        prompt = torch.randint(0, cfg.vocab_size, (1, T), device="cuda", generator=None)
        pos = torch.arange(T, device="cuda", dtype=torch.int32)
        n_lin = sum(int(m.weight.shape[0]) * int(m.weight.shape[1]) for m in model.modules() if isinstance(m, torch.nn.Linear))
        with torch.no_grad():
            for _ in range(3):
                prefill(model, prompt, pos, temperature=0.8, top_k=200)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(args.iters):
                prefill(model, prompt, pos, temperature=0.8, top_k=200)
            e1.record()
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.iters
        tfl = 2.0 * T * n_lin / (ms * 1e-3) / 1e12
        line = {"workload": wl, "tokens": T, "ms": round(ms, 3), "tok_per_s": round(T / (ms * 1e-3), 1),
                "linear_tflops": round(tfl, 1), "frac_tensor_sustained": round(tfl / sustained, 3),
                "frac_tensor_burst": round(tfl / burst, 3), "peaks_tflops": {"burst": burst, "sustained": sustained}}
        print(json.dumps(line), flush=True)
        lines.append(line)
        del model
        torch.cuda.empty_cache()
    if args.out:
        with open(args.out, "w") as f:
            for l in lines:
                f.write(json.dumps(l) + "\n")


if __name__ == "__main__":
    main()
