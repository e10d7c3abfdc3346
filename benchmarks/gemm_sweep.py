"""BASELINE.json config 5: int4 / int8 weight-only GEMM sweep, M in {1, 16, 128, 1024, 4096} over the ten Llama-3 8B / 70B
(N, K) shapes, through the public ops (torchao::b200_int4_weight_only_mm / b200_int8_weight_only_mm — what F.linear on a
quantized weight dispatches to).  One CSV row per (dtype, M, N, K):

    dtype,m,k,n,us,gbs,tflops,frac_hbm,frac_tensor

`gbs` = algorithmic bytes (packed weight + group parameters + activations in and out) / time, `tflops` = 2 M N K / time;
fractions against MEASURED_PEAKS.json (HBM copy GB/s, cuBLAS bf16 TFLOP/s) or the profiling guide's fallback.  Every
timed graph replays launches over DISTINCT weight copies totalling > 300 MB, so small-M rows measure DRAM, not the L2.

    python benchmarks/gemm_sweep.py --out profiles/r02_gemm_sweep.csv
"""
import argparse
import csv
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torchao_fork_b200 as ao  # noqa: E402,F401

T = torch.ops.torchao
# (N, K): wqkv, wo, w1||w3 (one launch in the decode engine: N = 2 I), w2, output — Llama-3-8B then Llama-3-70B
SHAPES = [(6144, 4096), (4096, 4096), (28672, 4096), (4096, 14336), (128256, 4096),
          (10240, 8192), (8192, 8192), (57344, 8192), (8192, 28672), (128256, 8192)]
MS = [1, 16, 128, 1024, 4096]


def time_graph(fns, iters):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for f in fns[:2]:
            f()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for f in fns:
            f()
    for _ in range(2):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / (iters * len(fns))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--group", type=int, default=128)
    ap.add_argument("--dtypes", nargs="+", default=["int4", "int8"])
    ap.add_argument("--m", type=int, nargs="+", default=MS)
    args = ap.parse_args()
    hbm, tensor = 6553.3, 1653.7
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        d = json.load(open(pk))
        hbm = float(d.get("hbm_gbs", hbm))
        tensor = float(d.get("bf16_tflops", d.get("tensor_tflops", tensor)))
    rows = []
    for kind in args.dtypes:
        for n, k in SHAPES:
            wbytes = n * k // 2 if kind == "int4" else n * k
            copies = max(2, min(48, int(300e6 // wbytes) + 1))
            torch.manual_seed(0)
            w = torch.randn(n, k, dtype=torch.bfloat16, device="cuda") * 0.02
            if kind == "int4":
                p, sz = T.b200_int4_quantize_pack(w, args.group, 1e-6, True)
                ws = [(p.clone(), sz.clone()) for _ in range(copies)]
            else:
                q, s = T.b200_int8_quantize(w, 1e-7, True)
                p8 = T.b200_int8_pack(q)
                ws = [(p8.clone(), s.clone()) for _ in range(copies)]
                del q
            del w
            for m in args.m:
                x = torch.randn(m, k, dtype=torch.bfloat16, device="cuda")
                if kind == "int4":
                    fns = [(lambda a=a, b=b: T.b200_int4_weight_only_mm(x, a, args.group, b, n)) for a, b in ws]
                    nbytes = n * k // 2 + (k // args.group) * n * 4 + m * k * 2 + m * n * 2
                else:
                    fns = [(lambda a=a, b=b: T.b200_int8_weight_only_mm(x, a, b)) for a, b in ws]
                    nbytes = n * k + n * 2 + m * k * 2 + m * n * 2
                if m >= 1024:
                    fns = fns[: max(2, len(fns) // 4)]
                t = time_graph(fns, 10 if m <= 128 else 4)
                gbs, tfl = nbytes / t / 1e9, 2.0 * m * n * k / t / 1e12
                row = dict(dtype=kind, m=m, k=k, n=n, us=round(t * 1e6, 2), gbs=round(gbs, 1), tflops=round(tfl, 2),
                           frac_hbm=round(gbs / hbm, 3), frac_tensor=round(tfl / tensor, 3))
                rows.append(row)
                print(json.dumps(row), flush=True)
                del x, fns
            del ws
            torch.cuda.empty_cache()
    if args.out:
        with open(args.out, "w", newline="") as f:
            wr = csv.DictWriter(f, fieldnames=list(rows[0].keys()))
            wr.writeheader()
            wr.writerows(rows)


if __name__ == "__main__":
    main()
