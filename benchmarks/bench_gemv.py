"""Micro-benchmark of the weight-only linear kernels: achieved algorithmic GB/s per (M, N, K).

Each timed graph replays `copies` launches over DISTINCT weight buffers whose total size exceeds
the 126 MB L2, so DRAM (not L2) is measured.  Also times aten._weight_int4pack_mm (the kernel the
reference calls) on the same shapes.  Usage: python benchmarks/bench_gemv.py [--int8] [--aten] [--csv path]
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torchao_fork_b200 as ao  # noqa: E402

T = torch.ops.torchao
SHAPES_8B = [(6144, 4096), (4096, 4096), (28672, 4096), (4096, 14336), (128256, 4096)]
SHAPES_70B = [(10240, 8192), (8192, 8192), (57344, 8192), (8192, 28672)]


def int4_bytes(m, n, k, g):
    return n * k // 2 + (k // g) * n * 4 + m * k * 2 + m * n * 2


def int8_bytes(m, n, k):
    return n * k + n * 2 + m * k * 2 + m * n * 2


def time_graph(fn_list, iters=20):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for f in fn_list[:2]:
            f()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for f in fn_list:
            f()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / (iters * len(fn_list))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--int8", action="store_true")
    ap.add_argument("--aten", action="store_true")
    ap.add_argument("--m", type=int, nargs="*", default=[1, 8, 16])
    ap.add_argument("--group", type=int, default=128)
    ap.add_argument("--shapes", default="8b")
    ap.add_argument("--out", default=None)
    ap.add_argument("--only", default=None, help="n,k")
    args = ap.parse_args()
    dev = "cuda"
    peak = 6550.7
    pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peak = json.load(open(pk))["hbm_gbs"]
    shapes = SHAPES_8B if args.shapes == "8b" else SHAPES_70B
    if args.only:
        shapes = [tuple(int(v) for v in args.only.split(","))]
    rows = []
    for (n, k) in shapes:
        wbytes = n * k // 2 if not args.int8 else n * k
        copies = max(2, min(64, int(300e6 // wbytes) + 1))
        torch.manual_seed(0)
        w = torch.randn(n, k, dtype=torch.bfloat16, device=dev)
        if args.int8:
            q, s = T.b200_int8_quantize(w, 1e-7, True)
            pk8 = T.b200_int8_pack(q)
            ws = [(pk8.clone(), s.clone()) for _ in range(copies)]
        else:
            p, sz = T.b200_int4_quantize_pack(w, args.group, 1e-6, True)
            ws = [(p.clone(), sz.clone()) for _ in range(copies)]
            if args.aten:
                qq = T.b200_int4_unpack(p)
                u8 = (qq[:, ::2] << 4 | qq[:, 1::2]).to(torch.uint8)
                kp = (k + 1023) // 1024 * 1024
                assert kp == k
                pa = torch.ops.aten._convert_weight_to_int4pack(u8, 8)
                from torchao_fork_b200.ops import sz_b200_to_tinygemm
                szr = sz_b200_to_tinygemm(sz)
                was = [(pa.clone(), szr.clone()) for _ in range(copies)]
        del w
        for m in args.m:
            x = torch.randn(m, k, dtype=torch.bfloat16, device=dev)
            if args.int8:
                fns = [(lambda a=a, b=b: T.b200_int8_weight_only_mm(x, a, b)) for a, b in ws]
                nbytes = int8_bytes(m, n, k)
            else:
                fns = [(lambda a=a, b=b: T.b200_int4_weight_only_mm(x, a, args.group, b, n)) for a, b in ws]
                nbytes = int4_bytes(m, n, k, args.group)
            t = time_graph(fns)
            row = {"kind": "int8" if args.int8 else "int4", "m": m, "n": n, "k": k, "us": t * 1e6,
                   "gbs": nbytes / t / 1e9, "frac_of_measured_peak": nbytes / t / 1e9 / peak}
            if args.aten and not args.int8:
                fa = [(lambda a=a, b=b: torch.ops.aten._weight_int4pack_mm(x, a, args.group, b)) for a, b in was]
                ta = time_graph(fa)
                row["aten_us"] = ta * 1e6
                row["aten_gbs"] = nbytes / ta / 1e9
            rows.append(row)
            print(json.dumps(row), flush=True)
        del ws
        torch.cuda.empty_cache()
    if args.out:
        json.dump(rows, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
