"""Prefill dequant-GEMM throughput: ours (tcgen05 / tensor memory) vs aten._weight_int4pack_mm (the kernel the
reference calls) vs a plain bf16 cuBLAS matmul on the dequantised weight.  TFLOP/s = 2 M N K / time.
Usage: python benchmarks/bench_gemm.py [--m 256 1024 4096]"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torchao_fork_b200 as ao  # noqa: E402,F401
from torchao_fork_b200 import ops as aops  # noqa: E402

T = torch.ops.torchao


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", type=int, nargs="+", default=[32, 128, 256, 1024, 2048, 4096, 8192])
    ap.add_argument("--gs", type=int, default=128)
    ap.add_argument("--int8", action="store_true", help="int8 weight-only: ours vs the reference's eager formula")
    args = ap.parse_args()
    shapes = [(6144, 4096), (4096, 4096), (28672, 4096), (4096, 14336)]
    rows = []
    if args.int8:
        for n, k in shapes:
            torch.manual_seed(0)
            w = torch.randn(n, k, dtype=torch.bfloat16, device="cuda")
            q, sc = T.b200_int8_quantize(w, 1e-7, True)
            packed = T.b200_int8_pack(q)
            wb = q.to(torch.bfloat16)
            for m in args.m:
                x = torch.randn(m, k, dtype=torch.bfloat16, device="cuda")
                t_ours = timeit(lambda: T.b200_int8_weight_only_mm(x, packed, sc))
                t_ref = timeit(lambda: torch.mm(x, q.t().to(torch.bfloat16)) * sc)   # plain_layout.py:256-262, eager
                t_bf16 = timeit(lambda: torch.nn.functional.linear(x, wb))
                fl = 2.0 * m * n * k
                rows.append({"kind": "int8", "N": n, "K": k, "M": m, "ours_us": round(t_ours, 1),
                             "ref_eager_us": round(t_ref, 1), "cublas_bf16_us": round(t_bf16, 1),
                             "ours_tflops": round(fl / t_ours * 1e-6, 1), "ref_eager_tflops": round(fl / t_ref * 1e-6, 1),
                             "cublas_bf16_tflops": round(fl / t_bf16 * 1e-6, 1)})
                print(json.dumps(rows[-1]), flush=True)
        return
    for n, k in shapes:
        torch.manual_seed(0)
        w = torch.randn(n, k, dtype=torch.bfloat16, device="cuda")
        packed, sz = T.b200_int4_quantize_pack(w, args.gs, 1e-6, True)
        q = T.b200_int4_unpack(packed)
        kp = (k + 1023) // 1024 * 1024
        qpad = torch.nn.functional.pad(q, (0, kp - k))
        u8 = (qpad[:, ::2] << 4 | qpad[:, 1::2]).to(torch.uint8)
        at_packed = torch.ops.aten._convert_weight_to_int4pack(u8.contiguous(), 8)
        sz_ref = aops.sz_b200_to_tinygemm(sz)
        szpad = torch.nn.functional.pad(sz_ref, (0, 0, 0, 0, 0, kp // args.gs - sz_ref.shape[0]))
        dq = T.b200_int4_dequantize(packed, sz, args.gs)[:n]
        for m in args.m:
            x = torch.randn(m, k, dtype=torch.bfloat16, device="cuda")
            xpad = torch.nn.functional.pad(x, (0, kp - k))
            t_ours = timeit(lambda: T.b200_int4_weight_only_mm(x, packed, args.gs, sz, n))
            t_aten = timeit(lambda: torch.ops.aten._weight_int4pack_mm(xpad, at_packed, args.gs, szpad))
            t_bf16 = timeit(lambda: torch.nn.functional.linear(x, dq))
            fl = 2.0 * m * n * k
            rows.append({"N": n, "K": k, "M": m, "ours_us": round(t_ours, 1), "aten_int4_us": round(t_aten, 1),
                         "cublas_bf16_us": round(t_bf16, 1), "ours_tflops": round(fl / t_ours * 1e-6, 1),
                         "aten_int4_tflops": round(fl / t_aten * 1e-6, 1), "cublas_bf16_tflops": round(fl / t_bf16 * 1e-6, 1)})
            print(json.dumps(rows[-1]), flush=True)


if __name__ == "__main__":
    main()
