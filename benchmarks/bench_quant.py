"""Throughput of the quantize / pack / unpack / dequant kernels against the
HBM roofline, next to the reference's eager op sequence on the same GPU.
Algorithmic bytes: int4 quantize+pack 2NK (read) + NK/2 + (K/G) N 4 (write); int8 quantize 2NK + NK + 2N;
unpack NK/2 -> 4NK; dequant NK/2 + qparams -> 2NK.
Usage: python benchmarks/bench_quant.py [--shapes 4096,4096 28672,4096]"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torchao_fork_b200 as ao  # noqa: E402,F401

T = torch.ops.torchao


def timeit(fn, iters=20, flush=None):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(iters):
        if flush is not None:
            flush.zero_()  # write a buffer larger than L2 between iterations
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters * 1e-3


def ref_eager_tinygemm(w, gs):  # quant_primitives.py:1287-1300, :569-570 + tensor_core_tiled_layout.py:276-279, utils.py:395-409
    n, k = w.shape
    wg = w.view(n, k // gs, gs)
    mn = torch.amin(wg, dim=-1)
    mx = torch.amax(wg, dim=-1)
    scale = torch.clamp((mx - mn) / float(15), min=1e-6)
    zero = mn + scale * 8.0
    min_val = zero - scale * 8.0
    q = torch.clamp(torch.round((wg - min_val.unsqueeze(-1)) / scale.unsqueeze(-1)), 0, 15).to(torch.int32).view(n, k)
    u8 = (q[:, ::2] << 4 | q[:, 1::2]).to(torch.uint8)
    packed = torch.ops.aten._convert_weight_to_int4pack(u8, 8)
    sz = torch.cat([scale.reshape(n, -1, 1), zero.reshape(n, -1, 1)], 2).transpose(0, 1).contiguous()
    return packed, sz


def ref_eager_int8(w):  # quant_primitives.py:1537-1565, :453-456
    mn = torch.amin(w, dim=-1)
    mx = torch.amax(w, dim=-1)
    mn = torch.min(mn, torch.zeros_like(mn))
    mx = torch.max(mx, torch.zeros_like(mx))
    mx = torch.max(-mn, mx)
    scale = torch.clamp(mx / 127.5, min=torch.finfo(torch.float32).eps)
    q = torch.clamp(torch.round(w * (1.0 / scale.unsqueeze(-1))), -128, 127).to(torch.int8)
    return scale, q


def ref_eager_hqq(w, gs):  # quant_primitives.py:1915-1973, 2035-2090 restated with torch ops (fp32), as the reference runs it
    W = w.float().reshape(-1, gs)
    mn, mx = W.min(dim=1, keepdim=True)[0], W.max(dim=1, keepdim=True)[0]
    scale = (15 / (mx - mn)).clamp(max=2e4)
    zero = torch.round(-mn * scale)
    beta, best = 10.0, 1e4
    for _ in range(20):
        Wq = torch.round(W * scale + zero).clamp(0, 15)
        Wr = (Wq - zero) / scale
        d = W - Wr
        We = torch.sign(d) * torch.relu(d.abs() - (1.0 / beta) * d.abs().pow(-0.3))
        zero = torch.mean(Wq - (W - We) * scale, dim=1, keepdim=True)
        beta *= 1.01
        err = float((W - Wr).abs().mean())
        if err < best:
            best = err
        else:
            break
    q = torch.round(W * scale + zero).clamp(0, 15).to(torch.int32).reshape(w.shape)
    return q, (1.0 / scale).to(torch.bfloat16), ((8 - zero) / scale).to(torch.bfloat16)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", nargs="*", default=["4096,4096", "28672,4096", "4096,14336"])
    ap.add_argument("--group", type=int, default=128)
    args = ap.parse_args()
    peak = 6550.7
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peak = json.load(open(pk))["hbm_gbs"]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    gs = args.group
    for sh in args.shapes:
        n, k = (int(v) for v in sh.split(","))
        torch.manual_seed(0)
        w = torch.randn(n, k, dtype=torch.bfloat16, device="cuda")
        rows = []

        def rec(name, fn, nbytes, ref_fn=None):
            t = timeit(fn, flush=flush)
            row = {"op": name, "n": n, "k": k, "us": round(t * 1e6, 2), "gbs": round(nbytes / t / 1e9, 1),
                   "frac_of_measured_peak": round(nbytes / t / 1e9 / peak, 3), "algorithmic_bytes": nbytes}
            if ref_fn is not None:
                tr = timeit(ref_fn, iters=5, flush=flush)
                row["reference_eager_us"] = round(tr * 1e6, 1)
                row["speedup_vs_reference_eager"] = round(tr / t, 1)
            rows.append(row)
            print(json.dumps(row), flush=True)

        rec("int4_quantize_pack", lambda: T.b200_int4_quantize_pack(w, gs, 1e-6, True),
            2 * n * k + n * k // 2 + (k // gs) * n * 4, lambda: ref_eager_tinygemm(w, gs))
        packed, sz = T.b200_int4_quantize_pack(w, gs, 1e-6, True)
        rec("int4_unpack", lambda: T.b200_int4_unpack(packed), n * k // 2 + 4 * n * k)
        rec("int4_dequantize", lambda: T.b200_int4_dequantize(packed, sz, gs), n * k // 2 + (k // gs) * n * 4 + 2 * n * k)
        if n * k <= 4096 * 14336:  # a solver iteration reads the weight once: (iterations + 2) * 2NK + 4NK written
            rec("int4_hqq_quantize (<= 20 iterations + init + finalize)", lambda: T.b200_int4_hqq_quantize(w, gs),
                22 * 2 * n * k + 4 * n * k, lambda: ref_eager_hqq(w, gs))
        rec("int8_quantize", lambda: T.b200_int8_quantize(w, 1.1920929e-07, True), 2 * n * k + n * k + 2 * n,
            lambda: ref_eager_int8(w))
        q8, _ = T.b200_int8_quantize(w, 1.1920929e-07, True)
        rec("int8_pack", lambda: T.b200_int8_pack(q8), 2 * n * k)
        p8 = T.b200_int8_pack(q8)
        rec("int8_unpack", lambda: T.b200_int8_unpack(p8, n, k), 2 * n * k)
        del w, packed, sz, q8, p8
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
