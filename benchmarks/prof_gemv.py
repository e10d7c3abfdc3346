"""Tiny driver for ncu: a handful of launches of one weight-only linear shape (no CUDA graph)."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torchao_fork_b200  # noqa: F401,E402

T = torch.ops.torchao
ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=28672)
ap.add_argument("--k", type=int, default=4096)
ap.add_argument("--m", type=int, default=1)
ap.add_argument("--group", type=int, default=128)
ap.add_argument("--int8", action="store_true")
ap.add_argument("--iters", type=int, default=6)
a = ap.parse_args()
torch.manual_seed(0)
w = torch.randn(a.n, a.k, dtype=torch.bfloat16, device="cuda")
x = torch.randn(a.m, a.k, dtype=torch.bfloat16, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
if a.int8:
    q, s = T.b200_int8_quantize(w, 1e-7, True)
    q = T.b200_int8_pack(q)
    f = lambda: T.b200_int8_weight_only_mm(x, q, s)
else:
    p, sz = T.b200_int4_quantize_pack(w, a.group, 1e-6, True)
    f = lambda: T.b200_int4_weight_only_mm(x, p, a.group, sz, a.n)
for _ in range(a.iters):
    flush.zero_()
    y = f()
torch.cuda.synchronize()
print("ok", float(y.float().abs().mean()))
