"""A few persistent-kernel decode steps of Llama-3-8B int4 for an ncu capture (run the same command without ncu first):

    ncu --set full --clock-control none --import-source on -k regex:decode_mega_kernel -s 3 -c 1 -o gpurun_out/r02_mega \
        python benchmarks/prof_mega.py [--workload llama3-8b-int4] [--pos 100]
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="llama3-8b-int4")
    ap.add_argument("--pos", type=int, default=100)
    ap.add_argument("--steps", type=int, default=6)
    args = ap.parse_args()
    from torchao_fork_b200._models.llama import Generator
    spec = bench.WORKLOADS[args.workload]
    model, cfg = bench.build_model(spec, torch.device("cuda"))
    gen = Generator(model, 1, args.pos + 64, temperature=0.8, top_k=200, use_graph=False, mega=True)
    eng = gen.engine
    assert eng.mega
    eng.set_state(torch.tensor([1], device="cuda"), args.pos)
    for _ in range(args.steps):
        eng.step()
    torch.cuda.synchronize()
    eng.check_status()
    print("ok", int(eng.pos.item()))


if __name__ == "__main__":
    main()
