"""Where a persistent decode step spends its time: per-phase clock stamps of every CTA (decode_mega.cu `trace`).

    python benchmarks/mega_trace.py [--workload llama3-8b-int4] [--pos 100] [--layers N] [--out gpurun_out/mega_trace.json]

Stamps (clock64 of the CTA's SM, so only differences inside a CTA are meaningful): 0 phase entered (previous main loop
done), 1 arrival counter satisfied, 6 statistics pass + barrier done, 2 staging done, 3 main loop done, 4 outputs
published, 5 producer issued the phase's last stage.  Reported per phase kind, averaged over layers (layer 0 skipped),
in microseconds at --mhz: [mean, max, min] over the CTAs."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="llama3-8b-int4")
    ap.add_argument("--pos", type=int, default=100)
    ap.add_argument("--out", default=None)
    ap.add_argument("--mhz", type=float, default=1965.0)
    ap.add_argument("--layers", type=int, default=0, help="override n_layer (short ncu runs)")
    args = ap.parse_args()
    from torchao_fork_b200._models.llama import Generator
    from torchao_fork_b200._models.llama.model import transformer_configs
    spec = bench.WORKLOADS[args.workload]
    if args.layers:
        transformer_configs[spec["model"]]["n_layer"] = args.layers
    # under torch.distributed.run: tensor parallel; every rank traces, rank 0 reports
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    tp = None
    if world > 1:
        import torch.distributed as dist
        from torchao_fork_b200._models.llama.tp import TPContext
        dist.init_process_group("nccl", device_id=device)
        tp = TPContext(rank, world, device)
    model, cfg = bench.build_model(spec, device, tp)
    gen = Generator(model, 1, args.pos + 64, temperature=0.8, top_k=200, use_graph=False, mega=True, tp=tp)
    eng = gen.engine
    assert eng.mega
    eng._setup_mega(trace=True)
    eng.set_state(torch.tensor([1], device=device), args.pos)
    for _ in range(5):
        eng.step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    eng.step()
    e1.record()
    torch.cuda.synchronize()
    eng.check_status()
    step_us = e0.elapsed_time(e1) * 1e3
    tr = eng._mega_trace.cpu().numpy().astype("float64") / args.mhz   # us, [cta][phase][8]
    n_cta, n_ph, _ = tr.shape
    nl = (n_ph - 1) // 5
    names = ["wqkv", "attn", "wo", "w13", "w2"]
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
        if rank != 0:
            return
    rep = {"step_us_with_sampling": step_us, "n_cta": n_cta, "world": world, "phases": {}}
    rep["kernel_us_mean_over_ctas"] = float((tr[:, n_ph - 1, 3] - tr[:, 0, 0]).mean())
    rep["slow_polls_total"] = int(eng._mega_state[16:20].view(torch.int32).item())  # sweeps that ran ahead of their packets (all steps)

    def mmm(x):
        return [round(float(x.mean()), 2), round(float(x.max()), 2), round(float(x.min()), 2)]

    def summarize(idx, gemv):
        d = {"wait": mmm(tr[:, idx, 1] - tr[:, idx, 0])}
        if gemv:
            d["first_packets"] = mmm(tr[:, idx, 7] - tr[:, idx, 1])   # thread 0's own sweep loads: one loaded L2 round trip
            d["sweep_stats"] = mmm(tr[:, idx, 6] - tr[:, idx, 1])
            d["limbs"] = mmm(tr[:, idx, 2] - tr[:, idx, 6])
            d["main"] = mmm(tr[:, idx, 3] - tr[:, idx, 2])
            d["publish_after_main"] = mmm(tr[:, idx, 4] - tr[:, idx, 3])
            d["producer_lead"] = mmm(tr[:, idx, 3] - tr[:, idx, 5])
            d["total"] = mmm(tr[:, idx, 3] - tr[:, idx, 0])
        else:
            act = (tr[:, idx, 4] - tr[:, idx, 1])
            d["attn"] = mmm(act)
            on = tr[:, idx, 2] > 0   # CTAs that own a slice of the attention
            if on.any():
                d["n_active"] = int(on.sum())
                d["q_ready"] = mmm((tr[:, idx, 2] - tr[:, idx, 1])[on])      # qkv packets polled, RoPE, q in shared memory
                d["tiles"] = mmm((tr[:, idx, 3] - tr[:, idx, 2])[on])        # K / V tiles out of the ring
                d["tiles_wait_first_round"] = mmm((tr[:, idx, 6] - tr[:, idx, 2])[on])   # of which: waiting for the tiles
                d["merge_publish"] = mmm((tr[:, idx, 4] - tr[:, idx, 3])[on])
            d["total"] = mmm(tr[:, idx, 4] - tr[:, idx, 0])
        return d

    for k, nm in enumerate(names):
        acc = {}
        for l in range(1 if nl > 1 else 0, nl):
            for kk, v in summarize(5 * l + k, nm != "attn").items():
                acc.setdefault(kk, []).append(v)
        rep["phases"][nm] = {kk: ([round(sum(x[j] for x in v) / len(v), 2) for j in range(3)] if isinstance(v[0], list) else v[0])
                             for kk, v in acc.items()}
    rep["phases"]["output"] = summarize(n_ph - 1, True)
    # timeline: when the CTAs enter each phase (mean over CTAs, relative to their own entry into phase 0)
    t_in = (tr[:, :, 0] - tr[:, :1, 0]).mean(axis=0)
    t_end = float((tr[:, n_ph - 1, 3] - tr[:, 0, 0]).mean())
    per_layer = [float(t_in[5 * (l + 1)] - t_in[5 * l]) if 5 * (l + 1) < n_ph else None for l in range(nl)]
    rep["layer_us_timeline"] = [round(x, 2) for x in per_layer if x is not None]
    rep["phase_entry_gaps_us"] = {nm: round(float(sum((t_in[5 * l + k + 1] - t_in[5 * l + k]) for l in range(1, nl)) / max(nl - 1, 1)), 2)
                                  for k, nm in enumerate(names)}   # entry of a phase to entry of the next one
    rep["output_phase_us_timeline"] = round(t_end - float(t_in[n_ph - 1]), 2)
    rep["layer_us"] = round(sum(rep["phases"][nm]["total"][0] for nm in names), 2)
    print(json.dumps({k: v for k, v in rep.items() if k != "phases"}))
    for nm, d in rep["phases"].items():
        print(nm, json.dumps(d))
    if args.out:
        open(args.out, "w").write(json.dumps(rep, indent=1))


if __name__ == "__main__":
    main()
