"""Staged diagnostic for the tensor-parallel decode path (run under torch.distributed.run with 2+ ranks):
each stage prints a line when it completes, faulthandler dumps the stack if a stage stalls."""
import faulthandler
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def stage(msg, t0=[time.time()]):
    torch.cuda.synchronize()
    print(f"[rank {os.environ.get('RANK')}] {msg}  (+{time.time() - t0[0]:.1f}s)", flush=True)
    t0[0] = time.time()
    faulthandler.cancel_dump_traceback_later()
    faulthandler.dump_traceback_later(40, exit=True)


def main():
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    faulthandler.dump_traceback_later(60, exit=True)
    dist.init_process_group("nccl", device_id=dev)
    stage("init_process_group")
    t = torch.ones(4096, device=dev, dtype=torch.bfloat16)
    dist.all_reduce(t)
    stage(f"eager all_reduce -> {t[0].item()}")
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        dist.all_reduce(t)
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    with torch.cuda.graph(g, capture_error_mode="thread_local"):
        dist.all_reduce(t)
    stage("captured all_reduce")
    g.replay()
    stage(f"replayed all_reduce -> {t[0].item()}")

    from torchao_fork_b200._models.llama import Generator, ModelArgs
    from torchao_fork_b200._models.llama.tp import TPContext, build_tp_transformer

    ctx = TPContext(rank, dist.get_world_size(), dev)
    name = sys.argv[1] if len(sys.argv) > 1 else "tiny-llama3"
    cfg = ModelArgs.from_name(name)
    model = build_tp_transformer(cfg, ctx, dev, dict(quant="int4", group_size=128), seed=11)
    stage("built model shard")
    gen = Generator(model, 1, 64, temperature=0.0, top_k=None, use_graph=False, tp=ctx)
    stage("engine (no graph)")
    prompt = torch.randint(0, cfg.vocab_size, (1, 8), generator=torch.Generator().manual_seed(1)).to(dev)
    logits = model(prompt, torch.arange(8, device=dev, dtype=torch.int32))
    stage(f"eager prefill {tuple(logits.shape)}")
    gen.engine.set_state(prompt[:, -1], 8)
    gen.engine.step()
    stage("one eager decode step")
    gen.engine.capture()
    stage("captured decode step")
    for _ in range(5):
        gen.engine.step()
    stage(f"5 graph steps, token {gen.engine.next_tok.tolist()}")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        gen.engine.step()
    e1.record()
    torch.cuda.synchronize()
    stage(f"20 graph steps: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us per step")
    gen.engine.graph = None
    gen.engine.graph_stream = None
    del gen, g
    import gc
    gc.collect()
    stage("graphs released")
    dist.barrier()
    stage("barrier")
    dist.destroy_process_group()
    print(f"[rank {rank}] destroyed process group", flush=True)
    faulthandler.cancel_dump_traceback_later()


if __name__ == "__main__":
    main()
