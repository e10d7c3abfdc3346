"""Where does the host-visible step (bench.py `e2e`) lose time against back-to-back graph replays?  Times, per token:
replay() on the host, the wait for the stream, and the Python bookkeeping.  Usage: python benchmarks/e2e_probe.py"""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

spec = bench.WORKLOADS["llama3-8b-int4"]
dev = torch.device("cuda", 0)
model, cfg = bench.build_model(spec, dev)
from torchao_fork_b200._models.llama import Generator  # noqa: E402

gen = Generator(model, 1, 512)
eng = gen.engine
eng.set_state(torch.tensor([1], device=dev), 8)
gen.decode_tokens(4, stream_to_host=True)
stream = torch.cuda.current_stream(dev)
N = 200
t_replay = t_sync = t_py = 0.0
h_tok, h_out = gen.h_tok.numpy(), gen.h_out.numpy()
torch.cuda.synchronize()
t_all0 = time.perf_counter()
for i in range(N):
    t0 = time.perf_counter()
    eng.graph_stream.replay()
    t1 = time.perf_counter()
    stream.synchronize()
    t2 = time.perf_counter()
    h_tok[:] = h_out
    t3 = time.perf_counter()
    t_replay += t1 - t0
    t_sync += t2 - t1
    t_py += t3 - t2
t_all = time.perf_counter() - t_all0
# device-only: back-to-back replays
eng.set_state(torch.tensor([1], device=dev), 8)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(N):
    eng.graph.replay()
e1.record()
torch.cuda.synchronize()
# one replay + sync, event-timed on the device
eng.set_state(torch.tensor([1], device=dev), 8)
dev_single = 0.0
for i in range(50):
    e0.record()
    eng.graph_stream.replay()
    e1.record()
    stream.synchronize()
    dev_single += e0.elapsed_time(e1)
print(json.dumps({"per_token_us": {"host_loop_total": t_all / N * 1e6, "replay_call": t_replay / N * 1e6,
                                   "wait_for_stream": t_sync / N * 1e6, "python": t_py / N * 1e6,
                                   "device_back_to_back": 0, "device_one_replay_between_events": dev_single / 50 * 1e3},
                  "nodes_per_step": eng.launches_per_step}))
