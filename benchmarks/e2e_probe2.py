"""Host-visible latency of ONE graph replay + stream sync for variants of the decode graph (which node costs the serving
loop its extra per-token latency?).  python benchmarks/e2e_probe2.py"""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from torchao_fork_b200._models.llama import Generator  # noqa: E402

spec = bench.WORKLOADS["llama3-8b-int4"]
dev = torch.device("cuda", 0)
model, cfg = bench.build_model(spec, dev)
stream = torch.cuda.current_stream(dev)
out = {}
for name, kw in [("mega_sample", dict(mega=True, temperature=0.8, top_k=200)), ("mega_greedy", dict(mega=True, temperature=0.0, top_k=None)),
                 ("proj_sample", dict(mega=False, temperature=0.8, top_k=200))]:
    gen = Generator(model, 1, 512, **kw)
    eng = gen.engine
    eng.set_state(torch.tensor([1], device=dev), 8)
    gen.decode_tokens(4, stream_to_host=True)
    res = {}
    for label, g in (("step_graph", eng.graph), ("streaming_graph", eng.graph_stream)):
        eng.set_state(torch.tensor([1], device=dev), 8)
        torch.cuda.synchronize()
        N = 100
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t_host = t_dev = 0.0
        for i in range(N):
            t0 = time.perf_counter()
            e0.record()
            g.replay()
            e1.record()
            stream.synchronize()
            t_host += time.perf_counter() - t0
            t_dev += e0.elapsed_time(e1)
        res[label] = {"host_us": t_host / N * 1e6, "device_events_us": t_dev / N * 1e3}
    out[name] = res
    del gen, eng
print(json.dumps(out, indent=1))
