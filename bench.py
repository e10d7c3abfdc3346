"""bench.py — Llama-3-8B int4 (group_size=128) weight-only decode on B200: tok/s + roofline of the GEMV.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload ...]

A "step" is one decoded token at batch size 1: one pass of the hot path (129 weight-only quantized GEMVs
+ the decode glue) over the whole random-init model.  Weights 3.99 GB >> the 126 MB L2, so every step
streams them from HBM (no L2 flush needed).  Prints ONE JSON line (see the task contract):
  value      decode tok/s with the token on the device (CUDA-graph replays, CUDA-event timed)
  e2e        the same through the public Generator.decode_tokens(stream_to_host=True): each step copies the
             token to feed from pinned host memory and reads the sampled token back to pinned host memory
  roofline   the dominant kernel (int4 GEMV of the fused w1|w3 projection, 28672x4096) timed live with
             CUDA events over a graph of that launch on the 32 layers' distinct weights
  cpu_baseline  the unmodified reference (oracle/_ref) Int4CPULayout path on the host cores, bounded sample
--impl reference times that CPU path alone (rank 0 only).  N > 1: tensor parallel (tp.py), strong scaling.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

# stdout carries ONE JSON line.  Libraries print to fd 1 too (NCCL's version banner under NCCL_DEBUG=VERSION, which the
# driver sets and this script leaves alone): keep the real stdout on a private descriptor for the JSON line and point
# fd 1 at stderr for everything else.
_JSON_OUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)


def emit(line: dict):
    _JSON_OUT.write(json.dumps(line) + "\n")
    _JSON_OUT.flush()


ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    "llama3-8b-int4": dict(model="Llama-3-8B", quant="int4", group_size=128),
    "llama3-8b-int8": dict(model="Llama-3-8B", quant="int8", group_size=None),
    "llama3-8b-bf16": dict(model="Llama-3-8B", quant="bf16", group_size=None),
    "llama3-70b-int4": dict(model="Llama-3-70B", quant="int4", group_size=128),
    "tiny-int4": dict(model="tiny-llama3", quant="int4", group_size=128),
}
# reference: torchao/_models/llama/model.py:104-113 (Llama-3-8B); Llama-3-70B = Llama-3.1-70B without scaled rope
MODEL_DIMS = {
    "Llama-3-8B": dict(dim=4096, intermediate_size=14336, n_layer=32, n_head=32, n_local_heads=8, head_dim=128, vocab_size=128256),
    "Llama-3-70B": dict(dim=8192, intermediate_size=28672, n_layer=80, n_head=64, n_local_heads=8, head_dim=128, vocab_size=128256),
    "tiny-llama3": dict(dim=1024, intermediate_size=2048, n_layer=2, n_head=8, n_local_heads=2, head_dim=128, vocab_size=1024),
}
PROMPT_LEN = 8  # reference default prompt "Hello, my name is" ~ 6-8 tokens (generate.py:260)


def int4_gemv_bytes(m, n, k, g):
    """algorithmic bytes of one int4 weight-only GEMM launch"""
    return n * k // 2 + (k // g) * n * 4 + m * k * 2 + m * n * 2


# --------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples SM clock + throttle reasons during the timed region (pynvml; nvidia-smi fallback)."""

    def __init__(self, index=0, period=0.05):
        self.index, self.period = index, period
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {}
        if nv is not None:
            for nm in dir(nv):
                if nm.startswith("nvmlClocksThrottleReason") or nm.startswith("nvmlClocksEventReason"):
                    v = getattr(nv, nm)
                    if isinstance(v, int) and v:
                        names[v] = nm.replace("nvmlClocksThrottleReason", "").replace("nvmlClocksEventReason", "")
        while not self._stop.is_set():
            try:
                if nv is not None:
                    self.samples.append(nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM))
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
                    for bit, nm in names.items():
                        if mask & bit and bin(bit).count("1") == 1:
                            self.reasons.add(nm)
                else:
                    out = subprocess.run(["nvidia-smi", f"--id={self.index}", "--query-gpu=clocks.sm,clocks.max.sm,"
                                          "clocks_event_reasons.active", "--format=csv,noheader,nounits"],
                                         capture_output=True, text=True, timeout=5).stdout.strip().split(",")
                    self.samples.append(int(out[0]))
                    self.max_mhz = int(out[1])
                    self.reasons.add(out[2].strip())
            except Exception:
                pass
            self._stop.wait(self.period)

    def __enter__(self):
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._thread.join(timeout=5)

    def summary(self):
        s = [x for x in self.samples if x]
        idle = {"GpuIdle", "None", "ApplicationsClocksSetting"}
        return {"sm_mhz": statistics.median(s) if s else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(r for r in self.reasons if r not in idle), "samples": len(s)}


# --------------------------------------------------------------------------------------- CPU reference arm
def cpu_reference(workload, steps, warmup, budget_s=25.0):
    """Times the UNMODIFIED reference (oracle/_ref = pip install --target of /root/reference) on the host
    cores: quantize_(Int4WeightOnlyConfig(group_size, layout=Int4CPULayout())) -> aten._weight_int4pack_mm_for_cpu
    (torchao/dtypes/uintx/int4_cpu_layout.py:281-316) for the five linears of ONE Llama transformer block at
    M=1, every layer of a token walked over distinct blocks, + the output projection: one step = one token's linears."""
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    kind = "reference"
    if not os.path.isdir(os.path.join(ref_dir, "torchao")):
        raise RuntimeError("oracle/_ref (the reference install) is missing; run __graft_entry__.build() where /root/reference exists")
    sys.path.insert(0, ref_dir)
    import torch
    import warnings
    warnings.filterwarnings("ignore")
    from torchao.dtypes import Int4CPULayout  # the reference, not this repo
    from torchao.quantization import Int4WeightOnlyConfig, Int8WeightOnlyConfig, quantize_

    from types import SimpleNamespace

    spec = WORKLOADS[workload]
    cfg = SimpleNamespace(**MODEL_DIMS[spec["model"]])  # no import of this repo's package in the reference arm
    cores = os.cpu_count() or 1
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:
        pass
    torch.set_num_threads(cores)
    torch.manual_seed(1234)
    D, I = cfg.dim, cfg.intermediate_size
    qkv = (cfg.n_head + 2 * cfg.n_local_heads) * cfg.head_dim
    shapes = [(qkv, D), (D, D), (I, D), (I, D), (D, I)]  # wqkv, wo, w1, w3, w2  (model.py:422-425,481-483)
    block_w = sum(n * k for n, k in shapes)
    wbytes = {"int4": 0.5, "int8": 1.0}.get(spec["quant"], 2.0)
    # distinct blocks: at least 1 GB of packed weights, so a token's walk over the layers misses the host's last-level
    # cache the way a real model does (one block of Llama-3-8B int4 is 109 MB: it would sit in a 105 MB L3)
    n_distinct = max(1, min(cfg.n_layer, int(-(-1.0e9 // (block_w * wbytes)))))

    def quantized(mods):
        if spec["quant"] == "int4":
            quantize_(mods, Int4WeightOnlyConfig(group_size=spec["group_size"], layout=Int4CPULayout()))
        elif spec["quant"] == "int8":
            quantize_(mods, Int8WeightOnlyConfig())
        return mods

    t0 = time.perf_counter()
    blocks = [quantized(torch.nn.ModuleList([torch.nn.Linear(k, n, bias=False, dtype=torch.bfloat16) for n, k in shapes]))
              for _ in range(n_distinct)]
    output = quantized(torch.nn.ModuleList([torch.nn.Linear(D, cfg.vocab_size, bias=False, dtype=torch.bfloat16)]))[0]
    t_quant = time.perf_counter() - t0
    xs = [torch.randn(1, k, dtype=torch.bfloat16) for n, k in shapes]

    def one_token():
        with torch.no_grad():
            for layer in range(cfg.n_layer):
                for lin, x in zip(blocks[layer % n_distinct], xs):
                    lin(x)
            output(xs[0])

    for _ in range(max(min(warmup, 5), 2)):
        one_token()
    times = []
    t_start = time.perf_counter()
    for _ in range(max(steps, 1)):
        t0 = time.perf_counter()
        one_token()
        times.append(time.perf_counter() - t0)
        if time.perf_counter() - t_start > budget_s:
            break
    t_token = statistics.median(times)
    path_name = {"int4": "Int4CPULayout path (_weight_int4pack_mm_for_cpu)",
                 "int8": "Int8WeightOnlyConfig PlainLayout path (mm(x, int8.to(bf16)) * scale on CPU)"}.get(
                     spec["quant"], "unquantized bf16 F.linear")
    return {"value": 1.0 / t_token, "unit": "tok/s", "cores": cores, "kind": kind,
            "sample": f"{len(times)} timed tokens ({sum(times):.1f} s of CPU work); a token = the 5 quantized linears of each "
                      f"of the {cfg.n_layer} transformer blocks + the output projection at M=1 through the reference's "
                      f"{path_name}, cycling over {n_distinct} distinct blocks ({n_distinct * block_w * wbytes / 1e9:.2f} GB "
                      f"packed, larger than the host's last-level cache); median {t_token * 1e3:.2f} ms/token; "
                      f"attention/norms/sampling excluded; reference quantize_ of the blocks took {t_quant:.1f} s",
            "ms_per_token": t_token * 1e3}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cb = cpu_reference(args.workload, args.steps, args.warmup, args.cpu_budget_s)
    spec = WORKLOADS[args.workload]
    line = {"impl": "reference", "metric": "llama3_8b_int4_decode_tok_per_s" if args.workload == "llama3-8b-int4" else args.workload + "_decode_tok_per_s",
            "value": cb["value"], "unit": "tok/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": cb["ms_per_token"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "bf16 activations x int4 weights, fp32 accumulate" if spec["quant"] == "int4" else spec["quant"],
            "data": "synthetic (random-init weights, random activations)",
            "config": {"workload": args.workload, "batch_size": 1, "group_size": spec["group_size"]},
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


# --------------------------------------------------------------------------------------- GPU arm
def build_model(spec, device, tp=None):
    import torch
    import torchao_fork_b200 as ao
    from torchao_fork_b200._models.llama import ModelArgs, Transformer

    cfg = ModelArgs.from_name(spec["model"])
    torch.manual_seed(1234)  # reference: generate.py:325
    old = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    try:
        if (tp is not None and tp.world_size > 1) or cfg.n_layer * cfg.dim * cfg.intermediate_size > 4e10:
            # sharded and / or too large to hold in bf16 next to its quantized copy: layer-streamed builder
            from torchao_fork_b200._models.llama.tp import build_tp_transformer
            model = build_tp_transformer(cfg, tp, device, spec)
        else:
            with torch.device(device):
                model = Transformer(cfg)
            model.eval()
            if spec["quant"] == "int4":
                ao.quantize_(model, ao.Int4WeightOnlyConfig(group_size=spec["group_size"]))
            elif spec["quant"] == "int8":
                ao.quantize_(model, ao.Int8WeightOnlyConfig())
    finally:
        torch.set_default_dtype(old)
    return model, cfg


def time_gemv_launch(engine, which, iters):
    """Average duration of ONE launch of the `which` linear of every layer (distinct weights, > L2 in total),
    replayed from a CUDA graph, CUDA events on the launching stream."""
    import torch
    lins = [l[which] for l in engine.layers if l[which] is not None]
    lin0 = lins[0]
    x = torch.randn(engine.B, lin0.k, dtype=torch.bfloat16, device=engine.device)
    y = torch.empty(engine.B, lin0.n, dtype=torch.bfloat16, device=engine.device)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for lin in lins[:2]:
            lin.launch(x, y, engine.ws, side.cuda_stream)
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        st = torch.cuda.current_stream().cuda_stream
        for lin in lins:
            lin.launch(x, y, engine.ws, st)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / (iters * len(lins)), lin0, sum(l.weight_bytes for l in lins)


def time_mega_kernel(eng, iters):
    """Average duration of the persistent decode kernel ALONE (no sampling): a CUDA graph of `iters` launches (each one
    advances the position, streams all the weights — 4 GB >> L2 — and appends to the caches), CUDA events around it."""
    import ctypes
    import torch
    from torchao_fork_b200 import _lib
    pos0 = int(eng.pos.item())
    iters = max(1, min(iters, eng.S - pos0 - 1))
    L = _lib.lib()

    def launch():
        _lib.check(L.ao_b200_decode_mega_step(ctypes.byref(eng._mega_model), eng._mega_plan.data_ptr(),
                                              torch.cuda.current_stream().cuda_stream), "decode_mega_step")
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        launch()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    eng.pos.fill_(pos0)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(iters):
            launch()
    eng.pos.fill_(pos0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    eng.check_status()
    eng.pos.fill_(pos0)
    return e0.elapsed_time(e1) * 1e-3 / iters, pos0, iters


def time_reference_gpu_kernels(quant, group_size, shapes, n_layer, iters, device):
    """The reference's own GPU kernel on the same shapes, same run: for int4
    aten._weight_int4pack_mm on weights packed by aten._convert_weight_to_int4pack(inner_k_tiles=8)
    (tensor_core_tiled_layout.py:104,279); for int8 the eager formula mm(x, w.t().to(bf16)) * scale (plain_layout.py:256-262).
    One CUDA graph per shape over `n_layer` DISTINCT weights (> L2 in total), M = 1, CUDA events."""
    import torch
    out = {}
    for name, (n, k) in shapes.items():
        layers = n_layer if name != "output" else 2
        if quant == "int8":
            layers = min(layers, 4)   # the eager formula materialises a bf16 copy of the weight per call
        x = torch.randn(1, k, dtype=torch.bfloat16, device=device)
        ws = []
        for _ in range(layers):
            if quant == "int4":
                q = torch.randint(0, 256, (n, k // 2), dtype=torch.uint8, device=device)
                ws.append((torch.ops.aten._convert_weight_to_int4pack(q, 8),
                           torch.randn(k // group_size, n, 2, dtype=torch.bfloat16, device=device)))
            else:
                ws.append((torch.randint(-128, 128, (n, k), dtype=torch.int8, device=device),
                           torch.rand(n, dtype=torch.bfloat16, device=device)))

        def run(w):
            if quant == "int4":
                return torch.ops.aten._weight_int4pack_mm(x, w[0], group_size, w[1])
            return torch.mm(x, w[0].t().to(torch.bfloat16)) * w[1]
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            run(ws[0])
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for w in ws:
                run(w)
        for _ in range(2):
            g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        out[name] = {"n": n, "k": k, "us_per_launch": e0.elapsed_time(e1) * 1e3 / (iters * layers)}
        del ws, g
        torch.cuda.empty_cache()
    return out


def run_b200(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world} (launch with torch.distributed.run)"
    assert torch.cuda.is_available(), "bench.py needs a B200 (there is no CPU fallback path)"
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    tp = None
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
        from torchao_fork_b200._models.llama.tp import TPContext
        tp = TPContext(rank, world, device)

    from torchao_fork_b200 import _lib
    from torchao_fork_b200._models.llama import Generator

    _lib.lib()  # fail loudly now if libao_b200.so is missing
    spec = WORKLOADS[args.workload]
    model, cfg = build_model(spec, device, tp)
    K, W = args.steps, max(args.warmup, 3)
    max_seq = PROMPT_LEN + 2 * (K + W) + 16
    gen = Generator(model, args.batch_size, max_seq, temperature=args.temperature, top_k=args.top_k, tp=tp)
    eng = gen.engine
    prompt = torch.randint(0, cfg.vocab_size, (PROMPT_LEN,), device=device)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- value: tokens stay on the device; K graph replays between CUDA events
    eng.set_state(prompt[-1:].repeat(args.batch_size), PROMPT_LEN)
    for _ in range(W):
        eng.step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clk:
        e0.record()
        for _ in range(K):
            eng.step()
        e1.record()
        barrier()
    t_dev = e0.elapsed_time(e1) * 1e-3
    # ---- e2e: public API, token fed from pinned host memory and read back every step; the SAME positions as `value`
    # (the attention of a step grows with the context: decoding on from where `value` stopped would time a different workload)
    eng.set_state(prompt[-1:].repeat(args.batch_size), PROMPT_LEN)
    for _ in range(W):
        gen.decode_tokens(1, stream_to_host=True)
    barrier()
    t0 = time.perf_counter()
    host_tokens = gen.decode_tokens(K, stream_to_host=True)
    barrier()
    t_e2e = time.perf_counter() - t0
    assert host_tokens.shape == (args.batch_size, K) and int(host_tokens.min()) >= 0
    if world > 1:
        t = torch.tensor([t_dev, t_e2e], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_dev, t_e2e = t.tolist()

    # the persistent kernel alone (every rank launches: with tensor parallelism the kernels wait on one another)
    mega_timing = None
    if eng.mega:
        eng.set_state(prompt[-1:].repeat(args.batch_size), PROMPT_LEN + (W + K) // 2)
        barrier()
        t_k, pos0_k, it_k = time_mega_kernel(eng, max(10, min(50, K // 4)))
        if world > 1:
            t = torch.tensor([t_k], device=device, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            t_k = float(t.item())
        mega_timing = (t_k, pos0_k, it_k)
    toks = K * args.batch_size
    pos_value = (PROMPT_LEN, PROMPT_LEN + W + K)                       # positions decoded in the `value` window
    pos_e2e = pos_value                                                # ... and in the `e2e` window: the same ones
    line = {"metric": "llama3_8b_int4_decode_tok_per_s" if args.workload == "llama3-8b-int4" else args.workload + "_decode_tok_per_s",
            "value": toks / t_dev, "unit": "tok/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": t_dev / K * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": {"int4": "bf16 activations x int4 weights, fp32 accumulate", "int8": "bf16 x int8 weights, fp32 accumulate",
                      "bf16": "bf16"}[spec["quant"]],
            "data": "synthetic (random-init weights, random tokens)",
            "config": {"workload": args.workload, "batch_size": args.batch_size, "group_size": spec["group_size"],
                       "prompt_len": PROMPT_LEN, "sampling": f"temperature={args.temperature}, top_k={args.top_k}",
                       "parallelism": f"tp{world}", "l2": "inputs larger than L2 (weights %.2f GB per step per rank)" % (eng.weight_bytes / 1e9),
                       "engine": "persistent decode kernel, 1 launch + sampling per token" if eng.mega else "per-projection CUDA graph",
                       "context": {"value_positions": list(pos_value), "e2e_positions": list(pos_e2e)},
                       "timing_windows_ms": {"value": t_dev * 1e3, "e2e": t_e2e * 1e3,
                                             "note": "value and e2e are timed one after the other over the SAME positions; "
                                                     "windows under ~100 ms are inside run-to-run noise of a few percent"}},
            "e2e": {"value": toks / t_e2e, "unit": "tok/s", "h2d_bytes_per_step": 8 * args.batch_size,
                    "d2h_bytes_per_step": 8 * args.batch_size},
            "gpu_launches": eng.launches_per_step * K,
            "model_bandwidth_gbs": eng.weight_bytes * world / (t_dev / K) / 1e9,
            "clocks": clk.summary()}

    if rank == 0:
        # ---- roofline of the dominant kernel
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "MEASURED_PEAKS.json hbm_gbs (of measured)"
        else:
            peak, peak_src = 6650.0, "B200_PROFILING.md fallback 6.65 TB/s (of fallback)"
        sm_mhz = line["clocks"].get("sm_mhz") or 1965.0
        shapes = {}
        for nm in ("wqkv", "wo", "w13", "w1", "w3", "w2"):
            if eng.layers[0][nm] is not None:
                shapes[nm] = (eng.layers[0][nm].n, eng.layers[0][nm].k)
        # ---- the per-projection kernels, each in a graph over the 32 layers' distinct weights (standalone launches:
        #      what `ao_b200_int4_mm` / the torchao:: ops deliver for one linear; config 1 = the 4096 x 4096 one)
        standalone, t_all, b_all = {}, 0.0, 0
        for nm in shapes:
            t1, l1, wb = time_gemv_launch(eng, nm, max(5, K // 20))
            alg1 = (int4_gemv_bytes(args.batch_size, l1.n, l1.k, l1.group_size) if l1.kind == "int4"
                    else l1.weight_bytes + args.batch_size * (l1.n + l1.k) * 2)
            standalone[nm] = {"n": l1.n, "k": l1.k, "us_per_launch": t1 * 1e6, "gbs": alg1 / t1 / 1e9, "frac": alg1 / t1 / 1e9 / peak}
            t_all += t1 * len(eng.layers)
            b_all += wb
        line["standalone_gemv_kernels"] = {"per_shape": standalone, "all_layers": {"gbs": b_all / t_all / 1e9, "frac": b_all / t_all / 1e9 / peak},
                                           "how": "each linear's own kernel launched back to back over the 32 layers' weights in a CUDA graph (PDL chain), CUDA events"}
        if "wo" in standalone and standalone["wo"]["n"] == 4096 and standalone["wo"]["k"] == 4096:
            line["config1"] = dict(standalone["wo"], workload="single Linear 4096x4096, Int4WeightOnlyConfig(group_size=128), M=1 (BASELINE.json configs[0])")
        traffic = None
        tp_path = os.path.join(ROOT, "profiles", "traffic.json")
        tr = json.load(open(tp_path)) if os.path.exists(tp_path) else {}
        if eng.mega:
            # ---- the dominant kernel IS the step: the persistent decode kernel, timed alone
            t_k, pos0, it_k = mega_timing
            kv_bytes = 2 * len(eng.layers) * eng.Hkv * 128 * 2 * (pos0 + it_k // 2)      # K and V rows read per token
            alg = eng.weight_bytes + kv_bytes
            if tr.get("kernel") == "decode_mega":
                traffic = tr.get("dram_bytes_per_launch")
            line["roofline"] = {"bound": "hbm", "achieved": alg / t_k / 1e9, "peak": peak, "unit": "GB/s",
                                "frac": alg / t_k / 1e9 / peak, "traffic": traffic,
                                "kernel": f"decode_mega_kernel<{eng.output.group_size}>: one launch = one decoded token (all 129 int4 linears, norms, RoPE, attention)",
                                "algorithmic_bytes_per_launch": alg, "us_per_launch": t_k * 1e6, "peak_source": peak_src,
                                "context_positions": [pos0, pos0 + it_k]}
            line["roofline"]["note"] = "per rank" if world > 1 else "single GPU"
            # ---- inside the kernel: the streaming loops of the projections (clock64 stamps of every CTA)
            cyc = eng.mega_phase_cycles() if world == 1 else None
            if cyc is not None:
                phases, t_sum, b_sum = {}, 0.0, 0
                for nm, key in (("wqkv", "wqkv"), ("wo", "wo"), ("w13", "w13"), ("w2", "w2")):
                    main = cyc[key]["main"] / sm_mhz                       # us [layers][ctas]
                    t_phase = float(main.max(axis=1).mean())               # a phase lasts until its slowest CTA is done
                    lin = eng.layers[0][nm]
                    alg1 = int4_gemv_bytes(1, lin.n, lin.k, lin.group_size)
                    phases[nm] = {"us": t_phase, "us_mean_cta": float(main.mean()), "gbs": alg1 / t_phase / 1e3, "frac": alg1 / t_phase / 1e3 / peak}
                    t_sum += t_phase
                    b_sum += alg1
                om = cyc["output"]["main"] / sm_mhz
                alg_o = int4_gemv_bytes(1, eng.output.n, eng.output.k, eng.output.group_size)
                phases["output"] = {"us": float(om.max()), "gbs": alg_o / float(om.max()) / 1e3, "frac": alg_o / float(om.max()) / 1e3 / peak}
                over = {nm: float((cyc[nm]["wait"] + cyc[nm].get("staging", cyc[nm].get("phase"))).mean() / sm_mhz) for nm in ("wqkv", "attn", "wo", "w13", "w2")}
                line["roofline"]["all_layer_gemvs"] = {
                    "gbs": b_sum / t_sum / 1e3, "frac": b_sum / t_sum / 1e3 / peak, "share_of_kernel": t_sum * len(eng.layers) / (t_k * 1e6),
                    "per_projection": phases, "hand_over_us_per_layer": over,
                    "how": "streaming (main) loops of the 4 projections of a layer inside the persistent kernel, per phase the "
                           "slowest CTA, mean over layers 1..31; clock64 stamps converted at the sampled SM clock "
                           f"({sm_mhz:.0f} MHz); the ring prefetches the next projection's weights during the hand-over, so a "
                           "loop can read above the HBM rate"}
        else:
            which = "w13" if eng.layers[0]["w13"] is not None else "w2"
            st = standalone[which]
            lin = eng.layers[0][which]
            for ent in (tr, tr.get("gemv_int4") or {}, tr.get("int8") or {}):   # per-kernel entries of profiles/traffic.json
                if ent.get("n") == lin.n and ent.get("k") == lin.k and ent.get("kind") == lin.kind:
                    traffic = ent.get("dram_bytes_per_launch")
            alg = st["gbs"] * 1e9 * st["us_per_launch"] * 1e-6
            line["roofline"] = {"bound": "hbm", "achieved": st["gbs"], "peak": peak, "unit": "GB/s", "frac": st["frac"], "traffic": traffic,
                                "kernel": f"{lin.kind} weight-only GEMV {lin.n}x{lin.k} (fused w1|w3), M={args.batch_size}",
                                "algorithmic_bytes_per_launch": alg, "us_per_launch": st["us_per_launch"], "peak_source": peak_src,
                                "all_layer_gemvs": {"gbs": b_all / t_all / 1e9, "frac": b_all / t_all / 1e9 / peak, "share_of_step": t_all / (t_dev / K)}}
        # BASELINE metric (ii) "decode GEMMs vs the HBM peak" over ALL layer GEMVs, next to the per-kernel number
        alg_all = line["roofline"].get("all_layer_gemvs")
        if alg_all is not None:
            line["roofline"]["frac_all_layer_gemvs"] = alg_all["frac"]
            line["all_layer_gemvs_frac"] = alg_all["frac"]
        # ---- the reference's GPU kernel on the same shapes, same run
        if spec["quant"] in ("int4", "int8") and world == 1 and not args.no_reference_gpu:
            try:
                ref_shapes = {"wqkv": shapes["wqkv"], "wo": shapes["wo"], "w1": (cfg.intermediate_size, cfg.dim), "w3": (cfg.intermediate_size, cfg.dim),
                              "w2": shapes["w2"], "output": (eng.output.n, eng.output.k)}
                rg = time_reference_gpu_kernels(spec["quant"], spec["group_size"], ref_shapes, len(eng.layers), 5, device)
                t_layer = sum(rg[nm]["us_per_launch"] for nm in ("wqkv", "wo", "w1", "w3", "w2"))
                t_tok = t_layer * len(eng.layers) + rg["output"]["us_per_launch"]
                line["reference_gpu"] = {
                    "kernel": "aten._weight_int4pack_mm (weights by aten._convert_weight_to_int4pack, inner_k_tiles=8)" if spec["quant"] == "int4"
                              else "eager mm(x, w.t().to(bf16)) * scale (plain_layout.py:256-262)",
                    "M": 1, "us_per_launch": {nm: v["us_per_launch"] for nm, v in rg.items()},
                    "all_linears_us_per_token": t_tok, "all_layer_gbs": eng.weight_bytes / (t_tok * 1e-6) / 1e9,
                    "implied_tok_s_linears_only": 1e6 / t_tok,
                    "this_repo_standalone_us_per_launch": {nm: v["us_per_launch"] for nm, v in standalone.items()}}
            except Exception as e:
                line["reference_gpu"] = {"unavailable": f"{type(e).__name__}: {e}"}
        if world == 1 and not args.no_cpu_baseline:
            # separate process: the reference's torchao and this repo's ops both own the `torchao::` op namespace
            try:
                out = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--workload",
                                      args.workload, "--steps", "1000", "--warmup", "3", "--cpu-budget-s", "12"], capture_output=True, text=True,
                                     timeout=600, env={k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")})
                ref_line = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][-1])
                line["cpu_baseline"] = ref_line["cpu_baseline"]
            except Exception as e:  # the checker is optional at run time; the product path is not
                line["cpu_baseline"] = {"value": None, "unit": "tok/s", "cores": 0, "kind": "reference",
                                        "sample": f"unavailable: {type(e).__name__}: {e}"}
        emit(line)
    if world > 1:
        # tear-down: the graph holds NCCL kernels, release it before the communicator; exit without
        # destroy_process_group (it can block on a communicator that captured graphs have used)
        eng.graph = None
        eng.graph_stream = None
        torch.cuda.synchronize()
        dist.barrier()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)  # reference: 200 new tokens per sample (generate.py:264)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="llama3-8b-int4", choices=sorted(WORKLOADS))
    ap.add_argument("--batch-size", type=int, default=1)
    ap.add_argument("--temperature", type=float, default=0.8)
    ap.add_argument("--top-k", type=int, default=200)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-budget-s", type=float, default=25.0, help="reference arm: stop timing tokens after this many seconds")
    ap.add_argument("--no-reference-gpu", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
