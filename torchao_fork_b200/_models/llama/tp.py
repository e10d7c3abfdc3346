"""Tensor-parallel wrapper for the gpt-fast Llama: column / row sharded weight-only-quantized linears over the GPUs of
one NVSwitch box, one process per GPU, an all-reduce only after the two row-parallel projections of a block.

The reference has no working equivalent for int4: its DTensor TP test of Int4WeightOnlyConfig is skipped
(test/dtypes/test_affine_quantized_tensor_parallel.py:153-158), and torchao/_models/llama has no tp.py.  The
partitioning is the Megatron / gpt-fast one, applied to the bf16 weights BEFORE quantize_, so every rank quantizes its
own shard and quantization groups never straddle a shard boundary:

  column-parallel (split output rows N): wqkv — by head groups: rank r takes q heads [r H/tp, (r+1) H/tp) and kv heads
      [r Hkv/tp, (r+1) Hkv/tp) out of the fused [q|k|v] rows (model.py:422-424); w1, w3; output (vocab rows, logits
      all-gathered)
  row-parallel (split input columns K): wo, w2 — partial sums all-reduced (sum) over the group

Everything else (embedding, norms, rope table) is replicated; the KV cache and attention are per-rank local over the
rank's own heads.  Results differ from tp=1 only by the bf16 rounding of the row-parallel partial sums before the
all-reduce.

`TPContext` is the collective plumbing (torch.distributed: NCCL on GPUs, gloo in the CPU tests), `apply_tp` shards an
existing bf16 model in place and installs the eager-mode hooks, `build_tp_transformer` builds a random-init model
shard by shard (layer-streamed: the full bf16 Llama-3-70B is never materialised) and quantizes as it goes.
"""
import math
import os
from typing import Optional

import torch
import torch.distributed as dist
import torch.nn as nn

from .model import ModelArgs, Transformer


class PeerExchange:
    """NVLink peer-memory exchange buffers of one rank (include/ao_b200.h, ao_b200_tp_*): allocated by the C library,
    IPC handles swapped over the process group's host channel, peers' buffers opened once.  `allreduce_add` is one
    kernel launch: h += sum over ranks of partial (csrc/tp_comm.cu)."""

    def __init__(self, rank: int, world_size: int, max_elems: int, device, group=None):
        import ctypes

        from ... import _lib

        self._lib, self._ct = _lib, ctypes
        L = _lib.lib()
        self.rank, self.world_size, self.max_elems = rank, world_size, int(max_elems)
        nbytes = int(L.ao_b200_tp_buffer_bytes(world_size, self.max_elems))
        if nbytes == 0:
            raise ValueError(f"peer exchange does not cover world={world_size}, max_elems={max_elems}")
        with torch.cuda.device(device):
            own, handle = ctypes.c_void_p(), ctypes.create_string_buffer(64)
            _lib.check(L.ao_b200_tp_alloc(nbytes, ctypes.byref(own), handle), "tp_alloc")
            handles = [None] * world_size
            dist.all_gather_object(handles, bytes(handle.raw), group=group)
            self.own = own.value
            self.bufs = (ctypes.c_void_p * world_size)()
            self._opened = []
            for r, hnd in enumerate(handles):
                if r == rank:
                    self.bufs[r] = self.own
                else:
                    peer = ctypes.c_void_p()
                    _lib.check(L.ao_b200_tp_open(ctypes.create_string_buffer(hnd, 64), ctypes.byref(peer)), f"tp_open(rank {r})")
                    self.bufs[r] = peer.value
                    self._opened.append(peer.value)
        dist.barrier(group=group)  # nobody pushes before everybody has zeroed and published its buffer

    def allreduce_add(self, partial: torch.Tensor, h: torch.Tensor):
        assert partial.is_contiguous() and h.is_contiguous() and partial.dtype == h.dtype == torch.bfloat16
        assert partial.numel() == h.numel()
        self._lib.check(self._lib.lib().ao_b200_tp_allreduce_add(
            partial.data_ptr(), h.data_ptr(), partial.numel(), self.bufs, self.rank, self.world_size, self.max_elems,
            torch.cuda.current_stream(partial.device).cuda_stream), "tp_allreduce_add")
        return h

    def status(self) -> int:
        """0 = ok, 1 = a peer never arrived within the kernel's spin bound (results are garbage)."""
        st = self._ct.c_int(0)
        self._lib.check(self._lib.lib().ao_b200_tp_status(self.own, self._ct.byref(st)), "tp_status")
        return st.value

    def close(self):
        L = self._lib.lib()
        for p in self._opened:
            L.ao_b200_tp_close(p)
        self._opened = []
        if self.own:
            L.ao_b200_tp_free(self.own)
            self.own = None


class TPContext:
    """Rank / world size of the tensor-parallel group and the three collectives the model needs."""

    def __init__(self, rank: Optional[int] = None, world_size: Optional[int] = None, device=None, group=None):
        self.group = group
        self.rank = dist.get_rank(group) if rank is None else int(rank)
        self.world_size = dist.get_world_size(group) if world_size is None else int(world_size)
        self.device = device
        self._gather = {}
        self.peer: Optional[PeerExchange] = None
        assert 0 <= self.rank < self.world_size

    def enable_peer_exchange(self, max_elems: int):
        """Decode path: replace NCCL all-reduce + add by the one-launch NVLink peer-memory kernel for tensors of up to
        max_elems bf16 elements.  AO_B200_TP_PEER=0 keeps NCCL (A/B comparison)."""
        if self.world_size > 1 and self.peer is None and os.environ.get("AO_B200_TP_PEER", "1") != "0":
            self.peer = PeerExchange(self.rank, self.world_size, max_elems, self.device, self.group)
        return self.peer

    @classmethod
    def from_env(cls, backend: Optional[str] = None, device=None):
        """Joins the process group described by RANK / WORLD_SIZE / MASTER_* (torch.distributed.run)."""
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29500")
            backend = backend or ("nccl" if torch.cuda.is_available() else "gloo")
            kw = {"device_id": device} if (backend == "nccl" and device is not None) else {}
            dist.init_process_group(backend, **kw)
        return cls(device=device)

    # -- collectives (all of them stream-ordered and CUDA-graph capturable under NCCL)
    def all_reduce(self, t: torch.Tensor) -> torch.Tensor:
        if self.world_size > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t

    def all_reduce_add(self, delta: torch.Tensor, h: torch.Tensor) -> torch.Tensor:
        """h += sum over ranks of delta  (the residual add that follows a row-parallel projection)."""
        if self.peer is not None and delta.is_cuda and delta.numel() <= self.peer.max_elems and delta.numel() % 8 == 0:
            return self.peer.allreduce_add(delta, h)
        self.all_reduce(delta)
        return h.add_(delta)

    def gather_logits(self, local: torch.Tensor) -> torch.Tensor:
        """[.., V / tp] on every rank -> [.., V] on every rank (vocab shards in rank order)."""
        if self.world_size == 1:
            return local
        lead, vl = local.shape[:-1], local.shape[-1]
        key = (tuple(lead), vl, local.dtype, local.device)
        bufs = self._gather.get(key)
        if bufs is None:  # fixed addresses: the decode graph replays over them
            bufs = (torch.empty((self.world_size,) + tuple(lead) + (vl,), dtype=local.dtype, device=local.device),
                    torch.empty(tuple(lead) + (self.world_size * vl,), dtype=local.dtype, device=local.device))
            self._gather[key] = bufs
        parts, full = bufs
        # (world * rows, vl) <- (rows, vl): the concatenated form every backend accepts
        dist.all_gather_into_tensor(parts.view(-1, vl), local.contiguous().view(-1, vl), group=self.group)
        full.view(tuple(lead) + (self.world_size, vl)).copy_(parts.movedim(0, -2))
        return full


# ----------------------------------------------------------------------------------------------------------------------
# shard arithmetic (pure tensor slicing: unit-tested on the CPU)
# ----------------------------------------------------------------------------------------------------------------------
def shard_rows(w: torch.Tensor, rank: int, tp: int) -> torch.Tensor:
    n = w.shape[0]
    assert n % tp == 0, f"{n} rows do not split over {tp} ranks"
    return w[rank * (n // tp):(rank + 1) * (n // tp)]


def shard_cols(w: torch.Tensor, rank: int, tp: int) -> torch.Tensor:
    k = w.shape[1]
    assert k % tp == 0, f"{k} columns do not split over {tp} ranks"
    return w[:, rank * (k // tp):(rank + 1) * (k // tp)]


def shard_wqkv(w: torch.Tensor, n_head: int, n_kv: int, head_dim: int, rank: int, tp: int) -> torch.Tensor:
    """Fused [q | k | v] rows -> this rank's [q_r | k_r | v_r] (whole heads; q heads stay with their kv head)."""
    assert n_head % tp == 0 and n_kv % tp == 0, f"{n_head} q / {n_kv} kv heads do not split over {tp} ranks"
    assert w.shape[0] == (n_head + 2 * n_kv) * head_dim
    q, k, v = w.split([n_head * head_dim, n_kv * head_dim, n_kv * head_dim], dim=0)
    return torch.cat([shard_rows(q, rank, tp), shard_rows(k, rank, tp), shard_rows(v, rank, tp)], dim=0)


def check_divisible(cfg: ModelArgs, tp: int, group_size: Optional[int] = None):
    assert cfg.n_head % tp == 0 and cfg.n_local_heads % tp == 0, "heads must split evenly over the ranks"
    assert cfg.intermediate_size % tp == 0 and cfg.vocab_size % tp == 0 and cfg.dim % tp == 0
    if group_size:
        # row-parallel shards keep whole quantization groups
        assert (cfg.intermediate_size // tp) % group_size == 0 and (cfg.n_head * cfg.head_dim // tp) % group_size == 0


def _set_linear(lin: nn.Linear, w: torch.Tensor):
    lin.weight = nn.Parameter(w.contiguous(), requires_grad=False)
    lin.out_features, lin.in_features = int(w.shape[0]), int(w.shape[1])


def _install_hooks(model: Transformer, ctx: TPContext):
    """Eager-mode collectives (prefill): all-reduce the outputs of attention and feed_forward (their last linear is
    row-parallel), all-gather the logits.  gpt-fast's tp.py does the same with forward hooks."""
    if ctx.world_size == 1:
        return
    for blk in model.layers:
        blk.attention.register_forward_hook(lambda _m, _i, out: ctx.all_reduce(out))
        blk.feed_forward.register_forward_hook(lambda _m, _i, out: ctx.all_reduce(out))
    model.output.register_forward_hook(lambda _m, _i, out: ctx.gather_logits(out))


def apply_tp(model: Transformer, ctx: TPContext) -> Transformer:
    """Shards an unquantized model in place (call BEFORE quantize_) and installs the eager collectives."""
    cfg, r, tp = model.config, ctx.rank, ctx.world_size
    check_divisible(cfg, tp)
    for blk in model.layers:
        att, ff = blk.attention, blk.feed_forward
        _set_linear(att.wqkv, shard_wqkv(att.wqkv.weight.detach(), cfg.n_head, cfg.n_local_heads, cfg.head_dim, r, tp))
        _set_linear(att.wo, shard_cols(att.wo.weight.detach(), r, tp))
        att.n_head, att.n_local_heads = cfg.n_head // tp, cfg.n_local_heads // tp
        _set_linear(ff.w1, shard_rows(ff.w1.weight.detach(), r, tp))
        _set_linear(ff.w3, shard_rows(ff.w3.weight.detach(), r, tp))
        _set_linear(ff.w2, shard_cols(ff.w2.weight.detach(), r, tp))
    _set_linear(model.output, shard_rows(model.output.weight.detach(), r, tp))
    _install_hooks(model, ctx)
    return model


# ----------------------------------------------------------------------------------------------------------------------
# layer-streamed random-init builder
# ----------------------------------------------------------------------------------------------------------------------
def _seed_of(seed: int, layer: int, name: str) -> int:
    return (seed * 1000003 + (layer + 1) * 7919 + sum(ord(c) * (i + 1) for i, c in enumerate(name))) % (2 ** 31 - 1)


def _full_linear_weight(n: int, k: int, seed: int, device, dtype) -> torch.Tensor:
    """nn.Linear's default init (kaiming_uniform(a = sqrt 5) = U(-1/sqrt K, 1/sqrt K)) from a per-tensor seed, so that
    every tp degree shards the SAME full weight."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    b = 1.0 / math.sqrt(k)
    return torch.empty(n, k, dtype=torch.float32, device=device).uniform_(-b, b, generator=g).to(dtype)


def _quantize(module: nn.Module, spec: Optional[dict]):
    if not spec or spec.get("quant") in (None, "bf16"):
        return
    from ... import Int4WeightOnlyConfig, Int8WeightOnlyConfig, quantize_

    if spec["quant"] == "int4":
        quantize_(module, Int4WeightOnlyConfig(group_size=spec.get("group_size", 128)))
    elif spec["quant"] == "int8":
        quantize_(module, Int8WeightOnlyConfig())
    else:
        raise ValueError(spec["quant"])


def build_tp_transformer(cfg, ctx: Optional[TPContext], device, spec: Optional[dict] = None, seed: int = 1234,
                         dtype=torch.bfloat16) -> Transformer:
    """Random-init Transformer shard of rank ctx.rank, quantized per `spec` ({"quant": "int4"|"int8"|"bf16",
    "group_size": G}).  One full bf16 tensor at a time is generated on the device, sharded and (per block) quantized,
    so the peak footprint is the quantized shard plus one block."""
    if not isinstance(cfg, ModelArgs):
        cfg = ModelArgs.from_name(cfg)
    r, tp = (0, 1) if ctx is None else (ctx.rank, ctx.world_size)
    check_divisible(cfg, tp, (spec or {}).get("group_size") if (spec or {}).get("quant") == "int4" else None)
    with torch.device("meta"):
        model = Transformer(cfg)
    model.eval()
    hd, H, Hkv, D, I, V = cfg.head_dim, cfg.n_head, cfg.n_local_heads, cfg.dim, cfg.intermediate_size, cfg.vocab_size

    def full(n, k, layer, name):
        return _full_linear_weight(n, k, _seed_of(seed, layer, name), device, dtype)

    g = torch.Generator(device=device)
    g.manual_seed(_seed_of(seed, -1, "tok_embeddings"))
    model.tok_embeddings.weight = nn.Parameter(
        torch.empty(V, D, dtype=torch.float32, device=device).normal_(generator=g).to(dtype), requires_grad=False)
    for li, blk in enumerate(model.layers):
        att, ff = blk.attention, blk.feed_forward
        _set_linear(att.wqkv, shard_wqkv(full((H + 2 * Hkv) * hd, D, li, "wqkv"), H, Hkv, hd, r, tp).clone())
        _set_linear(att.wo, shard_cols(full(D, H * hd, li, "wo"), r, tp).clone())
        att.n_head, att.n_local_heads = H // tp, Hkv // tp
        _set_linear(ff.w1, shard_rows(full(I, D, li, "w1"), r, tp).clone())
        _set_linear(ff.w3, shard_rows(full(I, D, li, "w3"), r, tp).clone())
        _set_linear(ff.w2, shard_cols(full(D, I, li, "w2"), r, tp).clone())
        for norm in (blk.attention_norm, blk.ffn_norm):
            norm.weight = nn.Parameter(torch.ones(D, dtype=dtype, device=device), requires_grad=False)
        _quantize(blk, spec)
    model.norm.weight = nn.Parameter(torch.ones(D, dtype=dtype, device=device), requires_grad=False)
    _set_linear(model.output, shard_rows(full(V, D, -2, "output"), r, tp).clone())
    _quantize(_Holder(model.output), spec)
    if ctx is not None:
        _install_hooks(model, ctx)
    return model


class _Holder(nn.Module):
    """quantize_ swaps the weights of the nn.Linear children of the module it is given; this gives a lone Linear a
    parent."""

    def __init__(self, lin: nn.Linear):
        super().__init__()
        self.lin = lin
