"""Single-token decode of the gpt-fast Llama through the B200 kernels, captured in one CUDA graph.

The reference gets its decode speed from `torch.compile(decode_one_token, mode="reduce-overhead")`
(torchao/_models/llama/generate.py:865-875), i.e. inductor + CUDA graphs.  Here there is no tracing compiler:
DecodeEngine walks the (already quantized) Transformer once, collects raw device pointers of every weight,
and issues the whole step — embedding gather, per layer [RMSNorm(+residual) -> wqkv -> RoPE+KV append ->
GQA attention -> wo -> RMSNorm(+residual) -> w1|w3 -> SiLU*mul -> w2], final norm, `output`, sampling —
as C-ABI kernel launches on the current stream; `capture()` records them into a CUDA graph which
`step()` replays.  The token position lives in device memory so the same graph serves every token.

w1 and w3 are fused by concatenating their packed rows (row-wise quantization: the fused weight is
exactly the two quantized weights stacked), which cuts one launch per layer.
"""
import math
import os
from typing import List, Optional

import torch
import torch.nn as nn

from ... import _lib
from ...dtypes.affine_quantized_tensor import AffineQuantizedTensor
from ...dtypes.layouts import PlainAQTTensorImpl, TensorCoreTiledAQTTensorImpl
from .model import Transformer


# persistent one-launch-per-token kernel (csrc/decode_mega.cu) is the batch-1 int4 default (it beats the per-projection
# graph of 161 launches: profiles/r02_*); AO_B200_MEGA=0 selects the per-projection graph
MEGA_DEFAULT = "1"


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


class _Linear:
    """One (possibly row-fused) weight-only linear: y[B][N] = x[B][K] @ W^T, launched through the C ABI."""

    def __init__(self, weights: List[torch.Tensor], interleave_gate_up: bool = False):
        w0 = weights[0]
        self.interleaved = False
        self.n = sum(int(w.shape[0]) for w in weights)
        self.k = int(w0.shape[1])
        if isinstance(w0, AffineQuantizedTensor) and isinstance(w0.tensor_impl, TensorCoreTiledAQTTensorImpl):
            self.kind = "int4"
            impls = [w.tensor_impl for w in weights]
            assert all(not i.transposed for i in impls)
            self.group_size = impls[0].group_size
            if interleave_gate_up:
                from ...ops import interleave_gate_up as _ilv
                assert len(impls) == 2 and int(weights[0].shape[0]) % 128 == 0
                self.packed, self.sz = _ilv(impls[0].packed_weight, impls[0].scale_and_zero,
                                            impls[1].packed_weight, impls[1].scale_and_zero)
                self.interleaved = True
            elif len(impls) > 1:
                assert all(int(w.shape[0]) % 128 == 0 for w in weights[:-1]), "row fusion needs N % 128 == 0"
                self.packed = torch.cat([i.packed_weight for i in impls], dim=0).contiguous()
                self.sz = torch.cat([i.scale_and_zero for i in impls], dim=0).contiguous()
            else:
                self.packed, self.sz = impls[0].packed_weight, impls[0].scale_and_zero
            self.np_, self.kp = self.packed.shape[0] * 16, self.packed.shape[1] * 64
            self.weight_bytes = self.packed.numel() * 4 + self.sz.numel() * 2
        elif isinstance(w0, AffineQuantizedTensor) and isinstance(w0.tensor_impl, PlainAQTTensorImpl):
            self.kind = "int8"
            impls = [w.tensor_impl for w in weights]
            assert all(not i.transposed for i in impls)
            if interleave_gate_up:
                from ...ops import interleave_gate_up_int8 as _ilv8
                assert len(impls) == 2 and int(weights[0].shape[0]) % 8 == 0
                q, self.scale = _ilv8(impls[0].int_data, impls[0].scale, impls[1].int_data, impls[1].scale)
                self.packed = torch.ops.torchao.b200_int8_pack(q)
                self.interleaved = True
            elif len(impls) > 1:
                self.packed = torch.ops.torchao.b200_int8_pack(torch.cat([i.int_data for i in impls], dim=0))
                self.scale = torch.cat([i.scale for i in impls], dim=0).contiguous()
            else:
                self.packed, self.scale = impls[0].packed, impls[0].scale.contiguous()
            self.np_, self.kp = self.packed.shape[0] * 16, self.packed.shape[1] * 64
            self.weight_bytes = self.n * self.k + self.scale.numel() * 2
        else:
            self.kind = "bf16"  # unquantized line: library GEMM (cuBLAS), for comparison only
            self.w = torch.cat([w.detach() for w in weights], dim=0).contiguous() if len(weights) > 1 else w0.detach()
            assert self.w.dtype == torch.bfloat16
            self.weight_bytes = self.w.numel() * 2

    def workspace_bytes(self, m):
        if self.kind == "int4":
            return int(_lib.lib().ao_b200_int4_mm_workspace_bytes(m, self.np_, self.kp))
        if self.kind == "int8":
            return int(_lib.lib().ao_b200_int8_mm_workspace_bytes(m, self.np_, self.kp))
        return 0

    def launch_fused(self, x, y, stream, norm_w=None, eps=0.0, residual=None, epilogue=0, ws=None):
        """ao_b200_int4_gemv_fused / ao_b200_int8_gemv_fused: RMSNorm prologue (norm_w), epilogue 0 plain / 1 residual
        add / 2 SiLU*mul."""
        if self.kind == "int8":
            _lib.check(_lib.lib().ao_b200_int8_gemv_fused(
                x.data_ptr(), x.shape[0], self.k, x.stride(0), self.packed.data_ptr(), self.scale.data_ptr(), self.np_,
                self.kp, y.data_ptr(), self.n, y.stride(0), _ptr(norm_w), eps, _ptr(residual),
                0 if residual is None else residual.stride(0), epilogue, stream), "int8_gemv_fused")
            return
        assert self.kind == "int4"
        _lib.check(_lib.lib().ao_b200_int4_gemv_fused(
            x.data_ptr(), x.shape[0], self.k, x.stride(0), self.packed.data_ptr(), self.sz.data_ptr(), self.np_,
            self.kp, self.group_size, y.data_ptr(), self.n, y.stride(0), _ptr(norm_w), eps, _ptr(residual),
            0 if residual is None else residual.stride(0), epilogue, _ptr(ws), 0 if ws is None else ws.numel(), stream),
            "int4_gemv_fused")

    def fused_fits(self, m):
        return self.kind in ("int4", "int8") and m <= 16 and m * (2 * self.kp + 16) <= 150 * 1024

    def launch(self, x: torch.Tensor, y: torch.Tensor, ws: Optional[torch.Tensor], stream: int):
        m = x.shape[0]
        L = _lib.lib()
        if self.kind == "int4":
            _lib.check(L.ao_b200_int4_mm(x.data_ptr(), m, self.k, x.stride(0), self.packed.data_ptr(),
                                         self.sz.data_ptr(), self.np_, self.kp, self.group_size, y.data_ptr(),
                                         self.n, y.stride(0), _ptr(ws), 0 if ws is None else ws.numel(), stream),
                       "int4_mm")
        elif self.kind == "int8":
            _lib.check(L.ao_b200_int8_mm(x.data_ptr(), m, self.k, x.stride(0), self.packed.data_ptr(),
                                         self.scale.data_ptr(), self.np_, self.kp, y.data_ptr(), self.n, y.stride(0),
                                         _ptr(ws), 0 if ws is None else ws.numel(), stream),
                       "int8_mm")
        else:
            torch.mm(x, self.w.t(), out=y)


class DecodeEngine:
    """decode_one_token (generate.py:111-117) + sample (generate.py:79-100) for a fixed batch size."""

    def __init__(self, model: Transformer, batch_size: int = 1, temperature: float = 0.8,
                 top_k: Optional[int] = 200, tp=None, fused: bool = True, mega: Optional[bool] = None):
        cfg = model.config
        assert model.freqs_cis is not None and model.max_batch_size >= batch_size, "call model.setup_caches first"
        assert cfg.head_dim == 128, "decode attention kernel is specialised for head_dim 128"
        assert batch_size <= 64, "the decode engine covers batch sizes up to 64 (M > 16 runs the prefill GEMM per linear)"
        self.model, self.cfg, self.B = model, cfg, batch_size
        self.temperature, self.top_k = temperature, top_k
        self.tp = tp  # tensor-parallel context (tp.py) or None
        dev = model.tok_embeddings.weight.device
        if tp is not None and tp.world_size > 1:
            assert batch_size <= 16, "the tensor-parallel exchange buffers are sized for B <= 16"
            tp.enable_peer_exchange(16 * cfg.dim)  # h / delta are [B <= 16][dim]
        self.device = dev
        bf = dict(dtype=torch.bfloat16, device=dev)
        B, D = batch_size, cfg.dim
        first = model.layers[0].attention
        self.H, self.Hkv, self.hd = first.n_head, first.n_local_heads, cfg.head_dim  # local (per-rank) head counts
        self.S = model.max_seq_length
        self.layers = []
        probe = [_Linear([model.layers[0].attention.wqkv.weight]), _Linear([model.layers[0].feed_forward.w2.weight])]
        self.fused = (fused and all(l.fused_fits(batch_size) for l in probe)
                      and int(model.layers[0].feed_forward.w1.weight.shape[0]) % 128 == 0
                      and type(model.layers[0].feed_forward.w1.weight) is type(model.layers[0].feed_forward.w3.weight))
        for blk in model.layers:
            ff = blk.feed_forward
            fuse = int(ff.w1.weight.shape[0]) % 128 == 0 and type(ff.w1.weight) is type(ff.w3.weight)
            self.layers.append(dict(
                attn_norm=blk.attention_norm.weight.detach(), ffn_norm=blk.ffn_norm.weight.detach(),
                wqkv=_Linear([blk.attention.wqkv.weight]), wo=_Linear([blk.attention.wo.weight]),
                w13=_Linear([ff.w1.weight, ff.w3.weight], interleave_gate_up=self.fused) if fuse else None,
                w1=None if fuse else _Linear([ff.w1.weight]), w3=None if fuse else _Linear([ff.w3.weight]),
                w2=_Linear([ff.w2.weight]),
                k_cache=blk.attention.kv_cache.k_cache, v_cache=blk.attention.kv_cache.v_cache,
                k_scale=getattr(blk.attention.kv_cache, "k_cache_scale", None),
                v_scale=getattr(blk.attention.kv_cache, "v_cache_scale", None)))
        # int8 KV cache (AffineQuantizedKVCache, model.py:198-240): the attention launch quantizes the new row and walks
        # the int8 rows (csrc/decode_attn_q8.cu)
        self.kv_int8 = self.layers[0]["k_cache"].dtype == torch.int8
        if self.kv_int8:
            assert self.layers[0]["k_scale"].dtype == torch.bfloat16, "int8 KV cache: bf16 scales"
        self.I = int(model.layers[0].feed_forward.w1.weight.shape[0])
        self.out_norm = model.norm.weight.detach()
        self.output = _Linear([model.output.weight])
        self.emb = model.tok_embeddings.weight.detach()
        self.eps = cfg.norm_eps
        self.freqs = model.freqs_cis
        for t in [self.out_norm, self.emb] + [l["attn_norm"] for l in self.layers]:
            assert t.dtype == torch.bfloat16, "model must be bf16"
        # activations (fixed addresses: the graph replays over them)
        self.tok = torch.zeros(B, dtype=torch.int64, device=dev)
        self.pos = torch.zeros(1, dtype=torch.int32, device=dev)
        self.h = torch.zeros(B, D, **bf)
        self.a = torch.zeros(B, D, **bf)
        self.qkv = torch.zeros(B, (self.H + 2 * self.Hkv) * self.hd, **bf)
        self.q = torch.zeros(B, self.H * self.hd, **bf)
        self.o = torch.zeros(B, self.H * self.hd, **bf)
        self.delta = torch.zeros(B, D, **bf)
        self.gu = torch.zeros(B, 2 * self.I, **bf)
        self.s = torch.zeros(B, self.I, **bf)
        self.logits = torch.zeros(B, int(model.output.weight.shape[0]), **bf)
        self.next_tok = torch.zeros(B, dtype=torch.int64, device=dev)
        lins = [self.output] + [l[k] for l in self.layers for k in ("wqkv", "wo", "w13", "w1", "w3", "w2") if l[k]]
        ws_bytes = max(l.workspace_bytes(B) for l in lins)
        self.ws = torch.zeros(max(ws_bytes, 16), dtype=torch.uint8, device=dev)
        # sequence splits of the decode attention: one 32-position block per split up to S = 1024 (a block is one DRAM
        # round trip; fewer splits = fewer CTAs and a cheaper merge), <= 32 (the merge keeps one split per lane)
        self.n_split = max(1, min(32, -(-self.S // 32)))
        if os.environ.get("AO_B200_ATTN_SPLIT"):  # tuning knob: sequence splits of the decode attention
            self.n_split = max(1, int(os.environ["AO_B200_ATTN_SPLIT"]))
        awb = int(_lib.lib().ao_b200_decode_attention_workspace_bytes(B, self.H, self.n_split))
        self.attn_ws = torch.zeros(max(awb, 16), dtype=torch.uint8, device=dev)
        fwb = int(_lib.lib().ao_b200_decode_attention_fused_workspace_bytes(B, self.H, self.Hkv, self.n_split))
        self.attn_fused_ws = torch.zeros(fwb, dtype=torch.uint8, device=dev)  # zeroed once: the kernel keeps its tickets 0
        self.weight_bytes = sum(l.weight_bytes for l in lins) + 2 * (
            self.out_norm.numel() + sum(l["attn_norm"].numel() + l["ffn_norm"].numel() for l in self.layers))
        self._noise = None
        self.graph = None
        self.graph_stream = None
        self.launches_per_step = 0
        self._host_pos = 0   # host mirror of *pos (set_state / step): decoding past the caches raises instead of corrupting them
        # one persistent launch per token (csrc/decode_mega.cu): batch 1, every linear int4 with one group size
        can_mega = (batch_size == 1 and self.fused and not self.kv_int8
                    and all(l[k] is not None and l[k].kind == "int4" for l in self.layers for k in ("wqkv", "wo", "w13", "w2"))
                    and self.output.kind == "int4"
                    and len({l[k].group_size for l in self.layers for k in ("wqkv", "wo", "w13", "w2")} | {self.output.group_size}) == 1
                    and D % 128 == 0 and self.I % 128 == 0 and self.H // self.Hkv <= 8)
        if mega is None:
            mega = can_mega and os.environ.get("AO_B200_MEGA", MEGA_DEFAULT) != "0"
        assert not mega or can_mega, "the persistent decode kernel needs batch 1 and int4 weight-only linears everywhere"
        self.mega = bool(mega)
        if self.mega:
            self._setup_mega()

    # ------------------------------------------------------------------ one persistent launch per token
    def _setup_mega(self, attn_blocks_per_cta: int = 0, timeout_ms: int = 0, trace: bool = False):
        import ctypes
        L = _lib.lib()
        dev = self.device
        nl = len(self.layers)
        arr = (_lib.MegaLayer * nl)()
        # the kernel's own K / V cache with padded rows (272 B: conflict-free ldmatrix): [layer][k | v][Hkv][S_pad][136]
        s_pad = -(-self.S // 32) * 32
        self._mega_kv = torch.zeros(nl, 2, self.Hkv, s_pad, 136, dtype=torch.bfloat16, device=dev)
        for i, l in enumerate(self.layers):
            arr[i].k_pad, arr[i].v_pad = self._mega_kv[i, 0].data_ptr(), self._mega_kv[i, 1].data_ptr()
        for i, l in enumerate(self.layers):
            a = arr[i]
            a.wqkv_packed, a.wqkv_sz = l["wqkv"].packed.data_ptr(), l["wqkv"].sz.data_ptr()
            a.wo_packed, a.wo_sz = l["wo"].packed.data_ptr(), l["wo"].sz.data_ptr()
            a.w13_packed, a.w13_sz = l["w13"].packed.data_ptr(), l["w13"].sz.data_ptr()
            a.w2_packed, a.w2_sz = l["w2"].packed.data_ptr(), l["w2"].sz.data_ptr()
            a.attn_norm, a.ffn_norm = l["attn_norm"].data_ptr(), l["ffn_norm"].data_ptr()
            a.k_cache, a.v_cache = l["k_cache"].data_ptr(), l["v_cache"].data_ptr()
        m = _lib.MegaModel()
        m.n_layer, m.dim, m.n_head, m.n_kv_head, m.head_dim = nl, self.cfg.dim, self.H, self.Hkv, self.hd
        m.inter, m.vocab, m.group_size, m.max_seq = self.I, self.output.n, self.output.group_size, self.S
        tp = self.tp if (self.tp is not None and self.tp.world_size > 1) else None
        m.world, m.rank = (1, 0) if tp is None else (tp.world_size, tp.rank)
        m.n_cta = torch.cuda.get_device_properties(dev).multi_processor_count
        if os.environ.get("AO_B200_MEGA_CTAS"):
            m.n_cta = int(os.environ["AO_B200_MEGA_CTAS"])
        if not attn_blocks_per_cta and os.environ.get("AO_B200_MEGA_ATTN_SPC"):  # tuning knob: 32-position tiles per CTA
            attn_blocks_per_cta = int(os.environ["AO_B200_MEGA_ATTN_SPC"])
        if not timeout_ms:
            # bound of every wait inside the kernel.  Tensor parallel: the ranks' launches are only loosely aligned (a busy
            # host on one rank delays its launch), so the bound is generous there; AO_B200_MEGA_TIMEOUT_MS overrides
            timeout_ms = int(os.environ.get("AO_B200_MEGA_TIMEOUT_MS", "10000" if tp is not None else "2000"))
        m.timeout_ms, m.attn_blocks_per_cta = timeout_ms, attn_blocks_per_cta
        m.eps = self.eps
        m.layers = ctypes.cast(arr, ctypes.POINTER(_lib.MegaLayer))
        m.out_packed, m.out_sz, m.out_norm = self.output.packed.data_ptr(), self.output.sz.data_ptr(), self.out_norm.data_ptr()
        m.emb, m.freqs = self.emb.data_ptr(), self.freqs.data_ptr()
        assert self.freqs.dtype == torch.bfloat16 and self.freqs.is_contiguous() and self.freqs.shape[0] >= self.S
        self.next_tok = self.tok   # the sampled token IS the token fed to the next launch
        m.tok, m.pos, m.logits = self.tok.data_ptr(), self.pos.data_ptr(), self.logits.data_ptr()
        self._mega_state = torch.zeros(int(L.ao_b200_decode_mega_state_bytes()), dtype=torch.uint8, device=dev)
        m.state = self._mega_state.data_ptr()
        self._mega_layers, self._mega_model = arr, m   # keep the host structs alive
        xb = int(L.ao_b200_decode_mega_xchg_bytes(ctypes.byref(m)))
        assert xb > 0, "ao_b200_decode_mega: unsupported model shape"
        if tp is None:
            self._mega_xchg = torch.zeros(xb, dtype=torch.uint8, device=dev)
            m.xchg[0] = self._mega_xchg.data_ptr()
        else:
            # tensor parallel: every rank's exchange region is mapped into every peer (CUDA IPC) — the wo / w2 epilogues
            # push their partial sums straight into the peers' regions over NVLink, the next projection's staging sums
            # the ranks in rank order.  All ranks call this together.
            import torch.distributed as dist
            with torch.cuda.device(dev):
                own, handle = ctypes.c_void_p(), ctypes.create_string_buffer(64)
                _lib.check(L.ao_b200_tp_alloc(xb, ctypes.byref(own), handle), "tp_alloc")   # cudaMalloc, zeroed
                handles = [None] * tp.world_size
                dist.all_gather_object(handles, bytes(handle.raw), group=tp.group)
                self._mega_xchg = own.value
                self._mega_peers = []
                for r, hnd in enumerate(handles):
                    if r == tp.rank:
                        m.xchg[r] = own.value
                    else:
                        peer = ctypes.c_void_p()
                        _lib.check(L.ao_b200_tp_open(ctypes.create_string_buffer(hnd, 64), ctypes.byref(peer)), f"tp_open(rank {r})")
                        m.xchg[r] = peer.value
                        self._mega_peers.append(peer.value)
            dist.barrier(group=tp.group)   # nobody pushes before everybody has zeroed and published its region
        pb = int(L.ao_b200_decode_mega_plan_bytes(ctypes.byref(m)))
        assert pb > 0, "ao_b200_decode_mega: unsupported model shape"
        self._mega_trace = None
        if trace:  # per-CTA, per-phase clock stamps (benchmarks/mega_trace.py)
            self._mega_trace = torch.zeros(m.n_cta, 5 * nl + 1, 8, dtype=torch.int64, device=dev)
            m.trace = self._mega_trace.data_ptr()
        self._mega_plan = torch.zeros(pb, dtype=torch.uint8, device=dev)
        torch.cuda.synchronize(dev)
        _lib.check(L.ao_b200_decode_mega_setup(ctypes.byref(m), self._mega_plan.data_ptr(), pb,
                                               torch.cuda.current_stream(dev).cuda_stream), "decode_mega_setup")

    def _sync_mega_kv(self):
        """Copies the model's KV caches (rows written by prefill / the eager model) into the kernel's padded copy."""
        for i, l in enumerate(self.layers):
            self._mega_kv[i, 0, :, : self.S, :128].copy_(l["k_cache"][0])
            self._mega_kv[i, 1, :, : self.S, :128].copy_(l["v_cache"][0])

    def mega_phase_cycles(self):
        """One traced launch of the persistent kernel at the current state -> {kind: {"main": cycles [layers][ctas], ...}}
        (SM clock cycles from the kernel's own stamps: 0 phase entered, 1 inputs published, 6 statistics done, 2 staging
        done, 3 main loop done, 4 outputs published).  Re-creates the kernel's buffers with a trace buffer attached: call
        it after the timed runs (graphs captured before keep the old buffers)."""
        assert self.mega
        self._keep_alive = getattr(self, "_keep_alive", []) + [self._mega_xchg, self._mega_state, self._mega_plan, self._mega_kv]
        self._setup_mega(trace=True)
        self._sync_mega_kv()
        for _ in range(2):
            self._launch_step_mega()
        torch.cuda.synchronize(self.device)
        self.check_status()
        tr = self._mega_trace.cpu().numpy().astype("float64")   # [cta][phase][8]
        nl = len(self.layers)
        kinds = ["wqkv", "attn", "wo", "w13", "w2"]
        out = {}
        for k, name in enumerate(kinds):
            idx = [5 * l + k for l in range(1 if nl > 1 else 0, nl)]   # layer 0 starts cold
            t = tr[:, idx, :]
            d = {"wait": t[:, :, 1] - t[:, :, 0]}
            if name == "attn":
                d["phase"] = t[:, :, 4] - t[:, :, 1]
            else:
                d["staging"] = t[:, :, 2] - t[:, :, 1]
                d["main"] = t[:, :, 3] - t[:, :, 2]
            out[name] = {kk: v.T for kk, v in d.items()}
        t = tr[:, [5 * nl], :]
        out["output"] = {"wait": (t[:, :, 1] - t[:, :, 0]).T, "staging": (t[:, :, 2] - t[:, :, 1]).T, "main": (t[:, :, 3] - t[:, :, 2]).T}
        out["_kernel_cycles"] = float((tr[:, 5 * nl, 3] - tr[:, 0, 0]).mean())
        return out

    def mega_status(self) -> int:
        """Sticky status word of the persistent kernel (0 = fine; non-zero = a wait inside it timed out)."""
        return int(self._mega_state[12:16].view(torch.int32).item()) if self.mega else 0

    def check_status(self):
        """Raises if a bounded wait on the device gave up since the last check (the persistent kernel's hand-overs, the
        tensor-parallel peer exchange): the tokens produced since then are garbage.  Called once per generate call."""
        st = self.mega_status()
        if st:
            raise RuntimeError(f"ao_b200_decode_mega: a wait inside the kernel timed out (status {st:#x}); "
                               "outputs since then are invalid — rebuild the engine")
        if self.tp is not None and self.tp.peer is not None and self.tp.peer.status() != 0:
            raise RuntimeError("tensor-parallel peer exchange: a rank never arrived within the kernel's spin bound; "
                               "outputs since then are invalid")

    def _launch_step_mega(self):
        import ctypes
        st = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(_lib.lib().ao_b200_decode_mega_step(ctypes.byref(self._mega_model), self._mega_plan.data_ptr(), st),
                   "decode_mega_step")
        logits = self.logits
        if self.tp is not None and self.tp.world_size > 1:
            logits = self.tp.gather_logits(self.logits)   # the output projection is sharded over the vocabulary
        self._sample(logits)
        self.launches_per_step = 2 + (1 if self.temperature != 0.0 else 0)

    # ------------------------------------------------------------------ one step, eager launches
    def _mm(self, lin, x, y, stream):
        lin.launch(x, y, self.ws, stream)
        self.launches_per_step += 1

    def _launch_step_fused(self):
        """4 GEMVs + one RoPE/KV-append/attention launch per layer: the two RMSNorms, both residual adds and SiLU*mul live in
        the GEMV prologues / epilogues (include/ao_b200.h, ao_b200_int4_gemv_fused)."""
        L = _lib.lib()
        st = torch.cuda.current_stream(self.device).cuda_stream
        B = self.B
        n = 0
        torch.index_select(self.emb, 0, self.tok, out=self.h)
        for l in self.layers:
            l["wqkv"].launch_fused(self.h, self.qkv, st, norm_w=l["attn_norm"], eps=self.eps, ws=self.ws)
            if self.kv_int8:
                _lib.check(L.ao_b200_decode_attention_fused_q8(
                    self.qkv.data_ptr(), self.freqs.data_ptr(), self.pos.data_ptr(), l["k_cache"].data_ptr(),
                    l["v_cache"].data_ptr(), l["k_scale"].data_ptr(), l["v_scale"].data_ptr(), self.o.data_ptr(), B, self.H,
                    self.Hkv, self.hd, self.S, 1.0 / math.sqrt(self.hd), self.n_split, self.attn_fused_ws.data_ptr(),
                    self.attn_fused_ws.numel(), st), "decode_attention_fused_q8")
            else:
                _lib.check(L.ao_b200_decode_attention_fused(
                    self.qkv.data_ptr(), self.freqs.data_ptr(), self.pos.data_ptr(), l["k_cache"].data_ptr(),
                    l["v_cache"].data_ptr(), self.o.data_ptr(), B, self.H, self.Hkv, self.hd, self.S,
                    1.0 / math.sqrt(self.hd), self.n_split, self.attn_fused_ws.data_ptr(), self.attn_fused_ws.numel(), st),
                    "decode_attention_fused")
            if self.tp is None:
                l["wo"].launch_fused(self.o, self.h, st, residual=self.h, epilogue=1, ws=self.ws)
            else:
                l["wo"].launch_fused(self.o, self.delta, st, ws=self.ws)
                self.tp.all_reduce_add(self.delta, self.h)
            l["w13"].launch_fused(self.h, self.s, st, norm_w=l["ffn_norm"], eps=self.eps, epilogue=2, ws=self.ws)
            if self.tp is None:
                l["w2"].launch_fused(self.s, self.h, st, residual=self.h, epilogue=1, ws=self.ws)
            else:
                l["w2"].launch_fused(self.s, self.delta, st, ws=self.ws)
                self.tp.all_reduce_add(self.delta, self.h)
            n += 5
        self.output.launch_fused(self.h, self.logits, st, norm_w=self.out_norm, eps=self.eps, ws=self.ws)
        self.launches_per_step = n + 1
        logits = self.logits
        if self.tp is not None:
            logits = self.tp.gather_logits(self.logits)
        self._sample(logits)
        self.tok.copy_(self.next_tok)
        self.pos.add_(1)

    def _launch_step(self):
        if self.mega:
            return self._launch_step_mega()
        if self.fused:
            return self._launch_step_fused()
        if self.kv_int8:
            raise NotImplementedError("int8 KV cache: only with the fused decode step (M * K must fit the GEMV staging)")
        L = _lib.lib()
        st = torch.cuda.current_stream(self.device).cuda_stream
        B, D = self.B, self.cfg.dim
        self.launches_per_step = 0
        torch.index_select(self.emb, 0, self.tok, out=self.h)
        residual = None
        for l in self.layers:
            _lib.check(L.ao_b200_rmsnorm(self.h.data_ptr(), _ptr(residual), l["attn_norm"].data_ptr(),
                                         self.h.data_ptr() if residual is not None else None, self.a.data_ptr(),
                                         B, D, self.eps, st), "rmsnorm")
            self._mm(l["wqkv"], self.a, self.qkv, st)
            _lib.check(L.ao_b200_rope_kv_append(self.qkv.data_ptr(), self.freqs.data_ptr(), self.pos.data_ptr(),
                                                self.q.data_ptr(), l["k_cache"].data_ptr(), l["v_cache"].data_ptr(),
                                                B, self.H, self.Hkv, self.hd, self.S, st), "rope_kv_append")
            _lib.check(L.ao_b200_decode_attention(self.q.data_ptr(), l["k_cache"].data_ptr(), l["v_cache"].data_ptr(),
                                                  self.pos.data_ptr(), self.o.data_ptr(), B, self.H, self.Hkv, self.hd,
                                                  self.S, 1.0 / math.sqrt(self.hd), self.n_split,
                                                  self.attn_ws.data_ptr(), self.attn_ws.numel(), st), "decode_attention")
            self._mm(l["wo"], self.o, self.delta, st)
            if self.tp is not None:
                self.tp.all_reduce(self.delta)
            _lib.check(L.ao_b200_rmsnorm(self.h.data_ptr(), self.delta.data_ptr(), l["ffn_norm"].data_ptr(),
                                         self.h.data_ptr(), self.a.data_ptr(), B, D, self.eps, st), "rmsnorm")
            if l["w13"] is not None:
                self._mm(l["w13"], self.a, self.gu, st)
            else:
                self._mm(l["w1"], self.a, self.gu[:, : self.I], st)
                self._mm(l["w3"], self.a, self.gu[:, self.I:], st)
            _lib.check(L.ao_b200_silu_mul(self.gu.data_ptr(), self.s.data_ptr(), B, self.I, st), "silu_mul")
            self._mm(l["w2"], self.s, self.delta, st)
            if self.tp is not None:
                self.tp.all_reduce(self.delta)
            residual = self.delta
            self.launches_per_step += 5 + (1 if self.n_split > 1 else 0)
        _lib.check(L.ao_b200_rmsnorm(self.h.data_ptr(), _ptr(residual), self.out_norm.data_ptr(), None,
                                     self.a.data_ptr(), B, D, self.eps, st), "rmsnorm")
        self._mm(self.output, self.a, self.logits, st)
        self.launches_per_step += 1
        logits = self.logits
        if self.tp is not None:
            logits = self.tp.gather_logits(self.logits)
        self._sample(logits)
        self.tok.copy_(self.next_tok)
        self.pos.add_(1)

    def _sample(self, logits):
        """reference: generate.py:79-100 (temperature, top-k, exponential-noise argmax).  The Exp(1) noise is drawn by
        torch's generator exactly as the reference draws it; everything else is one ao_b200_sample launch."""
        noise = None
        if self.temperature != 0.0:
            if self._noise is None or self._noise.shape != logits.shape:
                self._noise = torch.empty(logits.shape, dtype=torch.float32, device=logits.device)
            noise = self._noise.exponential_(1)
        _lib.check(_lib.lib().ao_b200_sample(
            logits.data_ptr(), logits.stride(0), logits.shape[0], logits.shape[1], float(self.temperature),
            0 if self.top_k is None else int(self.top_k), _ptr(noise), 0 if noise is None else noise.stride(0),
            self.next_tok.data_ptr(), torch.cuda.current_stream(self.device).cuda_stream), "sample")

    def _sample_torch(self, logits):
        """The reference's own op sequence (tests compare the kernel with it on the same noise)."""
        if self.temperature == 0.0:
            return torch.argmax(logits, dim=-1)
        lg = logits.float() / max(self.temperature, 1e-5)
        if self.top_k is not None:
            v, _ = torch.topk(lg, min(self.top_k, lg.size(-1)))
            lg = torch.where(lg < v[..., -1:], -float("inf"), lg)
        probs = torch.softmax(lg, dim=-1)
        return torch.argmax(probs / self._noise, dim=-1)

    # ------------------------------------------------------------------ graph
    def capture(self):
        tok0, pos0 = self.tok.clone(), self.pos.clone()
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            for _ in range(2):  # warm-up (lazy module loads, workspace growth) outside capture
                self._launch_step()
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        self.graph = torch.cuda.CUDAGraph()
        # with NCCL in the step, other threads (the process group's watchdog) may touch CUDA during the capture
        mode = {} if self.tp is None else {"capture_error_mode": "thread_local"}
        with torch.cuda.graph(self.graph, **mode):
            self._launch_step()
        self.tok.copy_(tok0)
        self.pos.copy_(pos0)
        torch.cuda.synchronize(self.device)
        return self

    def capture_streaming(self, h_tok: torch.Tensor, h_out: torch.Tensor):
        """A second graph for the serving loop: [H2D copy of the fed token from pinned `h_tok`] -> the step -> [D2H copy of
        the sampled token into pinned `h_out`], so one replay + one stream sync is a whole host-visible step (the copies
        are memcpy nodes; issuing them as separate calls costs the host ~20 us each per token)."""
        assert h_tok.is_pinned() and h_out.is_pinned() and self.graph is not None
        tok0, pos0 = self.tok.clone(), self.pos.clone()
        torch.cuda.synchronize(self.device)
        g = torch.cuda.CUDAGraph()
        mode = {} if self.tp is None else {"capture_error_mode": "thread_local"}
        with torch.cuda.graph(g, **mode):
            self.tok.copy_(h_tok, non_blocking=True)
            self._launch_step()
            h_out.copy_(self.next_tok, non_blocking=True)
        self.graph_stream = g
        self.tok.copy_(tok0)
        self.pos.copy_(pos0)
        torch.cuda.synchronize(self.device)
        return self

    def set_state(self, tokens: torch.Tensor, pos: int):
        """tokens int [B] = the token to feed next; pos = its position in the sequence."""
        if not 0 <= pos < self.S:
            raise ValueError(f"position {pos} outside the KV cache (max_seq_length {self.S})")
        self.tok.copy_(tokens.reshape(-1).to(torch.int64))
        self.pos.fill_(pos)
        self._host_pos = pos
        if self.mega:
            self._sync_mega_kv()

    def step(self) -> torch.Tensor:
        """Runs one decode step; returns the device buffer holding the sampled tokens [B] (overwritten by
        the next step).  The sampled token is already fed back for the following step."""
        if self._host_pos >= self.S:
            raise RuntimeError(f"decode step at position {self._host_pos} would run past the KV cache "
                               f"(max_seq_length {self.S}); the reference's KVCache.update raises here too")
        self._host_pos += 1
        if self.graph is not None:
            self.graph.replay()
        else:
            self._launch_step()
        return self.next_tok
