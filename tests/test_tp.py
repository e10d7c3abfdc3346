"""Tensor-parallel wrapper (torchao_fork_b200/_models/llama/tp.py), host-side logic on the CPU:
shard arithmetic, and — in two gloo processes — the collectives and the eager TP forward of an UNQUANTIZED tiny model
against the unsharded one (fp32: the only difference is the summation order of the row-parallel partial sums).
The quantized linears are CUDA-only (no CPU fallback), so the quantized TP path is covered by the gpu tests
(tests/test_tp_gpu.py) and bench.py --gpus N."""
import os
import socket
import sys

import pytest
import torch
import torch.multiprocessing as mp

from torchao_fork_b200._models.llama.model import ModelArgs, Transformer
from torchao_fork_b200._models.llama import tp as tpmod


def test_shard_rows_cols_cover_the_weight():
    w = torch.arange(24 * 16, dtype=torch.float32).view(24, 16)
    for tp in (1, 2, 4, 8):
        assert torch.equal(torch.cat([tpmod.shard_rows(w, r, tp) for r in range(tp)], 0), w)
        assert torch.equal(torch.cat([tpmod.shard_cols(w, r, tp) for r in range(tp)], 1), w)
    with pytest.raises(AssertionError):
        tpmod.shard_rows(w, 0, 5)


@pytest.mark.parametrize("tp", [1, 2, 4, 8])
def test_shard_wqkv_keeps_q_heads_with_their_kv_head(tp):
    H, Hkv, hd, D = 16, 8, 4, 6
    # row (head, lane) of q / k / v tagged with its kind and head index
    tag = torch.cat([torch.full((H * hd,), 0.0), torch.full((Hkv * hd,), 1.0), torch.full((Hkv * hd,), 2.0)])
    head = torch.cat([torch.arange(H).repeat_interleave(hd), torch.arange(Hkv).repeat_interleave(hd),
                      torch.arange(Hkv).repeat_interleave(hd)]).float()
    w = torch.stack([tag, head], dim=1).repeat(1, D // 2)
    for r in range(tp):
        s = tpmod.shard_wqkv(w, H, Hkv, hd, r, tp)
        hq, hk = H // tp, Hkv // tp
        assert s.shape == ((hq + 2 * hk) * hd, D)
        q, k, v = s.split([hq * hd, hk * hd, hk * hd])
        assert (q[:, 0] == 0).all() and (k[:, 0] == 1).all() and (v[:, 0] == 2).all()
        assert q[:, 1].unique().tolist() == list(range(r * hq, (r + 1) * hq))
        assert k[:, 1].unique().tolist() == list(range(r * hk, (r + 1) * hk)) == v[:, 1].unique().tolist()
        # GQA grouping survives: local q head j attends local kv head j // (hq / hk), the same pairing as globally
        rep = H // Hkv
        for j in range(hq):
            assert (r * hq + j) // rep == r * hk + j // (hq // hk)


def test_check_divisible_llama3_shapes():
    for name in ("Llama-3-8B", "Llama-3-70B"):
        cfg = ModelArgs.from_name(name)
        for tp in (1, 2, 4, 8):
            tpmod.check_divisible(cfg, tp, 128)
    with pytest.raises(AssertionError):
        tpmod.check_divisible(ModelArgs.from_name("Llama-3-8B"), 16, 128)   # 8 kv heads
    with pytest.raises(AssertionError):
        tpmod.check_divisible(ModelArgs.from_name("tiny-llama3"), 2, 2048)  # row-parallel shard smaller than a group


def test_streamed_builder_shards_the_same_full_model():
    cfg = ModelArgs.from_name("tiny-llama3")
    full = tpmod.build_tp_transformer(cfg, None, "cpu", None, seed=7, dtype=torch.float32)
    tp = 2
    shards = [tpmod.build_tp_transformer(cfg, tpmod.TPContext(r, tp), "cpu", None, seed=7, dtype=torch.float32)
              for r in range(tp)]
    b0, s0 = full.layers[1], [s.layers[1] for s in shards]
    assert torch.equal(torch.cat([s.feed_forward.w1.weight for s in s0], 0), b0.feed_forward.w1.weight)
    assert torch.equal(torch.cat([s.feed_forward.w2.weight for s in s0], 1), b0.feed_forward.w2.weight)
    assert torch.equal(torch.cat([s.attention.wo.weight for s in s0], 1), b0.attention.wo.weight)
    assert torch.equal(torch.cat([s.output.weight for s in shards], 0), full.output.weight)
    assert torch.equal(shards[1].tok_embeddings.weight, full.tok_embeddings.weight)
    assert s0[0].attention.n_head == cfg.n_head // tp and s0[0].attention.n_local_heads == cfg.n_local_heads // tp
    assert s0[0].attention.wqkv.out_features == (cfg.n_head + 2 * cfg.n_local_heads) * cfg.head_dim // tp
    # a different layer / tensor gets different numbers
    assert not torch.equal(full.layers[0].feed_forward.w1.weight, full.layers[1].feed_forward.w1.weight)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        torch.set_num_threads(2)
        ctx = tpmod.TPContext.from_env(backend="gloo")
        assert (ctx.rank, ctx.world_size) == (rank, world)
        # collectives
        t = torch.full((3, 4), float(rank + 1))
        assert torch.equal(ctx.all_reduce(t), torch.full((3, 4), float(sum(range(1, world + 1)))))
        h, d = torch.ones(2, 4), torch.full((2, 4), float(rank))
        ctx.all_reduce_add(d, h)
        assert torch.equal(h, torch.full((2, 4), 1.0 + sum(range(world))))
        local = torch.arange(2 * 3, dtype=torch.float32).view(2, 3) + 100 * rank
        g1 = ctx.gather_logits(local)
        expect = torch.cat([torch.arange(6, dtype=torch.float32).view(2, 3) + 100 * r for r in range(world)], dim=-1)
        assert torch.equal(g1, expect)
        assert ctx.gather_logits(local).data_ptr() == g1.data_ptr()      # fixed buffers (graph replay)
        # eager TP forward == unsharded forward
        cfg = ModelArgs.from_name("tiny-llama3")
        torch.manual_seed(5)                                              # the reference's TP test seed (:92)
        ref = Transformer(cfg).float().eval()
        sharded = Transformer(cfg).float().eval()
        sharded.load_state_dict(ref.state_dict())
        tpmod.apply_tp(sharded, ctx)
        for m in (ref, sharded):
            m.setup_caches(2, 32, dtype=torch.float32)
        idx = torch.randint(0, cfg.vocab_size, (2, 9))
        pos = torch.arange(9)
        with torch.no_grad():
            y_ref, y_tp = ref(idx, pos), sharded(idx, pos)
            assert y_tp.shape == y_ref.shape == (2, 9, cfg.vocab_size)
            assert torch.allclose(y_tp, y_ref, rtol=1e-4, atol=1e-4), (y_tp - y_ref).abs().max()
            # one decode step on top of the prefilled cache
            nxt, p1 = torch.randint(0, cfg.vocab_size, (2, 1)), torch.tensor([9])
            assert torch.allclose(sharded(nxt, p1), ref(nxt, p1), rtol=1e-4, atol=1e-4)
        # the streamed builder under a live group: logits of the shards agree with the tp=1 build
        one = tpmod.build_tp_transformer(cfg, None, "cpu", None, seed=3, dtype=torch.float32)
        mine = tpmod.build_tp_transformer(cfg, ctx, "cpu", None, seed=3, dtype=torch.float32)
        for m in (one, mine):
            m.setup_caches(2, 32, dtype=torch.float32)
        with torch.no_grad():
            assert torch.allclose(mine(idx, pos), one(idx, pos), rtol=1e-4, atol=1e-4)
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:  # surface the failure in the parent
        import traceback
        q.put((rank, "".join(traceback.format_exception(type(e), e, e.__traceback__))))


@pytest.mark.timeout(300)
def test_two_rank_gloo_collectives_and_eager_forward():
    world, port = 2, _free_port()
    mpctx = mp.get_context("spawn")
    q = mpctx.Queue()
    procs = [mpctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for rank, status in sorted(results):
        assert status == "ok", f"rank {rank}:\n{status}"
