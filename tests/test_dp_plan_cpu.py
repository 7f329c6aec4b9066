"""CPU tier: host-side logic of the data-parallel path (ownership map, gloo-backed plain path)."""
import os
import subprocess
import sys
import textwrap

import pytest
import torch

from pretraining_llm_b200.models.layout import ParamLayout
from pretraining_llm_b200.parallel.dp import plan_buckets

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_bucket_plan_partitions_every_element_once(world):
    lay = ParamLayout(12, 768, 1024, 50304, 3)
    plans = plan_buckets(lay, world)
    assert [p.name for p in plans] == ["head", "block2", "block1", "block0", "embed"]
    owner = torch.full((lay.total,), -1, dtype=torch.int8)
    for p in plans:
        assert p.chunk % 4 == 0 and p.start % 4 == 0
        for r in range(world):
            s, c = p.slice_of(r)
            assert c % 4 == 0 and p.start <= s <= p.end and s + c <= p.end
            assert (owner[s:s + c] == -1).all()
            owner[s:s + c] = r
        # slot table covers the bucket, ascending, flags: matrices bf16-only, the rest fp32
        assert p.slot_end[-1] == p.end and p.slot_end == sorted(p.slot_end)
    assert (owner >= 0).all()
    head = plans[0]
    names = [s.name for s in lay.slots.values() if head.start <= s.offset < head.end]
    flags = dict(zip(names, head.slot_fp32))
    assert flags == {"layer_norm.weight": 1, "layer_norm.bias": 1, "lm_head.weight": 0, "lm_head.bias": 1}
    emb = plans[-1]
    assert emb.slot_fp32 == [1, 1]


def test_two_process_gloo_training_keeps_replicas_identical(tmp_path):
    """world_size 2 on CPU (gloo): the plain-PyTorch DP path averages gradients every step, so the
    replicas stay bit-identical - what the reference's alternate-step sync does not do (defect D8)."""
    script = tmp_path / "run.py"
    script.write_text(textwrap.dedent(f"""
        import os, sys, torch
        sys.path.insert(0, {ROOT!r})
        from pretraining_llm_b200.config import default_config, model_args
        from pretraining_llm_b200.data.loader import get_batch_iterator
        from pretraining_llm_b200.models import Transformer
        from pretraining_llm_b200.parallel.bootstrap import init_distributed
        from pretraining_llm_b200.runtime.trainer import Trainer
        cfg = dict(default_config)
        cfg.update(dict(vocab_size=101, context_length=32, n_embed=64, n_head=4, n_blocks=2, t_batch_size=4,
                        t_train_steps=6, t_eval_steps=3, t_eval_iters=2, t_lr=1e-2, device="cpu", dtype="float32",
                        t_out_path={str(tmp_path / "m.pt")!r}, synthetic_data=True))
        ctx = init_distributed(cfg)
        torch.manual_seed(cfg["seed"])
        model = Transformer(**model_args(cfg))
        it = get_batch_iterator("synthetic:101", 4, 32, "cpu", ctx.ddp, ctx.rank, ctx.world, seed=1)
        vit = get_batch_iterator("synthetic:101", 4, 32, "cpu", ctx.ddp, ctx.rank, ctx.world, seed=2)
        x0 = next(get_batch_iterator("synthetic:101", 4, 32, "cpu", True, ctx.rank, ctx.world, seed=1))[0]
        Trainer(model, it, vit, None, config=cfg, dist_ctx=ctx).train()
        chk = model.flat_params.double().sum().item()
        import torch.distributed as dist
        outs = [None, None]
        dist.all_gather_object(outs, (chk, int(x0.sum())))
        if ctx.rank == 0:
            assert outs[0][0] == outs[1][0], outs          # identical replicas
            assert outs[0][1] != outs[1][1], outs          # but different data per rank
            print("OK", outs)
        ctx.shutdown()
    """))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29617", OMP_NUM_THREADS="2")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29617", str(script)],
                       capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert "OK" in r.stdout
    assert os.path.exists(tmp_path / "m.pt")


def test_bucket_plan_property_random_models():
    """Property test over random model shapes / world sizes / vocab-parallel: every element of every bucket has exactly one
    owner, slices are 16-byte aligned, and the per-slot table is consistent with the layout."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=40, deadline=None)
    @given(heads=st.sampled_from([1, 2, 4, 12]), hs=st.sampled_from([8, 64, 128]), layers=st.integers(1, 5),
           ctx=st.sampled_from([16, 100, 1024]), vocab=st.sampled_from([257, 1000, 50304]),
           world=st.sampled_from([1, 2, 4, 8]), vp=st.booleans())
    def check(heads, hs, layers, ctx, vocab, world, vp):
        lay = ParamLayout(heads, heads * hs, ctx, vocab, layers)
        plans = plan_buckets(lay, world, vocab_parallel=vp)
        assert plans[0].name == "head" and plans[-1].name == "embed" and len(plans) == layers + 2
        covered = 0
        for p in plans:
            assert p.start % 4 == 0 and p.chunk % 4 == 0
            prev_end = p.start
            for r in range(world):
                s, c = p.slice_of(r)
                assert s == min(p.start + r * p.chunk, p.end) and c >= 0 and c % 4 == 0
                assert s >= prev_end or c == 0
                prev_end = max(prev_end, s + c)
                covered += c
            assert prev_end == p.end
            assert len(p.slot_end) == len(p.slot_fp32) and p.slot_end[-1] == p.end
        excluded = 0
        if vp:   # the LM head leaves the head bucket (owned by vocabulary slice instead)
            excluded = lay.buckets()[0][2] - lay.slots["lm_head.weight"].offset
        assert covered == lay.total - excluded

    check()


# ---------------------------------------------------------------------------------------------- sharded mode plan
@pytest.mark.parametrize("world", [1, 2, 4, 8])
@pytest.mark.parametrize("shape", [(4, 256, 128, 1000, 3), (12, 768, 1024, 50304, 12), (16, 2048, 2048, 50304, 24)])
def test_zero_plan_partitions_every_bucket(world, shape):
    """parallel/zero.py ZeroPlan: the owned slices tile every bucket exactly, every slice is 16-byte aligned in fp32
    and bf16, shard offsets are the same on every rank, and each parameter lands either in the small replicated
    store or in a sharded bucket (never both for the big tables)."""
    from pretraining_llm_b200.models.layout import ParamLayout
    from pretraining_llm_b200.parallel.zero import ZeroPlan, _NOT_SMALL
    H, d, ctx, V, L = shape
    lay = ParamLayout(H, d, ctx, V, L)
    plan = ZeroPlan(lay, world)
    assert plan.names[0] == "head" and plan.names[-1] == "embed" and plan.nb == L + 2
    covered = 0
    for b in range(plan.nb):
        pos = 0
        for r in range(world):
            lo, cnt = plan.owned(b, r)
            assert lo == min(plan.size(b), pos) and lo % 8 == 0 and cnt % 4 == 0 and cnt <= plan.chunk[b]
            pos += cnt
        assert pos == plan.size(b)
        assert plan.chunk[b] % 128 == 0 and plan.chunk[b] * world >= plan.size(b)
        covered += plan.size(b)
        ends, smalls = plan.slot_tables(b)
        assert ends == sorted(ends) and ends[-1] == plan.size(b) and len(ends) <= 16
        inside = [s for s in lay.slots.values() if plan.start[b] <= s.offset < plan.end[b]]
        for sl, so in zip(inside, smalls):
            small = sl.name in plan.small_off
            assert (so != _NOT_SMALL) == small
            if small:      # bucket offset + so == offset in the small store
                assert (sl.offset - plan.start[b]) + so == plan.small_off[sl.name]
    assert covered == lay.total
    assert plan.shard_elems == sum(plan.chunk)
    assert "token_embed.weight" not in plan.small_off and "lm_head.weight" not in plan.small_off
    assert "lm_head.bias" in plan.small_off and "attn_blocks.0.ln1.weight" in plan.small_off
    offs = sorted((o, lay.slots[n].numel) for n, o in plan.small_off.items())
    for (o1, n1), (o2, _) in zip(offs, offs[1:]):
        assert o1 + n1 <= o2                     # no overlap in the small store


def test_zero_plan_memory_at_1p3b_on_8_ranks():
    from pretraining_llm_b200.models.layout import ParamLayout
    from pretraining_llm_b200.parallel.zero import ZeroPlan
    plan = ZeroPlan(ParamLayout(16, 2048, 2048, 50304, 24), 8)
    per_rank = plan.bytes_per_rank()
    assert sum(per_rank.values()) < 6e9
    replicated = 10 * plan.layout.total + 8 * plan.layout.total // 8      # what the replicated mode holds
    assert sum(per_rank.values()) < 0.35 * replicated
