"""CPU, world_size 2 over gloo: the exchange step of the data-parallel path (train_dp.BucketReducer) — bucket layout
covers the gradient arena exactly once, buckets are launched in the order backward completes them, and after
finish() every rank holds the element-wise SUM (AdamW folds 1/world into its gradient read)."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vision-transformer_b200"))
import train_dp  # noqa: E402

CFG = dict(C=3, P=4, E=32, H=2, L=5, T=17, B=2, classes=3)


def _layout():
    layout = train_dp.param_layout(**CFG)
    offs, total = train_dp.arena_offsets(layout)
    return layout, offs, total


@pytest.mark.parametrize("bucket_elems", [1, 5000, 30000, 10 ** 9])
def test_buckets_tile_the_arena_in_backward_order(bucket_elems):
    layout, offs, total = _layout()
    buckets = train_dp.make_buckets(layout, offs, total, bucket_elems)
    # contiguous, descending, covering [0, total) exactly once
    assert buckets[0][1] == total and buckets[-1][0] == 0
    for (a, b, _), (a2, b2, _) in zip(buckets, buckets[1:]):
        assert a < b and b2 == a
    # the stage that closes a bucket is the LAST one (in backward order) with parameters inside it
    order = ["head"] + list(range(CFG["L"] - 1, -1, -1)) + ["embed"]
    closing = [order.index(st) for _, _, st in buckets]
    assert closing == sorted(closing)
    for a, b, st in buckets:
        inside = {stage for (stage, _, _, _), (o, n) in zip(layout, offs) if a <= o < b}
        assert max(order.index(s) for s in inside) == order.index(st)
    if bucket_elems == 10 ** 9:
        assert len(buckets) == 1 and buckets[0][2] == "embed"
    if bucket_elems == 1:
        assert len(buckets) == CFG["L"] + 2


def test_arena_offsets_are_aligned_and_disjoint():
    layout, offs, total = _layout()
    end = 0
    for (o, n) in offs:
        assert o % train_dp.ALIGN == 0 and o >= end
        end = o + n
    assert total >= end and total % train_dp.ALIGN == 0


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, bucket_elems, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        layout, offs, total = _layout()
        buckets = train_dp.make_buckets(layout, offs, total, bucket_elems)
        gen = torch.Generator().manual_seed(100 + rank)
        g = torch.randn(total, generator=gen)
        mine = g.clone()
        red = train_dp.BucketReducer(g, buckets, dist.group.WORLD, world)
        handles, launched = [], []
        for stage in ["head"] + list(range(CFG["L"] - 1, -1, -1)) + ["embed"]:
            before = len(handles)
            red.ready(stage, handles)
            launched += [stage] * (len(handles) - before)
        red.finish(handles)
        others = [torch.randn(total, generator=torch.Generator().manual_seed(100 + r)) for r in range(world)]
        expect = torch.stack(others).sum(0)
        ok = torch.allclose(g, expect, rtol=0, atol=1e-6) and torch.equal(mine, others[rank])
        ok = ok and launched == [st for _, _, st in buckets]
        # parameters broadcast from rank 0 the way DPTrainer.broadcast_parameters does it
        p = torch.full((total,), float(rank))
        dist.broadcast(p, src=0)
        ok = ok and bool((p == 0).all())
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("bucket_elems", [5000, 10 ** 9])
def test_world2_bucketed_allreduce_sums_gradients(bucket_elems):
    world = 2
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker, args=(world, _free_port(), bucket_elems, out), nprocs=world, join=True)
        assert dict(out) == {0: True, 1: True}


def test_single_rank_reducer_is_a_no_op():
    layout, offs, total = _layout()
    g = torch.ones(total)
    red = train_dp.BucketReducer(g, train_dp.make_buckets(layout, offs, total, 5000), None, 1)
    handles = []
    red.ready("head", handles)
    red.finish(handles)
    assert handles == [] and bool((g == 1).all())


# ---------------------------------------------------------------------------------------------- ZeRO-1 exchange (CPU)
def _cpu_pack(g_slice, out_slice):
    """CPU stand-in for vitb200_grad_pack (test only): cast copy (if any) + clear."""
    if out_slice is not None:
        out_slice.copy_(g_slice.to(out_slice.dtype))
    g_slice.zero_()


def _zero1_worker(rank, world, port, comm, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        layout, offs, total = _layout()
        buckets = train_dp.make_buckets(layout, offs, total, 5000)
        comm_dtype = torch.bfloat16 if comm == "bf16" else torch.float32
        gs_all = [torch.randn(total, generator=torch.Generator().manual_seed(200 + r)) for r in range(world)]
        g = gs_all[rank].clone()
        red = train_dp.ShardedReducer(g, buckets, dist.group.WORLD, world, rank, comm_dtype, _cpu_pack)
        handles = []
        for stage in ["head"] + list(range(CFG["L"] - 1, -1, -1)) + ["embed"]:
            red.ready(stage, handles)
        red.finish(handles)
        ok = handles == list(range(len(buckets))) and int(torch.count_nonzero(g)) == 0
        # my shard of every bucket holds the sum over ranks (of the bf16-rounded values in bf16 mode)
        expect = torch.stack([t.to(comm_dtype).float() for t in gs_all]).sum(0)
        covered = torch.zeros(total, dtype=torch.bool)
        for k in range(len(buckets)):
            s0, s1 = red.shard_range(k)
            got = red.gs[red.shard_off[k]: red.shard_off[k] + red.shard_len[k]].float()
            tol = 1e-6 if comm == "fp32" else 4e-2
            ok = ok and torch.allclose(got, expect[s0:s1], rtol=0, atol=tol)
            ok = ok and s0 % train_dp.ALIGN == 0 and (s1 - s0) * world == buckets[k][1] - buckets[k][0]
            covered[s0:s1] = True
        # the shards of all ranks tile the arena exactly once
        cov = covered.to(torch.int32)
        dist.all_reduce(cov)
        ok = ok and bool((cov == 1).all())
        # "sharded optimizer": every rank rewrites ONLY its shards, gather() rebuilds the full arena everywhere
        p = torch.full((total,), -1.0)
        for k in range(len(buckets)):
            s0, s1 = red.shard_range(k)
            p[s0:s1] = float(rank + 1)
        seen = []
        red.gather(p, after=lambda k, a, b: seen.append(k))
        ok = ok and seen == list(range(len(buckets) - 1, -1, -1))      # forward order: embedding bucket first
        for k, (a, b, _) in enumerate(buckets):
            n = red.shard_len[k]
            for r in range(world):
                ok = ok and bool((p[a + r * n: a + (r + 1) * n] == float(r + 1)).all())
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("comm", ["fp32", "bf16"])
def test_world2_zero1_reduce_scatter_and_all_gather(comm):
    world = 2
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_zero1_worker, args=(world, _free_port(), comm, out), nprocs=world, join=True)
        assert dict(out) == {0: True, 1: True}


def test_every_stage_starts_on_a_shard_boundary():
    layout, offs, total = _layout()
    assert total % train_dp.SHARD_ALIGN == 0
    prev = None
    for (stage, _, _, _), (o, _) in zip(layout, offs):
        if stage != prev:
            assert o % train_dp.SHARD_ALIGN == 0, stage
            prev = stage
    for be in (1, 5000, 10 ** 9):
        for a, b, _ in train_dp.make_buckets(layout, offs, total, be):
            assert (b - a) % (8 * train_dp.ALIGN) == 0


def test_default_exchange_policy_is_two_buckets_that_shard_over_eight_ranks():
    """train_dp.auto_bucket_mb (profiles/r02_bucket_policy_n4.txt): 80 % of the gradient arena, then the rest — for
    every BASELINE config, each bucket a union of whole stages that splits into 8 aligned shards (ZeRO-1)."""
    for E, H, L, P, img in ((384, 6, 12, 16, 224), (768, 12, 12, 16, 224), (1024, 16, 24, 16, 224), (1280, 16, 32, 14, 336)):
        T = (img // P) ** 2 + 1
        layout = train_dp.param_layout(C=3, P=P, E=E, H=H, L=L, T=T, B=4, classes=2)
        offs, total = train_dp.arena_offsets(layout)
        mb = train_dp.auto_bucket_mb(total)
        buckets = train_dp.make_buckets(layout, offs, total, int(mb * (1 << 20) / 4))
        assert len(buckets) == 2, (E, [(b - a) for a, b, _ in buckets])
        (a0, b0, _), (a1, b1, st1) = buckets
        assert b0 == total and a0 == b1 and a1 == 0 and st1 == layout[0][0]      # contiguous, embedding stage closes
        assert 0.78 * total <= b0 - a0 <= 0.95 * total
        for a, b, _ in buckets:
            assert (b - a) % (8 * train_dp.ALIGN) == 0
