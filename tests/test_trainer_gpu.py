"""GPU: the flat-arena trainer (train_dp.DPTrainer) against the oracle: gradients, AdamW update, loss descent, and the
host-input (e2e) step."""
import pytest
import torch

from conftest import cosine, rel_err
from oracle import ref_vit

pytestmark = pytest.mark.gpu


def _make(dtype, E=64, H=2, L=2, B=4, P=8, img=32, classes=3, dropout_p=0.0, **kw):
    import train_dp
    tr = train_dp.DPTrainer(input_channels=3, num_classes=classes, image_size=img, patch_size=P, embedding_size=E,
                            num_heads=H, num_blocks=L, batch_size=B, dtype=dtype, dropout_p=dropout_p, lr=1e-3, **kw)
    cfg = ref_vit.Cfg(3, classes, (img // P) ** 2, E, P, H, L, B)
    sd = ref_vit.random_state_dict(cfg, seed=21)
    tr.load_state_dict(sd)
    gen = torch.Generator().manual_seed(22)
    x = torch.rand(B, 3, img, img, generator=gen)
    y = torch.randint(0, classes, (B,), generator=gen)
    return tr, cfg, sd, x, y


def test_trainer_state_dict_roundtrip(lib):
    tr, cfg, sd, x, y = _make(torch.float32)
    out = tr.state_dict()
    assert list(out.keys()) != [] and set(out) == set(sd)
    for k in sd:
        assert torch.equal(out[k].cpu(), sd[k]), k


def test_trainer_fp32_grads_and_adamw_match_oracle(lib):
    tr, cfg, sd, x, y = _make(torch.float32)
    d = tr._dims(False)
    tr.forward(x.cuda(), d)
    tr.backward(y.cuda(), d)
    ref_logits, ref_loss, ref_grads = ref_vit.loss_and_grads(sd, x, y, cfg)
    assert rel_err(tr.logits, ref_logits) < 1e-4
    assert abs(tr.loss.item() - ref_loss.item()) < 1e-5
    mine = tr.grads_by_ref_key()
    for k, g in ref_grads.items():
        assert rel_err(mine[k], g) < 1e-4, k
    # one AdamW step over the whole arena == oracle AdamW per tensor
    mine = {k: v.clone().cpu() for k, v in mine.items()}
    tr.optimizer_step()
    new = tr.state_dict()
    for k in sd:
        p = sd[k].clone()
        # first Adam step is lr * g / (|g| + eps): feed the oracle OUR gradient so near-zero entries do not flip
        ref_vit.adamw_step(p, mine[k], torch.zeros_like(p), torch.zeros_like(p), 1, lr=1e-3)
        assert torch.allclose(new[k].cpu(), p, rtol=1e-4, atol=2e-6), k
    assert torch.count_nonzero(tr.g) == 0  # gradients cleared for the next step


def test_trainer_bf16_grads_cosine(lib):
    tr, cfg, sd, x, y = _make(torch.bfloat16, E=128, H=2, B=32)   # 32 gradient-carrying rows, see test_model_gpu
    sd = {k: (v * 0.25 if k.endswith(("key.weight", "query.weight")) else v) for k, v in sd.items()}  # see _soften
    tr.load_state_dict(sd)
    d = tr._dims(False)
    tr.forward(x.cuda(), d)
    tr.backward(y.cuda(), d)
    ref_logits, _, ref_grads = ref_vit.loss_and_grads(sd, x, y, cfg)
    assert rel_err(tr.logits, ref_logits) < 2e-2
    mine = tr.grads_by_ref_key()
    bad = [(k, cosine(mine[k], g)) for k, g in ref_grads.items() if cosine(mine[k], g) < 0.999]
    assert not bad, bad[:5]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_trainer_training_reduces_loss_with_dropout(lib, dtype):
    tr, cfg, sd, x, y = _make(dtype, E=128, H=2, B=8, dropout_p=0.2)
    xs, ys = x.cuda(), y.cuda()
    losses = [tr.step(xs, ys).item() for _ in range(30)]
    assert losses[-1] < 0.6 * losses[0], losses
    if dtype == torch.bfloat16:
        assert torch.equal(tr.shadow, tr.p.bfloat16())  # AdamW keeps the bf16 shadow weights in sync


def test_trainer_step_host_matches_step(lib):
    tr_a, cfg, sd, x, y = _make(torch.float32)
    tr_b, *_ = _make(torch.float32)
    xh, yh = x.pin_memory(), y.pin_memory()
    for _ in range(3):
        tr_a.step(x.cuda(), y.cuda())
        li = tr_b.step_host(xh, yh)
    assert abs(tr_b.read_loss(li) - tr_a.loss.item()) < 1e-5
    # split-K weight gradients use fp32 atomics (order varies run to run) and Adam's g/(|g|+eps) amplifies that for
    # |g| ~ eps, so compare in bulk: the two trajectories agree to 1e-6 on >= 99.9% of the parameters
    diff = (tr_a.p - tr_b.p).abs()
    assert (diff > 1e-6).float().mean().item() < 1e-3
    assert diff.max().item() <= 3 * 2 * 1e-3
    with pytest.raises(RuntimeError, match="Sizes of tensors must match"):
        tr_a.step(x[:2].cuda(), y[:2].cuda())


@pytest.mark.parametrize("layout", [0, 1])
def test_trainer_step_host_u8_matches_float_step(lib, layout):
    """step_host_u8 (1 byte per pixel over PCIe, /255 fused into the patch gather) follows the same trajectory as
    step() on images_u8.float()/255."""
    tr_a, cfg, sd, x, y = _make(torch.float32)
    tr_b, *_ = _make(torch.float32)
    x8 = (x * 255).round().clamp(0, 255).to(torch.uint8)
    xf = (x8.float() / 255.0).cuda()
    xh = (x8.permute(0, 2, 3, 1).contiguous() if layout == 1 else x8).pin_memory()
    yh = y.pin_memory()
    for _ in range(3):
        tr_a.step(xf, y.cuda())
        li = tr_b.step_host_u8(xh, yh, layout=layout)
    assert abs(tr_b.read_loss(li) - tr_a.loss.item()) < 1e-5
    diff = (tr_a.p - tr_b.p).abs()
    assert (diff > 1e-6).float().mean().item() < 1e-3
    with pytest.raises(RuntimeError, match="uint8 pixels"):
        tr_b.step_host_u8(xh[:1], yh)


def test_trainer_optimizer_state_in_reference_format(lib):
    """DPTrainer.optimizer_state_dict() is a torch.optim.AdamW state_dict over the REFERENCE's parameter list
    (train.py:66,73,110): a real AdamW over per-head parameters loads it and, fed the trainer's gradients, takes the
    same next step; load_optimizer_state_dict resumes a second trainer on the same trajectory."""
    from VisionTransformer import checkpoint
    tr, cfg, sd, x, y = _make(torch.float32)
    xd, yd = x.cuda(), y.cuda()
    for _ in range(2):
        tr.step(xd, yd)
    osd, msd = tr.optimizer_state_dict(), tr.state_dict()
    keys = checkpoint.reference_param_keys(cfg.num_blocks, cfg.num_heads)
    assert osd["param_groups"][0]["params"] == list(range(len(keys))) and len(osd["state"]) == len(keys)
    ref_params = [msd[k].clone().cpu().requires_grad_(True) for k in keys]
    opt = torch.optim.AdamW(ref_params, lr=1e-3, weight_decay=1e-4)
    import copy
    opt.load_state_dict(copy.deepcopy(osd))        # what the reference does at src/train.py:73 (torch keeps the
    #                                                'step' tensors of the dict it is given, hence the copy)
    d = tr._dims(True)
    tr.forward(xd, d)
    tr.backward(yd, d)
    grads = {k: v.clone().cpu() for k, v in tr.grads_by_ref_key().items()}
    tr.optimizer_step()
    for k, p in zip(keys, ref_params):
        p.grad = grads[k]
    opt.step()
    new = tr.state_dict()
    for k, p in zip(keys, ref_params):
        assert torch.allclose(new[k].cpu(), p.detach(), rtol=1e-4, atol=2e-6), k
    # resume in a second trainer from (model, optimizer) state
    tr2, *_ = _make(torch.float32)
    tr2.load_state_dict(msd)
    tr2.load_optimizer_state_dict(osd)
    assert tr2.step_no == 2
    tr2.step(xd, yd)
    diff = (tr.p - tr2.p).abs()
    assert (diff > 1e-6).float().mean().item() < 1e-3
    with pytest.raises(ValueError, match="doesn't match the size"):
        tr2.load_optimizer_state_dict({"state": {}, "param_groups": [{"params": [0, 1]}]})


@pytest.mark.parametrize("dtype,shape", [(torch.float32, dict()), (torch.float32, dict(L=1)),
                                         (torch.bfloat16, dict(E=128, H=2, L=3, B=32, P=16, img=224)),
                                         (torch.float32, dict(E=96, H=3, L=2, B=5, P=8, img=40))])
def test_last_block_row0_mode_is_exact(lib, dtype, shape):
    """last_block_row0=True (only token 0 of the last block is computed past attention, reference vit.py:80) gives the
    logits and EVERY gradient of the full computation, and still matches the oracle in fp32."""
    tr_f, cfg, sd, x, y = _make(dtype, **shape)
    tr_r, *_ = _make(dtype, last_block_row0=True, **shape)
    xd, yd = x.cuda(), y.cuda()
    out = []
    for tr in (tr_f, tr_r):
        d = tr._dims(False)
        tr.forward(xd, d)
        tr.backward(yd, d)
        torch.cuda.synchronize()
        out.append((tr.logits.clone(), tr.loss.item(), {k: v.clone() for k, v in tr.grads_by_ref_key().items()}))
    assert rel_err(out[1][0], out[0][0]) < 1e-6
    assert abs(out[1][1] - out[0][1]) < 1e-6
    for k, g in out[0][2].items():
        assert rel_err(out[1][2][k], g) < 2e-5, k      # same per-row arithmetic; fp32 summation order differs
    if dtype == torch.float32:
        ref_logits, ref_loss, ref_grads = ref_vit.loss_and_grads(sd, x, y, cfg)
        assert rel_err(out[1][0], ref_logits) < 1e-4
        for k, g in ref_grads.items():
            assert rel_err(out[1][2][k], g) < 1e-4, k
    # with dropout the two modes draw different (equally valid) masks; training still descends
    tr_d, *_ = _make(dtype, dropout_p=0.2, last_block_row0=True, **shape)
    first = None
    for _ in range(30):
        tr_d.step(xd, yd)
        first = tr_d.loss.item() if first is None else first
    assert tr_d.loss.item() < first


# ---------------------------------------------------------------------------------------------- data parallel
def _dp_worker(rank, world, port, out):
    """One rank of a 2-process data-parallel step.  Both ranks share cuda:0 (gloo moves the buckets), so the real
    DPTrainer exchange path — bucket events, side stream, async all-reduce, 1/world inside AdamW — runs on one GPU."""
    import os
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import train_dp
        E, H, L, B, P, img, classes = 64, 2, 3, 2, 8, 32, 3
        tr = train_dp.DPTrainer(input_channels=3, num_classes=classes, image_size=img, patch_size=P, embedding_size=E,
                                num_heads=H, num_blocks=L, batch_size=B, dtype=torch.float32, dropout_p=0.0, lr=1e-3,
                                process_group=dist.group.WORLD, world_size=world, bucket_mb=0.05, zero1=False)
        cfg = ref_vit.Cfg(3, classes, (img // P) ** 2, E, P, H, L, B)
        sd = ref_vit.random_state_dict(cfg, seed=31)
        tr.load_state_dict(sd)
        tr.broadcast_parameters(0)
        gen = torch.Generator().manual_seed(32)
        x = torch.rand(world * B, 3, img, img, generator=gen)
        y = torch.randint(0, classes, (world * B,), generator=gen)
        xs, ys = x[rank * B:(rank + 1) * B], y[rank * B:(rank + 1) * B]
        d = tr._dims(False)
        tr.forward(xs.cuda(), d)
        tr.backward(ys.cuda(), d)
        torch.cuda.synchronize()
        # expectation: SUM over ranks of each rank's local-mean-loss gradient (oracle, fp32)
        expect = None
        for r in range(world):
            _, _, g = ref_vit.loss_and_grads(sd, x[r * B:(r + 1) * B], y[r * B:(r + 1) * B], cfg)
            expect = g if expect is None else {k: expect[k] + g[k] for k in g}
        mine = tr.grads_by_ref_key()
        worst = max(rel_err(mine[k], expect[k]) for k in expect)
        nb = len(tr.buckets)
        # the optimizer divides by world: parameters after one step == AdamW on the averaged gradient
        mine_c = {k: v.clone().cpu() for k, v in mine.items()}
        tr.optimizer_step()
        new = tr.state_dict()
        ok_step = True
        for k in sd:
            p = sd[k].clone()
            ref_vit.adamw_step(p, mine_c[k] / world, torch.zeros_like(p), torch.zeros_like(p), 1, lr=1e-3)
            ok_step = ok_step and torch.allclose(new[k].cpu(), p, rtol=1e-4, atol=2e-6)
        out[rank] = (float(worst), int(nb), bool(ok_step))
    finally:
        dist.destroy_process_group()


def test_two_rank_data_parallel_step_matches_oracle(lib):
    import socket
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_dp_worker, args=(2, port, out), nprocs=2, join=True)
        res = dict(out)
    assert set(res) == {0, 1}
    for r, (worst, nb, ok_step) in res.items():
        assert nb >= 3, "the test must exercise several buckets"
        assert worst < 1e-4, (r, worst)
        assert ok_step, r


def _zero1_worker(rank, world, port, backend, out):
    """One rank of a 2-process ZeRO-1 run (reduce-scatter -> sharded AdamW -> all-gather), 3 steps, against ONE
    all-reduce-free single-process trainer fed the concatenated batch semantics: the oracle's AdamW on the gradient
    averaged over ranks.  nccl when the box has two GPUs, else both ranks share cuda:0 over gloo."""
    import os
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dev = torch.device("cuda", rank if backend == "nccl" else 0)
    torch.cuda.set_device(dev)
    dist.init_process_group(backend, rank=rank, world_size=world)
    try:
        import train_dp
        E, H, L, B, P, img, classes = 64, 2, 3, 2, 8, 32, 3
        res = {}
        for comm, dtype in (("fp32", torch.float32), ("bf16", torch.float32)):
            kw = dict(input_channels=3, num_classes=classes, image_size=img, patch_size=P, embedding_size=E, num_heads=H,
                      num_blocks=L, batch_size=B, dtype=dtype, dropout_p=0.0, lr=1e-3, device=dev, bucket_mb=0.05)
            tr = train_dp.DPTrainer(process_group=dist.group.WORLD, world_size=world, zero1=True, grad_comm=comm, **kw)
            ar = train_dp.DPTrainer(process_group=dist.group.WORLD, world_size=world, zero1=False, **kw)
            cfg = ref_vit.Cfg(3, classes, (img // P) ** 2, E, P, H, L, B)
            sd = ref_vit.random_state_dict(cfg, seed=41)
            for t in (tr, ar):
                t.load_state_dict(sd)
                t.broadcast_parameters(0)
            assert tr.m.numel() * world == tr.total and len(tr.buckets) >= 3     # moments exist for the shard only
            gen = torch.Generator().manual_seed(42)
            worst = 0.0
            for step in range(3):
                x = torch.rand(world * B, 3, img, img, generator=gen)
                y = torch.randint(0, classes, (world * B,), generator=gen)
                xs, ys = x[rank * B:(rank + 1) * B].to(dev), y[rank * B:(rank + 1) * B].to(dev)
                la, lb = tr.step(xs, ys, train=False).item(), ar.step(xs, ys, train=False).item()
                # Adam's first steps move every weight by ~lr whatever the gradient's size: a gradient entry that is ~0
                # flips sign under ANY rounding change (bf16 exchange, or just a different summation order of the
                # reduce-scatter in fp32) and moves that one weight the other way; the loss follows to 1e-4 / 2 %
                assert abs(la - lb) < (1e-4 if comm == "fp32" else 2e-2) * abs(lb), (step, la, lb)
                if step == 0:
                    # after ONE step the first moment is 0.1 x the exchanged mean gradient: this isolates the exchange
                    # itself (bf16 rounding of each rank's contribution, 2^-9 relative) from trajectory divergence
                    o1, o2 = tr.optimizer_state_dict(), ar.optimizer_state_dict()
                    first = 1.0 - min(cosine(o1["state"][i]["exp_avg"], o2["state"][i]["exp_avg"]) for i in o1["state"])
                    assert first < (1e-10 if comm == "fp32" else 1e-4), (comm, first)
                    firstp = max(rel_err(o1["state"][i]["exp_avg"], o2["state"][i]["exp_avg"]) for i in o1["state"])
                    assert firstp < (1e-5 if comm == "fp32" else 1e-2), (comm, firstp)
            a, b = tr.state_dict(), ar.state_dict()
            # absolute: at most 2 lr per step on any one weight (3 steps, lr 1e-3), and nearly all weights agree
            worst = max(float((a[k] - b[k]).abs().max()) for k in a)
            mean_abs = sum(float((a[k] - b[k]).abs().sum()) for k in a) / sum(a[k].numel() for k in a)
            oa, ob = tr.optimizer_state_dict(), ar.optimizer_state_dict()
            worst_m = 1.0 - min(cosine(oa["state"][i]["exp_avg"], ob["state"][i]["exp_avg"]) for i in oa["state"])
            steps_ok = all(float(oa["state"][i]["step"]) == 3.0 for i in oa["state"])
            # resume: the gathered state loads back into shards
            tr.load_optimizer_state_dict(ob)
            oc = tr.optimizer_state_dict()
            resume = max(rel_err(oc["state"][i]["exp_avg_sq"], ob["state"][i]["exp_avg_sq"]) for i in oc["state"])
            res[comm] = (float(worst), float(mean_abs), float(worst_m), bool(steps_ok), float(resume),
                         int(torch.count_nonzero(tr.g)))
        out[rank] = res
    finally:
        dist.destroy_process_group()


def test_two_rank_zero1_matches_allreduce_trainer(lib):
    import socket
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    backend = "nccl" if torch.cuda.device_count() >= 2 else "gloo"
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_zero1_worker, args=(2, port, backend, out), nprocs=2, join=True)
        res = dict(out)
    assert set(res) == {0, 1}
    for r, by_comm in res.items():
        worst, mean_abs, worst_m, steps_ok, resume, g_nz = by_comm["fp32"]
        assert worst <= 2.1 * 3 * 1e-3 and mean_abs < 1e-6 and worst_m < 1e-3 and steps_ok and resume == 0.0 and g_nz == 0, \
            (r, by_comm["fp32"])
        worst, mean_abs, worst_m, steps_ok, resume, g_nz = by_comm["bf16"]
        # bf16 gradient exchange: the exchange itself is gated after the first step (inside the worker); after three
        # steps the trajectories have separated by sign flips of near-zero gradient entries
        assert worst <= 2.1 * 3 * 1e-3 and mean_abs < 1e-4 and worst_m < 2e-2 and steps_ok and g_nz == 0, (r, by_comm["bf16"])


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_graph_replay_equals_eager_steps_with_live_dropout(lib, dtype):
    """the whole step captured into ONE CUDA graph.  With dropout LIVE, replays must draw the masks and
    apply the AdamW update of the eager step with the same step number (the per-step seed term and the bias
    corrections live in device memory that the graph advances), also when eager steps are interleaved."""
    E = 128 if dtype == torch.bfloat16 else 64
    ta, cfg, sd, x, y = _make(dtype, E=E, H=2, L=2, B=8, dropout_p=0.2, seed=3)
    tb, *_ = _make(dtype, E=E, H=2, L=2, B=8, dropout_p=0.2, seed=3)
    gen = torch.Generator().manual_seed(5)
    xs = [torch.rand(x.shape, generator=gen).cuda() for _ in range(6)]
    ys = [torch.randint(0, 3, y.shape, generator=gen).cuda() for _ in range(6)]
    with pytest.raises(RuntimeError, match="at least one eager step"):
        tb.capture_graph()
    plan = ["eager", "graph", "graph", "eager", "graph", "graph"]
    for i, how in enumerate(plan):
        la = ta.step(xs[i], ys[i]).item()
        lb = (tb.step(xs[i], ys[i]) if how == "eager" else tb.step_graph(xs[i], ys[i])).item()
        # same masks, same update: only the fp32 summation order of the split-K dW atomics differs between two runs
        assert abs(la - lb) <= 2e-4 * abs(la), (i, how, la, lb)
    assert ta.step_no == tb.step_no == 6
    a, b = ta.state_dict(), tb.state_dict()
    mean_abs = sum(float((a[k] - b[k]).abs().sum()) for k in a) / sum(a[k].numel() for k in a)
    assert mean_abs < 2e-6, mean_abs
    # and a replay with the SAME inputs draws different masks each time (the seed really advances)
    l1 = tb.step_graph(xs[0], ys[0]).item()
    l2 = tb.step_graph(xs[0], ys[0]).item()
    assert l1 != l2
    oa, ob = ta.optimizer_state_dict(), tb.optimizer_state_dict()
    assert float(ob["state"][0]["step"]) == 8.0 and float(oa["state"][0]["step"]) == 6.0


def test_trainer_rejects_bad_labels_and_out_of_range_targets(lib):
    tr, cfg, sd, x, y = _make(torch.float32)
    with pytest.raises(RuntimeError, match="int64"):
        tr.step(x.cuda(), y.cuda().int())
    with pytest.raises(RuntimeError, match="int64"):
        tr.step(x.cuda(), y)                          # CPU labels
    bad = y.clone()
    bad[0] = 99                                       # out of range: torch device-asserts; here the loss turns NaN
    loss = tr.step(x.cuda(), bad.cuda())
    assert torch.isnan(loss).all()


def test_adamw_range_bf16_gradients_and_fused_zero(lib):
    """vitb200_adamw_range: a sub-range of the arena, gradients from a bf16 exchange buffer, fp32 gradient cleared."""
    import ctypes
    from VisionTransformer import _C
    n, a, b = 10007, 1024, 9220                      # odd sizes: vector body + scalar tail
    g = torch.Generator(device="cuda").manual_seed(0)
    p = torch.randn(n, device="cuda", generator=g)
    gr = torch.randn(n, device="cuda", generator=g)
    gb = gr.to(torch.bfloat16)
    m = torch.rand(n, device="cuda", generator=g) * 0.1
    v = torch.rand(n, device="cuda", generator=g) * 0.01
    sh = torch.zeros(n, dtype=torch.bfloat16, device="cuda")
    p0, m0, v0, g0 = p.clone(), m.clone(), v.clone(), gr.clone()
    h = _C.AdamWHyper(1e-3, 0.9, 0.999, 1e-8, 1e-2, 3, 0.5)
    L = _C.lib()
    _C.check(L.vitb200_adamw_range(p[a:].data_ptr(), gr[a:].data_ptr(), gb[a:].data_ptr(), m[a:].data_ptr(),
                                   v[a:].data_ptr(), sh[a:].data_ptr(), b - a, 1, h, _C.stream_ptr()))
    torch.cuda.synchronize()
    for t, t0 in ((p, p0), (m, m0), (v, v0)):
        assert torch.equal(t[:a], t0[:a]) and torch.equal(t[b:], t0[b:])
    assert torch.equal(gr[:a], g0[:a]) and torch.equal(gr[b:], g0[b:]) and torch.count_nonzero(gr[a:b]) == 0
    gg = gb[a:b].float() * 0.5
    b1, b2 = torch.tensor(0.9, dtype=torch.float32), torch.tensor(0.999, dtype=torch.float32)   # fp32 (1 - beta)
    em = b1 * m0[a:b] + (1 - b1) * gg
    ev = b2 * v0[a:b] + (1 - b2) * gg * gg
    ep = p0[a:b] * (1 - 1e-3 * 1e-2) - (1e-3 / (1 - 0.9 ** 3)) * em / (ev.sqrt() / (1 - 0.999 ** 3) ** 0.5 + 1e-8)
    assert torch.allclose(m[a:b], em.cuda(), rtol=1e-5, atol=1e-7) and torch.allclose(v[a:b], ev.cuda(), rtol=1e-5, atol=1e-9)
    assert torch.allclose(p[a:b], ep, rtol=1e-5, atol=1e-7)
    assert torch.equal(sh[a:b], p[a:b].to(torch.bfloat16)) and torch.count_nonzero(sh[:a]) == 0
