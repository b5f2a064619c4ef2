"""GPU: BASELINE.json's full bench configuration (ViT-B/16, 12 blocks, 256 images of 224x224x3 per GPU, bf16) through
size-independent properties of the reference's arithmetic: images are independent (permutation equivariance of logits),
the loss is a batch mean (gradients of the full batch = mean of the half-batch gradients), and the forward is
deterministic.  These complement — they do not replace — the direct comparison with the real reference modules on the
same B200 at full depth (12 / 24 / 32 blocks, batch 16-32), which lives in tests/test_reference_gpu.py; here the point
is the full BATCH of the bench line (M = 50 432 rows: every tile of every persistent kernel, the arenas, the 2-tile
attention items of all 3 072 (image, head) pairs)."""
import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu

CFG = dict(input_channels=3, num_classes=2, image_size=224, patch_size=16, embedding_size=768, num_heads=12,
           num_blocks=12)


def _trainer(B, seed=0):
    import train_dp
    tr = train_dp.DPTrainer(batch_size=B, dtype=torch.bfloat16, dropout_p=0.0, seed=seed, **CFG)
    tr.init_parameters(seed=5)
    # one CLS vector for every batch slot (the reference keeps one per slot, vit.py:31-33): makes images exchangeable
    cls = tr.view("embed", "cls")
    cls.copy_(cls[:1].expand_as(cls).clone())
    # trained-model regime for the logits (random init x sqrt(hs) scaling saturates the softmax)
    E = CFG["embedding_size"]
    for i in range(CFG["num_blocks"]):
        tr.view(i, "w_qkv")[:2 * E].mul_(0.25)
    tr.sync_shadow()
    return tr


def test_full_size_b16_invariants(lib):
    B = 256
    gen = torch.Generator().manual_seed(77)
    x = torch.rand(B, 3, 224, 224, generator=gen).cuda()
    y = torch.randint(0, 2, (B,), generator=gen).cuda()
    tr = _trainer(B)
    d = tr._dims(False)

    # forward determinism: same inputs, same bits
    l0 = tr.forward(x, d).clone()
    l1 = tr.forward(x, d).clone()
    assert torch.equal(l0, l1)
    assert torch.isfinite(l0).all()

    # images are independent: permuting the batch permutes the logits
    perm = torch.randperm(B, generator=gen).cuda()
    lp = tr.forward(x[perm].contiguous(), d).clone()
    assert rel_err(lp, l0[perm]) < 1e-6

    # gradients of the batch-mean loss
    tr.forward(x, d)
    tr.backward(y, d)
    torch.cuda.synchronize()
    full = {k: v.clone() for k, v in tr.grads_by_ref_key().items()}
    loss_full = tr.loss.item()
    assert all(torch.isfinite(v).all() for v in full.values())
    del tr
    torch.cuda.empty_cache()

    half = _trainer(B // 2)
    dh = half._dims(False)
    acc, losses = None, []
    for s in range(2):
        half.g.zero_()
        xs, ys = x[s * 128:(s + 1) * 128].contiguous(), y[s * 128:(s + 1) * 128].contiguous()
        logits_h = half.forward(xs, dh)
        assert rel_err(logits_h, l0[s * 128:(s + 1) * 128]) < 1e-6
        half.backward(ys, dh)
        torch.cuda.synchronize()
        losses.append(half.loss.item())
        g = {k: v.clone() for k, v in half.grads_by_ref_key().items()}
        acc = g if acc is None else {k: acc[k] + g[k] for k in g}
    assert abs(0.5 * sum(losses) - loss_full) < 1e-5
    worst = 0.0
    for k, v in full.items():
        if k == "emdeddings.cls_tkn_embd":
            continue  # per batch slot: 256 slots vs 128 slots
        e = rel_err(0.5 * acc[k], v)
        worst = max(worst, e)
        assert e < 2e-4, (k, e)   # identical per-image arithmetic; only fp32 summation order differs
    print(f"full-size half-batch gradient identity: worst rel-err {worst:.2e}")
