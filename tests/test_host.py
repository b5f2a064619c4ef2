"""CPU: host-side logic of the drop-in modules — reference state_dict keys/shapes, QKV packing rule, refusal of CPU
tensors (no fallback), batch-size contract, precision selection."""
import pytest
import torch

from conftest import load_golden
from VisionTransformer import config, transformer, vit


def _model(g, **kw):
    c = g["config"]
    cfg = config.ViTConfig(c["input_channels"], c["num_classes"], c["num_patches"], c["embedding_size"],
                           c["patch_size"], c["num_heads"], c["num_blocks"], "cpu", c["batch_size"], **kw)
    return vit.VisionTransformer(cfg)


@pytest.mark.parametrize("name", ["tiny", "odd", "c1"])
def test_state_dict_keys_and_shapes_match_reference(name):
    g = load_golden(name)
    m = _model(g)
    sd = m.state_dict()
    assert list(sd.keys()) == list(g["state_dict"].keys())
    for k, v in g["state_dict"].items():
        assert sd[k].shape == v.shape, k
    c = g["config"]
    assert len(sd) == 4 + c["num_blocks"] * (3 * c["num_heads"] + 10) + 6


def test_load_reference_state_dict_and_packing_rule():
    g = load_golden("tiny")
    m = _model(g)
    res = m.load_state_dict(g["state_dict"])
    assert not res.missing_keys and not res.unexpected_keys
    c = g["config"]
    E, H = c["embedding_size"], c["num_heads"]
    hs = E // H
    mh = m.transformer_encoder.blocks[1].multi_head
    for s, name in enumerate(("query", "key", "value")):
        for h in range(H):
            ref = g["state_dict"][f"transformer_encoder.blocks.1.multi_head.heads.{h}.{name}.weight"]
            assert torch.equal(mh.qkv_weight[s * E + h * hs: s * E + (h + 1) * hs], ref)
            assert torch.equal(getattr(mh.heads[h], name).weight, ref)
    out = m.state_dict()
    for k, v in g["state_dict"].items():
        assert torch.equal(out[k], v), k


def test_strict_load_reports_missing_head_key():
    g = load_golden("tiny")
    sd = dict(g["state_dict"])
    sd.pop("transformer_encoder.blocks.0.multi_head.heads.1.key.weight")
    with pytest.raises(RuntimeError, match="heads.1.key.weight"):
        _model(g).load_state_dict(sd)


def test_cpu_tensors_are_refused_everywhere():
    g = load_golden("tiny")
    m = _model(g)
    x = g["images_u8"].float() / 255.0
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(x)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m.transformer_encoder(torch.randn(2, 17, 16))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        transformer.FeedForward(16)(torch.randn(2, 17, 16))


def test_parameter_count_and_order():
    g = load_golden("tiny")
    m = _model(g)
    ps = m.all_param_tensors()
    assert len(ps) == 4 + 11 * g["config"]["num_blocks"] + 6
    assert {id(p) for p in ps} == {id(p) for p in m.parameters()}
    assert sum(p.numel() for p in ps) == sum(v.numel() for v in g["state_dict"].values())


def test_precision_selection():
    g = load_golden("tiny")
    m = _model(g)
    assert m._kernel_dtype() == torch.float32
    assert all(p.dtype == torch.float32 for p in m.parameters())
    mb = _model(g, precision=torch.bfloat16)
    assert mb._kernel_dtype() == torch.bfloat16
    assert mb.transformer_encoder.blocks[0].ffwd._kernel_dtype() == torch.bfloat16
    assert all(p.dtype == torch.float32 for p in mb.parameters())  # fp32 masters
    assert m.transformer_encoder.blocks[0].multi_head.attention_probs is None


def test_dropout_is_hard_coded_like_the_reference():
    g = load_golden("tiny")
    m = _model(g, dropout=0.7)   # stored, never passed down (reference config.py:17 / transformer.py:35,53)
    assert m.transformer_encoder.blocks[0].multi_head.dropout.p == 0.2
    assert m.transformer_encoder.blocks[0].ffwd.mlp[3].p == 0.2
    m.train()
    assert m._drop() == pytest.approx(0.2)
    m.eval()
    assert m._drop() == 0.0


# ---------------------------------------------------------------------------------------- checkpoint compatibility
def test_checkpoint_parameter_lists_match_reference_and_package():
    """reference_param_keys == the real reference's parameter order (its state_dict holds parameters only; fixture
    written by the reference itself); packed_param_keys == this package's named_parameters()."""
    from VisionTransformer import checkpoint
    for name in ("tiny", "odd", "c1"):
        g = load_golden(name)
        c = g["config"]
        assert checkpoint.reference_param_keys(c["num_blocks"], c["num_heads"]) == list(g["state_dict"].keys())
        m = _model(g)
        assert checkpoint.packed_param_keys(c["num_blocks"]) == [n for n, _ in m.named_parameters()]


def test_optimizer_state_translation_follows_real_adamw():
    """Run torch.optim.AdamW over the reference parameter list (per-head q/k/v) and over the packed list with the
    same gradients; the translated state equals the other optimizer's state, resumes identically, and the
    round trip is lossless (reference src/train.py:66,73,110)."""
    from VisionTransformer import checkpoint
    g = load_golden("tiny")
    c = g["config"]
    L, H, E = c["num_blocks"], c["num_heads"], c["embedding_size"]
    hs = E // H
    m = _model(g)
    m.load_state_dict(g["state_dict"])
    packed = [p.detach().clone().requires_grad_(True) for p in m.parameters()]
    ref_keys = checkpoint.reference_param_keys(L, H)
    ref = [g["state_dict"][k].detach().clone().requires_grad_(True) for k in ref_keys]
    opt_p = torch.optim.AdamW(packed, lr=1e-2, weight_decay=1e-4)
    opt_r = torch.optim.AdamW(ref, lr=1e-2, weight_decay=1e-4)
    pk_keys = checkpoint.packed_param_keys(L)
    gen = torch.Generator().manual_seed(3)

    def set_grads():
        for k, p in zip(pk_keys, packed):
            p.grad = torch.randn(p.shape, generator=gen)
        by_key = dict(zip(pk_keys, packed))
        for k, r in zip(ref_keys, ref):
            if "multi_head.heads." in k:
                pk, r0 = checkpoint._head_rows(k, H, hs)
                r.grad = by_key[pk].grad[r0:r0 + hs].clone()
            else:
                r.grad = by_key[k].grad.clone()

    for _ in range(3):
        set_grads()
        opt_p.step()
        opt_r.step()
    to_packed = checkpoint.optimizer_state_to_packed(opt_r.state_dict(), H)
    want = opt_p.state_dict()
    assert to_packed["param_groups"][0]["params"] == want["param_groups"][0]["params"]
    for j, st in want["state"].items():
        assert float(to_packed["state"][j]["step"]) == float(st["step"]) == 3.0
        assert torch.equal(to_packed["state"][j]["exp_avg"], st["exp_avg"]), pk_keys[j]
        assert torch.equal(to_packed["state"][j]["exp_avg_sq"], st["exp_avg_sq"]), pk_keys[j]
    back = checkpoint.optimizer_state_to_reference(to_packed, H)
    rs = opt_r.state_dict()
    assert set(back["state"]) == set(rs["state"])
    for i, st in rs["state"].items():
        assert torch.equal(back["state"][i]["exp_avg"], st["exp_avg"]), ref_keys[i]
        assert torch.equal(back["state"][i]["exp_avg_sq"], st["exp_avg_sq"]), ref_keys[i]
    # resume: a fresh packed optimizer loaded from the translated reference state takes the same next step
    fresh = [p.detach().clone().requires_grad_(True) for p in packed]
    opt_f = torch.optim.AdamW(fresh, lr=1e-2, weight_decay=1e-4)
    opt_f.load_state_dict(to_packed)
    set_grads()
    for f, p in zip(fresh, packed):
        f.grad = p.grad.clone()
    opt_p.step()
    opt_f.step()
    for f, p in zip(fresh, packed):
        assert torch.equal(f, p)
    # and the reference's own optimizer accepts the state translated back from the packed one
    opt_r2 = torch.optim.AdamW([r.detach().clone().requires_grad_(True) for r in ref], lr=1e-2, weight_decay=1e-4)
    opt_r2.load_state_dict(checkpoint.optimizer_state_to_reference(want, H))
    with pytest.raises(ValueError, match="not a reference ViT parameter list"):
        checkpoint.optimizer_state_to_packed({"state": {}, "param_groups": [{"params": list(range(13))}]}, H)


def test_convert_checkpoint_cli_roundtrip(tmp_path):
    from VisionTransformer import checkpoint
    g = load_golden("tiny")
    c = g["config"]
    H = c["num_heads"]
    keys = checkpoint.reference_param_keys(c["num_blocks"], H)
    params = [g["state_dict"][k].clone().requires_grad_(True) for k in keys]
    opt = torch.optim.AdamW(params, lr=1e-4, weight_decay=1e-4)
    for p in params:
        p.grad = torch.ones_like(p)
    opt.step()
    src, mid, dst = tmp_path / "ref.pt", tmp_path / "ours.pt", tmp_path / "back.pt"
    torch.save({"epoch": 3, "model_state_dict": g["state_dict"], "optimizer_state_dict": opt.state_dict(),
                "loss": 1.5, "step": 7}, src)
    checkpoint.main(["to-packed", str(src), str(mid), "--heads", str(H)])
    checkpoint.main(["to-reference", str(mid), str(dst), "--heads", str(H)])
    a, b = torch.load(src, weights_only=False), torch.load(dst, weights_only=False)
    assert a["epoch"] == b["epoch"] and a["step"] == b["step"]
    assert len(torch.load(mid, weights_only=False)["optimizer_state_dict"]["param_groups"][0]["params"]) == \
        4 + 11 * c["num_blocks"] + 6
    for i, st in a["optimizer_state_dict"]["state"].items():
        assert torch.equal(b["optimizer_state_dict"]["state"][i]["exp_avg"], st["exp_avg"])
