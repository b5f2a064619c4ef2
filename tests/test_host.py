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
