"""CPU: the oracle restatement (oracle/ref_vit.py) against fixtures produced by the real reference modules
(tests/golden/make_golden.py).  This is what pins the oracle."""
import pytest
import torch

from conftest import load_golden, rel_err
from oracle import ref_vit


def _cfg(g):
    c = g["config"]
    return ref_vit.Cfg(c["input_channels"], c["num_classes"], c["num_patches"], c["embedding_size"], c["patch_size"],
                       c["num_heads"], c["num_blocks"], c["batch_size"])


@pytest.mark.parametrize("name", ["tiny", "odd", "c1"])
def test_oracle_matches_reference_fixture(name):
    g = load_golden(name)
    cfg = _cfg(g)
    x = g["images_u8"].float() / 255.0
    logits, loss, grads = ref_vit.loss_and_grads(g["state_dict"], x, g["labels"], cfg)
    assert rel_err(logits, g["logits"]) < 2e-6
    assert abs(float(loss) - float(g["loss"])) < 1e-6
    assert set(grads) == set(g["grads"])
    for k, ref in g["grads"].items():
        assert grads[k].shape == ref.shape
        assert rel_err(grads[k], ref) < 2e-5, k


@pytest.mark.parametrize("name", ["tiny", "odd"])
def test_oracle_attention_probs(name):
    g = load_golden(name)
    _, weis = ref_vit.vit_forward(g["state_dict"], g["images_u8"].float() / 255.0, _cfg(g), return_probs=True)
    assert weis[0].shape == g["probs0"].shape
    assert rel_err(weis[0], g["probs0"]) < 1e-5


def test_oracle_im2col_layout_is_conv2d():
    """Bit-exact patch indexing: one-hot pixels light up exactly column c*P*P+ph*P+pw of row gy*gw+gx."""
    B, C, H, W, P = 2, 3, 12, 8, 4
    for (b, c, y, xx) in [(0, 0, 0, 0), (1, 2, 11, 7), (0, 1, 5, 6)]:
        x = torch.zeros(B, C, H, W)
        x[b, c, y, xx] = 1.0
        cols = ref_vit.im2col(x, P)
        nz = cols.nonzero()
        assert nz.shape[0] == 1
        assert nz[0].tolist() == [b, (y // P) * (W // P) + xx // P, c * P * P + (y % P) * P + xx % P]
    x = torch.randn(B, C, H, W)
    w = torch.randn(5, C, P, P)
    conv = torch.nn.functional.conv2d(x, w, stride=P).flatten(2).permute(0, 2, 1)
    assert torch.allclose(ref_vit.im2col(x, P) @ w.reshape(5, -1).t(), conv, atol=1e-5)


def test_oracle_batch_mismatch_raises():
    g = load_golden("tiny")
    x = g["images_u8"].float()[:1] / 255.0
    with pytest.raises(RuntimeError):
        ref_vit.vit_forward(g["state_dict"], x, _cfg(g))


def test_oracle_adamw_matches_torch():
    torch.manual_seed(0)
    p0 = torch.randn(1000)
    p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.AdamW([p], lr=1e-4, weight_decay=1e-4)
    q, m, v = p0.clone(), torch.zeros(1000), torch.zeros(1000)
    for step in range(1, 4):
        gr = torch.randn(1000)
        p.grad = gr.clone()
        opt.step()
        ref_vit.adamw_step(q, gr, m, v, step)
        assert torch.allclose(q, p.detach(), rtol=1e-6, atol=1e-7)


def test_vit_c_baseline_known_answer():
    """The C prototype timed as a CPU baseline (oracle/_ref/vit_c_time, built from the reference's csrc/vit.c where
    it lies) reproduces the reference's own known-answer notebook (tests/multihead-attention-test.ipynb cell 0:
    seed 69, 4x4x1 image of ones, P=2, D=4, 2 heads -> CLS+patch embeddings 5x4)."""
    import os
    import subprocess
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "vit_c_time")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/vit_c_time not built (reference sources absent)")
    out = subprocess.run([exe, "kat"], capture_output=True, text=True, timeout=60).stdout.splitlines()
    i = next(k for k, ln in enumerate(out) if ln.startswith("Matrix: CLS+Patch Embeddings"))
    got = torch.tensor([[float(v) for v in out[i + 1 + r].split()] for r in range(5)])
    want = torch.tensor([[0.7825, 1.0174, 1.6778, 1.2251], [3.3223, 2.8748, 3.0050, 3.2986],
                         [3.3452, 2.4302, 3.0276, 2.3968], [3.0843, 3.1493, 2.6917, 3.3506],
                         [3.5323, 3.3369, 2.9087, 2.9229]])
    assert torch.allclose(got, want, atol=1.5e-4)
