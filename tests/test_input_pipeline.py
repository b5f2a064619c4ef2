"""Device-side Resize (reference src/train.py:151-155 `transforms.Resize((img, img))` + `ToTensor()`).
CPU: the coefficient tables the host builds reproduce PIL's bilinear resample bit for bit when applied with PIL's own
two-pass fixed-point arithmetic (restated here in numpy: test infrastructure).  GPU: the kernel's output equals
`PIL.Image.resize(..., BILINEAR)` exactly, and feeding it to the model equals torchvision's Resize + ToTensor."""
import numpy as np
import pytest
import torch
from PIL import Image

from VisionTransformer import input_pipeline as ip

SHAPES = [(32, 32, 256, 256), (32, 32, 224, 224), (500, 375, 224, 224), (37, 53, 64, 48), (224, 224, 224, 224),
          (300, 300, 336, 336), (640, 480, 384, 384), (17, 19, 5, 7), (64, 48, 64, 96), (50, 96, 80, 96)]


def _apply_tables(img, hout, wout):
    """PIL ImagingResample restated: horizontal pass, then vertical pass, each rounded/clipped to uint8."""
    H, W = img.shape
    cur = img.astype(np.int64)
    if wout != W:
        b, k, _ = ip.resample_tables(W, wout)
        b, k = b.numpy(), k.numpy().astype(np.int64)
        out = np.zeros((H, wout), np.int64)
        for xx in range(wout):
            x0, n = b[xx]
            out[:, xx] = np.clip(((cur[:, x0:x0 + n] * k[xx, :n]).sum(1) + (1 << 21)) >> 22, 0, 255)
        cur = out
    if hout != H:
        b, k, _ = ip.resample_tables(H, hout)
        b, k = b.numpy(), k.numpy().astype(np.int64)
        out = np.zeros((hout, cur.shape[1]), np.int64)
        for yy in range(hout):
            y0, n = b[yy]
            out[yy] = np.clip(((cur[y0:y0 + n] * k[yy, :n, None]).sum(0) + (1 << 21)) >> 22, 0, 255)
        cur = out
    return cur.astype(np.uint8)


@pytest.mark.parametrize("h,w,ho,wo", SHAPES)
def test_tables_reproduce_pil_bilinear(h, w, ho, wo):
    a = np.random.default_rng(h * 1000 + w).integers(0, 256, (h, w), dtype=np.uint8)
    ref = np.asarray(Image.fromarray(a).resize((wo, ho), Image.BILINEAR))
    assert np.array_equal(_apply_tables(a, ho, wo), ref)


def test_tables_shape_and_normalisation():
    b, k, taps = ip.resample_tables(500, 224)
    assert b.shape == (224, 2) and k.shape == (224, taps) and taps == 7
    assert int(b[:, 0].min()) >= 0 and int((b[:, 0] + b[:, 1]).max()) <= 500
    s = k.sum(1)
    assert int((s - (1 << 22)).abs().max()) <= taps          # weights sum to 1 in fixed point (rounding only)


def test_resize_refuses_cpu_tensors():
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ip.Resize(8)(torch.zeros(1, 3, 4, 4, dtype=torch.uint8))


@pytest.mark.gpu
@pytest.mark.parametrize("h,w,ho,wo", SHAPES)
def test_gpu_resize_is_bit_identical_to_pil(lib, h, w, ho, wo):
    rng = np.random.default_rng(h + 7 * w)
    a = rng.integers(0, 256, (2, 3, h, w), dtype=np.uint8)
    out = ip.Resize(ho, wo)(torch.from_numpy(a).cuda()).cpu().numpy()
    for b in range(2):
        ref = np.asarray(Image.fromarray(a[b].transpose(1, 2, 0)).resize((wo, ho), Image.BILINEAR)).transpose(2, 0, 1)
        assert np.array_equal(out[b], ref), (b, int(np.abs(out[b].astype(int) - ref.astype(int)).max()))


@pytest.mark.gpu
def test_gpu_resize_plus_model_equals_torchvision_loader(lib):
    """CIFAR-sized 32x32 uint8 images -> device Resize(64) -> model(uint8) == model(torchvision Resize + ToTensor)."""
    import torchvision.transforms as T
    from VisionTransformer import config, vit
    B, img = 4, 64
    rng = np.random.default_rng(3)
    raw = rng.integers(0, 256, (B, 32, 32, 3), dtype=np.uint8)
    tf = T.Compose([T.Resize((img, img)), T.ToTensor()])                      # reference train.py:151-155
    x_ref = torch.stack([tf(Image.fromarray(r)) for r in raw]).cuda()
    x_u8 = ip.Resize(img)(torch.from_numpy(raw).permute(0, 3, 1, 2).contiguous().cuda())
    # ToTensor divides by 255 on the CPU (true division; a CUDA `x / 255` multiplies by a reciprocal and differs by 1 ulp)
    assert torch.equal(x_u8.cpu().float() / 255, x_ref.cpu())
    torch.manual_seed(0)
    m = vit.VisionTransformer(config.ViTConfig(3, 10, (img // 16) ** 2, 64, 16, 4, 2, "cuda", B)).cuda().eval()
    assert torch.equal(m(x_u8), m(x_ref))
