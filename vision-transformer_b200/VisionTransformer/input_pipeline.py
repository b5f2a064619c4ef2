"""Device-side input pipeline: the `transforms.Resize((img, img))` + `ToTensor()` of the reference's
loader (src/train.py:151-155) on the GPU, for raw uint8 batches.

torchvision's Resize on a PIL image is PIL's bilinear `ImagingResample`: a horizontal then a vertical pass, each a
normalised triangle filter (support widened by the scale factor when shrinking) in 22-bit fixed point, each pass
rounded and clipped to uint8.  `resample_tables` builds the per-output-pixel (first tap, tap count) bounds and the
fixed-point weights exactly as PIL does (double precision, round-half-away); `vitb200_resize_bilinear_u8` applies them.
The result is bit-identical to `PIL.Image.resize(..., BILINEAR)`; the `/255` of ToTensor is fused into the patch
gather (`model(uint8 batch)` / `DPTrainer.step_host_u8`).  No CPU fallback: CPU tensors raise.
"""
import math

import torch

from . import _C
from . import _engine as eng

PRECISION_BITS = 22  # PIL: 32 - 8 - 2


def resample_tables(in_size, out_size):
    """-> (bounds int32 [out, 2], coeffs int32 [out, taps], taps) of one bilinear pass in_size -> out_size."""
    scale = in_size / out_size
    filterscale = max(scale, 1.0)
    support = 1.0 * filterscale                      # bilinear filter support = 1
    taps = int(math.ceil(support)) * 2 + 1
    bounds = torch.zeros(out_size, 2, dtype=torch.int32)
    coeffs = torch.zeros(out_size, taps, dtype=torch.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        xmin = max(int(center - support + 0.5), 0)
        xmax = min(int(center + support + 0.5), in_size) - xmin
        w = []
        for x in range(xmax):
            a = abs((x + xmin - center + 0.5) * ss)
            w.append(1.0 - a if a < 1.0 else 0.0)
        ww = sum(w)
        if ww != 0.0:
            w = [v / ww for v in w]
        bounds[xx, 0], bounds[xx, 1] = xmin, xmax
        for x, v in enumerate(w):
            coeffs[xx, x] = int(-0.5 + v * (1 << PRECISION_BITS)) if v < 0 else int(0.5 + v * (1 << PRECISION_BITS))
    return bounds, coeffs, taps


class Resize:
    """`transforms.Resize((out_h, out_w))` for uint8 batches (B, C, H, W) resident on the GPU.  Tables are cached per
    input size; `__call__` enqueues one kernel on the current stream of the batch's device and allocates only the
    output (pass `out=` to reuse a buffer)."""

    def __init__(self, out_h, out_w=None):
        self.out_h, self.out_w = int(out_h), int(out_w if out_w is not None else out_h)
        self._tables = {}

    def _get(self, hin, win, device):
        key = (hin, win, str(device))
        if key not in self._tables:
            xb, xk, tx = resample_tables(win, self.out_w)
            yb, yk, ty = resample_tables(hin, self.out_h)
            self._tables[key] = (xb.to(device), xk.to(device), tx, yb.to(device), yk.to(device), ty)
        return self._tables[key]

    def __call__(self, images_u8, out=None):
        eng.require_cuda(images_u8, "image batch")
        if images_u8.dtype != torch.uint8 or images_u8.dim() != 4:
            raise RuntimeError(f"Resize expects a uint8 (B, C, H, W) batch, got {images_u8.dtype} {tuple(images_u8.shape)}")
        x = images_u8.contiguous()
        B, C, hin, win = x.shape
        if out is None:
            out = torch.empty(B, C, self.out_h, self.out_w, dtype=torch.uint8, device=x.device)
        elif tuple(out.shape) != (B, C, self.out_h, self.out_w) or out.dtype != torch.uint8 or not out.is_contiguous():
            raise RuntimeError("Resize: `out` must be a contiguous uint8 (B, C, out_h, out_w) tensor")
        xb, xk, tx, yb, yk, ty = self._get(hin, win, x.device)
        with torch.cuda.device(x.device):
            _C.check(_C.lib().vitb200_resize_bilinear_u8(x.data_ptr(), out.data_ptr(), B * C, hin, win, self.out_h,
                                                         self.out_w, xb.data_ptr(), xk.data_ptr(), tx, yb.data_ptr(),
                                                         yk.data_ptr(), ty, _C.stream_ptr(x.device)))
        return out
