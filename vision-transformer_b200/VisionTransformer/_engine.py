"""Host-side runtime over the stage-level C ABI (embed -> L x block -> head).

`Plan` fixes the shapes of one model instance, owns nothing but sizes, and turns torch tensors into the
pointer structs the C ABI takes.  It is shared by the nn.Module drop-in (autograd Functions in vit.py /
transformer.py) and by the flat-arena data-parallel trainer (trainer.py).  There is no CPU fallback.
"""
import ctypes

import torch

from . import _C

DROPOUT_P = 0.2  # hard-coded in the reference (transformer.py:35,53); config.dropout is never passed down


def require_cuda(t, what="input"):
    if not t.is_cuda:
        raise RuntimeError(f"VisionTransformer (B200): {what} must live on a CUDA device; there is no CPU fallback")


class Plan:
    def __init__(self, *, B, T, E, H, C, P, Hi, Wi, classes, L, dtype):
        self.B, self.T, self.E, self.H, self.C, self.P, self.Hi, self.Wi = B, T, E, H, C, P, Hi, Wi
        self.classes, self.L, self.dtype = classes, L, dtype
        self.code = _C.dtype_code(dtype)
        self.M = B * T
        d = self.dims(0.0, 0)
        L_ = _C.lib()
        self.embed_act_bytes = L_.vitb200_embed_act_bytes(ctypes.byref(d))
        self.block_act_bytes = L_.vitb200_block_act_bytes(ctypes.byref(d))
        self.block_scratch_bytes = L_.vitb200_block_scratch_bytes(ctypes.byref(d))
        self.head_act_bytes = L_.vitb200_head_act_bytes(ctypes.byref(d))

    def dims(self, dropout_p, seed, seed_dev=None):
        """seed_dev: optional device address of a uint64 added to `seed` on the device (captured steps)."""
        return _C.Dims(self.B, self.T, self.E, self.H, self.C, self.P, self.Hi, self.Wi, self.classes, self.code,
                       float(dropout_p), int(seed) & 0xFFFFFFFFFFFFFFFF, 1e-5, seed_dev)

    # ---- raw stage calls (tensors in, nothing returned; all async on the current stream) -------------------
    @staticmethod
    def _fill(struct_cls, fields, tensors):
        s = struct_cls()
        for f in fields:
            t = tensors.get(f)
            setattr(s, f, None if t is None else t.data_ptr())
        return s

    def embed_fwd(self, d, params, images, x_out, acts, u8_layout=0):
        p = self._fill(_C.EmbedParams, _C.EMBED_FIELDS_P, params)
        if images.dtype == torch.uint8:  # raw pixels, layout 0 = (B,C,H,W), 1 = (B,H,W,C); /255 fused on the device
            _C.check(_C.lib().vitb200_embed_fwd_u8(ctypes.byref(d), ctypes.byref(p), images.data_ptr(), int(u8_layout),
                                                   x_out.data_ptr(), acts.data_ptr(), _C.stream_ptr()))
            return
        if images.dtype != torch.float32:
            raise RuntimeError(f"images must be float32 or uint8, got {images.dtype}")
        _C.check(_C.lib().vitb200_embed_fwd(ctypes.byref(d), ctypes.byref(p), images.data_ptr(), x_out.data_ptr(),
                                            acts.data_ptr(), _C.stream_ptr()))

    def embed_bwd(self, d, grads, g, acts):
        gr = self._fill(_C.EmbedGrads, _C.EMBED_FIELDS_G, grads)
        _C.check(_C.lib().vitb200_embed_bwd(ctypes.byref(d), ctypes.byref(gr), g.data_ptr(), acts.data_ptr(),
                                            acts.data_ptr(), _C.stream_ptr()))

    def block_fwd(self, d, layer, params, x_in, x_out, acts, row0=False):
        p = self._fill(_C.BlockParams, _C.BLOCK_FIELDS_P, params)
        fn = _C.lib().vitb200_block_fwd_row0 if row0 else _C.lib().vitb200_block_fwd
        _C.check(fn(ctypes.byref(d), layer, ctypes.byref(p), x_in.data_ptr(), x_out.data_ptr(),
                                            acts.data_ptr(), _C.stream_ptr()))

    def block_bwd(self, d, layer, params, grads, x_in, acts, g, dy_in, dy_out, scratch, row0=False):
        p = self._fill(_C.BlockParams, _C.BLOCK_FIELDS_P, params)
        gr = self._fill(_C.BlockGrads, _C.BLOCK_GRADS_STRUCT, grads)
        fn = _C.lib().vitb200_block_bwd_row0 if row0 else _C.lib().vitb200_block_bwd
        _C.check(fn(ctypes.byref(d), layer, ctypes.byref(p), ctypes.byref(gr), x_in.data_ptr(),
                                            acts.data_ptr(), g.data_ptr(), dy_in.data_ptr(),
                                            None if dy_out is None else dy_out.data_ptr(), scratch.data_ptr(),
                                            _C.stream_ptr()))

    def head_fwd(self, d, params, x_last, logits, acts, row0=False):
        p = self._fill(_C.HeadParams, _C.HEAD_PARAMS_STRUCT, params)
        fn = _C.lib().vitb200_head_fwd_row0 if row0 else _C.lib().vitb200_head_fwd
        _C.check(fn(ctypes.byref(d), ctypes.byref(p), x_last.data_ptr(), logits.data_ptr(),
                                           acts.data_ptr(), _C.stream_ptr()))

    def head_bwd(self, d, params, grads, acts, dlogits, g, dy_out, scratch, row0=False):
        p = self._fill(_C.HeadParams, _C.HEAD_PARAMS_STRUCT, params)
        gr = self._fill(_C.HeadGrads, _C.HEAD_FIELDS, grads)
        fn = _C.lib().vitb200_head_bwd_row0 if row0 else _C.lib().vitb200_head_bwd
        _C.check(fn(ctypes.byref(d), self.L, ctypes.byref(p), ctypes.byref(gr), acts.data_ptr(),
                                           dlogits.data_ptr(), g.data_ptr(), dy_out.data_ptr(), scratch.data_ptr(),
                                           _C.stream_ptr()))

    # ---- buffers ------------------------------------------------------------------------------------------
    def new_bytes(self, n, device):
        return torch.empty(max(int(n), 256), dtype=torch.uint8, device=device)

    def new_stream_tensor(self, device):
        return torch.empty(self.M, self.E, dtype=torch.float32, device=device)

    def new_lp(self, device):
        """[M, E] buffer in the activation dtype (masked gradient operands)."""
        return torch.empty(self.M, self.E, dtype=self.dtype, device=device)


def chain_block_grads(block_grads):
    """Per-layer gradient dicts for a whole-model backward: the LayerNorm-1 backward of layer i+1 already sums the
    fc2 bias gradient of layer i (it produces the masked gradient that sum runs over), so layer i passes
    b_fc2 = None (skip) and layer i+1 passes b_fc2_prev = layer i's b_fc2.  The last layer keeps its own b_fc2:
    its incoming gradient is produced by the classifier head."""
    out = []
    n = len(block_grads)
    for i, g in enumerate(block_grads):
        c = dict(g)
        c["b_fc2_prev"] = block_grads[i - 1]["b_fc2"] if i > 0 else None
        if i < n - 1:
            c["b_fc2"] = None
        out.append(c)
    return out


# ---- small op wrappers used by the sub-module Functions and by tests -----------------------------------------
def cast_bf16(src, dst=None):
    if dst is None:
        dst = torch.empty(src.shape, dtype=torch.bfloat16, device=src.device)
    _C.check(_C.lib().vitb200_cast_f32_bf16(src.data_ptr(), dst.data_ptr(), src.numel(), _C.stream_ptr()))
    return dst


class ShadowCache:
    """bf16 copies of fp32 master parameters, refreshed when the parameter's version counter (or storage) moves.

    A refresh writes into a NEW buffer: an earlier forward may still hold the old shadow for its backward
    (forward, optimizer.step(), forward, backward-of-the-first must see the weights of ITS forward).  In-place
    updates made through `p.data` do not bump `p._version`; call `invalidate()` (or
    `model.invalidate_shadows()`) after such an update."""

    def __init__(self):
        self._c = {}

    def get(self, p):
        key = id(p)
        ent = self._c.get(key)
        tag = (p._version, p.data_ptr())
        if ent is None or ent[0] != tag or ent[1].device != p.device:
            buf = cast_bf16(p.detach().contiguous())
            self._c[key] = (tag, buf)
            return buf
        return ent[1]

    def invalidate(self):
        self._c.clear()


def mask_cast(g, dtype, p, seed, stream_id):
    out = torch.empty(g.shape, dtype=dtype, device=g.device)
    _C.check(_C.lib().vitb200_mask_cast(g.data_ptr(), out.data_ptr(), _C.dtype_code(dtype),
                                        _C.Dropout(float(p), int(seed) & 0xFFFFFFFFFFFFFFFF, int(stream_id)), g.numel(),
                                        _C.stream_ptr()))
    return out


_step_counter = [0]


def next_seed():
    """A fresh dropout seed per forward call, derived from torch's global seed."""
    _step_counter[0] += 1
    return (torch.initial_seed() * 0x9E3779B97F4A7C15 + _step_counter[0] * 0xD1B54A32D192ED03) & 0xFFFFFFFFFFFFFFFF
