"""PatchEmbedding and VisionTransformer with the reference's names, constructor signatures and state_dict keys
(reference src/VisionTransformer/vit.py:9-80), executing on libvitb200's sm_100a kernels.

Reference semantics reproduced (each verified against the reference modules, see tests/):
  * Conv2d(C,E,P,stride=P) patch embedding as an im2col GEMM, column order (c, ph, pw), token gy*(W/P)+gx  vit.py:20-30,40
  * one CLS vector per batch SLOT, parameter shape (batch_size, 1, E), concatenated LAST, then + pos_embd    vit.py:31-36,41-42
  * the classifier reads token 0 (the first patch, not CLS): Linear -> GELU(erf) -> LayerNorm(4E) -> Linear  vit.py:69-74,80
  * the attribute really is spelled `emdeddings`                                                              vit.py:52
The whole-model forward/backward is ONE autograd node (`_ViTFn`) that drives the stage-level C ABI
(embed -> L x block -> head); nothing in it falls back to torch ops, and CPU tensors are refused.
"""
import ctypes

import torch
import torch.nn as nn

from . import _C, config, transformer  # noqa: F401  (same sub-module imports as the reference vit.py:3)
from . import _engine as eng
from .transformer import _Precision, block_param_dict, block_param_tensors, export_block_probs


def _autocast_bf16():
    return torch.is_autocast_enabled("cuda") and torch.get_autocast_dtype("cuda") == torch.bfloat16


def _embed_param_dict(mod, w, b, cls, pos):
    E = w.shape[0]
    w2 = w.detach().reshape(E, -1)
    lp = mod._w(w).reshape(E, -1) if mod._kernel_dtype() == torch.bfloat16 else None
    return {"w_f32": w2, "w_lp": lp, "bias": b.detach(), "cls": cls.detach(), "pos": pos.detach()}


def _image_dims(mod_dtype, x, E, H, P, T, classes, p_drop, seed):
    B, C, Hi, Wi = x.shape
    return _C.Dims(B, T, E, H, C, P, Hi, Wi, classes, _C.dtype_code(mod_dtype), float(p_drop),
                   int(seed) & 0xFFFFFFFFFFFFFFFF, 1e-5)


def _check_images(x, C, batch_size):
    eng.require_cuda(x, "image batch")
    if x.dim() != 4 or x.shape[1] != C:
        raise RuntimeError(f"expected images of shape (B, {C}, H, W), got {tuple(x.shape)}")
    if x.shape[0] != batch_size:
        # the reference fails in torch.cat([x, cls_tkn_embd], dim=1) (vit.py:41) because cls is (batch_size,1,E)
        raise RuntimeError(f"Sizes of tensors must match except in dimension 1. Expected size {x.shape[0]} but got "
                           f"size {batch_size} for tensor number 1 in the list (cls_tkn_embd is per batch slot: "
                           f"the input batch must equal config.batch_size)")
    if x.dtype == torch.uint8:
        # raw pixels (B,C,H,W): the ToTensor /255 (train.py:151-155) is fused into the patch gather on the device
        return x.contiguous()
    return x.contiguous().float()


# ------------------------------------------------------------------------------------------------ PatchEmbedding
class _EmbedFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, cls, pos, mod):
        B = x.shape[0]
        E, T = w.shape[0], pos.shape[1]
        dt = mod._kernel_dtype()
        d = _image_dims(dt, x, E, 1, mod.patch_size, T, 1, 0.0, 0)
        L = _C.lib()
        acts = torch.empty(max(L.vitb200_embed_act_bytes(ctypes.byref(d)), 256), dtype=torch.uint8, device=x.device)
        out = torch.empty(B * T, E, dtype=torch.float32, device=x.device)
        plan = eng.Plan.__new__(eng.Plan)
        eng.Plan.embed_fwd(plan, d, _embed_param_dict(mod, w, b, cls, pos), x, out, acts)
        ctx.save_for_backward(acts)
        ctx.meta = (d, tuple(w.shape), tuple(cls.shape), tuple(pos.shape))
        return out.reshape(B, T, E)

    @staticmethod
    def backward(ctx, g):
        (acts,) = ctx.saved_tensors
        d, wshape, cshape, pshape = ctx.meta
        dev = g.device
        g = g.contiguous().float()
        grads = {"w": torch.zeros(wshape, dtype=torch.float32, device=dev),
                 "bias": torch.zeros(wshape[0], dtype=torch.float32, device=dev),
                 "cls": torch.zeros(cshape, dtype=torch.float32, device=dev),
                 "pos": torch.zeros(pshape, dtype=torch.float32, device=dev)}
        plan = eng.Plan.__new__(eng.Plan)
        eng.Plan.embed_bwd(plan, d, grads, g, acts)
        return None, grads["w"], grads["bias"], grads["cls"], grads["pos"], None


class PatchEmbedding(_Precision, nn.Module):
    """Reference vit.py:9-43.  `precision` picks the kernel family (fp32 CUDA-core / bf16 tcgen05); parameters
    are always fp32 masters."""

    def __init__(self, input_channels, embedding_size, patch_size, batch_size, num_patches, precision, device):
        super().__init__()
        self.sequence = nn.Sequential(
            nn.Conv2d(in_channels=input_channels, out_channels=embedding_size, kernel_size=patch_size,
                      stride=patch_size, device=device, dtype=torch.float32),
            nn.Flatten(2),
        )
        self.cls_tkn_embd = nn.Parameter(torch.randn(size=(batch_size, 1, embedding_size), device=device),
                                         requires_grad=True)
        self.pos_embd = nn.Parameter(torch.randn(size=(1, num_patches + 1, embedding_size), device=device),
                                     requires_grad=True)
        self.patch_size, self.batch_size, self.input_channels = patch_size, batch_size, input_channels
        self.precision = precision if precision in (torch.float32, torch.bfloat16) else torch.float32

    def forward(self, x):
        x = _check_images(x, self.input_channels, self.batch_size)
        conv = self.sequence[0]
        with torch.cuda.device(x.device):   # kernels go to torch's current stream OF THE TENSORS' device
            return _EmbedFn.apply(x, conv.weight, conv.bias, self.cls_tkn_embd, self.pos_embd, self)


# ------------------------------------------------------------------------------------------------ whole model
_HEAD_SHAPES = ("w1", "b1", "ln_w", "ln_b", "w3", "b3")


def head_param_tensors(model):
    m = model.mlp
    return (m[0].weight, m[0].bias, m[2].weight, m[2].bias, m[3].weight, m[3].bias)


def embed_param_tensors(model):
    e = model.emdeddings
    return (e.sequence[0].weight, e.sequence[0].bias, e.cls_tkn_embd, e.pos_embd)


class _ViTFn(torch.autograd.Function):
    """images -> logits as one autograd node over the stage ABI.  Parameter order of *params:
    4 embedding tensors, 11 per block (C-ABI order), 6 head tensors."""

    @staticmethod
    def forward(ctx, x, model, p_drop, seed, grad_mode, *params):
        emb, blocks = model.emdeddings, list(model.transformer_encoder.blocks)
        nL = len(blocks)
        dt = model._kernel_dtype()
        B = x.shape[0]
        E, T = emb.pos_embd.shape[2], emb.pos_embd.shape[1]
        H = blocks[0].multi_head.num_heads if nL else 1
        classes = model.mlp[3].out_features
        d = _image_dims(dt, x, E, H, emb.patch_size, T, classes, p_drop, seed)
        plan = eng.Plan.__new__(eng.Plan)
        plan.L = nL
        L = _C.lib()
        dev = x.device
        # `needs_input_grad` mirrors requires_grad of the inputs, NOT grad mode, and Function.forward itself always
        # runs with grad disabled: the caller captures torch.is_grad_enabled().  False under the reference's
        # `@torch.no_grad evaluate()` (train.py:29-44): no activation stash, one scratch arena, ping-pong residuals.
        need_grad = bool(grad_mode) and any(ctx.needs_input_grad)
        nb = lambda n: torch.empty(max(int(n), 256), dtype=torch.uint8, device=dev)  # noqa: E731

        pe = _embed_param_dict(emb, *params[:4])
        emb_acts = nb(L.vitb200_embed_act_bytes(ctypes.byref(d)))
        xs = [torch.empty(B * T, E, dtype=torch.float32, device=dev)]
        eng.Plan.embed_fwd(plan, d, pe, x, xs[0], emb_acts)

        blk_bytes = L.vitb200_block_act_bytes(ctypes.byref(d))
        blk_acts, blk_params = [], []
        shared = None if need_grad else nb(blk_bytes)   # eval / no_grad: one scratch arena, ping-pong residuals
        for i, blk in enumerate(blocks):
            pd = block_param_dict(blk, params[4 + 11 * i: 4 + 11 * (i + 1)])
            acts = nb(blk_bytes) if need_grad else shared
            if need_grad:
                x_out = torch.empty_like(xs[0])
            else:
                x_out = xs[-2] if len(xs) >= 2 else torch.empty_like(xs[0])
            eng.Plan.block_fwd(plan, d, i, pd, xs[-1], x_out, acts)
            export_block_probs(blk, d, acts)
            xs.append(x_out)
            if need_grad:
                blk_acts.append(acts)
                blk_params.append(pd)
            elif len(xs) > 2:
                xs = xs[-2:]

        hp = dict(zip(_C.HEAD_FIELDS, (t.detach() for t in params[4 + 11 * nL:])))
        hp["w1_lp"] = model._w(params[4 + 11 * nL]) if dt == torch.bfloat16 else None   # bf16 shadow of mlp.0.weight
        head_acts = nb(L.vitb200_head_act_bytes(ctypes.byref(d)))
        logits = torch.empty(B, classes, dtype=torch.float32, device=dev)
        eng.Plan.head_fwd(plan, d, hp, xs[-1], logits, head_acts)

        if need_grad:
            ctx.state = (d, nL, dt, emb_acts, xs, blk_acts, blk_params, hp, head_acts,
                         [tuple(p.shape) for p in params])
            # the kernels read the live fp32 parameters (and bf16 shadows of them) through raw pointers, outside
            # autograd's saved-tensor version check: remember the versions and check them in backward ourselves
            ctx.versions = [(p, p._version) for p in params]
        return logits

    @staticmethod
    def backward(ctx, dlogits):
        if getattr(ctx, "state", None) is None:
            raise RuntimeError("VisionTransformer (B200): the saved activations of this forward were already consumed "
                               "and released by a backward pass; retain_graph / a second backward is not supported")
        for p, ver in ctx.versions:
            if p._version != ver:
                raise RuntimeError("one of the variables needed for gradient computation has been modified by an "
                                   f"inplace operation: a parameter of shape {tuple(p.shape)} is at version "
                                   f"{p._version}; expected version {ver} (VisionTransformer B200 forward)")
        d, nL, dt, emb_acts, xs, blk_acts, blk_params, hp, head_acts, shapes = ctx.state
        ctx.state = None
        ctx.versions = None
        dev = dlogits.device
        L = _C.lib()
        plan = eng.Plan.__new__(eng.Plan)
        plan.L = nL
        # one zero-filled fp32 gradient arena; the kernels accumulate into it and autograd receives views
        numels = [int(torch.Size(s).numel()) for s in shapes]
        offs, tot = [], 0
        for n in numels:
            offs.append(tot)
            tot += (n + 63) // 64 * 64
        arena = torch.zeros(tot, dtype=torch.float32, device=dev)
        gv = [arena[o:o + n].view(s) for o, n, s in zip(offs, numels, shapes)]

        M, E = d.B * d.T, d.E
        g = torch.empty(M, E, dtype=torch.float32, device=dev)
        dy_a = torch.empty(M, E, dtype=dt, device=dev)
        dy_b = torch.empty(M, E, dtype=dt, device=dev)
        scratch = torch.empty(max(L.vitb200_block_scratch_bytes(ctypes.byref(d)),
                                  L.vitb200_head_act_bytes(ctypes.byref(d)), 256), dtype=torch.uint8, device=dev)
        hg = dict(zip(_C.HEAD_FIELDS, gv[4 + 11 * nL:]))
        eng.Plan.head_bwd(plan, d, hp, hg, head_acts, dlogits.contiguous().float(), g, dy_a, scratch)
        bgs = eng.chain_block_grads([dict(zip(_C.BLOCK_FIELDS_G, gv[4 + 11 * i: 4 + 11 * (i + 1)]))
                                     for i in range(nL)])
        for i in range(nL - 1, -1, -1):
            eng.Plan.block_bwd(plan, d, i, blk_params[i], bgs[i], xs[i], blk_acts[i], g, dy_a, dy_b, scratch)
            blk_acts[i] = None  # release this block's activations as soon as they are consumed
            dy_a, dy_b = dy_b, dy_a
        eg = dict(zip(_C.EMBED_FIELDS_G, gv[:4]))
        eng.Plan.embed_bwd(plan, d, eg, g, emb_acts)
        return (None, None, None, None, None) + tuple(gv)


class VisionTransformer(_Precision, nn.Module):
    """Reference vit.py:47-80: PatchEmbedding -> TransformerEncoder -> classifier MLP on token 0."""

    def __init__(self, config):
        super().__init__()
        self.emdeddings = PatchEmbedding(
            input_channels=config.input_channels,
            embedding_size=config.embedding_size,
            patch_size=config.patch_size,
            num_patches=config.num_patches,
            precision=config.precision,
            batch_size=config.batch_size,
            device=config.device,
        )
        self.transformer_encoder = transformer.TransformerEncoder(
            embedding_size=config.embedding_size,
            num_heads=config.num_heads,
            num_blocks=config.num_blocks,
            block_size=config.num_patches + 1,
        )
        self.mlp = nn.Sequential(
            nn.Linear(config.embedding_size, 4 * config.embedding_size),
            nn.GELU(),
            nn.LayerNorm(4 * config.embedding_size),
            nn.Linear(4 * config.embedding_size, config.num_classes),
        )
        self.set_precision(config.precision if config.precision in (torch.float32, torch.bfloat16)
                           else torch.float32)

    def all_param_tensors(self):
        """Parameters in the order `_ViTFn` and the flat-arena trainer use (embed, blocks in C-ABI order, head)."""
        out = list(embed_param_tensors(self))
        for blk in self.transformer_encoder.blocks:
            out.extend(block_param_tensors(blk))
        out.extend(head_param_tensors(self))
        return out

    def forward(self, x):
        x = _check_images(x, self.emdeddings.input_channels, self.emdeddings.batch_size)
        params = self.all_param_tensors()
        for p in params:
            eng.require_cuda(p, "model parameters (call model.to('cuda'))")
        with torch.cuda.device(x.device):   # kernels go to torch's current stream OF THE TENSORS' device
            return _ViTFn.apply(x, self, self._drop(), eng.next_seed(), torch.is_grad_enabled(), *params)
