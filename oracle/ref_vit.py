"""ORACLE — test infrastructure only.  Never imported by the product path (vision-transformer_b200/).

A functional CPU/any-device restatement, in plain torch ops, of the reference model's arithmetic
(SiddhantSKarki/Vision-Transformer, src/VisionTransformer/{vit,transformer}.py).  It consumes a state_dict
with the REFERENCE's key names and shapes, so weights copied out of either implementation can be fed to it.

Parity status: PINNED.  tests/golden/*.pt hold logits, attention probabilities and gradients produced by
importing the real reference modules in the build container (tests/golden/make_golden.py); tests/test_oracle.py
checks this file against them.  /root/reference does not exist on the GPU box, so GPU parity tests compare
the CUDA path with this oracle (and with the golden fixtures directly).

Reference semantics restated here (file:line under the reference root):
  * patch embedding = Conv2d(C,E,P,stride=P) -> flatten -> (B,N,E); written as an explicit im2col matmul with
    column order (c, ph, pw) and token order gy*(W/P)+gx, floor-cropping non-divisible images   vit.py:20-30,40
  * CLS vectors (one per batch slot, shape (B,1,E)) concatenated LAST, then + pos_embd          vit.py:31-36,41-42
  * per-head bias-free q/k/v projections; logits MULTIPLIED by sqrt(head_size)                 transformer.py:12-24
  * softmax over keys, out = wei @ v, heads concatenated, proj with bias, dropout              transformer.py:25-29,44-47
  * MLP = Linear(E,4E) -> ReLU -> Linear(4E,E) -> dropout                                      transformer.py:55-60
  * pre-LN residual blocks, LayerNorm eps 1e-5                                                  transformer.py:72-79
  * classifier reads token 0: Linear(E,4E) -> exact GELU -> LayerNorm(4E) -> Linear(4E,cls)     vit.py:69-74,80
"""
import math

import torch
import torch.nn.functional as F


class Cfg:
    """Shape bag mirroring reference config.py:6-29 (only the fields the arithmetic needs)."""

    def __init__(self, input_channels, num_classes, num_patches, embedding_size, patch_size, num_heads, num_blocks,
                 batch_size):
        self.input_channels = input_channels
        self.num_classes = num_classes
        self.num_patches = num_patches
        self.embedding_size = embedding_size
        self.patch_size = patch_size
        self.num_heads = num_heads
        self.num_blocks = num_blocks
        self.batch_size = batch_size


def im2col(x, P):
    """(B,C,H,W) -> (B, N, C*P*P): row gy*(W//P)+gx, column c*P*P + ph*P + pw (floor-cropped)."""
    B, C, H, W = x.shape
    gh, gw = H // P, W // P
    x = x[:, :, : gh * P, : gw * P]
    x = x.reshape(B, C, gh, P, gw, P).permute(0, 2, 4, 1, 3, 5)
    return x.reshape(B, gh * gw, C * P * P)


def patch_embed(sd, x, P, prefix="emdeddings."):
    w = sd[prefix + "sequence.0.weight"]
    b = sd[prefix + "sequence.0.bias"]
    E = w.shape[0]
    cols = im2col(x.to(w.dtype), P)
    tok = cols @ w.reshape(E, -1).t() + b
    cls = sd[prefix + "cls_tkn_embd"]
    if cls.shape[0] != tok.shape[0]:
        raise RuntimeError(f"batch {tok.shape[0]} != config.batch_size {cls.shape[0]} (reference vit.py:41 cat fails)")
    tok = torch.cat([tok, cls], dim=1)
    return tok + sd[prefix + "pos_embd"]


def head(sd, prefix, x):
    """Single attention head (transformer.py:20-31): returns (out (B,T,hs), wei (B,T,T))."""
    k = x @ sd[prefix + "key.weight"].t()
    q = x @ sd[prefix + "query.weight"].t()
    v = x @ sd[prefix + "value.weight"].t()
    hs = k.shape[-1]
    wei = torch.softmax((q @ k.transpose(-2, -1)) * math.sqrt(hs), dim=-1)
    return wei @ v, wei


def _drop(t, mask, p):
    return t if mask is None else t * mask.to(t.dtype) / (1.0 - p)


def multi_head(sd, prefix, x, num_heads, drop_mask=None, p=0.2):
    outs, weis = [], []
    for h in range(num_heads):
        o, w = head(sd, f"{prefix}heads.{h}.", x)
        outs.append(o)
        weis.append(w)
    y = torch.cat(outs, dim=-1) @ sd[prefix + "proj.weight"].t() + sd[prefix + "proj.bias"]
    return _drop(y, drop_mask, p), torch.stack(weis, dim=1)


def feed_forward(sd, prefix, x, drop_mask=None, p=0.2):
    h = torch.relu(x @ sd[prefix + "mlp.0.weight"].t() + sd[prefix + "mlp.0.bias"])
    y = h @ sd[prefix + "mlp.2.weight"].t() + sd[prefix + "mlp.2.bias"]
    return _drop(y, drop_mask, p)


def layer_norm(sd, prefix, x):
    return F.layer_norm(x, x.shape[-1:], sd[prefix + "weight"], sd[prefix + "bias"], 1e-5)


def block(sd, prefix, x, num_heads, masks=(None, None)):
    a, wei = multi_head(sd, prefix + "multi_head.", layer_norm(sd, prefix + "ln1.", x), num_heads, masks[0])
    x = x + a
    x = x + feed_forward(sd, prefix + "ffwd.", layer_norm(sd, prefix + "ln2.", x), masks[1])
    return x, wei


def encoder(sd, x, num_heads, num_blocks, masks=None, prefix="transformer_encoder.blocks."):
    weis = []
    for i in range(num_blocks):
        x, w = block(sd, f"{prefix}{i}.", x, num_heads, masks[i] if masks is not None else (None, None))
        weis.append(w)
    return x, weis


def classifier(sd, x0, prefix="mlp."):
    h = F.gelu(x0 @ sd[prefix + "0.weight"].t() + sd[prefix + "0.bias"])
    h = F.layer_norm(h, h.shape[-1:], sd[prefix + "2.weight"], sd[prefix + "2.bias"], 1e-5)
    return h @ sd[prefix + "3.weight"].t() + sd[prefix + "3.bias"]


def vit_forward(sd, x, cfg, masks=None, return_probs=False):
    """Full model (vit.py:77-80).  masks: optional per-block (proj_mask, fc2_mask) keep-masks for dropout."""
    t = patch_embed(sd, x, cfg.patch_size)
    t, weis = encoder(sd, t, cfg.num_heads, cfg.num_blocks, masks)
    logits = classifier(sd, t[:, 0, :])
    return (logits, weis) if return_probs else logits


def loss_and_grads(sd, x, labels, cfg, masks=None):
    """Cross-entropy (mean) + gradients w.r.t. every state_dict entry (train.py:81,93-95)."""
    leaf = {k: v.detach().clone().requires_grad_(True) for k, v in sd.items()}
    logits = vit_forward(leaf, x, cfg, masks)
    loss = F.cross_entropy(logits, labels)
    loss.backward()
    return logits.detach(), loss.detach(), {k: v.grad for k, v in leaf.items()}


def adamw_step(p, g, m, v, step, lr=1e-4, beta1=0.9, beta2=0.999, eps=1e-8, wd=1e-4):
    """torch.optim.AdamW semantics (train.py:66): decoupled decay, bias-corrected moments.  In place."""
    p.mul_(1.0 - lr * wd)
    m.mul_(beta1).add_(g, alpha=1.0 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1.0 - beta2)
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
    p.addcdiv_(m, denom, value=-lr / bc1)
    return p


def random_state_dict(cfg, seed=0, dtype=torch.float32, device="cpu"):
    """A state_dict with the reference's keys/shapes, filled deterministically (NOT the reference's init;
    parity tests copy the same tensors into both sides, so the distribution is irrelevant)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    E, H, C, P = cfg.embedding_size, cfg.num_heads, cfg.input_channels, cfg.patch_size
    hs, T = E // H, cfg.num_patches + 1

    def rn(*shape, scale=1.0):
        return (torch.randn(*shape, generator=g) * scale).to(dtype).to(device)

    sd = {
        "emdeddings.cls_tkn_embd": rn(cfg.batch_size, 1, E),
        "emdeddings.pos_embd": rn(1, T, E),
        "emdeddings.sequence.0.weight": rn(E, C, P, P, scale=(C * P * P) ** -0.5),
        "emdeddings.sequence.0.bias": rn(E, scale=0.1),
    }
    for i in range(cfg.num_blocks):
        b = f"transformer_encoder.blocks.{i}."
        for h in range(H):
            for n in ("key", "query", "value"):
                sd[f"{b}multi_head.heads.{h}.{n}.weight"] = rn(hs, E, scale=E ** -0.5 * (0.5 if n != "value" else 1.0))
        sd[b + "multi_head.proj.weight"] = rn(E, E, scale=E ** -0.5)
        sd[b + "multi_head.proj.bias"] = rn(E, scale=0.1)
        sd[b + "ffwd.mlp.0.weight"] = rn(4 * E, E, scale=E ** -0.5)
        sd[b + "ffwd.mlp.0.bias"] = rn(4 * E, scale=0.1)
        sd[b + "ffwd.mlp.2.weight"] = rn(E, 4 * E, scale=(4 * E) ** -0.5)
        sd[b + "ffwd.mlp.2.bias"] = rn(E, scale=0.1)
        for ln in ("ln1", "ln2"):
            sd[f"{b}{ln}.weight"] = 1.0 + rn(E, scale=0.1)
            sd[f"{b}{ln}.bias"] = rn(E, scale=0.1)
    sd["mlp.0.weight"] = rn(4 * E, E, scale=E ** -0.5)
    sd["mlp.0.bias"] = rn(4 * E, scale=0.1)
    sd["mlp.2.weight"] = 1.0 + rn(4 * E, scale=0.1)
    sd["mlp.2.bias"] = rn(4 * E, scale=0.1)
    sd["mlp.3.weight"] = rn(cfg.num_classes, 4 * E, scale=(4 * E) ** -0.5)
    sd["mlp.3.bias"] = rn(cfg.num_classes, scale=0.1)
    return sd
