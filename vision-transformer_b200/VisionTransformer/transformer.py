"""Transformer encoder modules with the reference's names, constructor signatures and state_dict keys
(reference src/VisionTransformer/transformer.py), executing on libvitb200's sm_100a kernels.

Differences in storage only: the 3*H per-head bias-free projections of a MultiHeadAttention are held as ONE packed
parameter `qkv_weight` [3E, E] (row s*E + h*hs + j = heads[h].{query,key,value}[s].weight[j]); state_dict()/
load_state_dict() translate to and from the reference's `heads.{h}.{key,query,value}.weight` keys.
"""
import ctypes
import math

import torch
import torch.nn as nn

from . import _C
from . import _engine as eng

_SECTIONS = ("query", "key", "value")  # packed order q | k | v


def _dims_for(x, E, H, dtype, p, seed):
    B, T, _ = x.shape
    return _C.Dims(B, T, E, H, 1, 1, 1, 1, 1, _C.dtype_code(dtype), float(p), int(seed) & 0xFFFFFFFFFFFFFFFF, 1e-5)


def _check_input(x, E):
    eng.require_cuda(x)
    if x.dim() != 3 or x.shape[-1] != E:
        raise RuntimeError(f"expected input of shape (B, T, {E}), got {tuple(x.shape)}")
    return x.contiguous().float()


def _lp(t, dtype):
    """Activation-dtype copy of an fp32 tensor (identity in fp32 mode)."""
    return t if dtype == torch.float32 else eng.cast_bf16(t)


class _Precision:
    """Mixin: `precision` (torch.float32 | torch.bfloat16) and the bf16 weight-shadow cache."""
    precision = torch.float32

    def _kernel_dtype(self):
        """bf16 (tcgen05 path) when configured so or when called under torch.autocast(cuda, bfloat16)."""
        if self.precision == torch.bfloat16:
            return torch.bfloat16
        if torch.is_autocast_enabled("cuda") and torch.get_autocast_dtype("cuda") == torch.bfloat16:
            return torch.bfloat16
        return torch.float32

    def _w(self, p):
        if self._kernel_dtype() == torch.float32:
            return p.detach()
        if not hasattr(self, "_shadow"):
            object.__setattr__(self, "_shadow", eng.ShadowCache())
        return self._shadow.get(p)

    def invalidate_shadows(self):
        """Drop every cached bf16 weight copy below this module (needed only after in-place updates through
        `param.data`, which do not move the version counter the cache keys on)."""
        for m in self.modules():
            if isinstance(m, _Precision) and hasattr(m, "_shadow"):
                m._shadow.invalidate()
        return self

    def set_precision(self, dtype):
        for m in self.modules():
            if isinstance(m, _Precision):
                m.precision = dtype
        return self

    def _drop(self):
        return eng.DROPOUT_P * float(getattr(self, "dropout_scale", 1.0)) if self.training else 0.0


# ------------------------------------------------------------------------------------------------ Head
class _HeadFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, wq, wk, wv, mod):
        B, T, E = x.shape
        hs = wq.shape[0]
        dt = mod._kernel_dtype()
        w = torch.cat([wq, wk, wv], dim=0).contiguous()      # packed [3hs, E]
        xl, wl = _lp(x.reshape(B * T, E), dt), _lp(w, dt)
        qkv = torch.empty(B * T, 3 * hs, dtype=dt, device=x.device)
        _C.gemm(xl, wl, qkv, M=B * T, N=3 * hs, K=E)
        out = torch.empty(B * T, hs, dtype=dt, device=x.device)
        lse = torch.empty(B, 1, T, dtype=torch.float32, device=x.device)
        scale = math.sqrt(hs)
        _C.check(_C.lib().vitb200_attn_fwd(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), _C.dtype_code(dt), B, T, 1, hs,
                                           scale, _C.stream_ptr()))
        probs = torch.empty(B, T, T, dtype=torch.float32, device=x.device)
        qkv32 = qkv if dt == torch.float32 else qkv.float()
        _C.check(_C.lib().vitb200_attn_probs(qkv32.data_ptr(), probs.data_ptr(), B, T, 1, hs, scale, _C.stream_ptr()))
        ctx.save_for_backward(xl, wl, qkv, out, lse)
        ctx.meta = (B, T, E, hs, dt)
        ctx.mark_non_differentiable(probs)
        return out.float().reshape(B, T, hs), probs

    @staticmethod
    def backward(ctx, g_out, _g_probs):
        xl, wl, qkv, out, lse = ctx.saved_tensors
        B, T, E, hs, dt = ctx.meta
        d_o = _lp(g_out.contiguous().float().reshape(B * T, hs), dt)
        dqkv = torch.empty_like(qkv)
        delta = torch.empty(B, 1, T, dtype=torch.float32, device=qkv.device)
        _C.check(_C.lib().vitb200_attn_bwd(qkv.data_ptr(), out.data_ptr(), d_o.data_ptr(), lse.data_ptr(),
                                           delta.data_ptr(), dqkv.data_ptr(), _C.dtype_code(dt), B, T, 1, hs,
                                           math.sqrt(hs), _C.stream_ptr()))
        dw = torch.zeros(3 * hs, E, dtype=torch.float32, device=qkv.device)
        _C.gemm(dqkv, xl, dw, a_major=_C.MAJOR_MN, b_major=_C.MAJOR_MN, M=3 * hs, N=E, K=B * T, accumulate=True)
        dx = torch.empty(B * T, E, dtype=torch.float32 if dt == torch.float32 else dt, device=qkv.device)
        _C.gemm(dqkv, wl, dx, b_major=_C.MAJOR_MN, M=B * T, N=E, K=3 * hs)
        return dx.float().reshape(B, T, E), dw[:hs], dw[hs:2 * hs], dw[2 * hs:], None


class Head(_Precision, nn.Module):
    """Single attention head (reference transformer.py:9-31): returns (out, wei)."""

    def __init__(self, head_size, n_embd, block_size):
        super().__init__()
        self.key = nn.Linear(n_embd, head_size, bias=False)
        self.query = nn.Linear(n_embd, head_size, bias=False)
        self.value = nn.Linear(n_embd, head_size, bias=False)

    def forward(self, x):
        x = _check_input(x, self.key.in_features)
        if self._kernel_dtype() == torch.bfloat16 and self.key.out_features % 16 != 0:
            raise RuntimeError("bf16 Head needs head_size % 16 == 0; use precision float32")
        with torch.cuda.device(x.device):
            return _HeadFn.apply(x, self.query.weight, self.key.weight, self.value.weight, self)


# ------------------------------------------------------------------------------------------------ MHA
class _ParamView:
    def __init__(self, weight):
        self.weight = weight


class _HeadView:
    """Read-only view of one head's slices of the packed qkv parameter (reference path heads[h].key.weight)."""

    def __init__(self, packed, h, hs, E):
        self.query = _ParamView(packed[0 * E + h * hs: 0 * E + (h + 1) * hs])
        self.key = _ParamView(packed[1 * E + h * hs: 1 * E + (h + 1) * hs])
        self.value = _ParamView(packed[2 * E + h * hs: 2 * E + (h + 1) * hs])


class _MHAFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, wqkv, wproj, bproj, mod, p_drop, seed):
        B, T, E = x.shape
        H = mod.num_heads
        Eh = wqkv.shape[0] // 3  # H * hs
        hs = Eh // H
        dt = mod._kernel_dtype()
        M = B * T
        xl = _lp(x.reshape(M, E), dt)
        wq, wp = mod._w(wqkv), mod._w(wproj)
        qkv = torch.empty(M, 3 * Eh, dtype=dt, device=x.device)
        _C.gemm(xl, wq, qkv, M=M, N=3 * Eh, K=E)
        att = torch.empty(M, Eh, dtype=dt, device=x.device)
        lse = torch.empty(B, H, T, dtype=torch.float32, device=x.device)
        _C.check(_C.lib().vitb200_attn_fwd(qkv.data_ptr(), att.data_ptr(), lse.data_ptr(), _C.dtype_code(dt), B, T, H, hs,
                                           math.sqrt(hs), _C.stream_ptr()))
        y = torch.empty(M, E, dtype=torch.float32, device=x.device)
        _C.gemm(att, wp, y, M=M, N=E, K=Eh, bias=bproj.detach(), dropout=(p_drop, seed, 0))
        ctx.save_for_backward(xl, wq, wp, qkv, att, lse)
        ctx.meta = (B, T, E, H, hs, dt, p_drop, seed)
        if mod.export_attention_probs:
            probs = torch.empty(B, H, T, T, dtype=torch.float32, device=x.device)
            qkv32 = qkv if dt == torch.float32 else qkv.float()
            _C.check(_C.lib().vitb200_attn_probs(qkv32.data_ptr(), probs.data_ptr(), B, T, H, hs, math.sqrt(hs),
                                                 _C.stream_ptr()))
            mod.attention_probs = probs
        return y.reshape(B, T, E)

    @staticmethod
    def backward(ctx, g):
        xl, wq, wp, qkv, att, lse = ctx.saved_tensors
        B, T, E, H, hs, dt, p_drop, seed = ctx.meta
        M, Eh = B * T, H * hs
        dev = g.device
        dy = eng.mask_cast(g.contiguous().float().reshape(M, E), dt, p_drop, seed, 0)
        dwp = torch.zeros(E, Eh, dtype=torch.float32, device=dev)
        dbp = torch.zeros(E, dtype=torch.float32, device=dev)
        _C.gemm(dy, att, dwp, a_major=_C.MAJOR_MN, b_major=_C.MAJOR_MN, M=E, N=Eh, K=M, accumulate=True)
        _C.check(_C.lib().vitb200_colsum(dy.data_ptr(), _C.dtype_code(dt), E, dbp.data_ptr(), M, E, _C.stream_ptr()))
        d_o = torch.empty(M, Eh, dtype=dt, device=dev)
        _C.gemm(dy, wp, d_o, b_major=_C.MAJOR_MN, M=M, N=Eh, K=E)
        dqkv = torch.empty_like(qkv)
        delta = torch.empty(B, H, T, dtype=torch.float32, device=dev)
        _C.check(_C.lib().vitb200_attn_bwd(qkv.data_ptr(), att.data_ptr(), d_o.data_ptr(), lse.data_ptr(),
                                           delta.data_ptr(), dqkv.data_ptr(), _C.dtype_code(dt), B, T, H, hs,
                                           math.sqrt(hs), _C.stream_ptr()))
        dwq = torch.zeros(3 * Eh, E, dtype=torch.float32, device=dev)
        _C.gemm(dqkv, xl, dwq, a_major=_C.MAJOR_MN, b_major=_C.MAJOR_MN, M=3 * Eh, N=E, K=M, accumulate=True)
        dx = torch.empty(M, E, dtype=dt, device=dev)
        _C.gemm(dqkv, wq, dx, b_major=_C.MAJOR_MN, M=M, N=E, K=3 * Eh)
        return dx.float().reshape(B, T, E), dwq, dwp, dbp, None, None, None


class MultiHeadAttention(_Precision, nn.Module):
    """Reference transformer.py:34-49.  `attention_probs` (B,H,T,T) is only materialised when
    `export_attention_probs` is set (the fused kernel never forms it); default None."""

    def __init__(self, num_heads, head_size, n_embd, block_size, dropout=0.2):
        super().__init__()
        self.num_heads, self.head_size, self.n_embd = num_heads, head_size, n_embd
        Eh = num_heads * head_size
        w = torch.empty(3 * Eh, n_embd)
        for s in range(3 * num_heads):  # same per-head init as 3H separate nn.Linear(n_embd, head_size, bias=False)
            nn.init.kaiming_uniform_(w[s * head_size:(s + 1) * head_size], a=math.sqrt(5))
        self.qkv_weight = nn.Parameter(w)
        self.proj = nn.Linear(n_embd, n_embd)
        self.attention_probs = None
        self.export_attention_probs = False
        self.dropout = nn.Dropout(dropout)

    @property
    def heads(self):
        Eh = self.num_heads * self.head_size
        return [_HeadView(self.qkv_weight, h, self.head_size, Eh) for h in range(self.num_heads)]

    # ---- reference-key translation ------------------------------------------------------------------------
    def _save_to_state_dict(self, destination, prefix, keep_vars):
        super()._save_to_state_dict(destination, prefix, keep_vars)
        packed = destination.pop(prefix + "qkv_weight")
        Eh, hs = self.num_heads * self.head_size, self.head_size
        for h in range(self.num_heads):
            for name in ("key", "query", "value"):  # the reference's registration order (transformer.py:12-17)
                s = _SECTIONS.index(name)
                destination[f"{prefix}heads.{h}.{name}.weight"] = packed[s * Eh + h * hs: s * Eh + (h + 1) * hs]

    def _load_from_state_dict(self, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys,
                              error_msgs):
        Eh, hs = self.num_heads * self.head_size, self.head_size
        keys = [(f"{prefix}heads.{h}.{name}.weight", s * Eh + h * hs)
                for h in range(self.num_heads) for s, name in enumerate(_SECTIONS)]
        if any(k in state_dict for k, _ in keys):
            packed = self.qkv_weight.detach().clone()
            for k, row in keys:
                if k in state_dict:
                    t = state_dict.pop(k)
                    if tuple(t.shape) != (hs, self.n_embd):
                        error_msgs.append(f"size mismatch for {k}: {tuple(t.shape)} vs {(hs, self.n_embd)}")
                        continue
                    packed[row:row + hs].copy_(t)
                elif strict:
                    missing_keys.append(k)
            state_dict[prefix + "qkv_weight"] = packed
        super()._load_from_state_dict(state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys,
                                      error_msgs)

    def forward(self, x):
        x = _check_input(x, self.n_embd)
        p = self._drop() and self.dropout.p
        with torch.cuda.device(x.device):
            return _MHAFn.apply(x, self.qkv_weight, self.proj.weight, self.proj.bias, self, float(p), eng.next_seed())


# ------------------------------------------------------------------------------------------------ FeedForward
class _FFNFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, mod, p_drop, seed):
        B, T, E = x.shape
        M, F = B * T, w1.shape[0]
        dt = mod._kernel_dtype()
        xl = _lp(x.reshape(M, E), dt)
        w1l, w2l = mod._w(w1), mod._w(w2)
        h = torch.empty(M, F, dtype=dt, device=x.device)
        _C.gemm(xl, w1l, h, M=M, N=F, K=E, bias=b1.detach(), relu=True)
        y = torch.empty(M, E, dtype=torch.float32, device=x.device)
        _C.gemm(h, w2l, y, M=M, N=E, K=F, bias=b2.detach(), dropout=(p_drop, seed, 1))
        ctx.save_for_backward(xl, w1l, w2l, h)
        ctx.meta = (B, T, E, F, dt, p_drop, seed)
        return y.reshape(B, T, E)

    @staticmethod
    def backward(ctx, g):
        xl, w1l, w2l, h = ctx.saved_tensors
        B, T, E, F, dt, p_drop, seed = ctx.meta
        M, dev, code = B * T, g.device, _C.dtype_code(dt)
        dy = eng.mask_cast(g.contiguous().float().reshape(M, E), dt, p_drop, seed, 1)
        dw2 = torch.zeros(E, F, dtype=torch.float32, device=dev)
        db2 = torch.zeros(E, dtype=torch.float32, device=dev)
        _C.gemm(dy, h, dw2, a_major=_C.MAJOR_MN, b_major=_C.MAJOR_MN, M=E, N=F, K=M, accumulate=True)
        _C.check(_C.lib().vitb200_colsum(dy.data_ptr(), code, E, db2.data_ptr(), M, E, _C.stream_ptr()))
        dh = torch.empty(M, F, dtype=dt, device=dev)
        _C.gemm(dy, w2l, dh, b_major=_C.MAJOR_MN, M=M, N=F, K=E, mask_aux=h)
        dw1 = torch.zeros(F, E, dtype=torch.float32, device=dev)
        db1 = torch.zeros(F, dtype=torch.float32, device=dev)
        _C.gemm(dh, xl, dw1, a_major=_C.MAJOR_MN, b_major=_C.MAJOR_MN, M=F, N=E, K=M, accumulate=True)
        _C.check(_C.lib().vitb200_colsum(dh.data_ptr(), code, F, db1.data_ptr(), M, F, _C.stream_ptr()))
        dx = torch.empty(M, E, dtype=dt, device=dev)
        _C.gemm(dh, w1l, dx, b_major=_C.MAJOR_MN, M=M, N=E, K=F)
        return dx.float().reshape(B, T, E), dw1, db1, dw2, db2, None, None, None


class FeedForward(_Precision, nn.Module):
    """Reference transformer.py:52-64: Linear(E,4E) -> ReLU -> Linear(4E,E) -> Dropout."""

    def __init__(self, n_embd, dropout=0.2):
        super().__init__()
        self.mlp = nn.Sequential(
            nn.Linear(n_embd, 4 * n_embd),
            nn.ReLU(),
            nn.Linear(4 * n_embd, n_embd),
            nn.Dropout(dropout),
        )

    def forward(self, x):
        x = _check_input(x, self.mlp[0].in_features)
        p = self._drop() and self.mlp[3].p
        with torch.cuda.device(x.device):
            return _FFNFn.apply(x, self.mlp[0].weight, self.mlp[0].bias, self.mlp[2].weight, self.mlp[2].bias, self,
                                float(p), eng.next_seed())


# ------------------------------------------------------------------------------------------------ Block
def block_param_tensors(blk):
    """The 11 parameter tensors of a Block in the order of the C ABI's vitb200_block_params."""
    return (blk.ln1.weight, blk.ln1.bias, blk.ln2.weight, blk.ln2.bias, blk.multi_head.qkv_weight,
            blk.multi_head.proj.weight, blk.multi_head.proj.bias, blk.ffwd.mlp[0].weight, blk.ffwd.mlp[0].bias,
            blk.ffwd.mlp[2].weight, blk.ffwd.mlp[2].bias)


_LP_WEIGHTS = ("w_qkv", "w_proj", "w_fc1", "w_fc2")


def block_param_dict(blk, tensors=None):
    """name -> tensor to hand to the kernels (bf16 shadows for the four matrices in bf16 mode)."""
    tensors = tensors if tensors is not None else block_param_tensors(blk)
    out = {}
    for name, t in zip(_C.BLOCK_FIELDS_P, tensors):
        out[name] = blk._w(t) if name in _LP_WEIGHTS else t.detach()
    return out


def export_block_probs(blk, d, acts):
    """MultiHeadAttention.attention_probs (B,H,T,T), which the reference sets on EVERY forward (transformer.py:48):
    recomputed from the q/k the block forward just stored in its activation arena, only when the module opted in
    (`export_attention_probs`) — the fused attention kernel never materialises it."""
    mh = blk.multi_head
    if not mh.export_attention_probs:
        mh.attention_probs = None
        return
    M, E, hs = d.B * d.T, d.E, d.E // d.H
    dt = torch.bfloat16 if d.dtype == _C.BF16 else torch.float32
    off = int(_C.lib().vitb200_block_act_offset(ctypes.byref(d), _C.ACT_QKV))
    qkv = acts[off: off + M * 3 * E * dt.itemsize].view(dt).view(M, 3 * E)
    qkv32 = qkv if dt == torch.float32 else qkv.float()
    probs = torch.empty(d.B, d.H, d.T, d.T, dtype=torch.float32, device=acts.device)
    _C.check(_C.lib().vitb200_attn_probs(qkv32.data_ptr(), probs.data_ptr(), d.B, d.T, d.H, hs, math.sqrt(hs),
                                         _C.stream_ptr()))
    mh.attention_probs = probs


class _BlockFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, blk, layer, p_drop, seed, *params):
        B, T, E = x.shape
        H = blk.multi_head.num_heads
        d = _dims_for(x, E, H, blk._kernel_dtype(), p_drop, seed)
        L = _C.lib()
        acts = torch.empty(L.vitb200_block_act_bytes(ctypes.byref(d)), dtype=torch.uint8, device=x.device)
        y = torch.empty(B * T, E, dtype=torch.float32, device=x.device)
        pd = block_param_dict(blk, params)
        plan = eng.Plan.__new__(eng.Plan)
        eng.Plan.block_fwd(plan, d, layer, pd, x, y, acts)
        export_block_probs(blk, d, acts)
        ctx.save_for_backward(x, acts, *[pd[n] for n in _C.BLOCK_FIELDS_P])
        ctx.meta = (d, layer, blk._kernel_dtype())
        return y.reshape(B, T, E)

    @staticmethod
    def backward(ctx, g):
        x, acts, *ptens = ctx.saved_tensors
        d, layer, dt = ctx.meta
        dev = x.device
        M, E = d.B * d.T, d.E
        g = g.contiguous().float().clone().reshape(M, E)
        pd = dict(zip(_C.BLOCK_FIELDS_P, ptens))
        shapes = {"ln1_w": (E,), "ln1_b": (E,), "ln2_w": (E,), "ln2_b": (E,), "w_qkv": (3 * E, E), "w_proj": (E, E),
                  "b_proj": (E,), "w_fc1": (4 * E, E), "b_fc1": (4 * E,), "w_fc2": (E, 4 * E), "b_fc2": (E,)}
        grads = {n: torch.zeros(shapes[n], dtype=torch.float32, device=dev) for n in _C.BLOCK_FIELDS_G}
        dy_in = eng.mask_cast(g, dt, d.dropout_p, d.seed, layer * 2 + 1)
        scratch = torch.empty(_C.lib().vitb200_block_scratch_bytes(ctypes.byref(d)), dtype=torch.uint8, device=dev)
        plan = eng.Plan.__new__(eng.Plan)
        eng.Plan.block_bwd(plan, d, layer, pd, grads, x, acts, g, dy_in, None, scratch)
        return (g.reshape(d.B, d.T, E), None, None, None, None) + tuple(grads[n] for n in _C.BLOCK_FIELDS_G)


class Block(_Precision, nn.Module):
    """Pre-LN encoder block (reference transformer.py:66-79): x + mha(ln1(x)); x + ffwd(ln2(x))."""

    def __init__(self, n_embd, n_head, block_size):
        super().__init__()
        head_size = n_embd // n_head
        self.multi_head = MultiHeadAttention(n_head, head_size, n_embd, block_size)
        self.ffwd = FeedForward(n_embd)
        self.ln1 = nn.LayerNorm(n_embd)
        self.ln2 = nn.LayerNorm(n_embd)
        self.layer_index = 0

    def forward(self, x):
        E = self.ln1.normalized_shape[0]
        x = _check_input(x, E)
        if self.multi_head.num_heads * self.multi_head.head_size != E:
            raise RuntimeError("Block requires n_embd divisible by n_head")
        with torch.cuda.device(x.device):
            return _BlockFn.apply(x, self, self.layer_index, self._drop(), eng.next_seed(),
                                  *block_param_tensors(self))


class TransformerEncoder(_Precision, nn.Module):
    """Reference transformer.py:82-90: nn.Sequential of Blocks, no final norm."""

    def __init__(self, embedding_size, num_heads, num_blocks, block_size):
        super().__init__()
        self.blocks = nn.Sequential(
            *[Block(embedding_size, num_heads, block_size) for _ in range(num_blocks)]
        )
        for i, b in enumerate(self.blocks):
            b.layer_index = i

    def forward(self, x):
        return self.blocks(x)
