"""GPU: every kernel family through the C ABI against a plain torch fp32 reference of the same op."""
import ctypes
import math

import pytest
import torch

from conftest import cosine, rel_err
from oracle import ref_vit

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _C():
    from VisionTransformer import _C as c
    return c


# ------------------------------------------------------------------------------------------------ contractions
@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("variant,M,N,K", [("nt", 300, 200, 104), ("nt", 1000, 768, 768), ("nn", 257, 264, 512),
                                           ("tn", 264, 136, 1000), ("nt", 128, 256, 64), ("nt", 1, 8, 8),
                                           ("tn", 520, 264, 6400)])      # deep K: several atomic splits per tile
def test_gemm_layouts(lib, dt, variant, M, N, K):
    C = _C()
    torch.manual_seed(1)
    if variant == "nt":
        A, B = torch.randn(M, K, device=DEV), torch.randn(N, K, device=DEV)
        am, bm = 0, 0
    elif variant == "nn":
        A, B = torch.randn(M, K, device=DEV), torch.randn(K, N, device=DEV)
        am, bm = 0, 1
    else:
        A, B = torch.randn(K, M, device=DEV), torch.randn(K, N, device=DEV)
        am, bm = 1, 1
    A, B = A.to(dt), B.to(dt)
    Af = A.float() if am == 0 else A.float().t()
    Bf = B.float() if bm == 0 else B.float().t()
    ref = Af @ Bf.t()
    out = torch.zeros(M, N, device=DEV, dtype=torch.float32)
    C.gemm(A, B, out, a_major=am, b_major=bm, M=M, N=N, K=K, accumulate=(variant == "tn"))
    tol = 1e-5 if dt == torch.float32 else 1e-5  # bf16 inputs are exact in the fp32 reference; accumulation is fp32
    assert rel_err(out, ref) < max(tol, 3e-6 * math.sqrt(K))


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
def test_gemm_epilogue_bias_relu_residual_rowmap(lib, dt):
    C = _C()
    torch.manual_seed(2)
    M, N, K = 6 * 16, 64, 72
    A, B = torch.randn(M, K, device=DEV).to(dt), torch.randn(N, K, device=DEV).to(dt)
    bias, res = torch.randn(N, device=DEV), torch.randn(M, N, device=DEV)
    out = torch.empty(M, N, device=DEV)
    C.gemm(A, B, out, M=M, N=N, K=K, bias=bias, relu=True, residual=res)
    ref = torch.relu(A.float() @ B.float().t() + bias) + res
    assert rel_err(out, ref) < 1e-5
    # row map used by the patch embedding: 6 groups of 16 rows land in groups of 17 rows (+ per-row add)
    pos = torch.randn(16, N, device=DEV)
    out2 = torch.zeros(6 * 17, N, device=DEV)
    C.gemm(A, B, out2, M=M, N=N, K=K, bias=bias, row_add=pos, row_period=16, row_group_stride=17)
    ref2 = (A.float() @ B.float().t() + bias).reshape(6, 16, N) + pos
    assert rel_err(out2.reshape(6, 17, N)[:, :16], ref2) < 1e-5
    assert torch.count_nonzero(out2.reshape(6, 17, N)[:, 16]) == 0
    # output in the operand dtype + aux mask (dReLU) + fused column sums (bias gradient)
    aux = torch.randn(M, N, device=DEV).to(dt)
    out3 = torch.empty(M, N, device=DEV, dtype=dt)
    cs = torch.ones(N, device=DEV)
    C.gemm(A, B, out3, M=M, N=N, K=K, mask_aux=aux, col_sum=cs)
    ref3 = (A.float() @ B.float().t()) * (aux.float() > 0)
    assert rel_err(out3, ref3) < (5e-3 if dt == torch.bfloat16 else 1e-5)
    assert rel_err(cs - 1, ref3.sum(0)) < 1e-5


@pytest.mark.parametrize("M,N,K", [(300, 96, 64), (1000, 3072, 128), (128, 32, 16)])
def test_gemm_relu_sign_bits_roundtrip(lib, M, N, K):
    """fc1 writes the ReLU mask as 1 bit per element (bits_out); the ReLU backward GEMM consumes it (mask_bits) and
    must give exactly what re-reading the bf16 activation (mask_aux) gives.  Word (n/32)*rows + m, bit n%32."""
    C = _C()
    torch.manual_seed(8)
    A, B = torch.randn(M, K, device=DEV).bfloat16(), torch.randn(N, K, device=DEV).bfloat16()
    bias = torch.randn(N, device=DEV)
    h = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
    bits = torch.zeros(N // 32, M, device=DEV, dtype=torch.int32)
    C.gemm(A, B, h, M=M, N=N, K=K, bias=bias, relu=True, bits_out=bits, bits_rows=M)
    ref = torch.relu(A.float() @ B.float().t() + bias)
    assert rel_err(h, ref) < 5e-3
    unpacked = ((bits.unsqueeze(-1) >> torch.arange(32, device=DEV, dtype=torch.int32)) & 1)   # [N/32, M, 32]
    unpacked = unpacked.permute(1, 0, 2).reshape(M, N).bool()
    assert torch.equal(unpacked, h.float() > 0)
    # backward-style product with an MN-major weight: mask by bits == mask by the activation
    dy = torch.randn(M, K, device=DEV).bfloat16()
    W = torch.randn(K, N, device=DEV).bfloat16()
    a, b = (torch.empty(M, N, device=DEV, dtype=torch.bfloat16) for _ in range(2))
    cs_a, cs_b = torch.zeros(N, device=DEV), torch.zeros(N, device=DEV)
    C.gemm(dy, W, a, M=M, N=N, K=K, b_major=C.MAJOR_MN, mask_aux=h, col_sum=cs_a)
    C.gemm(dy, W, b, M=M, N=N, K=K, b_major=C.MAJOR_MN, mask_bits=bits, bits_rows=M, col_sum=cs_b)
    assert torch.equal(a, b)
    assert rel_err(cs_b, cs_a) < 1e-5
    with pytest.raises(RuntimeError, match="bit masks"):
        C.gemm(A, B[:40], h[:, :40], M=M, N=40, K=K, relu=True, bits_out=bits, bits_rows=M, ldc=N)


def test_gemm_dropout_epilogue_matches_exported_mask(lib):
    C = _C()
    torch.manual_seed(3)
    M, N, K = 256, 128, 64
    A, B = torch.randn(M, K, device=DEV).bfloat16(), torch.randn(N, K, device=DEV).bfloat16()
    out = torch.empty(M, N, device=DEV)
    C.gemm(A, B, out, M=M, N=N, K=K, dropout=(0.2, 1234, 7))
    keep = torch.empty(M * N, dtype=torch.uint8, device=DEV)
    C.check(lib.vitb200_dropout_mask(keep.data_ptr(), M * N, C.Dropout(0.2, 1234, 7), C.stream_ptr()))
    ref = (A.float() @ B.float().t()) * keep.reshape(M, N).float() / 0.8
    assert rel_err(out, ref) < 1e-5
    frac = keep.float().mean().item()
    assert abs(frac - 0.8) < 0.01
    keep2 = torch.empty_like(keep)
    C.check(lib.vitb200_dropout_mask(keep2.data_ptr(), M * N, C.Dropout(0.2, 1234, 8), C.stream_ptr()))
    assert (keep != keep2).float().mean().item() > 0.2  # independent streams


@pytest.mark.parametrize("Bi,T,H,K", [(3, 197, 12, 768), (2, 65, 4, 256), (5, 130, 2, 64)])
def test_gemm_fused_head_dot_is_the_attention_backward_delta(lib, Bi, T, H, K):
    """kEpiDot: delta[b,h,t] = sum_j dO[b*T+t, 64h+j] * O[b*T+t, 64h+j] from the epilogue of the GEMM that produces dO
    (NN layout, as the block backward calls it), against torch on the STORED bf16 values and against the stand-alone
    delta kernel through vitb200_attn_bwd's scratch."""
    C = _C()
    torch.manual_seed(11)
    M, N = Bi * T, 64 * H
    A = (torch.randn(M, K, device=DEV) * 0.5).bfloat16()
    W = (torch.randn(K, N, device=DEV) / K ** 0.5).bfloat16()          # MN-major B operand: [K][N]
    O = torch.randn(M, N, device=DEV).bfloat16()
    dO = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
    delta = torch.full((Bi, H, T), float("nan"), device=DEV)
    C.gemm(A, W, dO, b_major=C.MAJOR_MN, M=M, N=N, K=K, dot_aux=O, dot_out=delta, dot_T=T, dot_H=H)
    plain = torch.empty_like(dO)
    C.gemm(A, W, plain, b_major=C.MAJOR_MN, M=M, N=N, K=K)
    assert torch.equal(dO, plain)                                       # the output itself is untouched by the fusion
    ref = (dO.float() * O.float()).reshape(Bi, T, H, 64).sum(-1).permute(0, 2, 1)
    assert torch.isfinite(delta).all()
    assert (delta - ref).abs().max().item() <= 2e-5 * (ref.abs().max().item() + 1.0)
    # bit-identical to the stand-alone delta kernel (same association of the 64 products): vitb200_attn_bwd fills its
    # `delta` scratch from (O, dO) before anything else
    qkv = (torch.randn(M, 3 * N, device=DEV) * 0.1).bfloat16()
    lse = torch.zeros(Bi, H, T, device=DEV)
    scratch = torch.full((Bi, H, T), float("nan"), device=DEV)
    dqkv = torch.empty_like(qkv)
    C.check(lib.vitb200_attn_bwd(qkv.data_ptr(), O.data_ptr(), dO.data_ptr(), lse.data_ptr(), scratch.data_ptr(),
                                 dqkv.data_ptr(), C.BF16, Bi, T, H, 64, 8.0, C.stream_ptr()))
    assert torch.equal(delta, scratch)
    with pytest.raises(RuntimeError, match="fused head dot"):
        C.gemm(A, W, dO, b_major=C.MAJOR_MN, M=M, N=N, K=K, dot_aux=O, dot_out=delta, dot_T=T, dot_H=H + 1)


def test_gemm_rejects_bad_arguments(lib):
    C = _C()
    A = torch.randn(16, 12, device=DEV).bfloat16()  # lda = 12: not a 16-byte row pitch
    B = torch.randn(16, 12, device=DEV).bfloat16()
    out = torch.empty(16, 16, device=DEV)
    with pytest.raises(RuntimeError, match="libvitb200"):
        C.gemm(A, B, out, M=16, N=16, K=12)


# ------------------------------------------------------------------------------------------------ LayerNorm
# M >= 4 x 148 rows take the register-resident forward kernel (one case per register-tile width: E <= 384, 768, 1024,
# 1536), fewer rows the shared-memory one
@pytest.mark.parametrize("M,E", [(197 * 3, 768), (50, 384), (7, 4), (33, 1280), (5, 3072), (3, 100), (1576, 384),
                                 (197 * 4, 768), (601, 1024), (1000, 1280), (597, 200)])
def test_layernorm_fwd_bwd(lib, M, E):
    C = _C()
    torch.manual_seed(4)
    x = (torch.randn(M, E, device=DEV) * 2 + 0.5).requires_grad_(True)
    w = (1 + 0.1 * torch.randn(E, device=DEV)).requires_grad_(True)
    b = (0.1 * torch.randn(E, device=DEV)).requires_grad_(True)
    y_ref = torch.nn.functional.layer_norm(x, (E,), w, b, 1e-5)
    dy = torch.randn(M, E, device=DEV)
    y_ref.backward(dy)
    y = torch.empty(M, E, device=DEV)
    mean, rstd = torch.empty(M, device=DEV), torch.empty(M, device=DEV)
    C.check(lib.vitb200_ln_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), C.F32, mean.data_ptr(),
                               rstd.data_ptr(), M, E, 1e-5, C.stream_ptr()))
    assert rel_err(y, y_ref) < 1e-5
    g0 = torch.randn(M, E, device=DEV)
    g = g0.clone()
    dgam, dbet = torch.zeros(E, device=DEV), torch.zeros(E, device=DEV)
    C.check(lib.vitb200_ln_bwd(dy.data_ptr(), C.F32, x.data_ptr(), mean.data_ptr(), rstd.data_ptr(), w.data_ptr(),
                               g.data_ptr(), 1, dgam.data_ptr(), dbet.data_ptr(), None, None, C.Dropout(0, 0, 0),
                               M, E, C.stream_ptr()))
    assert rel_err(g - g0, x.grad) < 2e-5
    assert rel_err(dgam, w.grad) < 2e-5
    assert rel_err(dbet, b.grad) < 2e-5
    if E % 8 == 0:
        yb = torch.empty(M, E, device=DEV, dtype=torch.bfloat16)
        C.check(lib.vitb200_ln_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), yb.data_ptr(), C.BF16, mean.data_ptr(),
                                   rstd.data_ptr(), M, E, 1e-5, C.stream_ptr()))
        assert rel_err(yb, y_ref) < 8e-3


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("M,E,add", [(197 * 7, 768, 1), (1000, 384, 1), (333, 1024, 0), (64, 8, 1), (4099, 192, 0),
                                     (50, 3072, 1), (9, 100, 1), (577 * 3, 1280, 1), (200, 1536, 0)])
def test_layernorm_bwd_masked_operand_and_bias_grad(lib, dt, M, E, add):
    """The fused outputs of the LayerNorm backward: g (+)= dx, dyn = cast(g_new * dropout factors) and
    dyn_colsum += column sums of dyn — bulk-copy ring kernel (E % 8 == 0, E <= 1024) and the generic one."""
    if dt == torch.bfloat16 and E % 8:
        pytest.skip("bf16 rows need 16-byte pitch")
    C = _C()
    torch.manual_seed(5)
    x = (torch.randn(M, E, device=DEV) * 1.5 - 0.3).requires_grad_(True)
    w = (1 + 0.1 * torch.randn(E, device=DEV)).requires_grad_(True)
    b = torch.zeros(E, device=DEV, requires_grad=True)
    dy = torch.randn(M, E, device=DEV).to(dt)
    torch.nn.functional.layer_norm(x, (E,), w, b, 1e-5).backward(dy.float())
    mean = x.detach().mean(1).contiguous()
    rstd = (x.detach().var(1, unbiased=False) + 1e-5).rsqrt().contiguous()
    g0 = torch.randn(M, E, device=DEV)
    g = g0.clone()
    dgam, dbet = torch.ones(E, device=DEV), torch.ones(E, device=DEV)   # accumulate on top of existing values
    csum = torch.full((E,), 2.0, device=DEV)
    dyn = torch.empty(M, E, device=DEV, dtype=dt)
    code = C.BF16 if dt == torch.bfloat16 else C.F32
    C.check(lib.vitb200_ln_bwd(dy.data_ptr(), code, x.data_ptr(), mean.data_ptr(), rstd.data_ptr(), w.data_ptr(),
                               g.data_ptr(), add, dgam.data_ptr(), dbet.data_ptr(), dyn.data_ptr(), csum.data_ptr(),
                               C.Dropout(0.2, 77, 5), M, E, C.stream_ptr()))
    g_ref = x.grad + (g0 if add else 0)
    assert rel_err(g, g_ref) < 2e-5
    assert rel_err(dgam - 1, w.grad) < 3e-5
    assert rel_err(dbet - 1, b.grad) < 3e-5
    keep = torch.empty(M * E, dtype=torch.uint8, device=DEV)
    C.check(lib.vitb200_dropout_mask(keep.data_ptr(), M * E, C.Dropout(0.2, 77, 5), C.stream_ptr()))
    dyn_ref = g_ref * keep.reshape(M, E).float() / 0.8
    assert rel_err(dyn, dyn_ref) < (5e-3 if dt == torch.bfloat16 else 2e-5)
    assert ((dyn == 0) == (keep.reshape(M, E) == 0)).float().mean().item() > 0.9999
    assert rel_err(csum - 2, dyn_ref.sum(0)) < (5e-3 if dt == torch.bfloat16 else 5e-5)


# ------------------------------------------------------------------------------------------------ attention
def _attn_ref(qkv, B, T, H, hs, scale):
    E = H * hs
    q, k, v = (qkv.reshape(B, T, 3, H, hs)[:, :, i].permute(0, 2, 1, 3) for i in range(3))
    p = torch.softmax((q @ k.transpose(-1, -2)) * scale, dim=-1)
    return (p @ v).permute(0, 2, 1, 3).reshape(B * T, E), p


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,T,H,hs", [(2, 197, 3, 64), (1, 257, 2, 64), (2, 65, 2, 80), (3, 17, 2, 32), (1, 577, 1, 80),
                                       (5, 256, 2, 64), (3, 128, 1, 64), (2, 100, 3, 64), (7, 5, 2, 64), (40, 129, 12, 64),
                                       # long rows / wide heads: the chunked tcgen05 kernels of attention_tc_gen.cu (bf16)
                                       (2, 300, 3, 64), (1, 1000, 2, 96), (2, 130, 2, 128), (3, 40, 2, 80),
                                       (2, 730, 2, 80), (3, 257, 1, 128), (20, 577, 16, 80),
                                       # chunk / tile boundaries of those kernels
                                       (2, 5, 2, 80), (1, 64, 1, 96), (1, 128, 1, 128), (2, 129, 1, 80), (1, 256, 2, 80),
                                       (1, 65, 1, 128), (2, 512, 1, 80)])
def test_attention_fwd_bwd(lib, dt, B, T, H, hs):
    C = _C()
    torch.manual_seed(5)
    E = H * hs
    scale = math.sqrt(hs)  # the reference MULTIPLIES by sqrt(hs): keep logits O(10) so softmax is peaked, not one-hot
    qkv = (torch.randn(B * T, 3 * E, device=DEV) * (0.45 / math.sqrt(hs))).to(dt)
    qkv[:, 2 * E:] = torch.randn(B * T, E, device=DEV).to(dt)
    ref_in = qkv.float().requires_grad_(True)
    o_ref, _ = _attn_ref(ref_in, B, T, H, hs, scale)
    d_o = torch.randn(B * T, E, device=DEV).to(dt)
    o_ref.backward(d_o.float())
    out = torch.empty(B * T, E, device=DEV, dtype=dt)
    lse = torch.empty(B, H, T, device=DEV)
    C.check(lib.vitb200_attn_fwd(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), C.dtype_code(dt), B, T, H, hs, scale,
                                 C.stream_ptr()))
    tol = 2e-5 if dt == torch.float32 else 1.5e-2
    assert rel_err(out, o_ref) < tol
    dqkv = torch.empty_like(qkv)
    delta = torch.empty(B, H, T, device=DEV)
    C.check(lib.vitb200_attn_bwd(qkv.data_ptr(), out.data_ptr(), d_o.data_ptr(), lse.data_ptr(), delta.data_ptr(),
                                 dqkv.data_ptr(), C.dtype_code(dt), B, T, H, hs, scale, C.stream_ptr()))
    if dt == torch.float32:
        assert rel_err(dqkv, ref_in.grad) < 1e-4
    else:
        for s in range(3):
            assert cosine(dqkv[:, s * E:(s + 1) * E], ref_in.grad[:, s * E:(s + 1) * E]) > 0.999


def test_attention_hs1_and_probs_export(lib):
    """hs = 1 is the repo-default config (E=4, H=4); probs export = MultiHeadAttention.attention_probs."""
    C = _C()
    torch.manual_seed(6)
    B, T, H, hs = 2, 197, 4, 1
    qkv = torch.randn(B * T, 3 * H * hs, device=DEV)
    out = torch.empty(B * T, H * hs, device=DEV)
    lse = torch.empty(B, H, T, device=DEV)
    C.check(lib.vitb200_attn_fwd(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), C.F32, B, T, H, hs, 1.0,
                                 C.stream_ptr()))
    o_ref, p_ref = _attn_ref(qkv, B, T, H, hs, 1.0)
    assert rel_err(out, o_ref) < 2e-5
    probs = torch.empty(B, H, T, T, device=DEV)
    C.check(lib.vitb200_attn_probs(qkv.data_ptr(), probs.data_ptr(), B, T, H, hs, 1.0, C.stream_ptr()))
    assert rel_err(probs, p_ref) < 1e-5


# ------------------------------------------------------------------------------------------------ patch gather (bit-exact)
@pytest.mark.parametrize("B,Cc,Hi,Wi,P", [(2, 3, 32, 48, 16), (1, 1, 40, 40, 10), (2, 3, 30, 30, 14)])
def test_im2col_bit_exact_layout(lib, B, Cc, Hi, Wi, P):
    C = _C()
    torch.manual_seed(7)
    x = torch.rand(B, Cc, Hi, Wi, device=DEV)
    N, K = (Hi // P) * (Wi // P), Cc * P * P
    A = torch.empty(B * N, K, device=DEV)
    C.check(lib.vitb200_im2col(x.data_ptr(), A.data_ptr(), C.F32, B, Cc, Hi, Wi, P, K, C.stream_ptr()))
    ref = ref_vit.im2col(x.cpu(), P).reshape(B * N, K)
    assert torch.equal(A.cpu(), ref)  # pure index permutation: must be bit-exact
    if K % 8 == 0:
        Ab = torch.empty(B * N, K, device=DEV, dtype=torch.bfloat16)
        C.check(lib.vitb200_im2col(x.data_ptr(), Ab.data_ptr(), C.BF16, B, Cc, Hi, Wi, P, K, C.stream_ptr()))
        assert torch.equal(Ab.cpu(), ref.bfloat16())


@pytest.mark.parametrize("B,Cc,Hi,Wi,P", [(2, 3, 32, 48, 16), (1, 1, 40, 40, 10), (2, 3, 30, 30, 14), (3, 3, 224, 224, 16)])
@pytest.mark.parametrize("layout", [0, 1])
def test_im2col_u8_bit_exact(lib, B, Cc, Hi, Wi, P, layout):
    """uint8 pixels with the ToTensor /255 fused in == the fp32 gather of x.float()/255, bit for bit,
    from (B,C,H,W) and (B,H,W,C) pixel orders; includes 0 and 255."""
    C = _C()
    gen = torch.Generator().manual_seed(11)
    x8 = torch.randint(0, 256, (B, Cc, Hi, Wi), dtype=torch.uint8, generator=gen)
    x8[0, 0, 0, 0], x8[0, 0, 0, 1] = 0, 255
    N, K = (Hi // P) * (Wi // P), Cc * P * P
    ref = ref_vit.im2col(x8.float() / 255.0, P).reshape(B * N, K)
    src = (x8.permute(0, 2, 3, 1).contiguous() if layout == 1 else x8).to(DEV)
    A = torch.empty(B * N, K, device=DEV)
    C.check(lib.vitb200_im2col_u8(src.data_ptr(), layout, A.data_ptr(), C.F32, B, Cc, Hi, Wi, P, K, C.stream_ptr()))
    assert torch.equal(A.cpu(), ref)
    if K % 8 == 0:
        Ab = torch.empty(B * N, K, device=DEV, dtype=torch.bfloat16)
        C.check(lib.vitb200_im2col_u8(src.data_ptr(), layout, Ab.data_ptr(), C.BF16, B, Cc, Hi, Wi, P, K, C.stream_ptr()))
        assert torch.equal(Ab.cpu(), ref.bfloat16())
    assert lib.vitb200_im2col_u8(src.data_ptr(), 2, A.data_ptr(), C.F32, B, Cc, Hi, Wi, P, K, C.stream_ptr()) != 0


# ------------------------------------------------------------------------------------------------ misc bandwidth kernels
def test_colsum_cross_entropy_cast(lib):
    C = _C()
    torch.manual_seed(8)
    M, N = 1000, 768
    a = torch.randn(M, N, device=DEV)
    out = torch.ones(N, device=DEV)
    C.check(lib.vitb200_colsum(a.data_ptr(), C.F32, N, out.data_ptr(), M, N, C.stream_ptr()))
    assert rel_err(out - 1, a.sum(0)) < 1e-5
    ab = a.bfloat16()
    out.zero_()
    C.check(lib.vitb200_colsum(ab.data_ptr(), C.BF16, N, out.data_ptr(), M, N, C.stream_ptr()))
    assert rel_err(out, ab.float().sum(0)) < 1e-5
    for Bc, cls in [(256, 2), (5, 100), (7, 1000)]:
        logits = (torch.randn(Bc, cls, device=DEV) * 3).requires_grad_(True)
        labels = torch.randint(0, cls, (Bc,), device=DEV)
        ref = torch.nn.functional.cross_entropy(logits, labels)
        ref.backward()
        loss, dl = torch.zeros(1, device=DEV), torch.empty(Bc, cls, device=DEV)
        C.check(lib.vitb200_cross_entropy(logits.data_ptr(), labels.data_ptr(), loss.data_ptr(), dl.data_ptr(), Bc, cls,
                                          1.0, C.stream_ptr()))
        assert abs(loss.item() - ref.item()) < 1e-5 * max(1.0, abs(ref.item()))
        assert rel_err(dl, logits.grad) < 1e-5
    src = torch.randn(1001 * 8, device=DEV)
    dst = torch.empty(1001 * 8, device=DEV, dtype=torch.bfloat16)
    C.check(lib.vitb200_cast_f32_bf16(src.data_ptr(), dst.data_ptr(), src.numel(), C.stream_ptr()))
    assert torch.equal(dst, src.bfloat16())


# ------------------------------------------------------------------------------------------------ AdamW
def test_adamw_flat_matches_torch_and_oracle(lib):
    C = _C()
    torch.manual_seed(9)
    n = 1_000_003
    p0 = torch.randn(n, device=DEV)
    p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.AdamW([p], lr=1e-4, weight_decay=1e-4)
    q, m, v = p0.clone(), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
    shadow = torch.empty(n, device=DEV, dtype=torch.bfloat16)
    qo, mo, vo = p0.clone().cpu(), torch.zeros(n), torch.zeros(n)
    for step in range(1, 4):
        g = torch.randn(n, device=DEV)
        p.grad = g.clone()
        opt.step()
        C.check(lib.vitb200_adamw_flat(q.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), shadow.data_ptr(), n,
                                       C.AdamWHyper(1e-4, 0.9, 0.999, 1e-8, 1e-4, step, 1.0), C.stream_ptr()))
        ref_vit.adamw_step(qo, g.cpu(), mo, vo, step)
        assert torch.allclose(q, p.detach(), rtol=2e-6, atol=1e-7)
        assert torch.allclose(q.cpu(), qo, rtol=2e-6, atol=1e-7)
        assert torch.equal(shadow, q.bfloat16())
