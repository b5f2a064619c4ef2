"""GPU: the allocation-free attention-backward entry point (vitb200_attn_bwd_noalloc
— no counterpart in the reference) against the regular entry point (vitb200_attn_bwd), through the C ABI.
Collected LAST on purpose (file name).

The scratch (`delta`, fp32 [B,H,T]) of the new entry point lives in a `static __device__` array inside the library;
the test makes two calls in a row (the second one smaller, so stale values of the first stay in the array), checks
both against the regular entry point, checks that no torch allocation happens inside the call, and that a problem
beyond the static limit is refused with a clear error."""
import math

import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu


def _case(B, T, H, hs, dtype, seed):
    from VisionTransformer import _C
    g = torch.Generator(device="cuda").manual_seed(seed)
    E = H * hs
    qkv = (torch.randn(B * T, 3 * E, device="cuda", generator=g) * (0.45 / math.sqrt(hs)))
    qkv[:, 2 * E:] = torch.randn(B * T, E, device="cuda", generator=g)
    qkv = qkv.to(dtype)
    d_o = torch.randn(B * T, E, device="cuda", generator=g).to(dtype)
    out = torch.empty(B * T, E, device="cuda", dtype=dtype)
    lse = torch.empty(B, H, T, device="cuda")
    L = _C.lib()
    _C.check(L.vitb200_attn_fwd(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), _C.dtype_code(dtype), B, T, H, hs,
                                math.sqrt(hs), _C.stream_ptr()))
    return qkv, out, d_o, lse


def _both(B, T, H, hs, dtype, seed):
    from VisionTransformer import _C
    L = _C.lib()
    qkv, out, d_o, lse = _case(B, T, H, hs, dtype, seed)
    code, scale = _C.dtype_code(dtype), math.sqrt(hs)
    delta = torch.empty(B, H, T, device="cuda")
    ref = torch.full_like(qkv, float("nan"))
    _C.check(L.vitb200_attn_bwd(qkv.data_ptr(), out.data_ptr(), d_o.data_ptr(), lse.data_ptr(), delta.data_ptr(),
                                ref.data_ptr(), code, B, T, H, hs, scale, _C.stream_ptr()))
    got = torch.full_like(qkv, float("nan"))
    torch.cuda.synchronize()
    allocs = torch.cuda.memory_stats()["allocation.all.allocated"]
    _C.check(L.vitb200_attn_bwd_noalloc(qkv.data_ptr(), out.data_ptr(), d_o.data_ptr(), lse.data_ptr(), got.data_ptr(),
                                        code, B, T, H, hs, scale, _C.stream_ptr()))
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats()["allocation.all.allocated"] == allocs      # no torch allocation inside the call
    return got, ref


@pytest.mark.parametrize("dtype,shapes", [
    (torch.bfloat16, [(8, 197, 12, 64), (3, 65, 4, 64)]),      # tcgen05 path (ViT-B/16 layer shape), then a smaller one
    (torch.bfloat16, [(2, 577, 16, 80), (2, 130, 2, 32)]),     # long-row tcgen05 path (ViT-H/14), then mma.sync path
    (torch.float32, [(2, 37, 4, 16), (1, 9, 2, 8)]),           # fp32 parity path
])
def test_noalloc_attention_backward_matches_regular_entry_point(lib, dtype, shapes):
    for i, (B, T, H, hs) in enumerate(shapes):                 # two calls in a row share the static scratch
        got, ref = _both(B, T, H, hs, dtype, seed=10 + i)
        assert torch.isfinite(got.float()).all() and torch.isfinite(ref.float()).all()
        # same kernels, same inputs: bit-identical in practice; gated at the tolerance of the kernel tests
        tol = 1e-5 if dtype == torch.float32 else 2e-2
        for s in range(3):
            E = H * hs
            assert rel_err(got[:, s * E:(s + 1) * E], ref[:, s * E:(s + 1) * E]) < tol, (i, s)
        assert torch.equal(got, ref), (i, "deterministic kernels: expected identical bits")


def test_noalloc_rejects_sizes_beyond_the_static_limit(lib):
    from VisionTransformer import _C
    L = _C.lib()
    B, T, H, hs = 4096, 197, 12, 64                            # B*H*T = 9.7 M > 1 Mi scratch values
    assert B * T * H > _C.NOALLOC_MAX_DELTA
    dummy = torch.zeros(16, device="cuda")
    rc = L.vitb200_attn_bwd_noalloc(dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(),
                                    dummy.data_ptr(), _C.BF16, B, T, H, hs, 8.0, _C.stream_ptr())
    assert rc != 0
    msg = L.vitb200_last_error().decode()
    assert "VITB200_NOALLOC_MAX_DELTA" in msg and "exceed" in msg
    # the largest bench shapes fit: ViT-B/16 and ViT-L/16 at 256 images, ViT-H/14 (T=577) at 64
    for b, t, h in ((256, 197, 12), (256, 197, 16), (64, 577, 16)):
        assert b * t * h <= _C.NOALLOC_MAX_DELTA
