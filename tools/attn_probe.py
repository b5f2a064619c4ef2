"""GPU probe: attention forward/backward correctness vs torch fp32 and timing at the ViT-B/16 shape.
   python tools/attn_probe.py [fwd|bwd|both] [B T H]      (VITB200_ATTN=legacy selects the mma.sync kernels)"""
import math, os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, os.path.join(ROOT, "vision-transformer_b200"))
import torch
from VisionTransformer import _C

mode = sys.argv[1] if len(sys.argv) > 1 else "both"
B, T, H = (int(v) for v in sys.argv[2:5]) if len(sys.argv) >= 5 else (256, 197, 12)
hs = int(os.environ.get("HS", 64))
E = H * hs
L = _C.lib()
torch.manual_seed(0)
scale = math.sqrt(hs)


def ref(qkv):
    q, k, v = (qkv.reshape(B, T, 3, H, hs)[:, :, i].permute(0, 2, 1, 3) for i in range(3))
    p = torch.softmax((q @ k.transpose(-1, -2)) * scale, dim=-1)
    return (p @ v).permute(0, 2, 1, 3).reshape(B * T, E)


def cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float(a @ b / (a.norm() * b.norm()))


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


qkv = (torch.randn(B * T, 3 * E, device="cuda") * (0.45 / math.sqrt(hs))).bfloat16()
qkv[:, 2 * E:] = torch.randn(B * T, E, device="cuda").bfloat16()
out = torch.empty(B * T, E, device="cuda", dtype=torch.bfloat16)
lse = torch.empty(B, H, T, device="cuda")
d_o = torch.randn(B * T, E, device="cuda").bfloat16()
dqkv = torch.empty_like(qkv)
delta = torch.empty(B, H, T, device="cuda")
fwd = lambda: _C.check(L.vitb200_attn_fwd(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), _C.BF16, B, T, H, hs, scale, _C.stream_ptr()))
bwd = lambda: _C.check(L.vitb200_attn_bwd(qkv.data_ptr(), out.data_ptr(), d_o.data_ptr(), lse.data_ptr(), delta.data_ptr(), dqkv.data_ptr(), _C.BF16, B, T, H, hs, scale, _C.stream_ptr()))
small = B * H <= 64
if mode in ("fwd", "both"):
    fwd(); torch.cuda.synchronize()
    if small or True:
        nb = min(B, 4)
        x = qkv[: nb * T].float().requires_grad_(True)
        Bs = B; B = nb
        o_ref = ref(x); B = Bs
        err = ((out[: nb * T].float() - o_ref).abs().max() / o_ref.abs().max()).item()
        print(f"fwd rel-err (first {nb} images) {err:.3e}  nan={torch.isnan(out.float()).any().item()}")
        # last image too (exercises the zero-filled tail rows)
        x2 = qkv[(B - 1) * T:].float()
        Bs = B; B = 1
        o2 = ref(x2); B = Bs
        print(f"fwd rel-err (last image) {((out[(B-1)*T:].float() - o2).abs().max() / o2.abs().max()).item():.3e}")
    ms = timeit(fwd)
    print(f"fwd B={B} T={T} H={H}: {ms*1e3:.1f} us  {4.0*B*H*T*T*hs/ms/1e9:.0f} TF/s")
if mode in ("bwd", "both"):
    fwd(); bwd(); torch.cuda.synchronize()
    nb = min(B, 2)
    x = qkv[: nb * T].float().requires_grad_(True)
    Bs = B; B = nb
    ref(x).backward(d_o[: nb * T].float()); B = Bs
    for s, nm in enumerate("qkv"):
        print(f"bwd d{nm} cosine {cos(dqkv[: nb*T, s*E:(s+1)*E], x.grad[:, s*E:(s+1)*E]):.6f}", end="  ")
    print("nan", torch.isnan(dqkv.float()).any().item())
    ms = timeit(bwd)
    print(f"bwd B={B} T={T} H={H}: {ms*1e3:.1f} us  {8.0*B*H*T*T*hs/ms/1e9:.0f} TF/s")
