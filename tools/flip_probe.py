import os, sys, ctypes
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, os.path.join(ROOT, "vision-transformer_b200")); sys.path.insert(0, ROOT)
import torch
from VisionTransformer import _C
torch.backends.cuda.matmul.allow_tf32 = False
torch.manual_seed(0)
M, E, F = 788, 768, 3072
x = torch.randn(M, E, device="cuda") * 1.5 + 0.3
gam = 1 + 0.1 * torch.randn(E, device="cuda"); bet = 0.1 * torch.randn(E, device="cuda")
W = torch.randn(F, E, device="cuda") * E ** -0.5; b = 0.1 * torch.randn(F, device="cuda")
L = _C.lib()
# fp32 truth
ln = torch.nn.functional.layer_norm(x, (E,), gam, bet, 1e-5)
pre = ln @ W.t() + b
# autocast-like
ln_b = ln.bfloat16(); W_b = W.bfloat16()
pre_a = (torch.nn.functional.linear(ln_b, W_b, b.bfloat16())).float()
# ours
y = torch.empty(M, E, device="cuda", dtype=torch.bfloat16); mean = torch.empty(M, device="cuda"); rstd = torch.empty(M, device="cuda")
_C.check(L.vitb200_ln_fwd(x.data_ptr(), gam.data_ptr(), bet.data_ptr(), y.data_ptr(), _C.BF16, mean.data_ptr(), rstd.data_ptr(), M, E, 1e-5, _C.stream_ptr()))
h = torch.empty(M, F, device="cuda", dtype=torch.bfloat16)
_C.gemm(y, W_b, h, M=M, N=F, K=E, bias=b, relu=True)
pre_o = torch.empty(M, F, device="cuda", dtype=torch.float32)
_C.gemm(y, W_b, pre_o, M=M, N=F, K=E, bias=b)
print("ln ours vs torch bf16: mismatched elems", (y != ln_b).float().mean().item(), "max abs", (y.float() - ln_b.float()).abs().max().item())
print("pre rel err (L2): autocast", ((pre_a - pre).norm() / pre.norm()).item(), "ours", ((pre_o - pre).norm() / pre.norm()).item())
print("flip fraction: autocast", ((pre_a > 0) != (pre > 0)).float().mean().item(), "ours(h>0)", ((h.float() > 0) != (pre > 0)).float().mean().item(),
      "ours(pre)", ((pre_o > 0) != (pre > 0)).float().mean().item())
