import os, sys, torch
sys.path.insert(0, "vision-transformer_b200")
from VisionTransformer import _C as C
lib = C.lib()
torch.manual_seed(0)
for M, E in [(600, 768), (1576, 384), (1000, 384), (700, 1024), (650, 1280), (4000, 1536), (50432, 768)]:
    x = (torch.randn(M, E, device="cuda") * 2 + 0.5)
    w, b = torch.randn(E, device="cuda"), torch.randn(E, device="cuda")
    for dt, code in ((torch.float32, C.F32), (torch.bfloat16, C.BF16)):
        y = torch.empty(M, E, device="cuda", dtype=dt)
        mean, rstd = torch.empty(M, device="cuda"), torch.empty(M, device="cuda")
        C.check(lib.vitb200_ln_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), code, mean.data_ptr(), rstd.data_ptr(), M, E, 1e-5, C.stream_ptr()))
        ref = torch.nn.functional.layer_norm(x.double(), (E,), w.double(), b.double(), 1e-5)
        err = ((y.double() - ref).abs().max() / ref.abs().max()).item()
        merr = (mean.double() - x.double().mean(1)).abs().max().item()
        rerr = ((rstd.double() - 1 / torch.sqrt(x.double().var(1, unbiased=False) + 1e-5)).abs().max()).item()
        print(M, E, dt, f"y rel {err:.2e} mean {merr:.2e} rstd {rerr:.2e}")
