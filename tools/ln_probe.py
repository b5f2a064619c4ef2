"""LayerNorm forward / backward timing at the ViT-B/16 training shape (M = 256*197, E = 768)."""
import os, sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "vision-transformer_b200"))
from VisionTransformer import _C as C

lib = C.lib()
DEV = "cuda"
M, E = int(os.environ.get("M", 256 * 197)), int(os.environ.get("E", 768))
x = torch.randn(M, E, device=DEV)
w, b = torch.ones(E, device=DEV), torch.zeros(E, device=DEV)
mean, rstd = torch.empty(M, device=DEV), torch.empty(M, device=DEV)
big = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)


def timeit(fn, n=20):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(n):
        big.zero_()
        a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); e.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2] * 1e3


for dt, code in ((torch.bfloat16, C.BF16), (torch.float32, C.F32)):
    es = 2 if dt == torch.bfloat16 else 4
    y = torch.empty(M, E, device=DEV, dtype=dt)
    us = timeit(lambda: C.check(lib.vitb200_ln_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), code,
                                                   mean.data_ptr(), rstd.data_ptr(), M, E, 1e-5, C.stream_ptr())))
    print(f"ln_fwd {dt}: {us:.1f} us  {M * E * (4 + es) / us / 1e6:.2f} TB/s")
    dy = torch.randn(M, E, device=DEV).to(dt)
    g = torch.randn(M, E, device=DEV)
    dyn = torch.empty(M, E, device=DEV, dtype=dt)
    dg, db, cs = (torch.zeros(E, device=DEV) for _ in range(3))
    for add, with_dyn, p in ((1, True, 0.2), (1, True, 0.0), (1, False, 0.0), (0, True, 0.2)):
        us = timeit(lambda: C.check(lib.vitb200_ln_bwd(dy.data_ptr(), code, x.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                                                       w.data_ptr(), g.data_ptr(), add, dg.data_ptr(), db.data_ptr(),
                                                       dyn.data_ptr() if with_dyn else None,
                                                       cs.data_ptr() if with_dyn else None, C.Dropout(p, 1, 2), M, E,
                                                       C.stream_ptr())))
        byts = M * E * (4 + es + 4 + (4 if add else 0) + (es if with_dyn else 0))
        print(f"ln_bwd {dt} add={add} dyn={with_dyn} p={p}: {us:.1f} us  {byts / us / 1e6:.2f} TB/s")
