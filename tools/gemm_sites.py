"""GPU probe: every GEMM call site of one ViT block (B/16 shapes by default) through vitb200_gemm with its real
epilogue, timed with CUDA events (inputs >> L2 are cycled), next to torch.matmul (cuBLAS) on the same shape."""
import os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, os.path.join(ROOT, "vision-transformer_b200"))
import torch
from VisionTransformer import _C

dev = "cuda"
B, T, E = int(os.environ.get("B", 256)), 197, int(os.environ.get("E", 768))
M, F = B * T, 4 * E
bf = torch.bfloat16


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


def rnd(*s, dt=bf):
    return (torch.randn(*s, device=dev) * 0.1).to(dt)


x_e, x_f, x_3e = rnd(M, E), rnd(M, F), rnd(M, 3 * E)
w_qkv, w_proj, w_fc1, w_fc2 = rnd(3 * E, E), rnd(E, E), rnd(F, E), rnd(E, F)
res, out_e32 = rnd(M, E, dt=torch.float32), torch.empty(M, E, device=dev)
out_e, out_f, out_3e = torch.empty(M, E, device=dev, dtype=bf), torch.empty(M, F, device=dev, dtype=bf), torch.empty(M, 3 * E, device=dev, dtype=bf)
b_e, b_f = rnd(E, dt=torch.float32), rnd(F, dt=torch.float32)
dw_qkv, dw_proj, dw_fc1, dw_fc2 = (torch.zeros(s, device=dev) for s in ((3 * E, E), (E, E), (F, E), (E, F)))
K0, K1 = _C.MAJOR_K, _C.MAJOR_MN
x_e2, delta_buf = rnd(M, E), torch.empty(B, E // 64, T, device=dev)
hbits = torch.zeros((F // 32) * M, device=dev, dtype=torch.int32)  # ReLU sign bits, word (n/32)*M + m
sites = [
    # name, flops, ours, torch
    ("fwd qkv   NT bf16", 2 * M * 3 * E * E, lambda: _C.gemm(x_e, w_qkv, out_3e, M=M, N=3 * E, K=E), lambda: x_e @ w_qkv.t()),
    ("fwd proj  NT f32 bias+res", 2 * M * E * E, lambda: _C.gemm(x_e, w_proj, out_e32, M=M, N=E, K=E, bias=b_e, residual=res), lambda: x_e @ w_proj.t()),
    ("fwd proj  NT f32 bias+drop+res", 2 * M * E * E, lambda: _C.gemm(x_e, w_proj, out_e32, M=M, N=E, K=E, bias=b_e, residual=res, dropout=(0.2, 1, 2)), lambda: x_e @ w_proj.t()),
    ("fwd fc1   NT bf16 bias+relu", 2 * M * F * E, lambda: _C.gemm(x_e, w_fc1, out_f, M=M, N=F, K=E, bias=b_f, relu=True), lambda: x_e @ w_fc1.t()),
    ("fwd fc1   NT bf16 bias+relu+bits", 2 * M * F * E, lambda: _C.gemm(x_e, w_fc1, out_f, M=M, N=F, K=E, bias=b_f, relu=True, bits_out=hbits, bits_rows=M), lambda: x_e @ w_fc1.t()),
    ("fwd fc2   NT f32 bias+res", 2 * M * F * E, lambda: _C.gemm(x_f, w_fc2, out_e32, M=M, N=E, K=F, bias=b_e, residual=res), lambda: x_f @ w_fc2.t()),
    ("fwd fc2   NT f32 bias+drop+res", 2 * M * F * E, lambda: _C.gemm(x_f, w_fc2, out_e32, M=M, N=E, K=F, bias=b_e, residual=res, dropout=(0.2, 1, 3)), lambda: x_f @ w_fc2.t()),
    ("bwd dW2   TN acc", 2 * M * F * E, lambda: _C.gemm(x_e, x_f, dw_fc2, a_major=K1, b_major=K1, M=E, N=F, K=M, accumulate=True), lambda: x_e.t() @ x_f),
    ("bwd dh    NN bf16 mask", 2 * M * F * E, lambda: _C.gemm(x_e, w_fc2, out_f, b_major=K1, M=M, N=F, K=E, mask_aux=x_f), lambda: x_e @ w_fc2),
    ("bwd dh    NN bf16 mask+colsum", 2 * M * F * E, lambda: _C.gemm(x_e, w_fc2, out_f, b_major=K1, M=M, N=F, K=E, mask_aux=x_f, col_sum=b_f), lambda: x_e @ w_fc2),
    ("bwd dh    NN bf16 bits+colsum", 2 * M * F * E, lambda: _C.gemm(x_e, w_fc2, out_f, b_major=K1, M=M, N=F, K=E, mask_bits=hbits, bits_rows=M, col_sum=b_f), lambda: x_e @ w_fc2),
    ("bwd dh    NN bf16 plain", 2 * M * F * E, lambda: _C.gemm(x_e, w_fc2, out_f, b_major=K1, M=M, N=F, K=E), lambda: x_e @ w_fc2),
    ("bwd dW1   TN acc", 2 * M * F * E, lambda: _C.gemm(x_f, x_e, dw_fc1, a_major=K1, b_major=K1, M=F, N=E, K=M, accumulate=True), lambda: x_f.t() @ x_e),
    ("bwd dln2  NN bf16", 2 * M * F * E, lambda: _C.gemm(x_f, w_fc1, out_e, b_major=K1, M=M, N=E, K=F), lambda: x_f @ w_fc1),
    ("bwd dWprj TN acc", 2 * M * E * E, lambda: _C.gemm(x_e, x_e, dw_proj, a_major=K1, b_major=K1, M=E, N=E, K=M, accumulate=True), lambda: x_e.t() @ x_e),
    ("bwd dO    NN bf16", 2 * M * E * E, lambda: _C.gemm(x_e, w_proj, out_e, b_major=K1, M=M, N=E, K=E), lambda: x_e @ w_proj),
    ("bwd dO    NN bf16 +delta", 2 * M * E * E, lambda: _C.gemm(x_e, w_proj, out_e, b_major=K1, M=M, N=E, K=E, dot_aux=x_e2, dot_out=delta_buf, dot_T=T, dot_H=E // 64), lambda: x_e @ w_proj),
    ("bwd dWqkv TN acc", 2 * M * 3 * E * E, lambda: _C.gemm(x_3e, x_e, dw_qkv, a_major=K1, b_major=K1, M=3 * E, N=E, K=M, accumulate=True), lambda: x_3e.t() @ x_e),
    ("bwd dln1  NN bf16", 2 * M * 3 * E * E, lambda: _C.gemm(x_3e, w_qkv, out_e, b_major=K1, M=M, N=E, K=3 * E), lambda: x_3e @ w_qkv),
]
tot_o = tot_t = 0.0
_only = os.environ.get("SITES")
if _only:
    sites = [x for x in sites if any(k in x[0] for k in _only.split(","))]
print(f"B={B} M={M} E={E}")
for name, fl, ours, ref in sites:
    to = timeit(ours)
    tt = to if os.environ.get("NOREF") else timeit(ref)
    if "drop" not in name or True:
        pass
    print(f"{name:<34} ours {to*1e3:7.1f} us {fl/to/1e9:7.0f} TF/s | cuBLAS {tt*1e3:7.1f} us {fl/tt/1e9:7.0f} TF/s | ratio {to/tt:4.2f}")
    if ("bias+res" in name and "drop" not in name) or "plain" in name or "mask" in name or name.endswith("relu") or "+delta" in name:
        continue  # the dropout variant is the one the training step uses
    tot_o += to; tot_t += tt
print(f"per-block GEMM total: ours {tot_o:.3f} ms, cuBLAS (no epilogues) {tot_t:.3f} ms; x12 blocks = {12*tot_o:.1f} / {12*tot_t:.1f} ms")
