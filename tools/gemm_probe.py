"""GPU probe for vitb200_gemm: correctness vs torch.matmul and CUDA-event timing.  Usage:
   python tools/gemm_probe.py <variant> [M N K]     variant in {nt, nn, tn, f32nt, f32tn, epi}
Run each variant in its own process under `timeout` (a barrier bug in a hand-written kernel hangs)."""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "vision-transformer_b200"))
import torch
from VisionTransformer import _C


def relerr(a, b):
    return ((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-30)).item()


def timeit(fn, iters=20):
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


def main():
    variant = sys.argv[1]
    dims = [int(v) for v in sys.argv[2:5]] if len(sys.argv) >= 5 else None
    torch.manual_seed(0)
    dev = "cuda"
    assert _C.lib().vitb200_device_ok() == 1
    f32 = variant.startswith("f32")
    dt = torch.float32 if f32 else torch.bfloat16
    v = variant[3:] if f32 else variant
    M, N, K = dims or (1024, 768, 768)
    if v in ("nt", "epi"):      # C = A[M,K] @ B[N,K]^T
        A = torch.randn(M, K, device=dev).to(dt); B = torch.randn(N, K, device=dev).to(dt)
        ref = A.float() @ B.float().t(); am, bm = 0, 0
    elif v == "nn":             # C = A[M,K] @ W[K,N]   (B stored [K][N]: MN-major)
        A = torch.randn(M, K, device=dev).to(dt); B = torch.randn(K, N, device=dev).to(dt)
        ref = A.float() @ B.float(); am, bm = 0, 1
    elif v == "tn":             # C = A[K,M]^T @ B[K,N]  (both MN-major) with split-K atomics
        A = torch.randn(K, M, device=dev).to(dt); B = torch.randn(K, N, device=dev).to(dt)
        ref = A.float().t() @ B.float(); am, bm = 1, 1
    else:
        raise SystemExit("bad variant")
    if v == "tn":
        C = torch.zeros(M, N, device=dev, dtype=torch.float32)
        _C.gemm(A, B, C, a_major=am, b_major=bm, M=M, N=N, K=K, accumulate=True, splits=0)
    elif v == "epi":
        bias = torch.randn(N, device=dev); res = torch.randn(M, N, device=dev)
        C = torch.empty(M, N, device=dev, dtype=torch.float32)
        _C.gemm(A, B, C, M=M, N=N, K=K, bias=bias, residual=res)
        ref = ref + bias + res
        torch.cuda.synchronize()
        print(f"{variant} bias+residual relerr {relerr(C, ref):.3e}")
        C2 = torch.empty(M, N, device=dev, dtype=dt)
        _C.gemm(A, B, C2, M=M, N=N, K=K, bias=bias, relu=True)
        ref = torch.relu(A.float() @ B.float().t() + bias)
    else:
        C = torch.empty(M, N, device=dev, dtype=torch.float32 if f32 else torch.bfloat16)
        _C.gemm(A, B, C, a_major=am, b_major=bm, M=M, N=N, K=K)
    torch.cuda.synchronize()
    out = C2 if v == "epi" else C
    err = relerr(out, ref)
    tol = 1e-5 if f32 else 1e-2
    print(f"{variant} M={M} N={N} K={K} relerr {err:.3e} {'OK' if err < tol else 'FAIL'}", flush=True)
    if err >= tol:
        d = (out.float() - ref).abs()
        bad = (d > tol * ref.abs().max()).nonzero()
        print(" n_bad", bad.shape[0], "first", bad[:8].tolist(), flush=True)
        print(" out[0,:8]", out[0, :8].tolist(), "\n ref[0,:8]", ref[0, :8].tolist())
        sys.exit(1)
    if v == "tn":
        fn = lambda: _C.gemm(A, B, C, a_major=am, b_major=bm, M=M, N=N, K=K, accumulate=True, splits=0)
    else:
        fn = lambda: _C.gemm(A, B, C, a_major=am, b_major=bm, M=M, N=N, K=K)
    ms = timeit(fn)
    tf = 2.0 * M * N * K / ms / 1e9
    A2, B2 = (A.t().contiguous() if am else A), (B if bm == 0 else B.t().contiguous())
    ms_t = timeit(lambda: torch.matmul(A2, B2.t()))
    print(f"{variant} M={M} N={N} K={K}: {ms:.4f} ms {tf:.1f} TFLOP/s | torch {ms_t:.4f} ms {2.0*M*N*K/ms_t/1e9:.1f} TFLOP/s", flush=True)


if __name__ == "__main__":
    main()
