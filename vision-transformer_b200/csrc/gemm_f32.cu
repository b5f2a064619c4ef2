// fp32 CUDA-core contraction used when the model runs in fp32 (the 1e-4 parity gate; tensor cores cannot
// meet it).  128x128x16 tiles double-buffered in shared memory, 8x8 register micro-tiles, arbitrary M/N/K/strides, same epilogue semantics as
// the tcgen05 kernel.  Replaces the fp32 cuBLAS calls behind nn.Linear / Conv2d of the reference
// (transformer.py:22-28,46,56-58; vit.py:21-28) and their autograd backward.
#include "common.cuh"
#include "gemm_epilogue.cuh"

namespace vb {

namespace {

constexpr int FBM = 128, FBN = 128, FBK = 16, FTHREADS = 256;

// first: this CTA holds the first K-split (the additive epilogue terms are applied exactly once per element)
__device__ __forceinline__ float epilogue_store1_f32(const GemmEpilogue& ep, int m, int n, int N, float v, bool first) {
  if (ep.bias != nullptr && first) v += __ldg(ep.bias + n);
  if (ep.relu) v = fmaxf(v, 0.f);
  const int64_t off = epi_out_row(ep, m) * ep.ldc + n;
  if (ep.mask_aux != nullptr) v = reinterpret_cast<const float*>(ep.mask_aux)[off] > 0.f ? v : 0.f;
  if (ep.drop.p > 0.f) {
    const uint64_t lin = static_cast<uint64_t>(m) * static_cast<uint64_t>(N) + static_cast<uint64_t>(n);
    v *= dropout_factor1(ep.drop, lin);
  }
  if (ep.row_add != nullptr && first) {
    const int pr = ep.row_period > 0 ? (m % ep.row_period) : m;
    v += __ldg(ep.row_add + static_cast<int64_t>(pr) * N + n);
  }
  if (ep.residual != nullptr && first) v += ep.residual[off];
  float* c = reinterpret_cast<float*>(ep.C) + off;
  if (ep.accumulate) atomicAdd(c, v);
  else *c = v;
  return v;
}

// Operand tile loader: 128 (rows) x 16 (k) elements of a K-major ([rows][ld]) or MN-major ([k][ld]) matrix into 8
// registers per thread.  Fast path: 16-byte loads when the whole tile is in bounds and 16-byte aligned; otherwise
// guarded scalar loads with zero fill.  store_tile() writes the registers to the [k][row] shared-memory tile.
template <bool KMAJOR>
__device__ __forceinline__ void load_tile_regs(const float* __restrict__ P, int64_t ld, int rows, int r0, int k0,
                                               int k_end, bool vec_ok, float (&v)[8]) {
  const int tid = threadIdx.x;
  if (vec_ok && r0 + FBM <= rows && k0 + FBK <= k_end) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int idx = tid + i * FTHREADS;   // float4 index inside the tile
      float4 t;
      if (KMAJOR) t = *reinterpret_cast<const float4*>(P + static_cast<int64_t>(r0 + (idx >> 2)) * ld + k0 + ((idx & 3) << 2));
      else        t = *reinterpret_cast<const float4*>(P + static_cast<int64_t>(k0 + (idx >> 5)) * ld + r0 + ((idx & 31) << 2));
      v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int idx = tid + i * FTHREADS;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        int r, kk;
        if (KMAJOR) { r = idx >> 2; kk = ((idx & 3) << 2) + c; } else { kk = idx >> 5; r = ((idx & 31) << 2) + c; }
        const int gr = r0 + r, gk = k0 + kk;
        float t = 0.f;
        if (gr < rows && gk < k_end) t = KMAJOR ? P[static_cast<int64_t>(gr) * ld + gk] : P[static_cast<int64_t>(gk) * ld + gr];
        v[4 * i + c] = t;
      }
    }
  }
}
template <bool KMAJOR>
__device__ __forceinline__ void store_tile(float (*S)[FBM + 4], const float (&v)[8]) {
  const int tid = threadIdx.x;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int idx = tid + i * FTHREADS;
    if (KMAJOR) {
      const int r = idx >> 2, kq = (idx & 3) << 2;
#pragma unroll
      for (int c = 0; c < 4; ++c) S[kq + c][r] = v[4 * i + c];
    } else {
      *reinterpret_cast<float4*>(&S[idx >> 5][(idx & 31) << 2]) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    }
  }
}

// A(m,k) = A[m*lda + k] (K-major) or A[k*lda + m]; B(n,k) likewise.  128 x 128 x 16 tiles, double-buffered in shared
// memory: the global loads of tile k+1 are issued into registers before tile k is multiplied (one barrier per tile).
template <bool A_KMAJOR, bool B_KMAJOR>
__global__ void __launch_bounds__(FTHREADS)
gemm_f32_kernel(const float* __restrict__ A, int64_t lda, const float* __restrict__ B, int64_t ldb, int M, int N,
                int K, int k_per_split, const GemmEpilogue ep, int vec_a, int vec_b, int vec_c) {
  __shared__ __align__(16) float As[2][FBK][FBM + 4];
  __shared__ __align__(16) float Bs[2][FBK][FBN + 4];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.y * FBM, n0 = blockIdx.x * FBN;
  const int k_begin = blockIdx.z * k_per_split;
  const int k_end = min(K, k_begin + k_per_split);
  const int tx = tid % 16, ty = tid / 16;  // micro-tile: rows ty*8.., cols tx*8..

  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  float ra[8], rb[8];
  load_tile_regs<A_KMAJOR>(A, lda, M, m0, k_begin, k_end, vec_a != 0, ra);
  load_tile_regs<B_KMAJOR>(B, ldb, N, n0, k_begin, k_end, vec_b != 0, rb);
  store_tile<A_KMAJOR>(As[0], ra);
  store_tile<B_KMAJOR>(Bs[0], rb);
  __syncthreads();
  int cur = 0;
  for (int k0 = k_begin; k0 < k_end; k0 += FBK) {
    const bool more = k0 + FBK < k_end;
    if (more) {
      load_tile_regs<A_KMAJOR>(A, lda, M, m0, k0 + FBK, k_end, vec_a != 0, ra);
      load_tile_regs<B_KMAJOR>(B, ldb, N, n0, k0 + FBK, k_end, vec_b != 0, rb);
    }
#pragma unroll
    for (int kk = 0; kk < FBK; ++kk) {
      float a[8], b[8];
      const float4 a0 = *reinterpret_cast<const float4*>(&As[cur][kk][ty * 8]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[cur][kk][ty * 8 + 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[cur][kk][tx * 8]);
      const float4 b1 = *reinterpret_cast<const float4*>(&Bs[cur][kk][tx * 8 + 4]);
      a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
      b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (more) {
      store_tile<A_KMAJOR>(As[cur ^ 1], ra);   // the other buffer was last read one iteration ago (barrier below)
      store_tile<B_KMAJOR>(Bs[cur ^ 1], rb);
    }
    __syncthreads();
    cur ^= 1;
  }

  float cs[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) cs[j] = 0.f;
  const bool first = blockIdx.z == 0;
  if (vec_c) {
    // 8 consecutive columns per row through the shared vector epilogue (16-byte loads / stores, one Philox call per
    // 8 elements); additive terms only on the first K-split.
    const int n = n0 + tx * 8;
    if (n < N) {
      GemmEpilogue e = ep;
      if (!first) { e.residual = nullptr; e.row_add = nullptr; }
      float bias8[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) bias8[j] = (e.bias != nullptr && first) ? __ldg(e.bias + n + j) : 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int m = m0 + ty * 8 + i;
        if (m >= M) continue;
        const int64_t off = epi_out_row(e, m) * e.ldc + n;
        EpiPre pre;
        epilogue_fetch8<float, kEpiGeneric>(e, off, m, n, pre);
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = acc[i][j];
        epilogue_finish8<float, kEpiGeneric>(e, m, n, N, off, v, pre, bias8);
#pragma unroll
        for (int j = 0; j < 8; ++j) cs[j] += v[j];
      }
    }
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int m = m0 + ty * 8 + i;
      if (m >= M) continue;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int n = n0 + tx * 8 + j;
        if (n < N) cs[j] += epilogue_store1_f32(ep, m, n, N, acc[i][j], first);
      }
    }
  }
  if (ep.col_sum != nullptr) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int n = n0 + tx * 8 + j;
      if (n < N) atomicAdd(ep.col_sum + n, cs[j]);
    }
  }
}

// Skinny output (N <= 16, both operands K-major): one warp per output row, lanes stride over K with 16-byte loads,
// fixed-order warp reduction => bit-reproducible.  The classifier's last Linear (N = num_classes) lands here.
__global__ void __launch_bounds__(256)
gemm_f32_skinny_kernel(const float* __restrict__ A, int64_t lda, const float* __restrict__ B, int64_t ldb, int M, int N,
                       int K, int vec, const GemmEpilogue ep) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= M) return;
  const float* a = A + static_cast<int64_t>(warp) * lda;
  for (int n = 0; n < N; ++n) {
    const float* b = B + static_cast<int64_t>(n) * ldb;
    float s = 0.f;
    if (vec) {
      for (int k = lane * 4; k < K; k += 128) {
        const float4 av = *reinterpret_cast<const float4*>(a + k), bv = __ldg(reinterpret_cast<const float4*>(b + k));
        s = fmaf(av.x, bv.x, s); s = fmaf(av.y, bv.y, s); s = fmaf(av.z, bv.z, s); s = fmaf(av.w, bv.w, s);
      }
    } else {
      for (int k = lane; k < K; k += 32) s = fmaf(a[k], __ldg(b + k), s);
    }
    s = warp_sum(s);
    if (lane == 0) epilogue_store1_f32(ep, warp, n, N, s, true);
  }
}

}  // namespace

int gemm_f32(int M, int N, int K, int a_major, int b_major, const float* A, int64_t lda, const float* B,
             int64_t ldb, int splits, const GemmEpilogue& ep_in, cudaStream_t stream) {
  VB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_f32: empty problem M=%d N=%d K=%d", M, N, K);
  VB_REQUIRE(ep_in.bits_out == nullptr && ep_in.mask_bits == nullptr, "gemm_f32: bit masks are a tensor-core-path option");
  GemmEpilogue ep = ep_in;
  if (N <= 16 && a_major == 0 && b_major == 0 && !ep.accumulate && ep.col_sum == nullptr && M >= 32) {
    ProfScope prof(kProfGemmF32, 2.0 * M * N * K, 1, stream);
    const int vec = (K & 3) == 0 && (lda & 3) == 0 && (ldb & 3) == 0 && aligned16(A) && aligned16(B);
    gemm_f32_skinny_kernel<<<ceil_div(M * 32, 256), 256, 0, stream>>>(A, lda, B, ldb, M, N, K, vec, ep);
    VB_CHECK_LAUNCH();
    return kOk;
  }
  int s = ep.accumulate ? splits : 1;
  // a K-split accumulates PARTIAL sums with atomics: an epilogue that is not linear in the accumulator (ReLU, ReLU-style
  // aux mask, dropout) would be applied to the partials — refuse to split then (one CTA owns the whole K range)
  if (ep.accumulate && (ep.relu || ep.mask_aux != nullptr || ep.drop.p > 0.f)) s = 1;
  const int tiles = ceil_div(M, FBM) * ceil_div(N, FBN);
  // Few output tiles but a long K (the classifier head, M = batch): split K across CTAs into a zeroed output.
  // Only for epilogues that are linear in the accumulator (bias / row_add / residual are added by split 0).
  // Atomic accumulation order is not reproducible, so this is opt-in (splits == -1; backward call sites).
  if (splits == -1 && !ep.accumulate && 2 * tiles <= num_sms() && K >= 256 && !ep.relu && ep.mask_aux == nullptr &&
      ep.drop.p <= 0.f && ep.col_sum == nullptr && ep.row_period == 0 && ep.residual == nullptr) {
    VB_CHECK_CUDA(cudaMemset2DAsync(ep.C, static_cast<size_t>(ep.ldc) * 4, 0, static_cast<size_t>(N) * 4, M, stream));
    ep.accumulate = 1;
    s = 0;
  }
  if (s <= 0) {
    s = ceil_div(2 * num_sms(), tiles);
    const int max_s = K / 64 > 0 ? K / 64 : 1;
    if (s > max_s) s = max_s;
    if (s < 1) s = 1;
  }
  int k_per_split = ceil_div(ceil_div(K, s), FBK) * FBK;
  s = ceil_div(K, k_per_split);
  dim3 grid(ceil_div(N, FBN), ceil_div(M, FBM), s);
  // 16-byte operand loads need 16-byte aligned rows; K-major tiles additionally start at multiples of 4 in k
  const int vec_a = aligned16(A) && lda % 4 == 0 && (a_major == 1 || k_per_split % 4 == 0);
  const int vec_b = aligned16(B) && ldb % 4 == 0 && (b_major == 1 || k_per_split % 4 == 0);
  const int vec_c = N % 8 == 0 && ep.ldc % 8 == 0 && aligned16(ep.C) && aligned16(ep.residual) && aligned16(ep.mask_aux) &&
                    aligned16(ep.row_add) && aligned16(ep.bias);
  ProfScope prof(kProfGemmF32, 2.0 * M * N * K, 1, stream);
  VB_REQUIRE(grid.y <= 65535 && grid.z <= 65535, "gemm_f32: grid too large");
  if (a_major == 0 && b_major == 0)
    gemm_f32_kernel<true, true><<<grid, FTHREADS, 0, stream>>>(A, lda, B, ldb, M, N, K, k_per_split, ep, vec_a, vec_b, vec_c);
  else if (a_major == 0 && b_major == 1)
    gemm_f32_kernel<true, false><<<grid, FTHREADS, 0, stream>>>(A, lda, B, ldb, M, N, K, k_per_split, ep, vec_a, vec_b, vec_c);
  else if (a_major == 1 && b_major == 1)
    gemm_f32_kernel<false, false><<<grid, FTHREADS, 0, stream>>>(A, lda, B, ldb, M, N, K, k_per_split, ep, vec_a, vec_b, vec_c);
  else if (a_major == 1 && b_major == 0)
    gemm_f32_kernel<false, true><<<grid, FTHREADS, 0, stream>>>(A, lda, B, ldb, M, N, K, k_per_split, ep, vec_a, vec_b, vec_c);
  else {
    set_error("gemm_f32: bad operand majors a=%d b=%d", a_major, b_major);
    return kErrInvalid;
  }
  VB_CHECK_LAUNCH();
  return kOk;
}

}  // namespace vb
