// Fused flash-style self-attention for the bf16 training path: forward, and backward as a dK/dV kernel plus a
// dQ kernel (both recompute P from the saved log-sum-exp, nothing of size T x T ever touches HBM).
// Tensor-core math via mma.sync.m16n8k16 (bf16 in, fp32 accumulate) with the online softmax kept in registers
// and reduced with warp shuffles; K/V (or Q/dO) tiles stream through a double-buffered cp.async pipeline.
// Semantics: reference transformer.py:20-31,44-45 — logits MULTIPLIED by sqrt(head_size) (scale is an
// argument), no mask, heads are column slices of the packed [M,3E] qkv buffer, output concatenated in [M,E].
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"

namespace vb {

static bool use_tc_gen_forward(int T, int hs);   // defined with the dispatchers at the end of this file

namespace {

constexpr int AT_THREADS = 128;  // 4 warps, 16 rows each
constexpr int AT_BLK = 64;       // rows per Q tile and per KV tile
constexpr float kLog2e = 1.4426950408889634f;

__device__ __forceinline__ uint32_t ptx_smem(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void cp_async16(uint32_t saddr, const void* g, bool valid) {
  const int sz = valid ? 16 : 0;  // src-size 0 => zero-fill
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(saddr), "l"(g), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

template <int HS>
struct Tile {
  static constexpr int LD = HS + 8;               // padded row (elements): conflict-free ldmatrix
  static constexpr int BYTES = AT_BLK * LD * 2;
};

// Stage a [64 x HS] bf16 tile: rows row0.. of a matrix with row stride ld (elements); rows >= nrows zero-filled.
template <int HS>
__device__ __forceinline__ void load_tile(uint32_t s_tile, const bf16* __restrict__ g, int64_t ld, int row0,
                                          int nrows, int tid) {
  constexpr int CH = HS / 8;  // 16-byte chunks per row
  for (int c = tid; c < AT_BLK * CH; c += AT_THREADS) {
    const int r = c / CH, k = c - r * CH;
    const bool ok = row0 + r < nrows;
    const bf16* src = g + static_cast<int64_t>(ok ? row0 + r : 0) * ld + k * 8;
    cp_async16(s_tile + (r * Tile<HS>::LD + k * 8) * 2, src, ok);
  }
}

// A-operand fragments for this warp's 16 rows (wr..wr+15) of a tile, all HS/16 k-steps.
template <int HS>
__device__ __forceinline__ void load_a_frags(uint32_t (&f)[HS / 16][4], uint32_t s_tile, int wr, int lane) {
  const int mid = lane >> 3, r = lane & 7;
#pragma unroll
  for (int kk = 0; kk < HS / 16; ++kk) {
    const int row = wr + (mid & 1) * 8 + r, col = kk * 16 + (mid >> 1) * 8;
    ldsm_x4(f[kk], s_tile + (row * Tile<HS>::LD + col) * 2);
  }
}

// acc[8][4] (16 x 64) = A(16 x HS) * Tile(64 x HS)^T   — tile rows are the n index.
template <int HS>
__device__ __forceinline__ void mma_a_tileT(float (&acc)[8][4], const uint32_t (&a)[HS / 16][4], uint32_t s_tile,
                                            int lane) {
  const int mid = lane >> 3, r = lane & 7;
#pragma unroll
  for (int nt = 0; nt < 8; ++nt)
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[nt][i] = 0.f;
#pragma unroll
  for (int kk = 0; kk < HS / 16; ++kk) {
#pragma unroll
    for (int ntp = 0; ntp < 4; ++ntp) {
      uint32_t b[4];
      const int row = (ntp * 2 + (mid >> 1)) * 8 + r, col = kk * 16 + (mid & 1) * 8;
      ldsm_x4(b, s_tile + (row * Tile<HS>::LD + col) * 2);
      mma_bf16(acc[ntp * 2], a[kk], b[0], b[1]);
      mma_bf16(acc[ntp * 2 + 1], a[kk], b[2], b[3]);
    }
  }
}

// out[HS/8][4] (16 x HS) += P(16 x 64, packed bf16 A-fragments) * Tile(64 x HS)  — tile rows are the k index.
template <int HS>
__device__ __forceinline__ void mma_p_tile(float (&out)[HS / 8][4], const uint32_t (&p)[4][4], uint32_t s_tile,
                                           int lane) {
  const int mid = lane >> 3, r = lane & 7;
#pragma unroll
  for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
    for (int ntp = 0; ntp < HS / 16; ++ntp) {
      uint32_t b[4];
      const int row = kk * 16 + (mid & 1) * 8 + r, col = (ntp * 2 + (mid >> 1)) * 8;
      ldsm_x4_t(b, s_tile + (row * Tile<HS>::LD + col) * 2);
      mma_bf16(out[ntp * 2], p[kk], b[0], b[1]);
      mma_bf16(out[ntp * 2 + 1], p[kk], b[2], b[3]);
    }
  }
}

// fp32 accumulator tile (16 x 64) -> bf16 A-operand fragments for 4 k-steps of 16.
__device__ __forceinline__ void pack_acc(uint32_t (&p)[4][4], const float (&acc)[8][4]) {
#pragma unroll
  for (int kk = 0; kk < 4; ++kk) {
    p[kk][0] = pack_bf16x2(acc[2 * kk][0], acc[2 * kk][1]);
    p[kk][1] = pack_bf16x2(acc[2 * kk][2], acc[2 * kk][3]);
    p[kk][2] = pack_bf16x2(acc[2 * kk + 1][0], acc[2 * kk + 1][1]);
    p[kk][3] = pack_bf16x2(acc[2 * kk + 1][2], acc[2 * kk + 1][3]);
  }
}

// Store a 16 x HS fp32 accumulator as bf16 rows (row0+g, row0+g+8) at column offset col0 of a matrix with stride ld.
template <int HS>
__device__ __forceinline__ void store_rows(bf16* __restrict__ dst, int64_t ld, int row_lo, int nrows, int64_t row_base,
                                           int col0, const float (&acc)[HS / 8][4], float s0, float s1, int lane) {
  const int g = lane >> 2, t = lane & 3;
#pragma unroll
  for (int nt = 0; nt < HS / 8; ++nt) {
    const int col = col0 + nt * 8 + 2 * t;
    if (row_lo + g < nrows)
      *reinterpret_cast<uint32_t*>(dst + (row_base + row_lo + g) * ld + col) = pack_bf16x2(acc[nt][0] * s0, acc[nt][1] * s0);
    if (row_lo + g + 8 < nrows)
      *reinterpret_cast<uint32_t*>(dst + (row_base + row_lo + g + 8) * ld + col) = pack_bf16x2(acc[nt][2] * s1, acc[nt][3] * s1);
  }
}

// ------------------------------------------------------------------------------------------------ forward
template <int HS>
__global__ void __launch_bounds__(AT_THREADS)
attn_bf16_fwd_kernel(const bf16* __restrict__ qkv, bf16* __restrict__ out, float* __restrict__ lse, int T, int H,
                     float scale) {
  extern __shared__ __align__(16) uint8_t at_smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, t = lane & 3;
  const int bh = blockIdx.y, b = bh / H, h = bh - b * H;
  const int E = H * HS;
  const int64_t ld3 = 3 * static_cast<int64_t>(E);
  const int q0 = blockIdx.x * AT_BLK;
  const bf16* qg = qkv + static_cast<int64_t>(b) * T * ld3 + h * HS;
  const bf16* kg = qg + E;
  const bf16* vg = kg + E;
  const uint32_t sQ = ptx_smem(at_smem);
  const uint32_t sK = sQ + Tile<HS>::BYTES;            // 2 buffers
  const uint32_t sV = sK + 2 * Tile<HS>::BYTES;        // 2 buffers
  const int nkv = ceil_div(T, AT_BLK);

  load_tile<HS>(sQ, qg, ld3, q0, T, tid);
  load_tile<HS>(sK, kg, ld3, 0, T, tid);
  load_tile<HS>(sV, vg, ld3, 0, T, tid);
  cp_async_commit();
  cp_async_wait_all();
  __syncthreads();

  uint32_t qf[HS / 16][4];
  load_a_frags<HS>(qf, sQ, warp * 16, lane);

  float o[HS / 8][4];
#pragma unroll
  for (int nt = 0; nt < HS / 8; ++nt)
#pragma unroll
    for (int i = 0; i < 4; ++i) o[nt][i] = 0.f;
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
  const float sl2 = scale * kLog2e;

  for (int kt = 0; kt < nkv; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nkv) {
      load_tile<HS>(sK + (buf ^ 1) * Tile<HS>::BYTES, kg, ld3, (kt + 1) * AT_BLK, T, tid);
      load_tile<HS>(sV + (buf ^ 1) * Tile<HS>::BYTES, vg, ld3, (kt + 1) * AT_BLK, T, tid);
      cp_async_commit();
    }
    float s[8][4];
    mma_a_tileT<HS>(s, qf, sK + buf * Tile<HS>::BYTES, lane);
    const int kv0 = kt * AT_BLK;
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const int key = kv0 + nt * 8 + 2 * t;
      s[nt][0] = key < T ? s[nt][0] * sl2 : -INFINITY;
      s[nt][1] = key + 1 < T ? s[nt][1] * sl2 : -INFINITY;
      s[nt][2] = key < T ? s[nt][2] * sl2 : -INFINITY;
      s[nt][3] = key + 1 < T ? s[nt][3] * sl2 : -INFINITY;
      mx0 = fmaxf(mx0, fmaxf(s[nt][0], s[nt][1]));
      mx1 = fmaxf(mx1, fmaxf(s[nt][2], s[nt][3]));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
    const float c0 = exp2f(m0 - mn0), c1 = exp2f(m1 - mn1);   // first tile: exp2(-inf) = 0
    m0 = mn0; m1 = mn1;
    float r0 = 0.f, r1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      s[nt][0] = exp2f(s[nt][0] - mn0); s[nt][1] = exp2f(s[nt][1] - mn0);
      s[nt][2] = exp2f(s[nt][2] - mn1); s[nt][3] = exp2f(s[nt][3] - mn1);
      r0 += s[nt][0] + s[nt][1];
      r1 += s[nt][2] + s[nt][3];
    }
    l0 = l0 * c0 + r0;
    l1 = l1 * c1 + r1;
#pragma unroll
    for (int nt = 0; nt < HS / 8; ++nt) { o[nt][0] *= c0; o[nt][1] *= c0; o[nt][2] *= c1; o[nt][3] *= c1; }
    uint32_t p[4][4];
    pack_acc(p, s);
    mma_p_tile<HS>(o, p, sV + buf * Tile<HS>::BYTES, lane);
    cp_async_wait_all();
    __syncthreads();
  }
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const int row_lo = q0 + warp * 16;
  store_rows<HS>(out, E, row_lo, T, static_cast<int64_t>(b) * T, h * HS, o, 1.f / l0, 1.f / l1, lane);
  if (t == 0 && lse != nullptr) {
    constexpr float kLn2 = 0.6931471805599453f;
    if (row_lo + g < T) lse[static_cast<int64_t>(bh) * T + row_lo + g] = m0 * kLn2 + logf(l0);
    if (row_lo + g + 8 < T) lse[static_cast<int64_t>(bh) * T + row_lo + g + 8] = m1 * kLn2 + logf(l1);
  }
}

// ------------------------------------------------------------------------------------------------ backward
// Stage 64 fp32 row statistics (lse * log2e, delta) for rows row0.. into shared memory; rows >= T get +inf lse
// (=> p = 0) and delta 0.
__device__ __forceinline__ void load_stats(float* s_lse2, float* s_delta, const float* __restrict__ lse,
                                           const float* __restrict__ delta, int row0, int T, int tid) {
  if (tid < AT_BLK) {
    const int r = row0 + tid;
    s_lse2[tid] = r < T ? lse[r] * kLog2e : INFINITY;
    s_delta[tid] = r < T ? delta[r] : 0.f;
  }
}

// dK, dV: the CTA owns 64 keys (16 per warp) of one (batch, head) and streams all query tiles.
//   S^T = K Q^T, P^T = exp(S^T*scale - lse[q]), dP^T = V dO^T, dS^T = P^T (dP^T - delta[q]) scale,
//   dV += P^T dO, dK += dS^T Q.
template <int HS>
__global__ void __launch_bounds__(AT_THREADS)
attn_bf16_bwd_dkv_kernel(const bf16* __restrict__ qkv, const bf16* __restrict__ d_o, const float* __restrict__ lse,
                         const float* __restrict__ delta, bf16* __restrict__ dqkv, int T, int H, float scale) {
  extern __shared__ __align__(16) uint8_t at_smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int t = lane & 3;
  const int bh = blockIdx.y, b = bh / H, h = bh - b * H;
  const int E = H * HS;
  const int64_t ld3 = 3 * static_cast<int64_t>(E);
  const int kv0 = blockIdx.x * AT_BLK;
  const bf16* qg = qkv + static_cast<int64_t>(b) * T * ld3 + h * HS;
  const bf16* kg = qg + E;
  const bf16* vg = kg + E;
  const bf16* dog = d_o + static_cast<int64_t>(b) * T * E + h * HS;
  const float* lse_bh = lse + static_cast<int64_t>(bh) * T;
  const float* delta_bh = delta + static_cast<int64_t>(bh) * T;
  const uint32_t sQ = ptx_smem(at_smem);                 // 2 buffers (first used for K at start)
  const uint32_t sDO = sQ + 2 * Tile<HS>::BYTES;         // 2 buffers (first used for V at start)
  float* s_stats = reinterpret_cast<float*>(at_smem + 4 * Tile<HS>::BYTES);  // [2][2][64]
  const int nq = ceil_div(T, AT_BLK);

  // K and V tiles -> persistent A fragments
  load_tile<HS>(sQ, kg, ld3, kv0, T, tid);
  load_tile<HS>(sDO, vg, ld3, kv0, T, tid);
  cp_async_commit();
  cp_async_wait_all();
  __syncthreads();
  uint32_t kf[HS / 16][4], vf[HS / 16][4];
  load_a_frags<HS>(kf, sQ, warp * 16, lane);
  load_a_frags<HS>(vf, sDO, warp * 16, lane);
  __syncthreads();

  load_tile<HS>(sQ, qg, ld3, 0, T, tid);
  load_tile<HS>(sDO, dog, E, 0, T, tid);
  load_stats(s_stats, s_stats + AT_BLK, lse_bh, delta_bh, 0, T, tid);
  cp_async_commit();
  cp_async_wait_all();
  __syncthreads();

  float dk[HS / 8][4], dv[HS / 8][4];
#pragma unroll
  for (int nt = 0; nt < HS / 8; ++nt)
#pragma unroll
    for (int i = 0; i < 4; ++i) { dk[nt][i] = 0.f; dv[nt][i] = 0.f; }
  const float sl2 = scale * kLog2e;

  for (int qt = 0; qt < nq; ++qt) {
    const int buf = qt & 1;
    if (qt + 1 < nq) {
      load_tile<HS>(sQ + (buf ^ 1) * Tile<HS>::BYTES, qg, ld3, (qt + 1) * AT_BLK, T, tid);
      load_tile<HS>(sDO + (buf ^ 1) * Tile<HS>::BYTES, dog, E, (qt + 1) * AT_BLK, T, tid);
      load_stats(s_stats + (buf ^ 1) * 2 * AT_BLK, s_stats + (buf ^ 1) * 2 * AT_BLK + AT_BLK, lse_bh, delta_bh,
                 (qt + 1) * AT_BLK, T, tid);
      cp_async_commit();
    }
    const float* s_lse2 = s_stats + buf * 2 * AT_BLK;
    const float* s_delta = s_lse2 + AT_BLK;
    float st[8][4], dp[8][4];
    mma_a_tileT<HS>(st, kf, sQ + buf * Tile<HS>::BYTES, lane);
    mma_a_tileT<HS>(dp, vf, sDO + buf * Tile<HS>::BYTES, lane);
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const int qc = nt * 8 + 2 * t;
      const float la = s_lse2[qc], lb = s_lse2[qc + 1];
      const float da = s_delta[qc], db = s_delta[qc + 1];
      const float p0 = exp2f(st[nt][0] * sl2 - la), p1 = exp2f(st[nt][1] * sl2 - lb);
      const float p2 = exp2f(st[nt][2] * sl2 - la), p3 = exp2f(st[nt][3] * sl2 - lb);
      st[nt][0] = p0; st[nt][1] = p1; st[nt][2] = p2; st[nt][3] = p3;
      dp[nt][0] = p0 * (dp[nt][0] - da) * scale; dp[nt][1] = p1 * (dp[nt][1] - db) * scale;
      dp[nt][2] = p2 * (dp[nt][2] - da) * scale; dp[nt][3] = p3 * (dp[nt][3] - db) * scale;
    }
    uint32_t pk[4][4];
    pack_acc(pk, st);
    mma_p_tile<HS>(dv, pk, sDO + buf * Tile<HS>::BYTES, lane);
    pack_acc(pk, dp);
    mma_p_tile<HS>(dk, pk, sQ + buf * Tile<HS>::BYTES, lane);
    cp_async_wait_all();
    __syncthreads();
  }
  const int row_lo = kv0 + warp * 16;
  store_rows<HS>(dqkv, ld3, row_lo, T, static_cast<int64_t>(b) * T, E + h * HS, dk, 1.f, 1.f, lane);
  store_rows<HS>(dqkv, ld3, row_lo, T, static_cast<int64_t>(b) * T, 2 * E + h * HS, dv, 1.f, 1.f, lane);
}

// dQ: the CTA owns 64 query rows and streams all key/value tiles.
//   S = Q K^T, P = exp(S*scale - lse[row]), dP = dO V^T, dS = P (dP - delta[row]) scale, dQ += dS K.
template <int HS>
__global__ void __launch_bounds__(AT_THREADS)
attn_bf16_bwd_dq_kernel(const bf16* __restrict__ qkv, const bf16* __restrict__ d_o, const float* __restrict__ lse,
                        const float* __restrict__ delta, bf16* __restrict__ dqkv, int T, int H, float scale) {
  extern __shared__ __align__(16) uint8_t at_smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, t = lane & 3;
  const int bh = blockIdx.y, b = bh / H, h = bh - b * H;
  const int E = H * HS;
  const int64_t ld3 = 3 * static_cast<int64_t>(E);
  const int q0 = blockIdx.x * AT_BLK;
  const bf16* qg = qkv + static_cast<int64_t>(b) * T * ld3 + h * HS;
  const bf16* kg = qg + E;
  const bf16* vg = kg + E;
  const bf16* dog = d_o + static_cast<int64_t>(b) * T * E + h * HS;
  const uint32_t sK = ptx_smem(at_smem);                 // 2 buffers (first used for Q at start)
  const uint32_t sV = sK + 2 * Tile<HS>::BYTES;          // 2 buffers (first used for dO at start)
  const int nkv = ceil_div(T, AT_BLK);

  load_tile<HS>(sK, qg, ld3, q0, T, tid);
  load_tile<HS>(sV, dog, E, q0, T, tid);
  cp_async_commit();
  cp_async_wait_all();
  __syncthreads();
  uint32_t qf[HS / 16][4], dof[HS / 16][4];
  load_a_frags<HS>(qf, sK, warp * 16, lane);
  load_a_frags<HS>(dof, sV, warp * 16, lane);
  __syncthreads();

  load_tile<HS>(sK, kg, ld3, 0, T, tid);
  load_tile<HS>(sV, vg, ld3, 0, T, tid);
  cp_async_commit();
  cp_async_wait_all();
  __syncthreads();

  const int r0 = q0 + warp * 16 + g, r1 = r0 + 8;
  const float* lse_bh = lse + static_cast<int64_t>(bh) * T;
  const float* delta_bh = delta + static_cast<int64_t>(bh) * T;
  const float lse0 = r0 < T ? lse_bh[r0] * kLog2e : INFINITY, lse1 = r1 < T ? lse_bh[r1] * kLog2e : INFINITY;
  const float dl0 = r0 < T ? delta_bh[r0] : 0.f, dl1 = r1 < T ? delta_bh[r1] : 0.f;
  float dq[HS / 8][4];
#pragma unroll
  for (int nt = 0; nt < HS / 8; ++nt)
#pragma unroll
    for (int i = 0; i < 4; ++i) dq[nt][i] = 0.f;
  const float sl2 = scale * kLog2e;

  for (int kt = 0; kt < nkv; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nkv) {
      load_tile<HS>(sK + (buf ^ 1) * Tile<HS>::BYTES, kg, ld3, (kt + 1) * AT_BLK, T, tid);
      load_tile<HS>(sV + (buf ^ 1) * Tile<HS>::BYTES, vg, ld3, (kt + 1) * AT_BLK, T, tid);
      cp_async_commit();
    }
    float s[8][4], dp[8][4];
    mma_a_tileT<HS>(s, qf, sK + buf * Tile<HS>::BYTES, lane);
    mma_a_tileT<HS>(dp, dof, sV + buf * Tile<HS>::BYTES, lane);
    const int kv0 = kt * AT_BLK;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const int key = kv0 + nt * 8 + 2 * t;
      const bool va = key < T, vb_ = key + 1 < T;
      const float p0 = va ? exp2f(s[nt][0] * sl2 - lse0) : 0.f, p1 = vb_ ? exp2f(s[nt][1] * sl2 - lse0) : 0.f;
      const float p2 = va ? exp2f(s[nt][2] * sl2 - lse1) : 0.f, p3 = vb_ ? exp2f(s[nt][3] * sl2 - lse1) : 0.f;
      dp[nt][0] = p0 * (dp[nt][0] - dl0) * scale; dp[nt][1] = p1 * (dp[nt][1] - dl0) * scale;
      dp[nt][2] = p2 * (dp[nt][2] - dl1) * scale; dp[nt][3] = p3 * (dp[nt][3] - dl1) * scale;
    }
    uint32_t pk[4][4];
    pack_acc(pk, dp);
    mma_p_tile<HS>(dq, pk, sK + buf * Tile<HS>::BYTES, lane);
    cp_async_wait_all();
    __syncthreads();
  }
  store_rows<HS>(dqkv, ld3, q0 + warp * 16, T, static_cast<int64_t>(b) * T, h * HS, dq, 1.f, 1.f, lane);
}

template <typename K>
int set_attn_smem(K kern, size_t smem) {
  if (smem > 48 * 1024) VB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  return kOk;
}

template <int HS>
int fwd_launch(const bf16* qkv, bf16* out, float* lse, int B, int T, int H, float scale, cudaStream_t stream) {
  const size_t smem = 5 * Tile<HS>::BYTES;
  VB_PROPAGATE(set_attn_smem(attn_bf16_fwd_kernel<HS>, smem));
  dim3 grid(ceil_div(T, AT_BLK), B * H);
  ProfScope prof(kProfAttnFwd, 4.0 * B * H * T * T * HS, 1, stream);
  attn_bf16_fwd_kernel<HS><<<grid, AT_THREADS, smem, stream>>>(qkv, out, lse, T, H, scale);
  VB_CHECK_LAUNCH();
  return kOk;
}

template <int HS>
int bwd_launch(const bf16* qkv, const bf16* d_o, const float* lse, const float* delta, bf16* dqkv, int B, int T, int H,
               float scale, cudaStream_t stream) {
  dim3 grid(ceil_div(T, AT_BLK), B * H);
  ProfScope prof(kProfAttnBwd, 8.0 * B * H * T * T * HS, 2, stream);
  if (use_tc_gen_forward(T, HS)) {   // long rows / wide heads: both backward kernels on tcgen05
    static const bool dkv_legacy = getenv("VITB200_ATTN_GEN_DKV") && getenv("VITB200_ATTN_GEN_DKV")[0] == '0';
    if (!dkv_legacy) {
      VB_PROPAGATE(attn_tc_gen_backward_dkv(qkv, d_o, lse, delta, dqkv, B, T, H, HS, scale, stream));
      return attn_tc_gen_backward_dq(qkv, d_o, lse, delta, dqkv, B, T, H, HS, scale, stream);
    }
  }
  const size_t smem_kv = 4 * Tile<HS>::BYTES + 4 * AT_BLK * sizeof(float);
  VB_PROPAGATE(set_attn_smem(attn_bf16_bwd_dkv_kernel<HS>, smem_kv));
  attn_bf16_bwd_dkv_kernel<HS><<<grid, AT_THREADS, smem_kv, stream>>>(qkv, d_o, lse, delta, dqkv, T, H, scale);
  VB_CHECK_LAUNCH();
  if (use_tc_gen_forward(T, HS))   // dQ on tcgen05 (long rows / wide heads); dK, dV above stay on mma.sync for now
    return attn_tc_gen_backward_dq(qkv, d_o, lse, delta, dqkv, B, T, H, HS, scale, stream);
  const size_t smem_q = 4 * Tile<HS>::BYTES;
  VB_PROPAGATE(set_attn_smem(attn_bf16_bwd_dq_kernel<HS>, smem_q));
  attn_bf16_bwd_dq_kernel<HS><<<grid, AT_THREADS, smem_q, stream>>>(qkv, d_o, lse, delta, dqkv, T, H, scale);
  VB_CHECK_LAUNCH();
  return kOk;
}

}  // namespace

int attn_delta(const void* o, const void* d_o, bool is_bf16, float* delta, int B, int T, int H, int hs,
               cudaStream_t stream);

// VITB200_ATTN=legacy forces the mma.sync kernels everywhere (A/B comparison and tests of the generic path).
static bool legacy_attention() {
  static const bool legacy = [] {
    const char* e = getenv("VITB200_ATTN");
    return e != nullptr && e[0] == 'l';
  }();
  return legacy;
}
static bool use_tc_attention(int T, int hs) { return !legacy_attention() && attn_tc_supported(T, hs); }
// forward only: long rows / wide heads (ViT-H/14) on tcgen05; VITB200_ATTN_GEN=0 keeps them on mma.sync
static bool use_tc_gen_forward(int T, int hs) {
  static const bool off = [] {
    const char* e = getenv("VITB200_ATTN_GEN");
    return e != nullptr && e[0] == '0';
  }();
  return !off && !legacy_attention() && attn_tc_gen_supported(T, hs);
}

int attn_bf16_forward(const bf16* qkv, bf16* out, float* lse, int B, int T, int H, int hs, float scale,
                      cudaStream_t stream) {
  if (use_tc_attention(T, hs)) return attn_tc_forward(qkv, out, lse, B, T, H, hs, scale, stream);
  if (use_tc_gen_forward(T, hs)) return attn_tc_gen_forward(qkv, out, lse, B, T, H, hs, scale, stream);
  switch (hs) {
    case 16: return fwd_launch<16>(qkv, out, lse, B, T, H, scale, stream);
    case 32: return fwd_launch<32>(qkv, out, lse, B, T, H, scale, stream);
    case 64: return fwd_launch<64>(qkv, out, lse, B, T, H, scale, stream);
    case 80: return fwd_launch<80>(qkv, out, lse, B, T, H, scale, stream);
    case 96: return fwd_launch<96>(qkv, out, lse, B, T, H, scale, stream);
    case 128: return fwd_launch<128>(qkv, out, lse, B, T, H, scale, stream);
    default:
      set_error("attention (bf16): head_size %d not instantiated (supported: 16, 32, 64, 80, 96, 128)", hs);
      return kErrUnsupported;
  }
}

int attn_bf16_backward(const bf16* qkv, const bf16* out, const bf16* d_o, const float* lse, float* delta, bf16* dqkv,
                       int B, int T, int H, int hs, float scale, cudaStream_t stream, bool delta_ready) {
  // delta_ready: the caller's dO GEMM already wrote delta = rowsum(dO o O) from its epilogue (kEpiDot)
  if (!delta_ready) VB_PROPAGATE(attn_delta(out, d_o, true, delta, B, T, H, hs, stream));
  if (use_tc_attention(T, hs)) return attn_tc_backward(qkv, d_o, lse, delta, dqkv, B, T, H, hs, scale, stream);
  switch (hs) {
    case 16: return bwd_launch<16>(qkv, d_o, lse, delta, dqkv, B, T, H, scale, stream);
    case 32: return bwd_launch<32>(qkv, d_o, lse, delta, dqkv, B, T, H, scale, stream);
    case 64: return bwd_launch<64>(qkv, d_o, lse, delta, dqkv, B, T, H, scale, stream);
    case 80: return bwd_launch<80>(qkv, d_o, lse, delta, dqkv, B, T, H, scale, stream);
    case 96: return bwd_launch<96>(qkv, d_o, lse, delta, dqkv, B, T, H, scale, stream);
    case 128: return bwd_launch<128>(qkv, d_o, lse, delta, dqkv, B, T, H, scale, stream);
    default:
      set_error("attention (bf16): head_size %d not instantiated (supported: 16, 32, 64, 80, 96, 128)", hs);
      return kErrUnsupported;
  }
}

}  // namespace vb
