// fp32 self-attention on CUDA cores, tiled flash-style: 64 x 64 score tiles held in registers (4 x 4 per thread),
// operands staged once per tile in shared memory, online softmax in registers with 16-lane shuffle reductions, nothing
// of size T x T written to HBM.  This is the fp32 training path (the 1e-4 parity gate, BASELINE configs[1]) for
// head sizes that are multiples of 4 up to 128; other head sizes (the reference's default config has hs = 1) use the
// row-per-warp kernels in attention_f32.cu.  Semantics: reference transformer.py:20-31 — logits = (q k^T) * scale with
// scale = +sqrt(head_size) MULTIPLIED, softmax over keys, out = wei v, heads = column slices of the packed [M,3E] qkv.
//
// Thread layout (256 threads): tm = tid / 16 owns tile rows 4*tm .. 4*tm+3; tn = tid % 16 owns score columns
// tn, tn+16, tn+32, tn+48 (strided, so the 8 threads of a quarter-warp read 8 consecutive shared-memory rows:
// conflict-free 16-byte loads with the +4 float row padding) and output columns 4*tn .. 4*tn+3 (+64 per d-chunk).
#include "common.cuh"
#include "kernels.h"

namespace vb {

namespace {

constexpr int TQ = 64;        // tile rows
constexpr int TKV = 64;       // tile columns
constexpr int TTHREADS = 256;
constexpr int PLD = TKV + 4;  // row pitch of the probability / dS tile

// rows [r0, r0+64) of one (image, head) slice -> smem tile [64][ld], rows >= T zero-filled.  src row pitch `pitch`.
__device__ __forceinline__ void load_tile(float* dst, int ld, const float* __restrict__ src, int64_t pitch, int r0, int T,
                                          int hs) {
  const int vec_per_row = hs >> 2;
  for (int i = threadIdx.x; i < TQ * vec_per_row; i += TTHREADS) {
    const int r = i / vec_per_row, c = (i - r * vec_per_row) << 2;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r0 + r < T) v = *reinterpret_cast<const float4*>(src + static_cast<int64_t>(r0 + r) * pitch + c);
    *reinterpret_cast<float4*>(dst + r * ld + c) = v;
  }
}

// acc[i][j] = sum_d A[4*tm+i][d] * B[tn+16*j][d]
__device__ __forceinline__ void tile_nt(const float* __restrict__ A, const float* __restrict__ Bm, int ld, int hs, int tm,
                                        int tn, float (&acc)[4][4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int d = 0; d < hs; d += 4) {
    float4 a[4], b[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) a[i] = *reinterpret_cast<const float4*>(A + (4 * tm + i) * ld + d);
#pragma unroll
    for (int j = 0; j < 4; ++j) b[j] = *reinterpret_cast<const float4*>(Bm + (tn + 16 * j) * ld + d);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        acc[i][j] = fmaf(a[i].x, b[j].x, acc[i][j]);
        acc[i][j] = fmaf(a[i].y, b[j].y, acc[i][j]);
        acc[i][j] = fmaf(a[i].z, b[j].z, acc[i][j]);
        acc[i][j] = fmaf(a[i].w, b[j].w, acc[i][j]);
      }
  }
}

// o[i][c*4+e] += sum_j P[4*tm+i][j] * V[j][64*c + 4*tn + e]        (P: [64][PLD], V: [64][ld])
template <int HC>
__device__ __forceinline__ void tile_pv(const float* __restrict__ P, const float* __restrict__ V, int ld, int hs, int tm,
                                        int tn, float (&o)[4][4 * HC]) {
#pragma unroll
  for (int c = 0; c < HC; ++c) {
    const int d0 = 64 * c + 4 * tn;
    if (d0 >= hs) break;
    for (int j = 0; j < TKV; j += 4) {
      float4 p[4], v[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) p[i] = *reinterpret_cast<const float4*>(P + (4 * tm + i) * PLD + j);
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) v[jj] = *reinterpret_cast<const float4*>(V + (j + jj) * ld + d0);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float pi[4] = {p[i].x, p[i].y, p[i].z, p[i].w};
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          o[i][c * 4 + 0] = fmaf(pi[jj], v[jj].x, o[i][c * 4 + 0]);
          o[i][c * 4 + 1] = fmaf(pi[jj], v[jj].y, o[i][c * 4 + 1]);
          o[i][c * 4 + 2] = fmaf(pi[jj], v[jj].z, o[i][c * 4 + 2]);
          o[i][c * 4 + 3] = fmaf(pi[jj], v[jj].w, o[i][c * 4 + 3]);
        }
      }
    }
  }
}

// same with the tile TRANSPOSED on the fly: o[i][..] += sum_j Pt[j][4*tm+i] * V[j][..]   (contraction over tile rows)
template <int HC>
__device__ __forceinline__ void tile_ptv(const float* __restrict__ Pt, const float* __restrict__ V, int ld, int hs, int tm,
                                         int tn, float (&o)[4][4 * HC]) {
#pragma unroll
  for (int c = 0; c < HC; ++c) {
    const int d0 = 64 * c + 4 * tn;
    if (d0 >= hs) break;
    for (int j = 0; j < TQ; ++j) {
      const float4 p = *reinterpret_cast<const float4*>(Pt + j * PLD + 4 * tm);
      const float4 v = *reinterpret_cast<const float4*>(V + j * ld + d0);
      const float pi[4] = {p.x, p.y, p.z, p.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        o[i][c * 4 + 0] = fmaf(pi[i], v.x, o[i][c * 4 + 0]);
        o[i][c * 4 + 1] = fmaf(pi[i], v.y, o[i][c * 4 + 1]);
        o[i][c * 4 + 2] = fmaf(pi[i], v.z, o[i][c * 4 + 2]);
        o[i][c * 4 + 3] = fmaf(pi[i], v.w, o[i][c * 4 + 3]);
      }
    }
  }
}

__device__ __forceinline__ float group16_max(float v) {
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float group16_sum(float v) {
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ------------------------------------------------------------------------------------------------ forward
// grid (B*H, ceil(T/64)).  PROBS: write the normalised probabilities [B,H,T,T] instead of the output (two sweeps).
template <int HC, bool PROBS>
__global__ void __launch_bounds__(TTHREADS)
attn_f32_tiled_fwd(const float* __restrict__ qkv, float* __restrict__ out, float* __restrict__ lse,
                   float* __restrict__ probs, int T, int H, int hs, float scale) {
  extern __shared__ __align__(16) float fsm[];
  const int ld = hs + 4;
  float* sQ = fsm;
  float* sK = sQ + TQ * ld;
  float* sV = sK + TKV * ld;
  float* sP = sV + TKV * ld;
  const int tid = threadIdx.x, tm = tid >> 4, tn = tid & 15;
  const int bh = blockIdx.x, b = bh / H, h = bh - b * H;
  const int q0 = blockIdx.y * TQ;
  const int E = H * hs;
  const int64_t row3 = 3 * static_cast<int64_t>(E);
  const float* qb = qkv + static_cast<int64_t>(b) * T * row3 + h * hs;
  load_tile(sQ, ld, qb, row3, q0, T, hs);
  float m[4], l[4], o[4][4 * HC];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    m[i] = -INFINITY;
    l[i] = 0.f;
#pragma unroll
    for (int e = 0; e < 4 * HC; ++e) o[i][e] = 0.f;
  }
  const int nkt = ceil_div(T, TKV);
  for (int sweep = 0; sweep < (PROBS ? 2 : 1); ++sweep) {
    for (int kt = 0; kt < nkt; ++kt) {
      const int k0 = kt * TKV;
      __syncthreads();                       // previous tile fully consumed
      load_tile(sK, ld, qb + E, row3, k0, T, hs);
      if (!PROBS) load_tile(sV, ld, qb + 2 * E, row3, k0, T, hs);
      __syncthreads();
      float s[4][4];
      tile_nt(sQ, sK, ld, hs, tm, tn, s);
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) s[i][j] = (k0 + tn + 16 * j < T) ? s[i][j] * scale : -INFINITY;
      if (PROBS && sweep == 1) {             // second sweep: statistics are final, emit p = exp(s - m) / l
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int r = q0 + 4 * tm + i;
          if (r >= T) continue;
          float* pr = probs + (static_cast<int64_t>(bh) * T + r) * T;
          const float inv = 1.f / l[i];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int c = k0 + tn + 16 * j;
            if (c < T) pr[c] = expf(s[i][j] - m[i]) * inv;
          }
        }
        continue;
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float mx = fmaxf(fmaxf(s[i][0], s[i][1]), fmaxf(s[i][2], s[i][3]));
        mx = group16_max(mx);
        const float mn = fmaxf(m[i], mx);    // finite: every key tile holds at least one valid column
        const float alpha = expf(m[i] - mn);
        float sum = 0.f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          s[i][j] = expf(s[i][j] - mn);
          sum += s[i][j];
        }
        sum = group16_sum(sum);
        l[i] = l[i] * alpha + sum;
        m[i] = mn;
        if (!PROBS) {
#pragma unroll
          for (int e = 0; e < 4 * HC; ++e) o[i][e] *= alpha;
#pragma unroll
          for (int j = 0; j < 4; ++j) sP[(4 * tm + i) * PLD + tn + 16 * j] = s[i][j];
        }
      }
      if (!PROBS) {
        __syncthreads();
        tile_pv<HC>(sP, sV, ld, hs, tm, tn, o);
      }
    }
  }
  if (PROBS) return;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = q0 + 4 * tm + i;
    if (r >= T) continue;
    const float inv = 1.f / l[i];
    float* orow = out + (static_cast<int64_t>(b) * T + r) * E + h * hs;
#pragma unroll
    for (int c = 0; c < HC; ++c) {
      const int d0 = 64 * c + 4 * tn;
      if (d0 < hs)
        *reinterpret_cast<float4*>(orow + d0) =
            make_float4(o[i][4 * c] * inv, o[i][4 * c + 1] * inv, o[i][4 * c + 2] * inv, o[i][4 * c + 3] * inv);
    }
    if (tn == 0 && lse != nullptr) lse[static_cast<int64_t>(bh) * T + r] = m[i] + logf(l[i]);
  }
}

// ------------------------------------------------------------------------------------------------ backward: dQ
// grid (B*H, ceil(T/64)): rows = a query tile, loop over key tiles.  ds = p (dp - delta) scale ; dq = ds k
template <int HC>
__global__ void __launch_bounds__(TTHREADS)
attn_f32_tiled_bwd_dq(const float* __restrict__ qkv, const float* __restrict__ d_o, const float* __restrict__ lse,
                      const float* __restrict__ delta, float* __restrict__ dqkv, int T, int H, int hs, float scale) {
  extern __shared__ __align__(16) float fsm[];
  const int ld = hs + 4;
  float* sQ = fsm;
  float* sDO = sQ + TQ * ld;
  float* sK = sDO + TQ * ld;
  float* sV = sK + TKV * ld;
  float* sS = sV + TKV * ld;
  const int tid = threadIdx.x, tm = tid >> 4, tn = tid & 15;
  const int bh = blockIdx.x, b = bh / H, h = bh - b * H;
  const int q0 = blockIdx.y * TQ;
  const int E = H * hs;
  const int64_t row3 = 3 * static_cast<int64_t>(E);
  const float* qb = qkv + static_cast<int64_t>(b) * T * row3 + h * hs;
  const float* dob = d_o + static_cast<int64_t>(b) * T * E + h * hs;
  load_tile(sQ, ld, qb, row3, q0, T, hs);
  load_tile(sDO, ld, dob, E, q0, T, hs);
  float lr[4], dl[4], dq[4][4 * HC];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = q0 + 4 * tm + i;
    lr[i] = r < T ? lse[static_cast<int64_t>(bh) * T + r] : INFINITY;     // +inf: p = 0 for rows that do not exist
    dl[i] = r < T ? delta[static_cast<int64_t>(bh) * T + r] : 0.f;
#pragma unroll
    for (int e = 0; e < 4 * HC; ++e) dq[i][e] = 0.f;
  }
  const int nkt = ceil_div(T, TKV);
  for (int kt = 0; kt < nkt; ++kt) {
    const int k0 = kt * TKV;
    __syncthreads();
    load_tile(sK, ld, qb + E, row3, k0, T, hs);
    load_tile(sV, ld, qb + 2 * E, row3, k0, T, hs);
    __syncthreads();
    float s[4][4], dp[4][4];
    tile_nt(sQ, sK, ld, hs, tm, tn, s);
    tile_nt(sDO, sV, ld, hs, tm, tn, dp);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const bool ok = k0 + tn + 16 * j < T;
        const float p = ok ? expf(s[i][j] * scale - lr[i]) : 0.f;
        sS[(4 * tm + i) * PLD + tn + 16 * j] = p * (dp[i][j] - dl[i]) * scale;
      }
    __syncthreads();
    tile_pv<HC>(sS, sK, ld, hs, tm, tn, dq);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = q0 + 4 * tm + i;
    if (r >= T) continue;
    float* dst = dqkv + (static_cast<int64_t>(b) * T + r) * row3 + h * hs;
#pragma unroll
    for (int c = 0; c < HC; ++c) {
      const int d0 = 64 * c + 4 * tn;
      if (d0 < hs) *reinterpret_cast<float4*>(dst + d0) = make_float4(dq[i][4 * c], dq[i][4 * c + 1], dq[i][4 * c + 2], dq[i][4 * c + 3]);
    }
  }
}

// ------------------------------------------------------------------------------------------------ backward: dK, dV
// grid (B*H, ceil(T/64)): COLUMNS of the score tile = this CTA's key tile, loop over query tiles (tile rows).
//   dv_j = sum_i p_ij dO_i ; dk_j = sum_i ds_ij q_i   — contractions over the tile ROWS (tile_ptv)
template <int HC>
__global__ void __launch_bounds__(TTHREADS)
attn_f32_tiled_bwd_dkv(const float* __restrict__ qkv, const float* __restrict__ d_o, const float* __restrict__ lse,
                       const float* __restrict__ delta, float* __restrict__ dqkv, int T, int H, int hs, float scale) {
  extern __shared__ __align__(16) float fsm[];
  const int ld = hs + 4;
  float* sK = fsm;
  float* sV = sK + TKV * ld;
  float* sQ = sV + TKV * ld;
  float* sDO = sQ + TQ * ld;
  float* sP = sDO + TQ * ld;
  float* sS = sP + TQ * PLD;
  const int tid = threadIdx.x, tm = tid >> 4, tn = tid & 15;
  const int bh = blockIdx.x, b = bh / H, h = bh - b * H;
  const int k0 = blockIdx.y * TKV;
  const int E = H * hs;
  const int64_t row3 = 3 * static_cast<int64_t>(E);
  const float* qb = qkv + static_cast<int64_t>(b) * T * row3 + h * hs;
  const float* dob = d_o + static_cast<int64_t>(b) * T * E + h * hs;
  load_tile(sK, ld, qb + E, row3, k0, T, hs);
  load_tile(sV, ld, qb + 2 * E, row3, k0, T, hs);
  float dk[4][4 * HC], dv[4][4 * HC];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int e = 0; e < 4 * HC; ++e) { dk[i][e] = 0.f; dv[i][e] = 0.f; }
  const int nqt = ceil_div(T, TQ);
  for (int qt = 0; qt < nqt; ++qt) {
    const int q0 = qt * TQ;
    __syncthreads();
    load_tile(sQ, ld, qb, row3, q0, T, hs);
    load_tile(sDO, ld, dob, E, q0, T, hs);
    __syncthreads();
    float s[4][4], dp[4][4];
    tile_nt(sQ, sK, ld, hs, tm, tn, s);       // rows = queries 4*tm+i, columns = keys tn+16*j
    tile_nt(sDO, sV, ld, hs, tm, tn, dp);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = q0 + 4 * tm + i;
      const float lrow = r < T ? lse[static_cast<int64_t>(bh) * T + r] : INFINITY;
      const float drow = r < T ? delta[static_cast<int64_t>(bh) * T + r] : 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const bool ok = k0 + tn + 16 * j < T;
        const float p = ok ? expf(s[i][j] * scale - lrow) : 0.f;
        sP[(4 * tm + i) * PLD + tn + 16 * j] = p;
        sS[(4 * tm + i) * PLD + tn + 16 * j] = p * (dp[i][j] - drow) * scale;
      }
    }
    __syncthreads();
    tile_ptv<HC>(sP, sDO, ld, hs, tm, tn, dv);   // output rows = keys 4*tm+i
    tile_ptv<HC>(sS, sQ, ld, hs, tm, tn, dk);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = k0 + 4 * tm + i;
    if (r >= T) continue;
    float* dst = dqkv + (static_cast<int64_t>(b) * T + r) * row3 + h * hs;
#pragma unroll
    for (int c = 0; c < HC; ++c) {
      const int d0 = 64 * c + 4 * tn;
      if (d0 < hs) {
        *reinterpret_cast<float4*>(dst + E + d0) = make_float4(dk[i][4 * c], dk[i][4 * c + 1], dk[i][4 * c + 2], dk[i][4 * c + 3]);
        *reinterpret_cast<float4*>(dst + 2 * E + d0) = make_float4(dv[i][4 * c], dv[i][4 * c + 1], dv[i][4 * c + 2], dv[i][4 * c + 3]);
      }
    }
  }
}

template <typename K>
int tiled_smem(K kern, size_t bytes) {
  if (bytes > 48 * 1024)
    VB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes)));
  return kOk;
}

}  // namespace

bool attn_f32_tiled_supported(int T, int hs, const void* a, const void* b) {
  return hs % 4 == 0 && hs >= 4 && hs <= 128 && T >= 1 && aligned16(a) && aligned16(b);
}

int attn_f32_tiled_forward(const float* qkv, float* out, float* lse, float* probs, int B, int T, int H, int hs,
                           float scale, cudaStream_t stream) {
  const int ld = hs + 4;
  const size_t smem = (static_cast<size_t>(3) * TQ * ld + TQ * PLD) * sizeof(float);
  const dim3 grid(B * H, ceil_div(T, TQ));
#define VB_FWD(HC, PR)                                                                                    \
  do {                                                                                                    \
    VB_PROPAGATE(tiled_smem(attn_f32_tiled_fwd<HC, PR>, smem));                                           \
    attn_f32_tiled_fwd<HC, PR><<<grid, TTHREADS, smem, stream>>>(qkv, out, lse, probs, T, H, hs, scale);  \
  } while (0)
  if (probs != nullptr) { if (hs <= 64) VB_FWD(1, true); else VB_FWD(2, true); }
  else { if (hs <= 64) VB_FWD(1, false); else VB_FWD(2, false); }
#undef VB_FWD
  VB_CHECK_LAUNCH();
  return kOk;
}

int attn_f32_tiled_backward(const float* qkv, const float* d_o, const float* lse, const float* delta, float* dqkv, int B,
                            int T, int H, int hs, float scale, cudaStream_t stream) {
  const int ld = hs + 4;
  const dim3 grid(B * H, ceil_div(T, TQ));
  const size_t smem_q = (static_cast<size_t>(4) * TQ * ld + TQ * PLD) * sizeof(float);
  const size_t smem_kv = (static_cast<size_t>(4) * TQ * ld + 2 * TQ * PLD) * sizeof(float);
  if (hs <= 64) {
    VB_PROPAGATE(tiled_smem(attn_f32_tiled_bwd_dq<1>, smem_q));
    attn_f32_tiled_bwd_dq<1><<<grid, TTHREADS, smem_q, stream>>>(qkv, d_o, lse, delta, dqkv, T, H, hs, scale);
    VB_CHECK_LAUNCH();
    VB_PROPAGATE(tiled_smem(attn_f32_tiled_bwd_dkv<1>, smem_kv));
    attn_f32_tiled_bwd_dkv<1><<<grid, TTHREADS, smem_kv, stream>>>(qkv, d_o, lse, delta, dqkv, T, H, hs, scale);
  } else {
    VB_PROPAGATE(tiled_smem(attn_f32_tiled_bwd_dq<2>, smem_q));
    attn_f32_tiled_bwd_dq<2><<<grid, TTHREADS, smem_q, stream>>>(qkv, d_o, lse, delta, dqkv, T, H, hs, scale);
    VB_CHECK_LAUNCH();
    VB_PROPAGATE(tiled_smem(attn_f32_tiled_bwd_dkv<2>, smem_kv));
    attn_f32_tiled_bwd_dkv<2><<<grid, TTHREADS, smem_kv, stream>>>(qkv, d_o, lse, delta, dqkv, T, H, hs, scale);
  }
  VB_CHECK_LAUNCH();
  return kOk;
}

}  // namespace vb
