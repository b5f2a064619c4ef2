// Fused self-attention FORWARD on the 5th-gen tensor cores for the shapes attention_tc.cu does not take: any number
// of tokens and head sizes 16..128 (ViT-H/14: T = 577, head_size = 80).  Same semantics (reference
// transformer.py:20-31,44-45: logits MULTIPLIED by sqrt(head_size), no mask, heads are column slices of the packed
// [M,3E] qkv buffer, output concatenated in [M,E]) and the same building blocks as attention_tc.cu (S/P/O in tensor
// memory, P written in place as bf16, TS-MMA for P.V), re-cut for long rows:
//
//   * a work item is (image, head, PAIR of 128-row query tiles); keys/values stream through a 2-stage TMA ring in
//     chunks of 128 keys;
//   * the head dimension is staged as 64-column swizzle atoms through a 3-D tensor map whose innermost extent is
//     head_size, so the second atom of an 80-wide head arrives as 16 real columns + 48 zeros (no neighbour-head data);
//     S = Q K^T uses only the head_size/16 real k-steps, O = P V runs with N = 64 * atoms;
//   * softmax is TWO-PASS over the item instead of online: steps 0..n-1 recompute nothing but the row maximum
//     (S chunk -> tcgen05.ld -> max), steps n..2n-1 recompute S, exponentiate against the final maximum and feed P.V.
//     The accumulator is never rescaled (no correction pass, no dependence of P.V on the previous chunk's statistics);
//     the price is a second Q K^T, which at these shapes is cheaper than the exponentials (16 ex2 / clk / SM).
//   * the MMA thread interleaves the two tiles: while the softmax warps of one tile work on step s, the tensor pipe
//     runs P.V of step s and S of step s+1 of the other tile.
#include <cuda.h>
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"
#include "ptx.cuh"

namespace vb {

namespace {

constexpr int GQT = 128;                 // query rows per tile = UMMA M
constexpr int GKC = 128;                 // keys per chunk = UMMA N of S, K extent of P.V
constexpr int kGenThreads = 64 + 8 * 32; // producer, MMA issuer, 2 x 4 softmax warps
constexpr float kGLog2e = 1.4426950408889634f;
constexpr float kGLn2 = 0.6931471805599453f;
constexpr uint32_t kDescHi = ptx::smem_desc_hi_sw128(1024);   // 128B-swizzle descriptors, 8-row groups 1024 B apart

template <int HS>
struct GenCfg {
  static constexpr int NA = (HS + 63) / 64;               // 64-column atoms per row
  static constexpr int HSK = HS / 16;                     // real k-steps of S = Q K^T
  static constexpr int ON = NA * 64;                      // UMMA N of O = P V (padded head dimension)
  static constexpr int ATOM_Q = 2 * GQT * 128;            // one atom of the Q pair: 256 rows x 128 B
  static constexpr int ATOM_KV = GKC * 128;               // one atom of a K or V chunk
  static constexpr int Q_BYTES = NA * ATOM_Q;
  static constexpr int KV_BYTES = NA * ATOM_KV;           // K (or V) chunk
  static constexpr int STAGE_BYTES = 2 * KV_BYTES;        // K | V
  static constexpr int SMEM_BYTES = Q_BYTES + 2 * STAGE_BYTES + 256 + 1024;
};

struct GenParams {
  int T, H, npair, nchunk, items;   // npair = ceil(T / 256), nchunk = ceil(T / 128), items = B * H * npair
  float scale_log2;                 // scale * log2(e)
  int dbg;                          // timing experiments only (VITB200_ATTN_DBG), 0 in production:
                                    // 1 = the max pass reads no tensor memory, 2 = no ex2, 4 = pass 2 reads/writes no
                                    // tensor memory (results are wrong in all three)
};

template <int HS>
__global__ void __launch_bounds__(kGenThreads, 1)
attn_tc_gen_fwd_kernel(const __grid_constant__ CUtensorMap tm_qkv, bf16* __restrict__ out, float* __restrict__ lse,
                       const GenParams p) {
  using Cfg = GenCfg<HS>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* s_q = smem;
  uint8_t* s_kv = smem + Cfg::Q_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_kv + 2 * Cfg::STAGE_BYTES);
  uint64_t* q_full = bars;              // [1] TMA -> MMA
  uint64_t* q_empty = bars + 1;         // [1] MMA -> TMA (all S MMAs of the item issued and retired)
  uint64_t* kv_full = bars + 2;         // [2] TMA -> MMA
  uint64_t* kv_empty = bars + 4;        // [2] MMA -> TMA
  uint64_t* s_full = bars + 6;          // [2] per tile: S chunk in tensor memory (MMA -> softmax)
  uint64_t* ready = bars + 8;           // [2] per tile: softmax done with this step's S (max read / P written)
  uint64_t* o_full = bars + 10;         // [2] per tile: O complete (MMA -> softmax)
  uint64_t* o_free = bars + 12;         // [2] per tile: O drained (softmax -> MMA)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 14);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int E = p.H * HS;
  const int nstep = 2 * p.nchunk;

  if (threadIdx.x == 0) {
    ptx::tma_prefetch_desc(&tm_qkv);
    ptx::mbar_init(q_full, 1);
    ptx::mbar_init(q_empty, 1);
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&kv_full[s], 1);
      ptx::mbar_init(&kv_empty[s], 1);
      ptx::mbar_init(&s_full[s], 1);
      ptx::mbar_init(&ready[s], 4);
      ptx::mbar_init(&o_full[s], 1);
      ptx::mbar_init(&o_free[s], 4);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_ptr, 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_wait();
  pdl_trigger();
  // tensor-memory map per query tile t: S chunk (fp32, 128 columns) -> P (bf16 pairs, 64 columns) at t*256,
  // O (fp32, ON <= 128 columns) at t*256 + 128

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      uint32_t n_item = 0, n_step = 0;
      for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++n_item) {
        const int pr = item % p.npair, bh = item / p.npair;
        const int b = bh / p.H, h = bh - b * p.H;
        const int row0 = b * p.T;
        ptx::mbar_wait(q_empty, (n_item & 1u) ^ 1u);
        ptx::mbar_arrive_expect_tx(q_full, Cfg::Q_BYTES);
#pragma unroll
        for (int j = 0; j < Cfg::NA; ++j)
#pragma unroll
          for (int t = 0; t < 2; ++t)
            ptx::tma_load_3d(s_q + j * Cfg::ATOM_Q + t * (GQT * 128), &tm_qkv, q_full, 64 * j, h,
                             row0 + pr * 2 * GQT + t * GQT);
        for (int s = 0; s < nstep; ++s, ++n_step) {
          const int st = n_step & 1;
          const int c = s < p.nchunk ? s : s - p.nchunk;
          const bool with_v = s >= p.nchunk;
          uint8_t* base = s_kv + st * Cfg::STAGE_BYTES;
          ptx::mbar_wait(&kv_empty[st], ((n_step >> 1) & 1u) ^ 1u);
          ptx::mbar_arrive_expect_tx(&kv_full[st], with_v ? Cfg::STAGE_BYTES : Cfg::KV_BYTES);
#pragma unroll
          for (int j = 0; j < Cfg::NA; ++j) {
            ptx::tma_load_3d(base + j * Cfg::ATOM_KV, &tm_qkv, &kv_full[st], 64 * j, p.H + h, row0 + c * GKC);
            if (with_v)
              ptx::tma_load_3d(base + Cfg::KV_BYTES + j * Cfg::ATOM_KV, &tm_qkv, &kv_full[st], 64 * j, 2 * p.H + h,
                               row0 + c * GKC);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (one thread)
    if (lane == 0) {
      const uint32_t idesc_s = ptx::make_idesc_bf16_f32(GQT, GKC, 0, 0);       // S: M=128, N=128 keys, A,B K-major
      const uint32_t idesc_o = ptx::make_idesc_bf16_f32(GQT, Cfg::ON, 0, 1);   // O: M=128, N=ON, B (=V) MN-major
      const uint32_t sq = ptx::smem_u32(s_q);
      const uint32_t skv = ptx::smem_u32(s_kv);
      uint32_t n_item = 0, n_step = 0;
      uint32_t n_ready[2] = {0u, 0u}, n_ofree[2] = {0u, 0u};
      for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++n_item) {
        const int pr = item % p.npair;
        const bool live[2] = {pr * 2 * GQT < p.T, pr * 2 * GQT + GQT < p.T};
        auto issue_s = [&](int t, uint32_t st) {
          // lean issue: descriptor low words advance by plain adds, the high word is a constant (ptx.cuh)
          const uint32_t d_s = tmem_base + static_cast<uint32_t>(t * 256);
          const uint32_t a0 = ptx::smem_desc_lo(sq + t * (GQT * 128), 16);
          const uint32_t b0 = ptx::smem_desc_lo(skv + st * Cfg::STAGE_BYTES, 16);
#pragma unroll
          for (int k = 0; k < Cfg::HSK; ++k) {
            const int j = k >> 2, kk = k & 3;
            ptx::umma_f16_lh(d_s, a0 + ((j * Cfg::ATOM_Q + kk * 32) >> 4), b0 + ((j * Cfg::ATOM_KV + kk * 32) >> 4),
                             kDescHi, idesc_s, k > 0 ? 1u : 0u);
          }
          ptx::umma_commit(&s_full[t]);
        };
        auto issue_pv = [&](int t, uint32_t st, bool first) {
          const uint32_t a_p = tmem_base + static_cast<uint32_t>(t * 256);
          const uint32_t d_o = a_p + 128u;
          // V as MN-major B: 16 key rows per k-step = 2048 B, 8-row groups 1024 B apart, next 64-wide atom ATOM_KV on
          const uint32_t b0 = ptx::smem_desc_lo(skv + st * Cfg::STAGE_BYTES + Cfg::KV_BYTES, Cfg::ATOM_KV);
#pragma unroll
          for (int k = 0; k < GKC / 16; ++k)
            ptx::umma_f16_ts_lh(d_o, a_p + static_cast<uint32_t>(k * 8), b0 + k * 128, kDescHi, idesc_o,
                                (!first || k > 0) ? 1u : 0u);
        };
        ptx::mbar_wait(q_full, n_item & 1u);
        ptx::mbar_wait(&kv_full[n_step & 1u], (n_step >> 1) & 1u);
        ptx::tc_fence_after();
        for (int t = 0; t < 2; ++t)
          if (live[t]) issue_s(t, n_step & 1u);
        for (int s = 0; s < nstep; ++s) {
          const uint32_t cur = n_step + static_cast<uint32_t>(s);
          const uint32_t st = cur & 1u;
          const bool phase_b = s >= p.nchunk;
          const int c = phase_b ? s - p.nchunk : s;
          for (int t = 0; t < 2; ++t) {
            if (live[t]) {
              ptx::mbar_wait(&ready[t], n_ready[t] & 1u);     // softmax of tile t is done with step s
              ++n_ready[t];
              ptx::tc_fence_after();
              if (phase_b) {
                if (c == 0) {                                // previous item's O must have been drained
                  ptx::mbar_wait(&o_free[t], (n_ofree[t] & 1u) ^ 1u);
                  ++n_ofree[t];
                  ptx::tc_fence_after();
                }
                issue_pv(t, st, c == 0);
                if (c == p.nchunk - 1) ptx::umma_commit(&o_full[t]);
              }
            }
            if (t == 1) ptx::umma_commit(&kv_empty[st]);      // both tiles have issued everything that reads stage st
            if (s + 1 < nstep) {
              const uint32_t nst = (cur + 1u) & 1u;
              if (t == 0) {
                ptx::mbar_wait(&kv_full[nst], ((cur + 1u) >> 1) & 1u);
                ptx::tc_fence_after();
              }
              if (live[t]) issue_s(t, nst);
              if (t == 1 && s + 2 == nstep) ptx::umma_commit(q_empty);   // last S of the item issued
            }
          }
        }
        n_step += static_cast<uint32_t>(nstep);
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax + epilogue: one thread per query row
    const int t = (warp - 2) >> 2;               // query tile of this warp within the pair
    const int q = warp & 3;                      // tensor-memory lane quarter
    const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(t * 256);
    const float sl2 = p.scale_log2;
    uint32_t n_sfull = 0, n_ofull = 0;
    for (int item = blockIdx.x; item < p.items; item += gridDim.x) {
      const int pr = item % p.npair, bh = item / p.npair;
      const int tile0 = pr * 2 * GQT + t * GQT;
      if (tile0 >= p.T) continue;                // dead tile: the MMA thread skips it too
      const int b = bh / p.H, h = bh - b * p.H;
      const int row = tile0 + q * 32 + lane;     // query index inside the image
      float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
      float l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f;
      float mx = 0.f, mneg = 0.f;
      for (int s = 0; s < nstep; ++s) {
        const bool phase_b = s >= p.nchunk;
        const int key0 = (phase_b ? s - p.nchunk : s) * GKC;
        ptx::mbar_wait(&s_full[t], n_sfull & 1u);
        ++n_sfull;
        ptx::tc_fence_after();
        if (!phase_b) {
          // pass 1: running row maximum of the raw logits over the valid keys of this chunk
#pragma unroll 1
          for (int sub = 0; sub < GKC / 32; ++sub) {
            const int kb = key0 + sub * 32;
            if (kb >= p.T) break;                // warp-uniform
            if (p.dbg & 1) { mx0 = 0.f; continue; }
            uint32_t r[32];
            ptx::tmem_ld_32x32(t_row + static_cast<uint32_t>(sub * 32), r);
            ptx::tmem_ld_wait();
            if (kb + 32 <= p.T) {
#pragma unroll
              for (int j = 0; j < 32; j += 4) {
                mx0 = fmaxf(mx0, __uint_as_float(r[j]));
                mx1 = fmaxf(mx1, __uint_as_float(r[j + 1]));
                mx2 = fmaxf(mx2, __uint_as_float(r[j + 2]));
                mx3 = fmaxf(mx3, __uint_as_float(r[j + 3]));
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (kb + j < p.T) mx0 = fmaxf(mx0, __uint_as_float(r[j]));
            }
          }
          if (s == p.nchunk - 1) {
            mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
            mneg = -mx * sl2;                    // scale > 0: max commutes with the scaling
          }
        } else {
          // pass 2: p = exp2(s*c - m), row sum in fp32, P written over S as packed bf16 (zeros past the last key)
#pragma unroll 1
          for (int sub = 0; sub < GKC / 32; ++sub) {
            const int kb = key0 + sub * 32;
            uint32_t pk[16];
            if (kb >= p.T) {
#pragma unroll
              for (int j = 0; j < 16; ++j) pk[j] = 0u;
            } else if (p.dbg & 4) {
#pragma unroll
              for (int j = 0; j < 16; ++j) pk[j] = 0x3c003c00u;
              l0 += 1.f;
              continue;
            } else {
              uint32_t r[32];
              ptx::tmem_ld_32x32(t_row + static_cast<uint32_t>(sub * 32), r);
              ptx::tmem_ld_wait();
              if (p.dbg & 2) {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                  const float p0 = fmaf(__uint_as_float(r[2 * j]), sl2, mneg), p1 = fmaf(__uint_as_float(r[2 * j + 1]), sl2, mneg);
                  l0 += p0; l1 += p1;
                  pk[j] = pack_bf16x2(p0, p1);
                }
              } else if (kb + 32 <= p.T) {
#pragma unroll
                for (int j = 0; j < 16; j += 2) {
                  const float p0 = ptx::ex2(fmaf(__uint_as_float(r[2 * j]), sl2, mneg));
                  const float p1 = ptx::ex2(fmaf(__uint_as_float(r[2 * j + 1]), sl2, mneg));
                  const float p2 = ptx::ex2(fmaf(__uint_as_float(r[2 * j + 2]), sl2, mneg));
                  const float p3 = ptx::ex2(fmaf(__uint_as_float(r[2 * j + 3]), sl2, mneg));
                  l0 += p0; l1 += p1; l2 += p2; l3 += p3;
                  pk[j] = pack_bf16x2(p0, p1);
                  pk[j + 1] = pack_bf16x2(p2, p3);
                }
              } else {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                  const int k0 = kb + 2 * j;
                  const float p0 = k0 < p.T ? ptx::ex2(fmaf(__uint_as_float(r[2 * j]), sl2, mneg)) : 0.f;
                  const float p1 = k0 + 1 < p.T ? ptx::ex2(fmaf(__uint_as_float(r[2 * j + 1]), sl2, mneg)) : 0.f;
                  l0 += p0; l1 += p1;
                  pk[j] = pack_bf16x2(p0, p1);
                }
              }
            }
            // the 16 packed columns of sub-chunk `sub` land on fp32 columns that sub-chunks <= sub/2 occupied:
            // all of them have been read by this thread already (ascending order)
            ptx::tmem_st_32x16(t_row + static_cast<uint32_t>(sub * 16), pk);
          }
          ptx::tmem_st_wait();
        }
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&ready[t]);
      }
      // epilogue: O / l -> bf16 row, log-sum-exp
      const float l = (l0 + l1) + (l2 + l3);
      ptx::mbar_wait(&o_full[t], n_ofull & 1u);
      ++n_ofull;
      ptx::tc_fence_after();
      const float inv = 1.f / l;
      const bool live_row = row < p.T;
      bf16* orow = out + (static_cast<int64_t>(b) * p.T + row) * E + h * HS;
#pragma unroll
      for (int c = 0; c < (HS + 31) / 32; ++c) {
        uint32_t r[32];
        ptx::tmem_ld_32x32(t_row + 128u + static_cast<uint32_t>(c * 32), r);
        ptx::tmem_ld_wait();
        if (live_row) {
          constexpr int kFull = HS / 32;         // loads whose 32 columns are all real
          const int ngrp = c < kFull ? 4 : (HS - kFull * 32) / 8;
#pragma unroll
          for (int g8 = 0; g8 < 4; ++g8) {
            if (g8 < ngrp) {
              uint4 w;
              w.x = pack_bf16x2(__uint_as_float(r[8 * g8]) * inv, __uint_as_float(r[8 * g8 + 1]) * inv);
              w.y = pack_bf16x2(__uint_as_float(r[8 * g8 + 2]) * inv, __uint_as_float(r[8 * g8 + 3]) * inv);
              w.z = pack_bf16x2(__uint_as_float(r[8 * g8 + 4]) * inv, __uint_as_float(r[8 * g8 + 5]) * inv);
              w.w = pack_bf16x2(__uint_as_float(r[8 * g8 + 6]) * inv, __uint_as_float(r[8 * g8 + 7]) * inv);
              *reinterpret_cast<uint4*>(orow + c * 32 + g8 * 8) = w;
            }
          }
        }
      }
      if (live_row && lse != nullptr)
        lse[(static_cast<int64_t>(b) * p.H + h) * p.T + row] = (mx * sl2) * kGLn2 + logf(l);
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&o_free[t]);
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

template <int HS>
int launch_gen_fwd(const bf16* qkv, bf16* out, float* lse, int B, int T, int H, float scale, cudaStream_t stream) {
  using Cfg = GenCfg<HS>;
  const int E = H * HS;
  const int64_t M = static_cast<int64_t>(B) * T;
  CUtensorMap tm;
  VB_PROPAGATE(make_tmap_bf16_heads(&tm, qkv, M, 3 * H, HS, 3 * static_cast<int64_t>(E), GQT));
  static bool attr_done = false;
  if (!attr_done) {
    VB_CHECK_CUDA(cudaFuncSetAttribute(attn_tc_gen_fwd_kernel<HS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       Cfg::SMEM_BYTES));
    attr_done = true;
  }
  GenParams p;
  p.T = T; p.H = H;
  p.npair = ceil_div(T, 2 * GQT);
  p.nchunk = ceil_div(T, GKC);
  p.items = B * H * p.npair;
  p.scale_log2 = scale * kGLog2e;
  static const int dbg = getenv("VITB200_ATTN_DBG") ? atoi(getenv("VITB200_ATTN_DBG")) : 0;
  p.dbg = dbg;
  ProfScope prof(kProfAttnFwd, 4.0 * B * H * T * static_cast<double>(T) * HS, 1, stream);
  const int grid = p.items < num_sms() ? p.items : num_sms();
  VB_CHECK_CUDA(launch_pdl(attn_tc_gen_fwd_kernel<HS>, dim3(grid), dim3(kGenThreads), Cfg::SMEM_BYTES, stream, 1, tm,
                           out, lse, p));
  VB_CHECK_LAUNCH();
  return kOk;
}


// ====================================================================================================== backward, dQ
// dQ_t = dS K with dS = P o (dP - delta) * scale, P = exp2(S*c - lse*log2e), S = Q_t K^T, dP = dO_t V^T — the same
// skeleton as the forward's second pass: a PAIR of 128-row query tiles per item, keys / values stream in chunks of 64
// (S and dP of a chunk are 64 fp32 columns each, dS goes back IN PLACE over S as packed bf16 and is the TS-MMA A
// operand of dQ += dS K, K chunk as MN-major B).  No row reductions: the statistics come from the forward (lse) and
// the delta kernel, one value per thread.
constexpr int DKC = 64;                  // keys per chunk
// The backward kernels are bound by the element-wise chain of ONE warp per 32 rows (~950 instructions per 64-column
// chunk), so they run 16 element-wise warps: two warps share each 32-row quarter of a tile / slot and take one
// 32-column half of every chunk each.
constexpr int kBwdGenThreads = 64 + 16 * 32;
// The packed bf16 half written by one warp of a pair lands on fp32 columns the OTHER warp still has to read, so the
// pair meets on its own named barrier between the tensor-memory loads and the stores.
__device__ __forceinline__ void pair_sync(int pair) {
  asm volatile("bar.sync %0, 64;" ::"r"(pair + 1) : "memory");
}

template <int HS>
struct DqCfg {
  static constexpr int NA = (HS + 63) / 64;
  static constexpr int HSK = HS / 16;
  static constexpr int ON = NA * 64;
  static constexpr int ATOM_Q = 2 * GQT * 128;            // 256 rows x 128 B
  static constexpr int ATOM_KV = DKC * 128;               // 64 rows x 128 B
  static constexpr int Q_BYTES = NA * ATOM_Q;             // Q pair (and dO pair)
  static constexpr int KV_BYTES = NA * ATOM_KV;
  static constexpr int STAGE_BYTES = 2 * KV_BYTES;        // K | V
  static constexpr int SMEM_BYTES = 2 * Q_BYTES + 2 * STAGE_BYTES + 256 + 1024;
};

struct DqParams {
  int T, H, npair, nchunk, items;   // nchunk = ceil(T / 64)
  float scale, scale_log2;
};

template <int HS>
__global__ void __launch_bounds__(kBwdGenThreads, 1)
attn_tc_gen_dq_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_kv,
                      const __grid_constant__ CUtensorMap tm_do, const float* __restrict__ lse,
                      const float* __restrict__ delta, bf16* __restrict__ dqkv, const DqParams p) {
  using Cfg = DqCfg<HS>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* s_q = smem;
  uint8_t* s_do = smem + Cfg::Q_BYTES;
  uint8_t* s_kv = smem + 2 * Cfg::Q_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_kv + 2 * Cfg::STAGE_BYTES);
  uint64_t* q_full = bars;              // [1] Q and dO pair landed
  uint64_t* q_empty = bars + 1;         // [1] last S/dP of the item retired
  uint64_t* kv_full = bars + 2;         // [2]
  uint64_t* kv_empty = bars + 4;        // [2]
  uint64_t* s_full = bars + 6;          // [2] per tile: S and dP of a chunk in tensor memory
  uint64_t* ready = bars + 8;           // [2] per tile: dS written
  uint64_t* o_full = bars + 10;         // [2] per tile: dQ complete
  uint64_t* o_free = bars + 12;         // [2] per tile: dQ drained
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 14);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int E = p.H * HS;
  const int nstep = p.nchunk;

  if (threadIdx.x == 0) {
    ptx::tma_prefetch_desc(&tm_q);
    ptx::tma_prefetch_desc(&tm_kv);
    ptx::tma_prefetch_desc(&tm_do);
    ptx::mbar_init(q_full, 1);
    ptx::mbar_init(q_empty, 1);
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&kv_full[s], 1);
      ptx::mbar_init(&kv_empty[s], 1);
      ptx::mbar_init(&s_full[s], 1);
      ptx::mbar_init(&ready[s], 8);
      ptx::mbar_init(&o_full[s], 1);
      ptx::mbar_init(&o_free[s], 8);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_ptr, 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_wait();
  pdl_trigger();
  // tensor memory per query tile t (base t*256): S (64 fp32 columns) -> dS (32 packed columns) at +0, dP at +64,
  // dQ accumulator (ON <= 128 columns) at +128

  if (warp == 0) {
    if (lane == 0) {
      uint32_t n_item = 0, n_step = 0;
      for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++n_item) {
        const int pr = item % p.npair, bh = item / p.npair;
        const int b = bh / p.H, h = bh - b * p.H;
        const int row0 = b * p.T;
        ptx::mbar_wait(q_empty, (n_item & 1u) ^ 1u);
        ptx::mbar_arrive_expect_tx(q_full, 2 * Cfg::Q_BYTES);
#pragma unroll
        for (int j = 0; j < Cfg::NA; ++j)
#pragma unroll
          for (int t = 0; t < 2; ++t) {
            const int r = row0 + pr * 2 * GQT + t * GQT;
            ptx::tma_load_3d(s_q + j * Cfg::ATOM_Q + t * (GQT * 128), &tm_q, q_full, 64 * j, h, r);
            ptx::tma_load_3d(s_do + j * Cfg::ATOM_Q + t * (GQT * 128), &tm_do, q_full, 64 * j, h, r);
          }
        for (int s = 0; s < nstep; ++s, ++n_step) {
          const int st = n_step & 1;
          uint8_t* base = s_kv + st * Cfg::STAGE_BYTES;
          ptx::mbar_wait(&kv_empty[st], ((n_step >> 1) & 1u) ^ 1u);
          ptx::mbar_arrive_expect_tx(&kv_full[st], Cfg::STAGE_BYTES);
#pragma unroll
          for (int j = 0; j < Cfg::NA; ++j) {
            ptx::tma_load_3d(base + j * Cfg::ATOM_KV, &tm_kv, &kv_full[st], 64 * j, p.H + h, row0 + s * DKC);
            ptx::tma_load_3d(base + Cfg::KV_BYTES + j * Cfg::ATOM_KV, &tm_kv, &kv_full[st], 64 * j, 2 * p.H + h,
                             row0 + s * DKC);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc_s = ptx::make_idesc_bf16_f32(GQT, DKC, 0, 0);       // S, dP: M=128, N=64 keys
      const uint32_t idesc_o = ptx::make_idesc_bf16_f32(GQT, Cfg::ON, 0, 1);   // dQ: N=ON, B (=K) MN-major
      const uint32_t sq = ptx::smem_u32(s_q), sdo = ptx::smem_u32(s_do), skv = ptx::smem_u32(s_kv);
      uint32_t n_item = 0, n_step = 0;
      uint32_t n_ready[2] = {0u, 0u}, n_ofree[2] = {0u, 0u};
      for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++n_item) {
        const int pr = item % p.npair;
        const bool live[2] = {pr * 2 * GQT < p.T, pr * 2 * GQT + GQT < p.T};
        auto issue_sdp = [&](int t, uint32_t st) {
          const uint32_t d_s = tmem_base + static_cast<uint32_t>(t * 256);
          const uint32_t sk = skv + st * Cfg::STAGE_BYTES;
          const uint32_t qa0 = ptx::smem_desc_lo(sq + t * (GQT * 128), 16), kb0 = ptx::smem_desc_lo(sk, 16);
          const uint32_t oa0 = ptx::smem_desc_lo(sdo + t * (GQT * 128), 16), vb0 = ptx::smem_desc_lo(sk + Cfg::KV_BYTES, 16);
          // S and dP are independent accumulator chains: alternate them so that no MMA waits on its predecessor
#pragma unroll
          for (int k = 0; k < Cfg::HSK; ++k) {
            const int j = k >> 2, kk = k & 3;
            const uint32_t oa = (j * Cfg::ATOM_Q + kk * 32) >> 4, ob = (j * Cfg::ATOM_KV + kk * 32) >> 4;
            ptx::umma_f16_lh(d_s, qa0 + oa, kb0 + ob, kDescHi, idesc_s, k > 0 ? 1u : 0u);
            ptx::umma_f16_lh(d_s + 64u, oa0 + oa, vb0 + ob, kDescHi, idesc_s, k > 0 ? 1u : 0u);
          }
          ptx::umma_commit(&s_full[t]);
        };
        auto issue_dq = [&](int t, uint32_t st, bool first) {
          const uint32_t a_ds = tmem_base + static_cast<uint32_t>(t * 256);
          const uint32_t b0 = ptx::smem_desc_lo(skv + st * Cfg::STAGE_BYTES, Cfg::ATOM_KV);
#pragma unroll
          for (int k = 0; k < DKC / 16; ++k)
            ptx::umma_f16_ts_lh(a_ds + 128u, a_ds + static_cast<uint32_t>(k * 8), b0 + k * 128, kDescHi, idesc_o,
                                (!first || k > 0) ? 1u : 0u);
        };
        ptx::mbar_wait(q_full, n_item & 1u);
        ptx::mbar_wait(&kv_full[n_step & 1u], (n_step >> 1) & 1u);
        ptx::tc_fence_after();
        for (int t = 0; t < 2; ++t)
          if (live[t]) issue_sdp(t, n_step & 1u);
        if (nstep == 1) ptx::umma_commit(q_empty);
        for (int s = 0; s < nstep; ++s) {
          const uint32_t cur = n_step + static_cast<uint32_t>(s);
          const uint32_t st = cur & 1u;
          for (int t = 0; t < 2; ++t) {
            if (live[t]) {
              ptx::mbar_wait(&ready[t], n_ready[t] & 1u);
              ++n_ready[t];
              ptx::tc_fence_after();
              if (s == 0) {
                ptx::mbar_wait(&o_free[t], (n_ofree[t] & 1u) ^ 1u);
                ++n_ofree[t];
                ptx::tc_fence_after();
              }
              issue_dq(t, st, s == 0);
              if (s == nstep - 1) ptx::umma_commit(&o_full[t]);
            }
            if (t == 1) ptx::umma_commit(&kv_empty[st]);
            if (s + 1 < nstep) {
              const uint32_t nst = (cur + 1u) & 1u;
              if (t == 0) {
                ptx::mbar_wait(&kv_full[nst], ((cur + 1u) >> 1) & 1u);
                ptx::tc_fence_after();
              }
              if (live[t]) issue_sdp(t, nst);
              if (t == 1 && s + 2 == nstep) ptx::umma_commit(q_empty);
            }
          }
        }
        n_step += static_cast<uint32_t>(nstep);
      }
    }
  } else {
    // ------------------------------------------------------------------ element-wise + drain: one thread per query row
    const int t = ((warp - 2) >> 2) & 1;         // query tile
    const int half = (warp - 2) >> 3;            // which 32-key half of every chunk this warp converts
    const int q = warp & 3;
    const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(t * 256);
    const float sl2 = p.scale_log2, scale = p.scale;
    uint32_t n_sfull = 0, n_ofull = 0;
    for (int item = blockIdx.x; item < p.items; item += gridDim.x) {
      const int pr = item % p.npair, bh = item / p.npair;
      const int tile0 = pr * 2 * GQT + t * GQT;
      if (tile0 >= p.T) continue;
      const int b = bh / p.H, h = bh - b * p.H;
      const int row = tile0 + q * 32 + lane;
      const bool live_row = row < p.T;
      // dead rows: lse = +inf => P = 0 => dS = 0
      const float lse2 = live_row ? lse[static_cast<int64_t>(bh) * p.T + row] * kGLog2e : INFINITY;
      const float dlt = live_row ? delta[static_cast<int64_t>(bh) * p.T + row] : 0.f;
      for (int s = 0; s < nstep; ++s) {
        const int key0 = s * DKC;
        ptx::mbar_wait(&s_full[t], n_sfull & 1u);
        ++n_sfull;
        ptx::tc_fence_after();
        {
          const int sub = half;
          const int kb = key0 + sub * 32;
          uint32_t pk[16];
          if (kb >= p.T) {
#pragma unroll
            for (int j = 0; j < 16; ++j) pk[j] = 0u;
            pair_sync(t * 4 + q);
          } else {
            uint32_t rs[32], rp[32];
            ptx::tmem_ld_32x32(t_row + static_cast<uint32_t>(sub * 32), rs);
            ptx::tmem_ld_32x32(t_row + 64u + static_cast<uint32_t>(sub * 32), rp);
            ptx::tmem_ld_wait();
            pair_sync(t * 4 + q);
            const bool full = kb + 32 <= p.T;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float p0 = ptx::ex2(fmaf(__uint_as_float(rs[2 * j]), sl2, -lse2));
              const float p1 = ptx::ex2(fmaf(__uint_as_float(rs[2 * j + 1]), sl2, -lse2));
              float d0 = p0 * (__uint_as_float(rp[2 * j]) - dlt) * scale;
              float d1 = p1 * (__uint_as_float(rp[2 * j + 1]) - dlt) * scale;
              if (!full) {
                if (kb + 2 * j >= p.T) d0 = 0.f;
                if (kb + 2 * j + 1 >= p.T) d1 = 0.f;
              }
              pk[j] = pack_bf16x2(d0, d1);
            }
          }
          ptx::tmem_st_32x16(t_row + static_cast<uint32_t>(sub * 16), pk);
        }
        ptx::tmem_st_wait();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&ready[t]);
      }
      ptx::mbar_wait(&o_full[t], n_ofull & 1u);
      ++n_ofull;
      ptx::tc_fence_after();
      bf16* orow = dqkv + (static_cast<int64_t>(b) * p.T + row) * (3 * static_cast<int64_t>(E)) + h * HS;
#pragma unroll
      for (int c = half; c < (HS + 31) / 32; c += 2) {   // the two warps of a quarter split the 32-column groups
        uint32_t r[32];
        ptx::tmem_ld_32x32(t_row + 128u + static_cast<uint32_t>(c * 32), r);
        ptx::tmem_ld_wait();
        if (live_row) {
          constexpr int kFull = HS / 32;
          const int ngrp = c < kFull ? 4 : (HS - kFull * 32) / 8;
#pragma unroll
          for (int g8 = 0; g8 < 4; ++g8) {
            if (g8 < ngrp) {
              uint4 w;
              w.x = pack_bf16x2(__uint_as_float(r[8 * g8]), __uint_as_float(r[8 * g8 + 1]));
              w.y = pack_bf16x2(__uint_as_float(r[8 * g8 + 2]), __uint_as_float(r[8 * g8 + 3]));
              w.z = pack_bf16x2(__uint_as_float(r[8 * g8 + 4]), __uint_as_float(r[8 * g8 + 5]));
              w.w = pack_bf16x2(__uint_as_float(r[8 * g8 + 6]), __uint_as_float(r[8 * g8 + 7]));
              *reinterpret_cast<uint4*>(orow + c * 32 + g8 * 8) = w;
            }
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&o_free[t]);
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

template <int HS>
int launch_gen_dq(const bf16* qkv, const bf16* d_o, const float* lse, const float* delta, bf16* dqkv, int B, int T,
                  int H, float scale, cudaStream_t stream) {
  using Cfg = DqCfg<HS>;
  const int E = H * HS;
  const int64_t M = static_cast<int64_t>(B) * T;
  CUtensorMap tq, tkv, tdo;
  VB_PROPAGATE(make_tmap_bf16_heads(&tq, qkv, M, 3 * H, HS, 3 * static_cast<int64_t>(E), GQT));
  VB_PROPAGATE(make_tmap_bf16_heads(&tkv, qkv, M, 3 * H, HS, 3 * static_cast<int64_t>(E), DKC));
  VB_PROPAGATE(make_tmap_bf16_heads(&tdo, d_o, M, H, HS, E, GQT));
  static bool attr_done = false;
  if (!attr_done) {
    VB_CHECK_CUDA(cudaFuncSetAttribute(attn_tc_gen_dq_kernel<HS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       Cfg::SMEM_BYTES));
    attr_done = true;
  }
  DqParams p;
  p.T = T; p.H = H;
  p.npair = ceil_div(T, 2 * GQT);
  p.nchunk = ceil_div(T, DKC);
  p.items = B * H * p.npair;
  p.scale = scale;
  p.scale_log2 = scale * kGLog2e;
  const int grid = p.items < num_sms() ? p.items : num_sms();
  VB_CHECK_CUDA(launch_pdl(attn_tc_gen_dq_kernel<HS>, dim3(grid), dim3(kBwdGenThreads), Cfg::SMEM_BYTES, stream, 1, tq,
                           tkv, tdo, lse, delta, dqkv, p));
  VB_CHECK_LAUNCH();
  return kOk;
}


// ====================================================================================================== backward, dK / dV
// Transposed orientation (a tensor-memory A operand cannot be transposed): rows = keys.  An item is (image, head, ONE
// 128-key tile); queries stream in chunks of 64 through a 3-stage TMA ring (Q chunk | dO chunk):
//   S^T = K_t Q_c^T, dP^T = V_t dO_c^T  (64 fp32 columns each)  ->  P^T = exp2(S^T*c - lse[q]*log2e),
//   dS^T = P^T o (dP^T - delta[q]) * scale, both written back in place as packed bf16
//   dV_t += P^T dO_c, dK_t += dS^T Q_c   (TS-MMA, the chunk operands as MN-major B)
// Chunks alternate between two tensor-memory slots, each owned by one group of four element-wise warps, so the S/dP
// MMAs of chunk c+2 and the output MMAs of chunk c+1 run while a group works on chunk c.  The per-QUERY statistics are
// per-column here: each lane holds them for two columns of the chunk and they are broadcast with warp shuffles.
constexpr int KQC = 64;                  // queries per chunk
constexpr int kKvStages = 3;             // 5 stages measured slower (958 vs 888 us per ViT-H/14 layer backward)

template <int HS>
struct DkvCfg {
  static constexpr int NA = (HS + 63) / 64;
  static constexpr int HSK = HS / 16;
  static constexpr int ON = NA * 64;
  static constexpr int ATOM_T = GQT * 128;                // key tile atom: 128 rows x 128 B
  static constexpr int ATOM_C = KQC * 128;                // chunk atom: 64 rows x 128 B
  static constexpr int T_BYTES = NA * ATOM_T;             // K tile (and V tile)
  static constexpr int C_BYTES = NA * ATOM_C;             // Q chunk (and dO chunk)
  static constexpr int STAGE_BYTES = 2 * C_BYTES;         // Q | dO
  static constexpr int SMEM_BYTES = 2 * T_BYTES + kKvStages * STAGE_BYTES + 256 + 1024;
};

struct DkvParams {
  int T, H, ntile, nchunk, items;   // ntile = ceil(T / 128), nchunk = ceil(T / 64), items = B * H * ntile
  float scale, scale_log2;
  int dbg;                          // timing experiments only (VITB200_ATTN_DBG): 1 = element-wise pass does nothing
};

template <int HS>
__global__ void __launch_bounds__(kBwdGenThreads, 1)
attn_tc_gen_dkv_kernel(const __grid_constant__ CUtensorMap tm_kv, const __grid_constant__ CUtensorMap tm_q,
                       const __grid_constant__ CUtensorMap tm_do, const float* __restrict__ lse,
                       const float* __restrict__ delta, bf16* __restrict__ dqkv, const DkvParams p) {
  using Cfg = DkvCfg<HS>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* s_k = smem;
  uint8_t* s_v = smem + Cfg::T_BYTES;
  uint8_t* s_c = smem + 2 * Cfg::T_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_c + kKvStages * Cfg::STAGE_BYTES);
  uint64_t* kv_full = bars;             // [1] K and V tile landed
  uint64_t* kv_empty = bars + 1;        // [1] last S/dP of the item retired
  uint64_t* c_full = bars + 2;                     // [kKvStages] chunk ring
  uint64_t* c_empty = c_full + kKvStages;          // [kKvStages]
  uint64_t* s_full = c_empty + kKvStages;          // [2] per slot: S^T and dP^T in tensor memory
  uint64_t* ready = s_full + 2;                    // [2] per slot: P^T and dS^T written
  uint64_t* o_full = ready + 2;                    // [1] dV and dK complete
  uint64_t* o_free = o_full + 1;                   // [1] both accumulators drained (16 warps)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(o_free + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int E = p.H * HS;
  const int n = p.nchunk;

  if (threadIdx.x == 0) {
    ptx::tma_prefetch_desc(&tm_kv);
    ptx::tma_prefetch_desc(&tm_q);
    ptx::tma_prefetch_desc(&tm_do);
    ptx::mbar_init(kv_full, 1);
    ptx::mbar_init(kv_empty, 1);
    for (int s = 0; s < kKvStages; ++s) {
      ptx::mbar_init(&c_full[s], 1);
      ptx::mbar_init(&c_empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&s_full[s], 1);
      ptx::mbar_init(&ready[s], 8);
    }
    ptx::mbar_init(o_full, 1);
    ptx::mbar_init(o_free, 16);
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_ptr, 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_wait();
  pdl_trigger();
  // tensor memory: slot g at g*128: S^T (64 fp32 columns) -> P^T (32 packed) at +0, dP^T -> dS^T at +64;
  // dV accumulator at 256, dK accumulator at 384 (ON <= 128 columns each)

  if (warp == 0) {
    if (lane == 0) {
      uint32_t n_item = 0, n_chunk = 0;
      for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++n_item) {
        const int tile = item % p.ntile, bh = item / p.ntile;
        const int b = bh / p.H, h = bh - b * p.H;
        const int row0 = b * p.T;
        ptx::mbar_wait(kv_empty, (n_item & 1u) ^ 1u);
        ptx::mbar_arrive_expect_tx(kv_full, 2 * Cfg::T_BYTES);
#pragma unroll
        for (int j = 0; j < Cfg::NA; ++j) {
          ptx::tma_load_3d(s_k + j * Cfg::ATOM_T, &tm_kv, kv_full, 64 * j, p.H + h, row0 + tile * GQT);
          ptx::tma_load_3d(s_v + j * Cfg::ATOM_T, &tm_kv, kv_full, 64 * j, 2 * p.H + h, row0 + tile * GQT);
        }
        for (int c = 0; c < n; ++c, ++n_chunk) {
          const uint32_t st = n_chunk % kKvStages;
          uint8_t* base = s_c + st * Cfg::STAGE_BYTES;
          ptx::mbar_wait(&c_empty[st], ((n_chunk / kKvStages) & 1u) ^ 1u);
          ptx::mbar_arrive_expect_tx(&c_full[st], Cfg::STAGE_BYTES);
#pragma unroll
          for (int j = 0; j < Cfg::NA; ++j) {
            ptx::tma_load_3d(base + j * Cfg::ATOM_C, &tm_q, &c_full[st], 64 * j, h, row0 + c * KQC);
            ptx::tma_load_3d(base + Cfg::C_BYTES + j * Cfg::ATOM_C, &tm_do, &c_full[st], 64 * j, h, row0 + c * KQC);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc_s = ptx::make_idesc_bf16_f32(GQT, KQC, 0, 0);       // S^T, dP^T: M=128 keys, N=64 queries
      const uint32_t idesc_o = ptx::make_idesc_bf16_f32(GQT, Cfg::ON, 0, 1);   // dV, dK: N=ON, B MN-major
      const uint32_t sk = ptx::smem_u32(s_k), sv = ptx::smem_u32(s_v), sc = ptx::smem_u32(s_c);
      uint32_t n_item = 0, n_chunk = 0;
      uint32_t n_ready[2] = {0u, 0u};
      for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++n_item) {
        auto issue_s = [&](int c, int slot) {
          const uint32_t g = n_chunk + static_cast<uint32_t>(c);
          const uint32_t st = g % kKvStages;
          ptx::mbar_wait(&c_full[st], (g / kKvStages) & 1u);
          ptx::tc_fence_after();
          const uint32_t cq = sc + st * Cfg::STAGE_BYTES, cdo = cq + Cfg::C_BYTES;
          const uint32_t d = tmem_base + static_cast<uint32_t>(slot * 128);
          // independent accumulator chains, alternated; lean descriptor arithmetic
          const uint32_t ka0 = ptx::smem_desc_lo(sk, 16), qb0 = ptx::smem_desc_lo(cq, 16);
          const uint32_t va0 = ptx::smem_desc_lo(sv, 16), ob0 = ptx::smem_desc_lo(cdo, 16);
#pragma unroll
          for (int k = 0; k < Cfg::HSK; ++k) {
            const int j = k >> 2, kk = k & 3;
            const uint32_t oa = (j * Cfg::ATOM_T + kk * 32) >> 4, ob = (j * Cfg::ATOM_C + kk * 32) >> 4;
            ptx::umma_f16_lh(d, ka0 + oa, qb0 + ob, kDescHi, idesc_s, k > 0 ? 1u : 0u);
            ptx::umma_f16_lh(d + 64u, va0 + oa, ob0 + ob, kDescHi, idesc_s, k > 0 ? 1u : 0u);
          }
          ptx::umma_commit(&s_full[slot]);
        };
        auto issue_out = [&](int c, int slot, bool first) {
          const uint32_t g = n_chunk + static_cast<uint32_t>(c);
          const uint32_t st = g % kKvStages;
          const uint32_t cq = sc + st * Cfg::STAGE_BYTES, cdo = cq + Cfg::C_BYTES;
          const uint32_t a = tmem_base + static_cast<uint32_t>(slot * 128);
          const uint32_t db0 = ptx::smem_desc_lo(cdo, Cfg::ATOM_C), qb0 = ptx::smem_desc_lo(cq, Cfg::ATOM_C);
#pragma unroll
          for (int k = 0; k < KQC / 16; ++k) {   // dV += P^T dO_c and dK += dS^T Q_c, alternated
            ptx::umma_f16_ts_lh(tmem_base + 256u, a + static_cast<uint32_t>(k * 8), db0 + k * 128, kDescHi, idesc_o,
                                (!first || k > 0) ? 1u : 0u);
            ptx::umma_f16_ts_lh(tmem_base + 384u, a + 64u + static_cast<uint32_t>(k * 8), qb0 + k * 128, kDescHi,
                                idesc_o, (!first || k > 0) ? 1u : 0u);
          }
          ptx::umma_commit(&c_empty[st]);
        };
        ptx::mbar_wait(kv_full, n_item & 1u);
        ptx::tc_fence_after();
        issue_s(0, 0);
        if (n > 1) issue_s(1, 1);
        if (n <= 2) ptx::umma_commit(kv_empty);
        for (int c = 0; c < n; ++c) {
          const int slot = c & 1;
          ptx::mbar_wait(&ready[slot], n_ready[slot] & 1u);
          ++n_ready[slot];
          ptx::tc_fence_after();
          if (c == 0) {
            ptx::mbar_wait(o_free, (n_item & 1u) ^ 1u);
            ptx::tc_fence_after();
          }
          issue_out(c, slot, c == 0);
          if (c == n - 1) ptx::umma_commit(o_full);
          if (c + 2 < n) {
            issue_s(c + 2, slot);
            if (c + 2 == n - 1) ptx::umma_commit(kv_empty);
          }
        }
        n_chunk += static_cast<uint32_t>(n);
      }
    }
  } else {
    // ------------------------------------------------------------------ element-wise groups: one thread per key row
    const int g = ((warp - 2) >> 2) & 1;         // group = tensor-memory slot = chunk parity
    const int half = (warp - 2) >> 3;            // which 32-query half of every chunk this warp converts
    const int q = warp & 3;
    const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const uint32_t t_slot = t_lane + static_cast<uint32_t>(g * 128);
    const float sl2 = p.scale_log2, scale = p.scale;
    uint32_t n_sfull = 0, n_item = 0;
    for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++n_item) {
      const int tile = item % p.ntile, bh = item / p.ntile;
      const int b = bh / p.H, h = bh - b * p.H;
      const int key = tile * GQT + q * 32 + lane;
      const float* lse_bh = lse + static_cast<int64_t>(bh) * p.T;
      const float* dl_bh = delta + static_cast<int64_t>(bh) * p.T;
      for (int c = g; c < n; c += 2) {
        const int q0 = c * KQC;
        // statistics of query column q0 + 32*half + lane; columns past the last query: lse = +inf => P = 0
        const int qa = q0 + half * 32 + lane;
        const float nl_mine = qa < p.T ? -__ldg(lse_bh + qa) * kGLog2e : -INFINITY;
        const float dl_mine = qa < p.T ? __ldg(dl_bh + qa) : 0.f;
        ptx::mbar_wait(&s_full[g], n_sfull & 1u);
        ++n_sfull;
        ptx::tc_fence_after();
        {
          const int sub = half;
          uint32_t pkp[16], pkd[16];
          if (p.dbg & 1) {
            // nothing
          } else
          if (q0 + sub * 32 >= p.T) {            // warp-uniform
#pragma unroll
            for (int j = 0; j < 16; ++j) { pkp[j] = 0u; pkd[j] = 0u; }
            pair_sync(g * 4 + q);
          } else {
            uint32_t rs[32], rp[32];
            ptx::tmem_ld_32x32(t_slot + static_cast<uint32_t>(sub * 32), rs);
            ptx::tmem_ld_32x32(t_slot + 64u + static_cast<uint32_t>(sub * 32), rp);
            ptx::tmem_ld_wait();
            pair_sync(g * 4 + q);
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float nl0 = __shfl_sync(0xffffffffu, nl_mine, 2 * j), nl1 = __shfl_sync(0xffffffffu, nl_mine, 2 * j + 1);
              const float dl0 = __shfl_sync(0xffffffffu, dl_mine, 2 * j), dl1 = __shfl_sync(0xffffffffu, dl_mine, 2 * j + 1);
              const float p0 = ptx::ex2(fmaf(__uint_as_float(rs[2 * j]), sl2, nl0));
              const float p1 = ptx::ex2(fmaf(__uint_as_float(rs[2 * j + 1]), sl2, nl1));
              const float d0 = p0 * (__uint_as_float(rp[2 * j]) - dl0) * scale;
              const float d1 = p1 * (__uint_as_float(rp[2 * j + 1]) - dl1) * scale;
              pkp[j] = pack_bf16x2(p0, p1);
              pkd[j] = pack_bf16x2(d0, d1);
            }
          }
          if (!(p.dbg & 1)) {
            ptx::tmem_st_32x16(t_slot + static_cast<uint32_t>(sub * 16), pkp);
            ptx::tmem_st_32x16(t_slot + 64u + static_cast<uint32_t>(sub * 16), pkd);
          }
        }
        ptx::tmem_st_wait();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&ready[g]);
      }
      // drain: group 0 stores dV (section 2 of dqkv), group 1 stores dK (section 1)
      ptx::mbar_wait(o_full, n_item & 1u);
      ptx::tc_fence_after();
      const bool live_key = key < p.T;
      bf16* orow = dqkv + (static_cast<int64_t>(b) * p.T + key) * (3 * static_cast<int64_t>(E)) + (g == 0 ? 2 : 1) * E +
                   h * HS;
      const uint32_t t_acc = t_lane + (g == 0 ? 256u : 384u);
#pragma unroll
      for (int c = half; c < (HS + 31) / 32; c += 2) {   // the two warps of a quarter split the 32-column groups
        uint32_t r[32];
        ptx::tmem_ld_32x32(t_acc + static_cast<uint32_t>(c * 32), r);
        ptx::tmem_ld_wait();
        if (live_key) {
          constexpr int kFull = HS / 32;
          const int ngrp = c < kFull ? 4 : (HS - kFull * 32) / 8;
#pragma unroll
          for (int g8 = 0; g8 < 4; ++g8) {
            if (g8 < ngrp) {
              uint4 w;
              w.x = pack_bf16x2(__uint_as_float(r[8 * g8]), __uint_as_float(r[8 * g8 + 1]));
              w.y = pack_bf16x2(__uint_as_float(r[8 * g8 + 2]), __uint_as_float(r[8 * g8 + 3]));
              w.z = pack_bf16x2(__uint_as_float(r[8 * g8 + 4]), __uint_as_float(r[8 * g8 + 5]));
              w.w = pack_bf16x2(__uint_as_float(r[8 * g8 + 6]), __uint_as_float(r[8 * g8 + 7]));
              *reinterpret_cast<uint4*>(orow + c * 32 + g8 * 8) = w;
            }
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(o_free);
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

template <int HS>
int launch_gen_dkv(const bf16* qkv, const bf16* d_o, const float* lse, const float* delta, bf16* dqkv, int B, int T,
                   int H, float scale, cudaStream_t stream) {
  using Cfg = DkvCfg<HS>;
  const int E = H * HS;
  const int64_t M = static_cast<int64_t>(B) * T;
  CUtensorMap tkv, tq, tdo;
  VB_PROPAGATE(make_tmap_bf16_heads(&tkv, qkv, M, 3 * H, HS, 3 * static_cast<int64_t>(E), GQT));
  VB_PROPAGATE(make_tmap_bf16_heads(&tq, qkv, M, 3 * H, HS, 3 * static_cast<int64_t>(E), KQC));
  VB_PROPAGATE(make_tmap_bf16_heads(&tdo, d_o, M, H, HS, E, KQC));
  static bool attr_done = false;
  if (!attr_done) {
    VB_CHECK_CUDA(cudaFuncSetAttribute(attn_tc_gen_dkv_kernel<HS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       Cfg::SMEM_BYTES));
    attr_done = true;
  }
  DkvParams p;
  p.T = T; p.H = H;
  p.ntile = ceil_div(T, GQT);
  p.nchunk = ceil_div(T, KQC);
  p.items = B * H * p.ntile;
  p.scale = scale;
  p.scale_log2 = scale * kGLog2e;
  static const int dbg = getenv("VITB200_ATTN_DBG") ? atoi(getenv("VITB200_ATTN_DBG")) : 0;
  p.dbg = dbg;
  const int grid = p.items < num_sms() ? p.items : num_sms();
  VB_CHECK_CUDA(launch_pdl(attn_tc_gen_dkv_kernel<HS>, dim3(grid), dim3(kBwdGenThreads), Cfg::SMEM_BYTES, stream, 1, tkv,
                           tq, tdo, lse, delta, dqkv, p));
  VB_CHECK_LAUNCH();
  return kOk;
}

}  // namespace

bool attn_tc_gen_supported(int T, int hs) { return T >= 1 && (hs == 64 || hs == 80 || hs == 96 || hs == 128); }

int attn_tc_gen_forward(const bf16* qkv, bf16* out, float* lse, int B, int T, int H, int hs, float scale,
                        cudaStream_t stream) {
  if (!attn_tc_gen_supported(T, hs)) {
    set_error("attention (tcgen05, long rows): head_size %d not instantiated (64, 80, 96, 128)", hs);
    return kErrUnsupported;
  }
  VB_REQUIRE(scale > 0.f, "attention: scale must be positive");
  VB_REQUIRE(aligned16(qkv) && aligned16(out), "attention: qkv/out must be 16-byte aligned");
  VB_REQUIRE(static_cast<int64_t>(B) * H * ceil_div(T, 2 * GQT) < (1ll << 31), "attention: too many work items");
  switch (hs) {
    case 64: return launch_gen_fwd<64>(qkv, out, lse, B, T, H, scale, stream);
    case 80: return launch_gen_fwd<80>(qkv, out, lse, B, T, H, scale, stream);
    case 96: return launch_gen_fwd<96>(qkv, out, lse, B, T, H, scale, stream);
    default: return launch_gen_fwd<128>(qkv, out, lse, B, T, H, scale, stream);
  }
}

// dK and dV (the k and v sections of dqkv).
int attn_tc_gen_backward_dkv(const bf16* qkv, const bf16* d_o, const float* lse, const float* delta, bf16* dqkv, int B,
                             int T, int H, int hs, float scale, cudaStream_t stream) {
  if (!attn_tc_gen_supported(T, hs)) {
    set_error("attention dK/dV (tcgen05, long rows): head_size %d not instantiated (64, 80, 96, 128)", hs);
    return kErrUnsupported;
  }
  VB_REQUIRE(scale > 0.f, "attention: scale must be positive");
  VB_REQUIRE(aligned16(qkv) && aligned16(d_o) && aligned16(dqkv), "attention: operands must be 16-byte aligned");
  switch (hs) {
    case 64: return launch_gen_dkv<64>(qkv, d_o, lse, delta, dqkv, B, T, H, scale, stream);
    case 80: return launch_gen_dkv<80>(qkv, d_o, lse, delta, dqkv, B, T, H, scale, stream);
    case 96: return launch_gen_dkv<96>(qkv, d_o, lse, delta, dqkv, B, T, H, scale, stream);
    default: return launch_gen_dkv<128>(qkv, d_o, lse, delta, dqkv, B, T, H, scale, stream);
  }
}

// dQ only (the q section of dqkv).
int attn_tc_gen_backward_dq(const bf16* qkv, const bf16* d_o, const float* lse, const float* delta, bf16* dqkv, int B,
                            int T, int H, int hs, float scale, cudaStream_t stream) {
  if (!attn_tc_gen_supported(T, hs)) {
    set_error("attention dQ (tcgen05, long rows): head_size %d not instantiated (64, 80, 96, 128)", hs);
    return kErrUnsupported;
  }
  VB_REQUIRE(scale > 0.f, "attention: scale must be positive");
  VB_REQUIRE(aligned16(qkv) && aligned16(d_o) && aligned16(dqkv), "attention: operands must be 16-byte aligned");
  switch (hs) {
    case 64: return launch_gen_dq<64>(qkv, d_o, lse, delta, dqkv, B, T, H, scale, stream);
    case 80: return launch_gen_dq<80>(qkv, d_o, lse, delta, dqkv, B, T, H, scale, stream);
    case 96: return launch_gen_dq<96>(qkv, d_o, lse, delta, dqkv, B, T, H, scale, stream);
    default: return launch_gen_dq<128>(qkv, d_o, lse, delta, dqkv, B, T, H, scale, stream);
  }
}

}  // namespace vb
