// Fused self-attention on the 5th-gen tensor cores for the ViT shapes (head_size 64, T <= 256 tokens: the whole key
// axis of one (image, head) fits ONE tcgen05.mma in N, so softmax is single-pass, never online).
//
//   S = Q K^T            tcgen05.mma SS: A = Q tile [128 x 64] (smem, K-major), B = K [TP x 64] (smem, K-major),
//                        D = S [128 lanes x TP fp32 columns] in TENSOR MEMORY            (TP = T rounded up to 16)
//   P = exp2(S*c - max)  softmax warps: one thread per query row reads its row of S with tcgen05.ld, finds the
//                        max, exponentiates, and writes P back IN PLACE as packed bf16 (tcgen05.st) over S
//   O = P V              tcgen05.mma TS: A = P straight from tensor memory, B = V [TP x 64] (smem, MN-major),
//                        D = O [128 x 64] in the columns of the S region that P no longer needs
//   out = O / rowsum     same softmax warps drain O, normalise, write bf16 rows + the log-sum-exp
//
// One persistent CTA per SM walks over (image, head) items; warp 0 = TMA producer (2-deep Q/K/V ring), warp 1 = MMA
// issuer, warps 2-5 / 6-9 = softmax + epilogue for query rows 0-127 / 128-255.  Nothing of size T x T touches HBM.
// Semantics: reference transformer.py:20-31,44-45 — logits MULTIPLIED by sqrt(head_size) (scale is an argument),
// no mask, heads are column slices of the packed [M,3E] qkv buffer, output concatenated in [M,E].
#include <cuda.h>
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"
#include "ptx.cuh"

namespace vb {

namespace {

constexpr int HS = 64;                  // head size (one 128-byte swizzle row of bf16)
constexpr int QT = 128;                 // query rows per tile = UMMA M
constexpr int KV_ROWS = 256;            // rows staged for K / V / Q per item (2 TMA boxes of 128)
constexpr int TILE_BYTES = KV_ROWS * HS * 2;        // 32 KB
constexpr int STAGE_BYTES = 3 * TILE_BYTES;         // Q | K | V
constexpr int kStages = 2;
constexpr int kThreads = 64 + 8 * 32;
constexpr int SMEM_BYTES = kStages * STAGE_BYTES + 256 + 1024;
// EWG = 2 (round 2, T in 193..256): TWO softmax groups per query tile, each owning half of the key columns of every row
// (chunks 0-3 | 4-7) — the per-row chain of tcgen05.ld -> ex2 -> tcgen05.st, which bounds the kernel, is half as long.
// The groups exchange their partial row maxima / sums through 4 KB of shared memory and a 256-thread named barrier.
constexpr int kThreads2 = 64 + 16 * 32;
constexpr int XCH_BYTES = 2 * 2 * 2 * 128 * 4;      // [max | sum][tile][group][row]
constexpr int SMEM_BYTES2 = SMEM_BYTES + XCH_BYTES;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

#define VB_TR(tag)                                                                      \
  do {                                                                                  \
    if (trbuf != nullptr && trn < 512) trbuf[trn++] = (clock64() << 8) | (tag);         \
  } while (0)

struct FwdParams {
  int T, TP, H, items;   // TP = T rounded up to a multiple of 16 (UMMA N and PV K extent)
  float scale_log2;      // scale * log2(e)
  int st256;             // output rows are 32-byte aligned: store whole sectors per thread
  int order;             // MMA issue order: 1 skewed (default), 0 lock-step (VITB200_ATTN_FWD_ORDER=0, A/B timing)
  long long* trace;      // optional event trace of CTA 0 (tools/attn_trace.py fwd), else nullptr
};

template <int EWG>
__global__ void __launch_bounds__(EWG == 2 ? kThreads2 : kThreads, 1)
attn_tc_fwd_kernel(const __grid_constant__ CUtensorMap tm_qkv, bf16* __restrict__ out, float* __restrict__ lse,
                   const FwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * STAGE_BYTES);
  uint64_t* kv_full = bars;             // [2]  TMA -> MMA
  uint64_t* kv_empty = bars + 2;        // [2]  MMA -> TMA
  uint64_t* s_full = bars + 4;          // [2]  per query tile: S ready (MMA -> softmax)
  uint64_t* p_ready = bars + 6;         // [2]  P written (softmax -> MMA)
  uint64_t* o_full = bars + 8;          // [2]  O ready (MMA -> softmax)
  uint64_t* t_free = bars + 10;         // [2]  tile's tensor-memory region drained (softmax -> MMA)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 12);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int E = p.H * HS;

  if (threadIdx.x == 0) {
    ptx::tma_prefetch_desc(&tm_qkv);
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&kv_full[s], 1);
      ptx::mbar_init(&kv_empty[s], 1);
      ptx::mbar_init(&s_full[s], 1);
      ptx::mbar_init(&p_ready[s], 4 * EWG);
      ptx::mbar_init(&o_full[s], 1);
      ptx::mbar_init(&t_free[s], 4 * EWG);
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
  // EWG = 2: group 0 packs its P (keys 0..127) at columns [0,64), group 1 (keys 128..) at [128, ...): each group only
  // overwrites columns it has already read.  The O accumulators sit in columns that are dead once P is complete:
  // [64,128) and the 64 columns behind group 1's P.
  const int nchunk_all = (p.TP + 31) / 32;
  const uint32_t kO0 = EWG == 2 ? 64u : 128u;
  const uint32_t kO1 = EWG == 2 ? static_cast<uint32_t>(128 + 16 * (nchunk_all - 4)) : 192u;
  // tensor-memory map per query tile t: [t*256, t*256+TP) S (fp32) -> [t*256, +TP/2) P (bf16 pairs), O at t*256+128
  // (TP <= 256: P needs <= 128 columns, O 64 columns; S is dead once P is written)

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int it = 0;
      for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++it) {
        const int st = it & 1;
        const uint32_t ph = (it >> 1) & 1;
        const int b = item / p.H, h = item - b * p.H;
        ptx::mbar_wait(&kv_empty[st], ph ^ 1u);
        ptx::mbar_arrive_expect_tx(&kv_full[st], STAGE_BYTES);
        uint8_t* base = smem + st * STAGE_BYTES;
        const int row0 = b * p.T;
#pragma unroll
        for (int sec = 0; sec < 3; ++sec)
#pragma unroll
          for (int half = 0; half < 2; ++half)
            ptx::tma_load_2d(base + sec * TILE_BYTES + half * (QT * HS * 2), &tm_qkv, &kv_full[st],
                             sec * E + h * HS, row0 + half * QT);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (one thread)
    if (lane == 0) {
      const uint32_t idesc_s = ptx::make_idesc_bf16_f32(QT, p.TP, 0, 0);   // S: M=128, N=TP, A,B K-major
      const uint32_t idesc_o = ptx::make_idesc_bf16_f32(QT, HS, 0, 1);     // O: M=128, N=64, B (=V) MN-major
      const int ntile = p.T > QT ? 2 : 1;
      const int nks = p.TP / 16;
      constexpr uint32_t kHiSw = ptx::smem_desc_hi_sw128(1024);
      const bool dual = nks >= 2;              // both O accumulators get initialised by the first two k-steps
      // S = Q_t K^T of one (item, tile) unit into the tile's tensor-memory slot
      long long* trbuf = (p.trace != nullptr && blockIdx.x == 0) ? p.trace : nullptr;
      int trn = 0;
      auto issue_s = [&](int it2, int t) {
        const uint32_t sq = ptx::smem_u32(smem + (it2 & 1) * STAGE_BYTES), sk = sq + TILE_BYTES;
        VB_TR(1);
        ptx::mbar_wait(&t_free[t], (static_cast<uint32_t>(it2) & 1u) ^ 1u);
        ptx::tc_fence_after();
        VB_TR(2);
        const uint32_t d_s = tmem_base + static_cast<uint32_t>(t * 256);
        // lean issue (ptx.cuh): descriptor low words advance by plain adds, the high word is a constant
        const uint32_t a_lo = ptx::smem_desc_lo(sq + t * (QT * HS * 2), 16), b_lo = ptx::smem_desc_lo(sk, 16);
#pragma unroll
        for (int k = 0; k < HS / 16; ++k)
          ptx::umma_f16_lh(d_s, a_lo + 2u * k, b_lo + 2u * k, kHiSw, idesc_s, k > 0 ? 1u : 0u);
        ptx::umma_commit(&s_full[t]);
        VB_TR(3);
      };
      // O = P V of one unit.  Back-to-back MMAs into ONE accumulator serialise on its latency (~114 clk each against
      // ~55 clk for independent ones, tools/probe): the k-steps alternate between two accumulators (columns 128..191
      // and 192..255 of the slot; S is dead by then) which the epilogue adds.
      auto issue_pv = [&](int it2, int t) {
        const uint32_t sv = ptx::smem_u32(smem + (it2 & 1) * STAGE_BYTES) + 2 * TILE_BYTES;
        VB_TR(4);
        ptx::mbar_wait(&p_ready[t], static_cast<uint32_t>(it2) & 1u);
        ptx::tc_fence_after();
        VB_TR(5);
        const uint32_t a_p = tmem_base + static_cast<uint32_t>(t * 256);
        // V as MN-major B: 16 key rows per k-step = 16 * 128 B (low word + 128); 8-row groups 1024 B apart
        const uint32_t b_lo = ptx::smem_desc_lo(sv, 64 * 128);
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          if (k < nks) {
            const uint32_t d_o = a_p + ((dual && (k & 1)) ? kO1 : kO0);
            const uint32_t a_k = (EWG == 2 && k >= 8) ? a_p + 128u + static_cast<uint32_t>((k - 8) * 8)
                                                      : a_p + static_cast<uint32_t>(k * 8);
            ptx::umma_f16_ts_lh(d_o, a_k, b_lo + static_cast<uint32_t>(k * 128), kHiSw, idesc_o,
                                k > (dual ? 1 : 0) ? 1u : 0u);
          }
        }
        ptx::umma_commit(&o_full[t]);
        VB_TR(6);
      };
      const int my_items = blockIdx.x < p.items ? (p.items - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) /
                                                    static_cast<int>(gridDim.x) : 0;
      if (p.order == 0) {
        // lock-step order (round 1): S of both tiles, then P.V of both tiles, item by item
        for (int it = 0; it < my_items; ++it) {
          ptx::mbar_wait(&kv_full[it & 1], (it >> 1) & 1);
          ptx::tc_fence_after();
          for (int t = 0; t < ntile; ++t) issue_s(it, t);
          for (int t = 0; t < ntile; ++t) issue_pv(it, t);
          ptx::umma_commit(&kv_empty[it & 1]);
        }
      } else if (my_items > 0) {
        // skewed order: P.V(it, t) is followed at once by S(it+1, t), so that a tile's softmax warps get their next
        // logits while the OTHER tile's are still exponentiating — the two element-wise groups stop running in
        // lock-step, and the SFU works on one tile while the tensor pipe and the epilogue serve the other
        ptx::mbar_wait(&kv_full[0], 0);
        ptx::tc_fence_after();
        for (int t = 0; t < ntile; ++t) issue_s(0, t);
        for (int it = 0; it < my_items; ++it) {
          const bool more = it + 1 < my_items;
          for (int t = 0; t < ntile; ++t) {
            issue_pv(it, t);
            if (t == ntile - 1) ptx::umma_commit(&kv_empty[it & 1]);
            if (more) {
              if (t == 0) {
                VB_TR(7);
                ptx::mbar_wait(&kv_full[(it + 1) & 1], ((it + 1) >> 1) & 1);
                ptx::tc_fence_after();
                VB_TR(8);
              }
              issue_s(it + 1, t);
            }
          }
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax + epilogue: one thread per query row
    // (EWG = 2: per row TWO threads in two warps of the same lane quarter, each owning half of the key columns)
    const int grp4 = (warp - 2) >> 2;            // group of four warps covering the four tensor-memory lane quarters
    const int t = EWG == 2 ? grp4 >> 1 : grp4;   // query tile of this warp
    const int hg = EWG == 2 ? grp4 & 1 : 0;      // column half owned by this warp
    const int q = warp & 3;                      // tensor-memory lane quarter
    const int row = t * QT + q * 32 + lane;      // query index inside the item
    const bool tile_live = t * QT < p.T;
    const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(t * 256);
    const int nchunk = nchunk_all;
    const int cb = EWG == 2 ? (hg == 0 ? 0 : 4) : 0;                    // my chunks of 32 key columns: [cb, ce)
    const int ce = EWG == 2 ? (hg == 0 ? 4 : nchunk) : nchunk;
    // packed bf16 P of chunk c lands at column pdst(c): always at or below the chunk's own first column
    auto pdst = [&](int c) { return EWG == 2 && hg == 1 ? 128 + 16 * (c - 4) : 16 * c; };
    float* xch = reinterpret_cast<float*>(smem + kStages * STAGE_BYTES + 256);   // EWG = 2 only (SMEM_BYTES2)
    const int xi = (t * 2 + hg) * 128 + q * 32 + lane, xo = (t * 2 + (hg ^ 1)) * 128 + q * 32 + lane;
    long long* trbuf = (p.trace != nullptr && blockIdx.x == 0 && q == 0 && lane == 0) ? p.trace + 512 * (1 + t * 2 + hg) : nullptr;
    int trn = 0;
    int it = 0;
    for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++it) {
      if (!tile_live) continue;
      const uint32_t tp = it & 1;
      const int b = item / p.H, h = item - b * p.H;
      VB_TR(10);
      ptx::mbar_wait(&s_full[t], tp);
      ptx::tc_fence_after();
      VB_TR(11);
      // pass 1: row maximum of the raw logits over my share of the T valid keys (tail chunk masked)
      float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
      for (int c = cb; c < ce; ++c) {
        uint32_t r[32];
        ptx::tmem_ld_32x32(t_row + static_cast<uint32_t>(c * 32), r);
        ptx::tmem_ld_wait();
        if (c * 32 + 32 <= p.T) {
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
            if (c * 32 + j < p.T) mx0 = fmaxf(mx0, __uint_as_float(r[j]));
        }
      }
      float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
      if (EWG == 2) {                            // the row maximum is over BOTH halves
        xch[xi] = mx;
        asm volatile("bar.sync %0, 256;" ::"r"(1 + t) : "memory");
        mx = fmaxf(mx, xch[xo]);
      }
      const float sl2 = p.scale_log2;
      const float mneg = -mx * sl2;              // scale > 0: max commutes with the scaling
      VB_TR(12);
      // pass 2: p = exp2(s*c - m), row sum in fp32, P written over S as packed bf16
      float l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f;
      for (int c = cb; c < ce; ++c) {
        uint32_t r[32];
        ptx::tmem_ld_32x32(t_row + static_cast<uint32_t>(c * 32), r);
        ptx::tmem_ld_wait();
        uint32_t pk[16];
        if (c * 32 + 32 <= p.T) {
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
            const int k0 = c * 32 + 2 * j;
            const float p0 = k0 < p.T ? ptx::ex2(fmaf(__uint_as_float(r[2 * j]), sl2, mneg)) : 0.f;
            const float p1 = k0 + 1 < p.T ? ptx::ex2(fmaf(__uint_as_float(r[2 * j + 1]), sl2, mneg)) : 0.f;
            l0 += p0; l1 += p1;
            pk[j] = pack_bf16x2(p0, p1);
          }
        }
        ptx::tmem_st_32x16(t_row + static_cast<uint32_t>(pdst(c)), pk);
      }
      float l = (l0 + l1) + (l2 + l3);
      ptx::tmem_st_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&p_ready[t]);
      VB_TR(13);
      if (EWG == 2) {                            // the row sum is over BOTH halves
        xch[512 + xi] = l;
        asm volatile("bar.sync %0, 256;" ::"r"(1 + t) : "memory");
        l += xch[512 + xo];
      }
      // epilogue: O / l -> bf16 row, log-sum-exp.  EWG = 1: all 64 columns; EWG = 2: my 32 of them.
      ptx::mbar_wait(&o_full[t], tp);
      ptx::tc_fence_after();
      VB_TR(14);
      const float inv = 1.f / l;
      const bool live = row < p.T;
      bf16* orow = out + (static_cast<int64_t>(b) * p.T + row) * E + h * HS;
#pragma unroll
      for (int c = (EWG == 2 ? hg : 0); c < (EWG == 2 ? hg + 1 : 2); ++c) {
        uint32_t r[32];
        ptx::tmem_ld_32x32(t_row + kO0 + static_cast<uint32_t>(c * 32), r);
        if (p.TP >= 32) {                       // second accumulator (odd k-steps of P.V), see the MMA issuer
          uint32_t r2[32];
          ptx::tmem_ld_32x32(t_row + kO1 + static_cast<uint32_t>(c * 32), r2);
          ptx::tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) r[j] = __float_as_uint(__uint_as_float(r[j]) + __uint_as_float(r2[j]));
        }
        ptx::tmem_ld_wait();
        if (c == (EWG == 2 ? hg : 1)) {          // my last read of the slot: release it BEFORE the scaling and the stores,
          ptx::tc_fence_before();                //  the next item's S = Q K^T of this tile can be issued meanwhile
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(&t_free[t]);
        }
        if (live) {
          uint32_t w[16];
#pragma unroll
          for (int j = 0; j < 16; ++j)
            w[j] = pack_bf16x2(__uint_as_float(r[2 * j]) * inv, __uint_as_float(r[2 * j + 1]) * inv);
          if (p.st256) {                            // whole 32-byte sectors per thread (rows are 32-byte aligned)
            const uint32_t lo[8] = {w[0], w[1], w[2], w[3], w[4], w[5], w[6], w[7]};
            const uint32_t hi[8] = {w[8], w[9], w[10], w[11], w[12], w[13], w[14], w[15]};
            ptx::stg256(orow + c * 32, lo);
            ptx::stg256(orow + c * 32 + 16, hi);
          } else {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              *reinterpret_cast<uint4*>(orow + c * 32 + j * 8) = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
          }
        }
      }
      if (live && hg == 0 && lse != nullptr)
        lse[(static_cast<int64_t>(b) * p.H + h) * p.T + row] = (mx * sl2) * kLn2 + logf(l);
      VB_TR(15);
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}


// ====================================================================================================== backward
// Two kernels with one skeleton (MODE 0: dQ, rows = queries, columns = keys; MODE 1: dK/dV, rows = keys, columns =
// queries — the transposed orientation, because a tensor-memory A operand cannot be transposed):
//   MODE 0:  S = Q_t K^T, dP = dO_t V^T            -> dS = P o (dP - delta) * scale         -> dQ_t = dS K
//   MODE 1:  S^T = K_t Q^T, dP^T = V_t dO^T        -> P^T, dS^T                              -> dV_t = P^T dO, dK_t = dS^T Q
// with P = exp2(S*c - lse*log2e) recomputed from the saved log-sum-exp (no row reductions in backward).
//
// A unit = one (image, head) item x one 128-row tile.  Its columns are cut into nchk chunks of <= W columns; chunks
// flow through TWO tensor-memory slots ([S | dP], 2W columns each) so that the three engines overlap:
//   MMA thread    : S/dP of chunk g+1 is issued BEFORE the output MMAs of chunk g (tensor pipe never waits on math)
//   element-wise  : warps 2-5 own slot 0 (even chunks), warps 6-9 own slot 1 (odd chunks); one thread per row reads
//                   fp32 S/dP, writes bf16 P/dS back IN PLACE (ascending, so it only overwrites what it already read)
//   output MMAs   : TS form, A = the bf16 chunk straight from tensor memory, accumulate over the unit's chunks into
//                   two accumulators at columns 384..447 | 448..511 (MODE 0: dQ split by k-step parity, MODE 1: dV | dK)
// Both groups drain half of each accumulator (32 columns) once a unit's last chunk is committed.
constexpr int kBwdThreads = 64 + 8 * 32 + 32;   // producer, S/dP issuer, 8 element-wise warps, output-MMA issuer
constexpr int kOutWarp = 10;
// DW (round 2): four more warps (11-14, one per tensor-memory lane quarter) drain the accumulators, so the element-wise
// warps never stop for the read-out and its uncoalesced row stores (35 of 319 us per ViT-B/16 layer with the stores
// removed, profiles/r02_attn_bwd_bound_experiments.txt) and the accumulators are released as early as possible.
constexpr int kBwdThreadsDW = kBwdThreads + 4 * 32;
constexpr int kStatBytes = 2048;            // lse*log2e [256] | delta*scale [256] per stage

struct BwdParams {
  int T, TP, H, items, stages;
  int W, nchk;                              // chunk width (multiple of 16) and chunks per unit
  int tile_bytes, stage_bytes;              // TP*128, 4*tile_bytes + kStatBytes
  float scale, scale_log2;
  int st256;                                // dqkv rows are 32-byte aligned: drain with 32-byte stores
  int full;                                 // 1: ONE full-width chunk per unit in ONE slot (see attn_tc_backward), 0: chunked, 2 slots
  int hc;                                   // full: column split between the two element-wise groups (0: balanced)
  int acc0, acc1;                           // tensor-memory columns of the two accumulators
  int dbg;                                  // timing experiments only (VITB200_ATTN_DBG), 0 in production
  long long* trace;                         // optional event trace of CTA 0 (tools/attn_trace.py), else nullptr
};


template <int N>
__device__ __forceinline__ void tmem_ld_n(uint32_t taddr, uint32_t (&r)[N]) {
  if constexpr (N == 32) ptx::tmem_ld_32x32(taddr, r); else ptx::tmem_ld_32x16(taddr, r);
}
template <int N>
__device__ __forceinline__ void tmem_st_n(uint32_t taddr, const uint32_t (&r)[N]) {
  if constexpr (N == 16) ptx::tmem_st_32x16(taddr, r); else ptx::tmem_st_32x8(taddr, r);
}

// Column split of a chunk of wc columns (a multiple of 16) between the two element-wise groups, in whole k-steps.
__device__ __forceinline__ int bwd_split(int wc, int hc = 0) {
  return hc > 0 ? (hc < wc ? hc : wc) : ((wc >> 4) + 1) >> 1 << 4;
}

// N columns [c0, c0+N) of one chunk (global column index gcol0 .. gcol0+N-1) for this thread's row.
template <int MODE, int N>
__device__ __forceinline__ void bwd_piece(uint32_t tS, uint32_t tP, int c0, int d0, int gcol0, int T, float sl2,
                                          float scale, float lse2r, float dscr, uint32_t stat_s) {
  uint32_t rs[N], rp[N];
  tmem_ld_n<N>(tS + static_cast<uint32_t>(c0), rs);
  tmem_ld_n<N>(tP + static_cast<uint32_t>(c0), rp);
  ptx::tmem_ld_wait();
  uint32_t pk0[N / 2], pk1[N / 2];
  if (MODE == 0) {
    if (gcol0 + N <= T) {                  // all keys valid (warp-uniform): no per-element predicates
#pragma unroll
      for (int j = 0; j < N / 2; ++j) {
        const float e0 = ptx::ex2(fmaf(__uint_as_float(rs[2 * j]), sl2, -lse2r)) *
                         fmaf(__uint_as_float(rp[2 * j]), scale, -dscr);
        const float e1 = ptx::ex2(fmaf(__uint_as_float(rs[2 * j + 1]), sl2, -lse2r)) *
                         fmaf(__uint_as_float(rp[2 * j + 1]), scale, -dscr);
        pk0[j] = pack_bf16x2(e0, e1);
      }
    } else {
#pragma unroll
      for (int j = 0; j < N / 2; ++j) {
        const int k0 = gcol0 + 2 * j;
        float e0 = ptx::ex2(fmaf(__uint_as_float(rs[2 * j]), sl2, -lse2r));
        float e1 = ptx::ex2(fmaf(__uint_as_float(rs[2 * j + 1]), sl2, -lse2r));
        e0 = k0 < T ? e0 * fmaf(__uint_as_float(rp[2 * j]), scale, -dscr) : 0.f;
        e1 = k0 + 1 < T ? e1 * fmaf(__uint_as_float(rp[2 * j + 1]), scale, -dscr) : 0.f;
        pk0[j] = pack_bf16x2(e0, e1);
      }
    }
  } else {
    const uint32_t l4 = stat_s + static_cast<uint32_t>(gcol0) * 4u, d4 = l4 + 1024u;
#pragma unroll
    for (int j4 = 0; j4 < N / 4; ++j4) {
      const float4 lv = ptx::lds128(l4 + 16u * j4), dv = ptx::lds128(d4 + 16u * j4);
      const float lj[4] = {lv.x, lv.y, lv.z, lv.w}, dj[4] = {dv.x, dv.y, dv.z, dv.w};
      float pe[4], de[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        pe[i] = ptx::ex2(fmaf(__uint_as_float(rs[4 * j4 + i]), sl2, -lj[i]));
        de[i] = pe[i] * fmaf(__uint_as_float(rp[4 * j4 + i]), scale, -dj[i]);
      }
      pk0[2 * j4] = pack_bf16x2(pe[0], pe[1]); pk0[2 * j4 + 1] = pack_bf16x2(pe[2], pe[3]);
      pk1[2 * j4] = pack_bf16x2(de[0], de[1]); pk1[2 * j4 + 1] = pack_bf16x2(de[2], de[3]);
    }
  }
  tmem_st_n<N / 2>(tS + static_cast<uint32_t>(d0), pk0);
  if (MODE == 1) tmem_st_n<N / 2>(tP + static_cast<uint32_t>(d0), pk1);
}

template <int MODE, bool DW>
__global__ void __launch_bounds__(DW ? kBwdThreadsDW : kBwdThreads, 1)
attn_tc_bwd_kernel(const __grid_constant__ CUtensorMap tm_qkv, const __grid_constant__ CUtensorMap tm_do,
                   const float* __restrict__ lse, const float* __restrict__ delta, bf16* __restrict__ dqkv,
                   const BwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  // barriers live AFTER the stages and the over-read pad (see host side)
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.stages * p.stage_bytes + (32768 - p.tile_bytes));
  uint64_t* kv_full = bars;          // [2] TMA bytes landed
  uint64_t* st_full = bars + 2;      // [2] row statistics staged (32 producer lanes)
  uint64_t* kv_empty = bars + 4;     // [2]
  uint64_t* sd_full = bars + 6;      // [2] S and dP of the chunk in slot s are ready
  uint64_t* ew_done = bars + 8;      // [2][2] bf16 operands of slot s, column half h written (4 warps each)
  uint64_t* slot_free = bars + 12;   // [2] output MMAs that read slot s have completed
  uint64_t* acc_full = bars + 14;    // unit's accumulators complete
  uint64_t* acc_free = bars + 15;    // accumulators drained (8 warps)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 16);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int E = p.H * HS;
  const int ntile = p.T > QT ? 2 : 1;
  const int nchk = p.nchk, W = p.W;
  const bool dual = p.TP >= 32;      // MODE 0: a unit's first chunk then has >= 2 k-steps, so both accumulators get initialised
  const uint32_t kAcc0 = static_cast<uint32_t>(p.acc0);   // MODE 0: dQ (even k-steps) | MODE 1: dV
  const uint32_t kAcc1 = static_cast<uint32_t>(p.acc1);   // MODE 0: dQ (odd k-steps)  | MODE 1: dK
  const bool full = p.full != 0;

  if (threadIdx.x == 0) {
    ptx::tma_prefetch_desc(&tm_qkv);
    ptx::tma_prefetch_desc(&tm_do);
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&kv_full[s], 1);
      ptx::mbar_init(&st_full[s], 32);
      ptx::mbar_init(&kv_empty[s], 1);
      ptx::mbar_init(&sd_full[s], 1);
      ptx::mbar_init(&ew_done[2 * s], 4);
      ptx::mbar_init(&ew_done[2 * s + 1], 4);
      ptx::mbar_init(&slot_free[s], 1);
    }
    ptx::mbar_init(acc_full, 1);
    ptx::mbar_init(acc_free, DW ? 4 : 8);
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

  if (warp == 0) {
    // ------------------------------------------------------------------ producer: TMA tiles + row statistics
    int it = 0;
    for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++it) {
      const int st = p.stages == 2 ? (it & 1) : 0;
      const uint32_t ph = p.stages == 2 ? ((it >> 1) & 1) : (it & 1);
      const int b = item / p.H, h = item - b * p.H;
      ptx::mbar_wait(&kv_empty[st], ph ^ 1u);
      uint8_t* base = smem + st * p.stage_bytes;
      if (lane == 0) {
        ptx::mbar_arrive_expect_tx(&kv_full[st], 4 * p.tile_bytes);
        const int row0 = b * p.T;
        ptx::tma_load_2d(base + 0 * p.tile_bytes, &tm_qkv, &kv_full[st], 0 * E + h * HS, row0);
        ptx::tma_load_2d(base + 1 * p.tile_bytes, &tm_qkv, &kv_full[st], 1 * E + h * HS, row0);
        ptx::tma_load_2d(base + 2 * p.tile_bytes, &tm_qkv, &kv_full[st], 2 * E + h * HS, row0);
        ptx::tma_load_2d(base + 3 * p.tile_bytes, &tm_do, &kv_full[st], h * HS, row0);
      }
      float* stat = reinterpret_cast<float*>(base + 4 * p.tile_bytes);
      const float* lrow = lse + static_cast<int64_t>(item) * p.T;
      const float* drow = delta + static_cast<int64_t>(item) * p.T;
      for (int j = lane; j < 256; j += 32) {
        stat[j] = j < p.T ? lrow[j] * kLog2e : INFINITY;        // +inf => p = 0 for padded rows / columns
        stat[256 + j] = j < p.T ? drow[j] * p.scale : 0.f;
      }
      ptx::mbar_arrive(&st_full[st]);
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer 1 (one thread): S / dP of every chunk
    // Two issuing threads in two warps (this one and warp kOutWarp below).  A single thread that alternates between
    // the S/dP batch of chunk g+1 and the output batch of chunk g spends more clocks on its own waits, fences and
    // bookkeeping than the tensor pipe needs for the MMAs (2.9K clk per chunk against ~1.1K of MMA time, measured
    // with the clock64 trace); as two independent state machines the two streams overlap.  The issue path itself is
    // kept lean: descriptor low words are computed once per item / chunk and advanced with one 32-bit add per MMA.
    if (lane == 0) {
      long long* trbuf = (p.trace != nullptr && blockIdx.x == 0) ? p.trace + MODE * 2048 : nullptr;
      int trn = 0;
      constexpr uint32_t kHi = ptx::smem_desc_hi_sw128(1024);
      constexpr uint32_t kIdescS0 = ptx::make_idesc_bf16_f32(QT, 0, 0, 0);
      int it = 0, g = 0;
      for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++it) {
        const int st = p.stages == 2 ? (it & 1) : 0;
        const uint32_t ph = p.stages == 2 ? ((it >> 1) & 1) : (it & 1);
        const uint32_t sq = ptx::smem_u32(smem + st * p.stage_bytes);
        const uint32_t sk = sq + p.tile_bytes, sv = sk + p.tile_bytes, sdo = sv + p.tile_bytes;
        // MODE 0: rows from Q / dO, columns from K / V.   MODE 1: rows from K / V, columns from Q / dO.
        const uint32_t a1 = ptx::smem_desc_lo(MODE == 0 ? sq : sk, 16), b1 = ptx::smem_desc_lo(MODE == 0 ? sk : sq, 16);
        const uint32_t a2 = ptx::smem_desc_lo(MODE == 0 ? sdo : sv, 16), b2 = ptx::smem_desc_lo(MODE == 0 ? sv : sdo, 16);
        VB_TR(7);
        ptx::mbar_wait(&kv_full[st], ph);
        ptx::tc_fence_after();
        VB_TR(8);
        for (int t = 0; t < ntile; ++t) {
          const uint32_t a1t = a1 + static_cast<uint32_t>(t * (QT * 128 / 16)), a2t = a2 + static_cast<uint32_t>(t * (QT * 128 / 16));
          for (int j = 0; j < nchk; ++j, ++g) {
            const int slot = full ? 0 : (g & 1);
            const uint32_t gph = full ? (static_cast<uint32_t>(g) & 1u) : ((static_cast<uint32_t>(g) >> 1) & 1u);
            const int col0 = j * W, wc = min(W, p.TP - col0);
            const uint32_t idesc_s = kIdescS0 | (static_cast<uint32_t>(wc >> 3) << 17);
            const uint32_t tS = tmem_base + static_cast<uint32_t>(slot * 2 * W), tP = tS + static_cast<uint32_t>(W);
            const uint32_t b1c = b1 + static_cast<uint32_t>(col0 * 8), b2c = b2 + static_cast<uint32_t>(col0 * 8);
            VB_TR(1);
            ptx::mbar_wait(&slot_free[slot], gph ^ 1u);
            // full width: the accumulators live INSIDE the slot (columns the packed operands do not need), so the
            // previous unit must have been drained before its slot is overwritten
            if (full) ptx::mbar_wait(acc_free, (static_cast<uint32_t>(g) & 1u) ^ 1u);
            ptx::tc_fence_after();
            VB_TR(2);
#pragma unroll
            for (int k = 0; k < HS / 16; ++k) {   // S and dP interleaved: two independent accumulator chains
              ptx::umma_f16_lh(tS, a1t + 2u * k, b1c + 2u * k, kHi, idesc_s, k > 0 ? 1u : 0u);
              ptx::umma_f16_lh(tP, a2t + 2u * k, b2c + 2u * k, kHi, idesc_s, k > 0 ? 1u : 0u);
            }
            ptx::umma_commit(&sd_full[slot]);
            VB_TR(3);
          }
        }
      }
    }
  } else if (warp == kOutWarp) {
    // ------------------------------------------------------------------ MMA issuer 2 (one thread): output MMAs
    if (lane == 0) {
      long long* trbuf = (p.trace != nullptr && blockIdx.x == 0) ? p.trace + MODE * 2048 + 1536 : nullptr;
      int trn = 0;
      constexpr uint32_t kHi = ptx::smem_desc_hi_sw128(1024);
      const uint32_t idesc_o = ptx::make_idesc_bf16_f32(QT, HS, 0, 1);
      const uint32_t acc_odd = tmem_base + (dual ? kAcc1 : kAcc0);
      int it = 0, u = 0, g = 0;
      for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++it) {
        const int st = p.stages == 2 ? (it & 1) : 0;
        const uint32_t sq = ptx::smem_u32(smem + st * p.stage_bytes);
        const uint32_t sk = sq + p.tile_bytes, sdo = sk + 2 * p.tile_bytes;
        // B operands (MN-major, 16 rows = 2048 B per k-step): MODE 0 K | MODE 1 dO and Q
        const uint32_t o0 = ptx::smem_desc_lo(MODE == 0 ? sk : sdo, 64 * 128), o1 = ptx::smem_desc_lo(sq, 64 * 128);
        for (int t = 0; t < ntile; ++t, ++u) {
          for (int j = 0; j < nchk; ++j, ++g) {
            const int slot = full ? 0 : (g & 1);
            const uint32_t gph = full ? (static_cast<uint32_t>(g) & 1u) : ((static_cast<uint32_t>(g) >> 1) & 1u);
            const int col0 = j * W, nk = min(W, p.TP - col0) >> 4;
            const uint32_t tS = tmem_base + static_cast<uint32_t>(slot * 2 * W), tP = tS + static_cast<uint32_t>(W);
            const uint32_t b0 = o0 + static_cast<uint32_t>(col0 * 8), b1 = o1 + static_cast<uint32_t>(col0 * 8);
            const uint32_t keep = j == 0 ? 0u : 1u;                    // first chunk of a unit overwrites the accumulators
            // bf16 operand of k-step k: columns 8k.. of the region for the first group's k-steps, shifted by hc/2 for
            // the second group's (it packs in place from column hc on)
            const int kh = bwd_split(nk << 4, p.hc) >> 4;
            const uint32_t sh = static_cast<uint32_t>(kh * 8);
            VB_TR(4);
            ptx::mbar_wait(&ew_done[2 * slot], gph);
            VB_TR(5);
            if (j == 0) ptx::mbar_wait(acc_free, (static_cast<uint32_t>(u) & 1u) ^ 1u);
            ptx::tc_fence_after();
            // Back-to-back MMAs into ONE accumulator serialise on its latency (~114 clk each, measured); two
            // independent accumulators interleaved run at ~55 clk each.  MODE 0 therefore splits dQ over two
            // accumulators by k-step parity (summed in the drain); MODE 1 alternates dV and dK.
            const uint32_t keep1 = dual ? keep : 1u;
#pragma unroll
            for (int k = 0; k < 16; ++k) {
              if (k < nk) {
                if (k == kh) {
                  ptx::mbar_wait(&ew_done[2 * slot + 1], gph);
                  ptx::tc_fence_after();
                }
                const uint32_t ao = static_cast<uint32_t>(k * 8) + (k >= kh ? sh : 0u);
                if (MODE == 0) {
                  ptx::umma_f16_ts_lh((k & 1) ? acc_odd : tmem_base + kAcc0, tS + ao,               // dQ(k parity) += dS K
                                      b0 + static_cast<uint32_t>(k * 128), kHi, idesc_o, k == 0 ? keep : (k == 1 ? keep1 : 1u));
                } else {
                  ptx::umma_f16_ts_lh(tmem_base + kAcc0, tS + ao, b0 + static_cast<uint32_t>(k * 128), kHi, idesc_o,
                                      k == 0 ? keep : 1u);                                         // dV += P^T dO
                  ptx::umma_f16_ts_lh(tmem_base + kAcc1, tP + ao, b1 + static_cast<uint32_t>(k * 128), kHi, idesc_o,
                                      k == 0 ? keep : 1u);                                         // dK += dS^T Q
                }
              }
            }
            if (nk <= kh) ptx::mbar_wait(&ew_done[2 * slot + 1], gph);   // one-k-step chunk: keep the barrier's phase in step
            ptx::umma_commit(&slot_free[slot]);
            VB_TR(6);
            if (j == nchk - 1) {
              ptx::umma_commit(acc_full);
              if (t == ntile - 1) ptx::umma_commit(&kv_empty[st]);
            }
          }
        }
      }
    }
  } else if (DW && warp > kOutWarp) {
    // ------------------------------------------------------------------ drain warps (DW): accumulators -> bf16 rows of dqkv
    const int q = warp & 3;
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const int my_items = blockIdx.x < p.items ? (p.items - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const int total_units = my_items * ntile;
    for (int u = 0; u < total_units; ++u) {
      const int it_u = u / ntile, t = u - it_u * ntile;
      const int item = blockIdx.x + it_u * gridDim.x;
      const int b = item / p.H, h = item - b * p.H;
      const int row = t * QT + q * 32 + lane;
      const bool live = row < p.T && !(p.dbg & 2);
      bf16* grow = dqkv + (static_cast<int64_t>(b) * p.T + (live ? row : 0)) * (3 * static_cast<int64_t>(E)) + h * HS;
      ptx::mbar_wait(acc_full, static_cast<uint32_t>(u) & 1u);
      ptx::tc_fence_after();
#pragma unroll
      for (int half = 0; half < 2; ++half) {     // 32 columns of each accumulator at a time (register budget)
        uint32_t r0[32], r1[32];
        ptx::tmem_ld_32x32(lane_base + kAcc0 + static_cast<uint32_t>(half * 32), r0);
        if (MODE == 1 || dual) ptx::tmem_ld_32x32(lane_base + kAcc1 + static_cast<uint32_t>(half * 32), r1);
        ptx::tmem_ld_wait();
        if (half == 1) {                         // everything is in registers: the next unit may overwrite the accumulators
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(acc_free);
        }
        if (MODE == 0 && dual) {
#pragma unroll
          for (int j = 0; j < 32; ++j) r0[j] = __float_as_uint(__uint_as_float(r0[j]) + __uint_as_float(r1[j]));
        }
        auto put = [&](bf16* dst, const uint32_t (&r)[32]) {     // 32 columns = 64 contiguous bytes of my row
          uint32_t w[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) w[j] = pack_bf16x2(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
          if (p.st256) {
            const uint32_t lo[8] = {w[0], w[1], w[2], w[3], w[4], w[5], w[6], w[7]};
            const uint32_t hi[8] = {w[8], w[9], w[10], w[11], w[12], w[13], w[14], w[15]};
            ptx::stg256(dst, lo);
            ptx::stg256(dst + 16, hi);
          } else {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              *reinterpret_cast<uint4*>(dst + j * 8) = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
          }
        };
        if (live) {
          // MODE 0: dQ -> section 0.  MODE 1: accumulator 0 = dV -> section 2, accumulator 1 = dK -> section 1.
          put(grow + (MODE == 0 ? 0 : 2) * static_cast<int64_t>(E) + half * 32, r0);
          if (MODE == 1) put(grow + static_cast<int64_t>(E) + half * 32, r1);
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ element-wise warps: one thread per row
    const int q = warp & 3;                      // tensor-memory lane quarter
    const int grp = (warp - 2) >> 2;             // slot / chunk parity owned by this warp; half of the drain columns
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const float sl2 = p.scale_log2, scale = p.scale;
    const int my_items = blockIdx.x < p.items ? (p.items - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const int total_units = my_items * ntile;
    int next_drain = 0;
    long long* trbuf = (p.trace != nullptr && blockIdx.x == 0 && q == 0 && lane == 0) ? p.trace + MODE * 2048 + 512 * (1 + grp) : nullptr;
    int trn = 0;

    auto drain = [&](int u) {
      const int it_u = u / ntile, t = u - it_u * ntile;
      const int item = blockIdx.x + it_u * gridDim.x;
      const int b = item / p.H, h = item - b * p.H;
      const int row = t * QT + q * 32 + lane;
      VB_TR(14);
      ptx::mbar_wait(acc_full, static_cast<uint32_t>(u) & 1u);
      ptx::tc_fence_after();
      VB_TR(15);
      uint32_t r0[32], r1[32];
      ptx::tmem_ld_32x32(lane_base + kAcc0 + static_cast<uint32_t>(grp * 32), r0);
      if (MODE == 1 || dual) ptx::tmem_ld_32x32(lane_base + kAcc1 + static_cast<uint32_t>(grp * 32), r1);
      ptx::tmem_ld_wait();
      if (MODE == 0 && dual) {
#pragma unroll
        for (int j = 0; j < 32; ++j) r0[j] = __float_as_uint(__uint_as_float(r0[j]) + __uint_as_float(r1[j]));
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(acc_free);
      if (row < p.T && !(p.dbg & 2)) {
        bf16* grow = dqkv + (static_cast<int64_t>(b) * p.T + row) * (3 * static_cast<int64_t>(E)) + h * HS + grp * 32;
        // MODE 0: dQ -> section 0.  MODE 1: accumulator 0 = dV -> section 2, accumulator 1 = dK -> section 1.
        bf16* d0 = grow + (MODE == 0 ? 0 : 2) * static_cast<int64_t>(E);
        auto put = [&](bf16* dst, const uint32_t (&r)[32]) {     // my 32 columns = 64 contiguous bytes of the row
          uint32_t w[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) w[j] = pack_bf16x2(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
          if (p.st256) {
            const uint32_t lo[8] = {w[0], w[1], w[2], w[3], w[4], w[5], w[6], w[7]};
            const uint32_t hi[8] = {w[8], w[9], w[10], w[11], w[12], w[13], w[14], w[15]};
            ptx::stg256(dst, lo);
            ptx::stg256(dst + 16, hi);
          } else {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              *reinterpret_cast<uint4*>(dst + j * 8) = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
          }
        };
        put(d0, r0);
        if (MODE == 1) put(grow + static_cast<int64_t>(E), r1);
      }
    };

    int it = 0, u = 0, g = 0;
    for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++it) {
      const int st = p.stages == 2 ? (it & 1) : 0;
      const uint32_t ph = p.stages == 2 ? ((it >> 1) & 1) : (it & 1);
      const float* stat = reinterpret_cast<const float*>(smem + st * p.stage_bytes + 4 * p.tile_bytes);
      const uint32_t stat_s = ptx::smem_u32(stat);
      ptx::mbar_wait(&st_full[st], ph);
      for (int t = 0; t < ntile; ++t, ++u) {
        const int row = t * QT + q * 32 + lane;  // query (MODE 0) or key (MODE 1) index of this thread
        float lse2r = 0.f, dscr = 0.f;
        if (MODE == 0) { lse2r = stat[row]; dscr = stat[256 + row]; }
        for (int j = 0; j < nchk; ++j, ++g) {
          // BOTH groups work on every chunk: group 0 takes columns [0, hc), group 1 [hc, wc).  That halves the
          // S/dP -> operand latency of a chunk, which is what bounds the slot cycle (S/dP MMAs -> element-wise ->
          // output MMAs -> slot free).  Each group packs its bf16 results in place at the start of ITS OWN column
          // range (ascending, so it only overwrites what it has already read); the output-MMA issuer knows the layout.
          const int slot = full ? 0 : (g & 1);
          const int col0 = j * W, wc = min(W, p.TP - col0);
          const int hc = bwd_split(wc, p.hc);
          const int cbeg = grp == 0 ? 0 : hc, cend = grp == 0 ? hc : wc;
          const uint32_t tS = lane_base + static_cast<uint32_t>(slot * 2 * W), tP = tS + static_cast<uint32_t>(W);
          VB_TR(10);
          ptx::mbar_wait(&sd_full[slot], full ? (static_cast<uint32_t>(g) & 1u) : ((static_cast<uint32_t>(g) >> 1) & 1u));
          ptx::tc_fence_after();
          VB_TR(11);
          int c0 = cbeg;
          if (p.dbg & 1) c0 = cend;               // experiment: no element-wise work
          for (; c0 + 32 <= cend; c0 += 32)
            bwd_piece<MODE, 32>(tS, tP, c0, cbeg + ((c0 - cbeg) >> 1), col0 + c0, p.T, sl2, scale, lse2r, dscr, stat_s);
          if (c0 < cend)
            bwd_piece<MODE, 16>(tS, tP, c0, cbeg + ((c0 - cbeg) >> 1), col0 + c0, p.T, sl2, scale, lse2r, dscr, stat_s);
          ptx::tmem_st_wait();
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(&ew_done[2 * slot + grp]);
          VB_TR(12);
          // Drain a unit's accumulators one chunk LATE (after the first chunk of the next unit): right after the unit's
          // own last chunk its output MMAs have only just been issued and both groups would sit on acc_full.
          // (full width: at once — the next unit's S/dP cannot be issued before this unit's accumulators are drained)
          if constexpr (!DW) {
            while (next_drain < total_units && (next_drain + 1) * nchk - 1 < g + (full ? 1 : 0)) drain(next_drain++);
          }
        }
      }
    }
    if constexpr (!DW) {
      while (next_drain < total_units) drain(next_drain++);
    }
    VB_TR(16);
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

bool attn_tc_supported(int T, int hs) { return hs == HS && T >= 1 && T <= 256; }

int attn_tc_forward(const bf16* qkv, bf16* out, float* lse, int B, int T, int H, int hs, float scale,
                    cudaStream_t stream) {
  if (!attn_tc_supported(T, hs)) {
    set_error("attention (tcgen05): needs head_size 64 and T <= 256 (got hs=%d T=%d)", hs, T);
    return kErrUnsupported;
  }
  VB_REQUIRE(scale > 0.f, "attention: scale must be positive");
  VB_REQUIRE(aligned16(qkv) && aligned16(out), "attention: qkv/out must be 16-byte aligned");
  const int E = H * hs;
  const int64_t M = static_cast<int64_t>(B) * T;
  CUtensorMap tm;
  VB_PROPAGATE(make_tmap_bf16(&tm, qkv, M, 3 * static_cast<int64_t>(E), 3 * static_cast<int64_t>(E), HS, QT));
  static bool attr_done = false;
  if (!attr_done) {
    VB_CHECK_CUDA(cudaFuncSetAttribute(attn_tc_fwd_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    VB_CHECK_CUDA(cudaFuncSetAttribute(attn_tc_fwd_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES2));
    attr_done = true;
  }
  FwdParams p;
  p.T = T; p.TP = (T + 15) / 16 * 16; p.H = H; p.items = B * H;
  p.scale_log2 = scale * kLog2e;
  p.st256 = (reinterpret_cast<uintptr_t>(out) & 31u) == 0 && (E * 2) % 32 == 0;
  static const int order = getenv("VITB200_ATTN_FWD_ORDER") ? atoi(getenv("VITB200_ATTN_FWD_ORDER")) : 1;
  p.order = order;
  p.trace = getenv("VITB200_ATTN_TRACE") ? reinterpret_cast<long long*>(strtoull(getenv("VITB200_ATTN_TRACE"), nullptr, 0)) : nullptr;
  ProfScope prof(kProfAttnFwd, 4.0 * B * H * T * static_cast<double>(T) * hs, 1, stream);
  const int grid = p.items < num_sms() ? p.items : num_sms();
  // two softmax groups per tile when there are 7 or 8 chunks of 32 key columns (T in 193..256); VITB200_ATTN_FWD_SPLIT=0/1
  static const int split_env = getenv("VITB200_ATTN_FWD_SPLIT") ? atoi(getenv("VITB200_ATTN_FWD_SPLIT")) : 1;
  if (split_env && p.TP > 192)
    VB_CHECK_CUDA(launch_pdl(attn_tc_fwd_kernel<2>, dim3(grid), dim3(kThreads2), SMEM_BYTES2, stream, 1, tm, out, lse, p));
  else
    VB_CHECK_CUDA(launch_pdl(attn_tc_fwd_kernel<1>, dim3(grid), dim3(kThreads), SMEM_BYTES, stream, 1, tm, out, lse, p));
  VB_CHECK_LAUNCH();
  return kOk;
}


int attn_tc_backward(const bf16* qkv, const bf16* d_o, const float* lse, const float* delta, bf16* dqkv, int B, int T,
                     int H, int hs, float scale, cudaStream_t stream) {
  if (!attn_tc_supported(T, hs)) {
    set_error("attention (tcgen05): needs head_size 64 and T <= 256 (got hs=%d T=%d)", hs, T);
    return kErrUnsupported;
  }
  VB_REQUIRE(scale > 0.f, "attention: scale must be positive");
  VB_REQUIRE(aligned16(qkv) && aligned16(d_o) && aligned16(dqkv), "attention: operands must be 16-byte aligned");
  const int E = H * hs;
  const int64_t M = static_cast<int64_t>(B) * T;
  BwdParams p;
  p.T = T; p.TP = (T + 15) / 16 * 16; p.H = H; p.items = B * H;
  p.tile_bytes = p.TP * 128;
  p.stage_bytes = 4 * p.tile_bytes + kStatBytes;
  const int pad = 32768 - p.tile_bytes;      // an A tile of rows [128,256) may over-read past a TP-row operand
  p.stages = (2 * p.stage_bytes + pad + 256 + 1024 <= 227 * 1024) ? 2 : 1;
  p.scale = scale; p.scale_log2 = scale * kLog2e;
  p.st256 = (reinterpret_cast<uintptr_t>(dqkv) & 31u) == 0 && (3 * E * 2) % 32 == 0;
  static const int dbg = getenv("VITB200_ATTN_DBG") ? atoi(getenv("VITB200_ATTN_DBG")) : 0;
  p.dbg = dbg;
  p.trace = getenv("VITB200_ATTN_TRACE") ? reinterpret_cast<long long*>(strtoull(getenv("VITB200_ATTN_TRACE"), nullptr, 0)) : nullptr;
  const int smem = p.stages * p.stage_bytes + pad + 256 + 1024;
  // chunk width: two [S | dP] slots of 2W columns + the accumulators must fit 512 tensor-memory columns
  auto chunking = [&](int wmax) {
    p.nchk = ceil_div(p.TP, wmax);
    p.W = ceil_div(ceil_div(p.TP, p.nchk), 16) * 16;
  };
  CUtensorMap tq, td;
  VB_PROPAGATE(make_tmap_bf16(&tq, qkv, M, 3 * static_cast<int64_t>(E), 3 * static_cast<int64_t>(E), HS, p.TP));
  VB_PROPAGATE(make_tmap_bf16(&td, d_o, M, E, E, HS, p.TP));
  VB_CHECK_CUDA(cudaFuncSetAttribute(attn_tc_bwd_kernel<0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  VB_CHECK_CUDA(cudaFuncSetAttribute(attn_tc_bwd_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  VB_CHECK_CUDA(cudaFuncSetAttribute(attn_tc_bwd_kernel<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  VB_CHECK_CUDA(cudaFuncSetAttribute(attn_tc_bwd_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  // VITB200_ATTN_BWD_DRAIN: bit 0 = dQ kernel, bit 1 = dK/dV kernel run with dedicated drain warps.  Measured at the
  // ViT-B/16 layer shape (profiles/r02_attn_bwd_drain_warps_ab.txt): 342.1 us with 0, 327.1 with 1, 332.4 with 2, 317.4-317.8 with 3.
  static const int drain_mask = getenv("VITB200_ATTN_BWD_DRAIN") ? atoi(getenv("VITB200_ATTN_BWD_DRAIN")) : 3;
  ProfScope prof(kProfAttnBwd, 8.0 * B * H * T * static_cast<double>(T) * hs, 2, stream);
  const int grid = p.items < num_sms() ? p.items : num_sms();
  // Two tensor-memory plans.  CHUNKED (round 1): chunks of W <= 96 columns flow through two [S | dP] slots, accumulators
  // at columns 384 / 448 (4W + 128 <= 512).  FULL WIDTH (round 2): ONE chunk of all TP columns per unit in ONE slot
  // (S at 0, dP at TP), the accumulators inside columns that are dead by the time the output MMAs start.  These
  // kernels are bound by the NUMBER of MMAs (each 128 x N x 16 instruction costs 55-110 clk whatever N <= 112 is:
  // profiles/r02_attn_bwd_bound_experiments.txt — without any element-wise work the two kernels still take 190 of
  // 290 us), and full width needs 8 instead of 24 S/dP MMAs per 128-row tile at T = 197, at the price of no overlap
  // between a unit's element-wise pass and the next unit's MMAs.  VITB200_ATTN_BWD_FULL: bit 0 = dQ kernel, bit 1 =
  // dK/dV kernel (A/B switch; default set from the measurement).
  static const int full_mask = getenv("VITB200_ATTN_BWD_FULL") ? atoi(getenv("VITB200_ATTN_BWD_FULL")) : 0;
  const bool can_full = p.TP >= 144 && p.TP <= 224;
  auto plan = [&](int mode) {
    if (can_full && ((full_mask >> mode) & 1)) {
      p.full = 1; p.nchk = 1; p.W = p.TP;
      if (mode == 0) { p.hc = 0; p.acc0 = 2 * p.TP; p.acc1 = p.TP; }      // dQ halves: spare columns | group 0's dP columns
      else { p.hc = 128; p.acc0 = 64; p.acc1 = p.TP + 64; }              // dV | dK: the unpacked upper half of group 0's columns
    } else {
      p.full = 0; p.hc = 0; p.acc0 = 384; p.acc1 = 448;
      chunking(96);    // two slots of 2W columns + two 64-column accumulators: 4W + 128 <= 512
    }
  };
  plan(1);
  if ((drain_mask & 2) && !p.full)
    VB_CHECK_CUDA(launch_pdl(attn_tc_bwd_kernel<1, true>, dim3(grid), dim3(kBwdThreadsDW), smem, stream, 1, tq, td, lse, delta, dqkv, p));
  else
    VB_CHECK_CUDA(launch_pdl(attn_tc_bwd_kernel<1, false>, dim3(grid), dim3(kBwdThreads), smem, stream, 1, tq, td, lse, delta, dqkv, p));
  VB_CHECK_LAUNCH();
  plan(0);
  if ((drain_mask & 1) && !p.full)
    VB_CHECK_CUDA(launch_pdl(attn_tc_bwd_kernel<0, true>, dim3(grid), dim3(kBwdThreadsDW), smem, stream, 1, tq, td, lse, delta, dqkv, p));
  else
    VB_CHECK_CUDA(launch_pdl(attn_tc_bwd_kernel<0, false>, dim3(grid), dim3(kBwdThreads), smem, stream, 1, tq, td, lse, delta, dqkv, p));
  VB_CHECK_LAUNCH();
  return kOk;
}

}  // namespace vb
