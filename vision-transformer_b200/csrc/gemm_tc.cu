// bf16 x bf16 -> fp32 contraction on the 5th-gen tensor cores (tcgen05.mma, accumulators in TMEM),
// operands staged by TMA (128B swizzle) through a 4-deep mbarrier ring, persistent over output tiles,
// warp-specialised: warp 0 = TMA producer, warp 1 = MMA issuer (+TMEM allocator), warps 2-5 = epilogue.
// Two TMEM accumulator stages (2 x BN columns) let the epilogue of tile i overlap the mainloop of tile i+1.
//
// Replaces the cuBLAS calls behind nn.Linear at reference transformer.py:22-28,46,56-58 and the patch
// Conv2d at vit.py:21-28 (as an im2col GEMM), plus their autograd backward GEMMs.
#include <cuda.h>
#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "gemm_epilogue.cuh"
#include "ptx.cuh"

namespace vb {

namespace {

constexpr int BM = 128;          // UMMA M (cta_group::1)
constexpr int BK = 64;           // 64 bf16 = 128 B = one swizzle row
constexpr int UMMA_K = 16;
// Epilogue warps per CTA: EW / 4 per TMEM lane quarter, each draining BN * 4 / EW accumulator columns.  8 everywhere.
// 16 warps for the 2-byte-output CTA-pair call sites (64 columns each, <= 96 registers per thread, one ring stage
// traded for the extra transpose buffers) was built and measured on B200: fc1+bits 200 -> 194 us, but qkv / dh / dln2
// 3-6 us slower each (5-stage ring), per-block total unchanged (1.703 vs 1.706 ms) — kept behind
// -DVITB200_EPI_WARPS_LP=16.
#ifndef VITB200_EPI_WARPS_LP
#define VITB200_EPI_WARPS_LP 8
#endif
template <typename OutT, int CG>
constexpr int epi_warps() { return (sizeof(OutT) == 2 && CG == 2) ? VITB200_EPI_WARPS_LP : 8; }
template <typename OutT, int CG>
constexpr int gemm_threads() { return 64 + epi_warps<OutT, CG>() * 32; }

// CG = 1: one CTA owns a 128 x BN tile.  CG = 2: a CTA PAIR (cluster of 2, tcgen05 cta_group::2) owns a 256 x BN tile;
// each CTA stages its own 128 rows of A and HALF of the B tile, the leader issues M = 256 MMAs that read both halves,
// so every SM pulls a third less operand traffic through L2 and shared memory per output element.  CG = 4: a cluster
// of TWO pairs works on two vertically adjacent 256 x BN tiles; they need the same B tile, so each CTA fetches only a
// QUARTER of it and TMA-multicasts that quarter to its counterpart in the other pair (24 KB instead of 32 KB per CTA
// and k-block: the K-major call sites are bound by the ~43 B/clk/SM L2->SM feed, not by the tensor pipe).
template <int BN, int CG, int EW>
struct TileCfg {
  static constexpr int STAGES = CG >= 2 ? (EW > 8 ? 5 : 6) : 4;   // 16 transpose buffers cost one ring stage
  static constexpr int A_BYTES = BM * BK * 2;            // 16 KB
  static constexpr int B_BYTES = BN * BK * 2 / (CG >= 2 ? 2 : 1);   // 32 KB @ BN=256 (16 KB per CTA of a pair)
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int BAR_BYTES = 256;
  static constexpr int STG_BYTES = EW * 32 * 128;   // per epilogue warp: 32 rows x 32 fp32 transpose buffer
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + BAR_BYTES + STG_BYTES + 1024;  // +1024 alignment slack
  static constexpr int TMEM_COLS = 2 * BN;
};

struct TcParams {
  int M, N, K;
  int m_tiles, n_tiles;
  int splits, kb_per_split, kb_total;
  int flags;   // bit 0: epilogue prefetches its read-side operands one tile ahead
};

// EXACT: M is a multiple of the tile rows (128 * CG) and N of BN — the epilogue carries no bounds tests, so its four row
// passes per chunk are one straight-line region the compiler can interleave (loads first, stores last).
template <int AMAJ, int BMAJ, int BN, typename OutT, uint32_t EPI, int CG, bool EXACT>
__global__ void __launch_bounds__((gemm_threads<OutT, CG>()), 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
               const TcParams p, const GemmEpilogue ep) {
  constexpr int EW = epi_warps<OutT, CG>();
  constexpr int CW = BN * 4 / EW;          // accumulator columns drained by one epilogue warp
  using Cfg = TileCfg<BN, CG, EW>;
  constexpr int kStages = Cfg::STAGES;
  // pair rank / work-unit walk: a unit is one (tile, split); CTAs of a pair walk the same units
  const uint32_t rank = CG >= 2 ? ptx::cluster_ctarank() : 0u;   // rank in the cluster: pair = rank >> 1, half = rank & 1
  const uint32_t pair = rank >> 1, half_id = rank & 1u;
  const int walker = static_cast<int>(blockIdx.x) / CG;
  const int nwalkers = static_cast<int>(gridDim.x) / CG;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kStages * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full = empty_bar + kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int tiles = p.m_tiles * p.n_tiles;
  const int total_work = tiles * p.splits;

  if (threadIdx.x == 0) {
    ptx::tma_prefetch_desc(&tma_a);
    ptx::tma_prefetch_desc(&tma_b);
    for (int s = 0; s < kStages; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], CG == 4 ? 2 : 1);   // CG 4: both pairs read what lands in this slot
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&tmem_full[s], 1);
      ptx::mbar_init(&tmem_empty[s], EW * (CG >= 2 ? 2 : 1));   // pair: the leader collects both CTAs' epilogue warps
    }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    if constexpr (CG >= 2) {
      ptx::tmem_alloc_pair(tmem_ptr, Cfg::TMEM_COLS);
      ptx::tmem_relinquish_pair();
    } else {
      ptx::tmem_alloc(tmem_ptr, Cfg::TMEM_COLS);
      ptx::tmem_relinquish();
    }
  }
  ptx::tc_fence_before();
  if constexpr (CG >= 2) ptx::cluster_sync(); else __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_wait();      // everything above overlapped the previous kernel's tail; global memory is touched only below
  pdl_trigger();

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int w = walker; w < total_work; w += nwalkers) {
        const int split = w / tiles;
        const int tile = w - split * tiles;
        const int m_blk = tile / p.n_tiles;
        const int n_blk = tile - m_blk * p.n_tiles;
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(p.kb_total, kb0 + p.kb_per_split);
        const int m_row = (m_blk * CG + static_cast<int>(rank)) * BM;          // my 128 rows of A
        // my share of the B tile: the whole tile, my half, or (CG 4) the quarter I fetch for both pairs
        const int n_row = n_blk * BN + (CG >= 2 ? static_cast<int>(half_id) * (BN / 2) : 0) +
                          (CG == 4 ? static_cast<int>(pair) * (BN / 4) : 0);
        const uint16_t mc_mask = static_cast<uint16_t>((1u << half_id) | (1u << (2u + half_id)));
        for (int kb = kb0; kb < kb1; ++kb) {
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1u);
          // pair: the leader's barrier expects the bytes of BOTH CTAs; the peer's copies are credited to it
          if (CG == 1 || half_id == 0) ptx::mbar_arrive_expect_tx(&full_bar[stage], (CG >= 2 ? 2 : 1) * Cfg::STAGE_BYTES);
          uint8_t* sA = smem + stage * Cfg::STAGE_BYTES;
          uint8_t* sB = sA + Cfg::A_BYTES;
          auto load = [&](uint8_t* dst, const CUtensorMap* tm, int c0, int c1) {
            if constexpr (CG >= 2) ptx::tma_load_2d_pair(dst, tm, &full_bar[stage], c0, c1);
            else ptx::tma_load_2d(dst, tm, &full_bar[stage], c0, c1);
          };
          auto load_b = [&](uint8_t* dst, int c0, int c1) {
            if constexpr (CG == 4) ptx::tma_load_2d_pair_mc(dst, &tma_b, &full_bar[stage], c0, c1, mc_mask);
            else load(dst, &tma_b, c0, c1);
          };
          if constexpr (AMAJ == 0) {
            load(sA, &tma_a, kb * BK, m_row);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j) load(sA + j * (BK * 128), &tma_a, m_row + j * 64, kb * BK);
          }
          // CG 4: my quarter goes to offset pair * 8 KB of the half tile (64 K-major rows, or one 64-wide MN slab)
          uint8_t* sBq = sB + (CG == 4 ? pair * (Cfg::B_BYTES / 2) : 0u);
          if constexpr (BMAJ == 0) {
            load_b(sBq, kb * BK, n_row);
          } else {
#pragma unroll
            for (int j = 0; j < BN / CG / 64; ++j) load_b(sBq + j * (BK * 128), n_row + j * 64, kb * BK);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (one thread)
    if (lane == 0 && half_id == 0) {
      constexpr uint32_t idesc = ptx::make_idesc_bf16_f32(BM * (CG >= 2 ? 2 : 1), BN, AMAJ, BMAJ);
      const uint16_t pair_mask = static_cast<uint16_t>(3u << (2u * pair));   // both CTAs of my pair
      const uint16_t slot_mask = CG == 4 ? uint16_t(0xF) : pair_mask;          // everyone whose loads land in the slot
      // K-major: 8-row groups 1024 B apart, k-step = 32 B inside the swizzle row.
      // MN-major: 64-wide MN slabs BK*128 B apart (LBO), 8 k-rows = 1024 B (SBO), k-step = 16 rows * 128 B.
      constexpr uint32_t a_lbo = AMAJ == 0 ? 16 : BK * 128, a_sbo = 1024, a_kstep = AMAJ == 0 ? 32 : 2048;
      constexpr uint32_t b_lbo = BMAJ == 0 ? 16 : BK * 128, b_sbo = 1024, b_kstep = BMAJ == 0 ? 32 : 2048;
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int w = walker; w < total_work; w += nwalkers) {
        const int split = w / tiles;
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(p.kb_total, kb0 + p.kb_per_split);
        ptx::mbar_wait(&tmem_empty[acc], acc_phase ^ 1u);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * BN);
        for (int kb = kb0; kb < kb1; ++kb) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after();
          const uint32_t a_addr = ptx::smem_u32(smem + stage * Cfg::STAGE_BYTES);
          const uint32_t b_addr = a_addr + Cfg::A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t da = ptx::make_smem_desc_sw128(a_addr + k * a_kstep, a_lbo, a_sbo);
            const uint64_t db = ptx::make_smem_desc_sw128(b_addr + k * b_kstep, b_lbo, b_sbo);
            if constexpr (CG >= 2) ptx::umma_f16_pair(d_tmem, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            else ptx::umma_f16(d_tmem, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          // smem slot reusable once these MMAs retire (pair: in both CTAs)
          if constexpr (CG >= 2) ptx::umma_commit_pair(&empty_bar[stage], slot_mask); else ptx::umma_commit(&empty_bar[stage]);
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
        // accumulator complete -> epilogue (pair: each CTA drains its own 128 lanes)
        if constexpr (CG >= 2) ptx::umma_commit_pair(&tmem_full[acc], pair_mask); else ptx::umma_commit(&tmem_full[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (EW warps)
    // Warp (q, half) drains TMEM lanes [32q, 32q+32) x accumulator columns [half*CW, (half+1)*CW) in chunks of
    // 32 columns: tcgen05.ld gives each thread ONE row, a swizzled warp-private shared-memory transpose turns that
    // into "4 lanes cover 32 consecutive columns of a row", so residual / mask loads and the C stores are coalesced.
    const int ew = warp - 2;
    const int q = warp & 3;              // TMEM lane quarter this warp may read
    const int half = ew >> 2;
    const uint32_t stg_s = ptx::smem_u32(smem + kStages * Cfg::STAGE_BYTES + Cfg::BAR_BYTES) + static_cast<uint32_t>(ew * 4096);
    const int rsub = lane >> 2;          // row within a pass of 8 rows
    const int cg = lane & 3;             // 8-column group within the chunk
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int w = walker; w < total_work; w += nwalkers) {
      const int split = w / tiles;
      const int tile = w - split * tiles;
      const int m_blk = tile / p.n_tiles;
      const int n_blk = tile - m_blk * p.n_tiles;
      const int m0 = (m_blk * CG + static_cast<int>(rank)) * BM + q * 32;
      // Pull the epilogue's read-side operands (residual rows, dReLU mask rows) of my 32 x (BN/2) sub-tile into L2, one
      // row segment per lane: for THIS tile while its mainloop may still be running and, when the epilogue is the
      // slower engine (short K: the mainloop is done long before), for the walker's NEXT tile (p.flags bit 0).
      auto prefetch_unit = [&](int wu) {
        const int tl = wu - (wu / tiles) * tiles;
        const int mb = tl / p.n_tiles;
        const int m = (mb * CG + static_cast<int>(rank)) * BM + q * 32 + lane;
        const int nb = (tl - mb * p.n_tiles) * BN + half * CW;
        if (EXACT || (m < p.M && nb < p.N)) {
          const int ncols = EXACT ? CW : min(CW, p.N - nb);
          const int64_t orow = VB_EPI(kEpiRowMap, ep.row_period > 0) ? epi_out_row(ep, m) : static_cast<int64_t>(m);
          const int64_t off = orow * ep.ldc + nb;
          // bit 1: one prefetch.global.L2 per 128-byte line and lane (the bulk form is a uniform-datapath instruction:
          // the compiler serialises it into a 32-iteration loop over the lanes, ~10 % of the epilogue's instructions)
          if (VB_EPI(kEpiResidual, ep.residual != nullptr)) {
            if (p.flags & 2) {
              for (int b = 0; b < ncols * 4; b += 128) ptx::prefetch_l2_line(reinterpret_cast<const char*>(ep.residual + off) + b);
            } else {
              ptx::prefetch_l2_bulk(ep.residual + off, static_cast<uint32_t>(ncols) * 4u);
            }
          }
          if (VB_EPI(kEpiMask | kEpiDot, ep.mask_aux != nullptr || ep.dot_out != nullptr)) {
            const OutT* aux = reinterpret_cast<const OutT*>(VB_EPI(kEpiDot, ep.dot_out != nullptr) ? ep.dot_aux : ep.mask_aux) + off;
            if (p.flags & 2) {
              for (int b = 0; b < ncols * static_cast<int>(sizeof(OutT)); b += 128)
                ptx::prefetch_l2_line(reinterpret_cast<const char*>(aux) + b);
            } else {
              ptx::prefetch_l2_bulk(aux, static_cast<uint32_t>(ncols) * static_cast<uint32_t>(sizeof(OutT)));
            }
          }
        }
      };
      if (VB_EPI(kEpiResidual | kEpiMask | kEpiDot, ep.residual != nullptr || ep.mask_aux != nullptr || ep.dot_out != nullptr)) {
        if (!(p.flags & 1) || w == walker) prefetch_unit(w);
        if ((p.flags & 1) && w + nwalkers < total_work) prefetch_unit(w + nwalkers);
      }
      ptx::mbar_wait(&tmem_full[acc], acc_phase);
      ptx::tc_fence_after();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(acc * BN);
      const int nb = n_blk * BN + half * CW;     // first column of my share
      constexpr int NCH = CW / 32;
      uint32_t r[32];
      if (EXACT || nb < p.N) ptx::tmem_ld_32x32(t_row + static_cast<uint32_t>(half * CW), r);
      // Read-side operands (residual / dReLU mask) are fetched ONE CHUNK AHEAD into registers: their L2 latency is
      // covered by the previous chunk's transpose + stores instead of stalling every chunk.
      EpiPre pre_nx[4];
      float dacc[4] = {0.f, 0.f, 0.f, 0.f};   // kEpiDot: per-row partial sums over the two 32-column chunks of a 64-wide head
      auto fetch_chunk = [&](int cc, EpiPre (&dst)[4]) {
        const int nn = nb + cc * 32 + cg * 8;
        if (!EXACT && nn >= p.N) return;
#pragma unroll
        for (int pass = 0; pass < 4; ++pass) {
          const int m = m0 + pass * 8 + rsub;
          const int64_t orow = VB_EPI(kEpiRowMap, ep.row_period > 0) ? epi_out_row(ep, m) : static_cast<int64_t>(m);
          if (EXACT || m < p.M) epilogue_fetch8<OutT, EPI>(ep, orow * ep.ldc + nn, m, nn, dst[pass], p.flags);
        }
      };
      fetch_chunk(0, pre_nx);
#pragma unroll 1
      for (int c = 0; c < NCH; ++c) {
        const int n0 = nb + c * 32;
        if (!EXACT && n0 >= p.N) break;                // warp-uniform
        const int n = n0 + cg * 8;
        const bool ncol_ok = EXACT || n < p.N;
        EpiPre pre[4];
#pragma unroll
        for (int pass = 0; pass < 4; ++pass) pre[pass] = pre_nx[pass];
        if (c + 1 < NCH) fetch_chunk(c + 1, pre_nx);
        int64_t off[4];
        float bias8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) bias8[i] = 0.f;
        if (VB_EPI(kEpiBias, ep.bias != nullptr) && ncol_ok) {
          const float4 b0 = __ldg(reinterpret_cast<const float4*>(ep.bias + n));
          const float4 b1 = __ldg(reinterpret_cast<const float4*>(ep.bias + n + 4));
          bias8[0] = b0.x; bias8[1] = b0.y; bias8[2] = b0.z; bias8[3] = b0.w;
          bias8[4] = b1.x; bias8[5] = b1.y; bias8[6] = b1.z; bias8[7] = b1.w;
        }
#pragma unroll
        for (int pass = 0; pass < 4; ++pass) {
          const int m = m0 + pass * 8 + rsub;
          const int64_t orow = VB_EPI(kEpiRowMap, ep.row_period > 0) ? epi_out_row(ep, m) : static_cast<int64_t>(m);
          off[pass] = orow * ep.ldc + n;
        }
        // (2) accumulator chunk: TMEM -> registers (one row per thread) -> swizzled shared-memory transpose
        ptx::tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 8; ++j)                    // my row, 16-byte chunk j -> physical chunk j ^ (row & 7)
          ptx::sts128(stg_s + static_cast<uint32_t>(lane * 128 + ((j ^ (lane & 7)) << 4)), r[4 * j], r[4 * j + 1],
                      r[4 * j + 2], r[4 * j + 3]);
        __syncwarp();
        // (3) next chunk's TMEM load flies while this chunk is finished
        if (c + 1 < NCH && (EXACT || n0 + 32 < p.N))
          ptx::tmem_ld_32x32(t_row + static_cast<uint32_t>(half * CW + (c + 1) * 32), r);
        float cs[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) cs[i] = 0.f;
#pragma unroll
        for (int pass = 0; pass < 4; ++pass) {
          const int R = pass * 8 + rsub;
          const int m = m0 + R;
          const float4 lo = ptx::lds128(stg_s + static_cast<uint32_t>(R * 128 + (((2 * cg) ^ (R & 7)) << 4)));
          const float4 hi = ptx::lds128(stg_s + static_cast<uint32_t>(R * 128 + (((2 * cg + 1) ^ (R & 7)) << 4)));
          uint32_t sign8 = 0;
          float dpart = 0.f;
          if (EXACT || (m < p.M && ncol_ok)) {
            float v[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
            epilogue_finish8<OutT, EPI>(ep, m, n, p.N, off[pass], v, pre[pass], bias8, p.flags);
            if constexpr (sizeof(OutT) == 2) {
              if (VB_EPI(kEpiDot, ep.dot_out != nullptr)) {   // stored (bf16-rounded) values x the aux operand
                const uint32_t ov[4] = {pre[pass].a0.x, pre[pass].a0.y, pre[pass].a0.z, pre[pass].a0.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  const uint32_t pk = pack_bf16x2(v[2 * j], v[2 * j + 1]);
                  dpart = fmaf(__uint_as_float(pk << 16), __uint_as_float(ov[j] << 16), dpart);
                  dpart = fmaf(__uint_as_float(pk & 0xffff0000u), __uint_as_float(ov[j] & 0xffff0000u), dpart);
                }
              }
            }
            if (VB_EPI(kEpiColSum, ep.col_sum != nullptr)) {
#pragma unroll
              for (int i = 0; i < 8; ++i) cs[i] += v[i];
            }
            if (VB_EPI(kEpiBitsOut, ep.bits_out != nullptr)) {
#pragma unroll
              for (int i = 0; i < 8; ++i) sign8 |= (v[i] > 0.f ? 1u : 0u) << i;
            }
          }
          if constexpr (sizeof(OutT) == 2) {
            if (VB_EPI(kEpiDot, ep.dot_out != nullptr)) {   // warp-uniform
              // Same association as attn_delta_vec_kernel (8 sequential FMAs per 16-byte piece, then a butterfly over
              // the head's 8 pieces with strides 4, 2, 1), so the fused statistic is BIT-IDENTICAL to the stand-alone
              // kernel's: lane cg holds piece cg of the even chunk and piece 4 + cg of the odd one.
              if (c & 1) {                                   // second chunk of the head: n0 % 64 == 32
                float t = dacc[pass] + dpart;                //  stride 4
                t += __shfl_xor_sync(0xffffffffu, t, 2);
                t += __shfl_xor_sync(0xffffffffu, t, 1);
                if (cg == 0 && (EXACT || (m < p.M && ncol_ok))) {
                  const int bi = m / ep.dot_T, ti = m - bi * ep.dot_T;
                  ep.dot_out[(static_cast<int64_t>(bi) * ep.dot_H + (n0 >> 6)) * ep.dot_T + ti] = t;
                }
              } else {
                dacc[pass] = dpart;
              }
            }
          }
          if (VB_EPI(kEpiBitsOut, ep.bits_out != nullptr)) {   // warp-uniform: the 4 lanes of a row merge their bytes
            uint32_t w = sign8 << (8 * cg);
            w |= __shfl_xor_sync(0xffffffffu, w, 1);
            w |= __shfl_xor_sync(0xffffffffu, w, 2);
            if (cg == 0 && (EXACT || m < p.M)) ep.bits_out[static_cast<int64_t>(n0 >> 5) * ep.bits_rows + m] = w;
          }
        }
        if (VB_EPI(kEpiColSum, ep.col_sum != nullptr)) {   // warp-uniform
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            cs[i] += __shfl_xor_sync(0xffffffffu, cs[i], 4);
            cs[i] += __shfl_xor_sync(0xffffffffu, cs[i], 8);
            cs[i] += __shfl_xor_sync(0xffffffffu, cs[i], 16);
          }
          if (rsub == 0 && ncol_ok) {
            red_add_f32x4(ep.col_sum + n, cs[0], cs[1], cs[2], cs[3]);
            red_add_f32x4(ep.col_sum + n + 4, cs[4], cs[5], cs[6], cs[7]);
          }
        }
        __syncwarp();
      }
      ptx::tmem_ld_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (CG >= 2) ptx::mbar_arrive_cluster(&tmem_empty[acc], 2u * pair); else ptx::mbar_arrive(&tmem_empty[acc]);
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
  }

  ptx::tc_fence_before();
  if constexpr (CG >= 2) ptx::cluster_sync(); else __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    if constexpr (CG >= 2) ptx::tmem_dealloc_pair(tmem_base, Cfg::TMEM_COLS); else ptx::tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

}  // namespace

// ----------------------------------------------------------------------------------------------------
// Host side: tensor maps + launch
// ----------------------------------------------------------------------------------------------------
namespace {

using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  });
  return fn;
}

}  // namespace

// 2-D bf16 tensor map over a row-major matrix [rows][ld] whose logical width is `cols`;
// box = box_cols (inner, <= 64 elements = 128 B) x box_rows, 128B swizzle, OOB reads return zero.
int make_tmap_bf16(void* map_out, const void* base, int64_t rows, int64_t cols, int64_t ld, int box_cols,
                   int box_rows) {
  CUtensorMap* map = static_cast<CUtensorMap*>(map_out);
  EncodeTiledFn enc = get_encode_fn();
  if (enc == nullptr) {
    set_error("cuTensorMapEncodeTiled entry point not available");
    return kErrCuda;
  }
  const cuuint64_t gdim[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  const cuuint64_t gstr[1] = {static_cast<cuuint64_t>(ld) * 2};
  const cuuint32_t box[2] = {static_cast<cuuint32_t>(box_cols), static_cast<cuuint32_t>(box_rows)};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstr, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d): rows=%lld cols=%lld ld=%lld box=%dx%d", static_cast<int>(r),
              static_cast<long long>(rows), static_cast<long long>(cols), static_cast<long long>(ld), box_cols, box_rows);
    return kErrCuda;
  }
  return kOk;
}

// 3-D bf16 tensor map over the packed qkv buffer [rows][3H][hs] (dim0 = hs elements of one head section, dim1 = the 3H
// head sections of a row, dim2 = rows; row stride ld elements): box = 64 x 1 x box_rows, 128B swizzle.  Because dim0
// is BOUNDED at hs, a 64-wide box starting at column 64 of an hs = 80 head delivers 16 real columns and 48 zeros —
// the head dimension is split into 64-column swizzle atoms without touching the neighbouring head.
int make_tmap_bf16_heads(void* map_out, const void* base, int64_t rows, int sections, int hs, int64_t ld,
                         int box_rows) {
  CUtensorMap* map = static_cast<CUtensorMap*>(map_out);
  EncodeTiledFn enc = get_encode_fn();
  if (enc == nullptr) {
    set_error("cuTensorMapEncodeTiled entry point not available");
    return kErrCuda;
  }
  const cuuint64_t gdim[3] = {static_cast<cuuint64_t>(hs), static_cast<cuuint64_t>(sections),
                              static_cast<cuuint64_t>(rows)};
  const cuuint64_t gstr[2] = {static_cast<cuuint64_t>(hs) * 2, static_cast<cuuint64_t>(ld) * 2};
  const cuuint32_t box[3] = {64, 1, static_cast<cuuint32_t>(box_rows)};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), gdim, gstr, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (heads) failed (%d): rows=%lld sections=%d hs=%d ld=%lld", static_cast<int>(r),
              static_cast<long long>(rows), sections, hs, static_cast<long long>(ld));
    return kErrCuda;
  }
  return kOk;
}

namespace {

template <int AMAJ, int BMAJ, int BN, typename OutT, uint32_t EPI, int CG, bool EXACT = false>
int launch_tc_cg(const CUtensorMap& ta, const CUtensorMap& tb, const TcParams& p, const GemmEpilogue& ep,
                 cudaStream_t stream) {
  using Cfg = TileCfg<BN, CG, epi_warps<OutT, CG>()>;
  constexpr int kThreads = gemm_threads<OutT, CG>();
  auto kern = gemm_tc_kernel<AMAJ, BMAJ, BN, OutT, EPI, CG, EXACT>;
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [&] {
    attr_err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
  });
  VB_CHECK_CUDA(attr_err);
  ProfScope prof(kProfGemmTc, 2.0 * p.M * p.N * p.K, 1, stream);
  const int total = p.m_tiles * p.n_tiles * p.splits;     // work units (one per CTA, or per CTA pair)
  if constexpr (CG == 1) {
    const int grid = total < num_sms() ? total : num_sms();
    VB_CHECK_CUDA(launch_pdl(kern, dim3(grid), dim3(kThreads), Cfg::SMEM_BYTES, stream, 1, ta, tb, p, ep));
  } else {
    // persistent walk: only as many clusters as can be co-resident (clusters of 4 do not tile all 148 SMs)
    static int max_clusters = 0;
    static std::once_flag occ_once;
    std::call_once(occ_once, [&] {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(num_sms() / CG * CG);
      cfg.blockDim = dim3(kThreads);
      cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = CG;
      attr[0].val.clusterDim.y = 1;
      attr[0].val.clusterDim.z = 1;
      cfg.attrs = attr;
      cfg.numAttrs = 1;
      int n = 0;
      if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess || n <= 0) n = num_sms() / CG;
      max_clusters = n;
    });
    const int clusters = total < max_clusters ? total : max_clusters;
    VB_CHECK_CUDA(launch_pdl(kern, dim3(CG * clusters), dim3(kThreads), Cfg::SMEM_BYTES, stream, CG, ta, tb, p, ep));
  }
  VB_CHECK_LAUNCH();
  return kOk;
}

// cg: CTAs per cluster the operands / TcParams were prepared for (128*cg-row tiles, BN/cg-row B boxes).  The cluster
// of 4 is instantiated for the forward-type call sites only (K-major x K-major and K-major x MN-major operands).
template <int AMAJ, int BMAJ, int BN, typename OutT, uint32_t EPI>
int launch_tc(int cg, const CUtensorMap& ta, const CUtensorMap& tb, const TcParams& p, const GemmEpilogue& ep,
              cudaStream_t stream) {
#ifdef VITB200_CG4   // measured 4-8 % SLOWER than pairs on B200 (only 33 clusters of 4 fit the 148 SMs): off by default
  if constexpr (AMAJ == 0) {
    if (cg == 4) return launch_tc_cg<AMAJ, BMAJ, BN, OutT, EPI, 4>(ta, tb, p, ep, stream);
  }
#endif
  if (cg == 4) cg = 2;
  if (cg == 2) {
    // the hot call sites of the training step tile exactly (ViT-B/16 at batch 256: M = 50432 = 197 x 256, N = 768 .. 3072)
    if constexpr (EPI != kEpiGeneric) {
      static const bool exact_on = !(getenv("VITB200_GEMM_EXACT") && atoi(getenv("VITB200_GEMM_EXACT")) == 0);
      if (exact_on && p.M % (2 * BM) == 0 && p.N % BN == 0)
        return launch_tc_cg<AMAJ, BMAJ, BN, OutT, EPI, 2, true>(ta, tb, p, ep, stream);
    }
    return launch_tc_cg<AMAJ, BMAJ, BN, OutT, EPI, 2>(ta, tb, p, ep, stream);
  }
  return launch_tc_cg<AMAJ, BMAJ, BN, OutT, EPI, 1>(ta, tb, p, ep, stream);
}

}  // namespace

// Split count of an atomically accumulated contraction on the static persistent walk: `tiles` output tiles, `kb_total`
// 64-deep k-blocks, `walkers` CTAs (or CTA pairs).  Minimises rounds x (k-blocks per split + hand-over); a larger count
// must win by more than 1 % (fewer splits = less atomic traffic).  Pure host arithmetic (vitb200_gemm_auto_splits).
int gemm_tc_auto_splits(int tiles, int kb_total, int walkers) {
  constexpr int kHandOver = 6;   // k-block equivalents per unit: accumulator hand-over + its share of the atomics
  const int max_s = kb_total / 4 > 0 ? kb_total / 4 : 1;
  long best = -1;
  int s = 1;
  for (int c = 1; c <= max_s && c <= 64; ++c) {
    const int per = ceil_div(kb_total, c);
    const int units = tiles * ceil_div(kb_total, per);
    const long cost = static_cast<long>(ceil_div(units, walkers)) * (per + kHandOver);
    if (best < 0 || cost * 100 < best * 99) { best = cost; s = c; }
  }
  return s;
}

// Entry used by vitb200_gemm() for in_dtype == BF16.  out_f32 selects the C dtype.
int gemm_tc_bf16(int M, int N, int K, int a_major, int b_major, const void* A, int64_t lda, const void* B,
                 int64_t ldb, bool out_f32, int splits, const GemmEpilogue& ep, cudaStream_t stream) {
  VB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_tc: empty problem M=%d N=%d K=%d", M, N, K);
  VB_REQUIRE(N % 8 == 0 && ep.ldc % 8 == 0, "gemm_tc: N (%d) and ldc (%lld) must be multiples of 8", N,
             static_cast<long long>(ep.ldc));
  VB_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "gemm_tc: lda/ldb must be multiples of 8 elements (16 B rows for TMA)");
  VB_REQUIRE(aligned16(A) && aligned16(B) && aligned16(ep.C), "gemm_tc: operands must be 16-byte aligned");
  VB_REQUIRE(!ep.accumulate || out_f32, "gemm_tc: atomic accumulation needs an fp32 output");
  VB_REQUIRE(ep.residual == nullptr || out_f32, "gemm_tc: residual add needs an fp32 output");
  VB_REQUIRE((ep.bits_out == nullptr && ep.mask_bits == nullptr) || (N % 32 == 0 && ep.bits_rows >= M),
             "gemm_tc: bit masks need N %% 32 == 0 and bits_rows >= M (N=%d, bits_rows=%d)", N, ep.bits_rows);
  VB_REQUIRE(ep.dot_out == nullptr || (!out_f32 && ep.dot_aux != nullptr && N % 64 == 0 && ep.dot_T > 0 &&
                                       ep.dot_H * 64 == N && M % ep.dot_T == 0 && ep.mask_aux == nullptr &&
                                       ep.row_period == 0 && aligned16(ep.dot_aux)),
             "gemm_tc: the fused head dot needs a 2-byte output, N = 64 * dot_H, M a multiple of dot_T and no aux mask / row map");
  constexpr int BN = 256;

  // CTA pairs (256-row tiles) whenever there is more than one 128-row tile of output; VITB200_GEMM_CG=1|2 forces.
  static const int force_cg = getenv("VITB200_GEMM_CG") ? atoi(getenv("VITB200_GEMM_CG")) : 0;   // 1 | 2 | 4 forces
  int cg = M > BM ? 2 : 1;
#ifdef VITB200_CG4
  if (force_cg == 4 && a_major == 0 && !ep.accumulate) cg = 4;   // experiment only, see launch_tc
#endif
  if (force_cg == 1 || force_cg == 2) cg = force_cg;

  CUtensorMap ta, tb;
  if (a_major == 0) VB_PROPAGATE(make_tmap_bf16(&ta, A, M, K, lda, BK, BM));
  else              VB_PROPAGATE(make_tmap_bf16(&ta, A, K, M, lda, 64, BK));
  if (b_major == 0) VB_PROPAGATE(make_tmap_bf16(&tb, B, N, K, ldb, BK, BN / cg));
  else              VB_PROPAGATE(make_tmap_bf16(&tb, B, K, N, ldb, 64, BK));

  TcParams p;
  p.M = M; p.N = N; p.K = K;
  p.m_tiles = ceil_div(M, BM * cg);
  p.n_tiles = ceil_div(N, BN);
  p.kb_total = ceil_div(K, BK);
  int s = splits;
  if (!ep.accumulate) s = 1;
  if (s <= 0) {
    // auto.  The persistent walk is static: walker i takes units i, i + W, ... (W = CTAs or CTA pairs), all units the
    // same length, so the launch lasts rounds * (k-blocks per split + hand-over).  Round 1 aimed at "about two waves"
    // (s = ceil(2W / tiles)), which left the weight-gradient sites of ViT-B/16 at 180 / 162 / 153 units on 74 pairs =
    // THREE rounds at 81 / 73 / 69 % occupancy — exactly their measured deficit against cuBLAS.  Now: the split count
    // with the lowest modelled duration, fewest splits on ties (less atomic traffic into the fp32 gradient arena).
    static const bool old_rule = getenv("VITB200_GEMM_SPLITS_OLD") && atoi(getenv("VITB200_GEMM_SPLITS_OLD")) != 0;
    const int tiles = p.m_tiles * p.n_tiles;
    const int walkers = num_sms() / cg;
    const int max_s = p.kb_total / 4 > 0 ? p.kb_total / 4 : 1;
    if (old_rule) {
      s = ceil_div(2 * walkers, tiles);
      if (s > max_s) s = max_s;
      if (s < 1) s = 1;
    } else {
      s = gemm_tc_auto_splits(tiles, p.kb_total, walkers);
    }
  }
  // Epilogue memory-side switches (VITB200_GEMM_EPI_FLAGS overrides; measured on B200 at the ViT-B/16 proj site, fp32 out
  // + bias + dropout + residual, 92.3 us with 0): 1 = L2 prefetch one tile ahead (+10 us), 2 = per-lane L2 prefetch
  // (-5.5 us; aux-mask site 252 -> 237 us), 4 = residual loads bypass L1 (+12 us: the two 16-byte halves of a sector
  // are then fetched twice), 8 = 256-bit residual loads / output stores (-2.5 us more).  Default 2 | 8.
  static const int epi_flags_env = getenv("VITB200_GEMM_EPI_FLAGS") ? atoi(getenv("VITB200_GEMM_EPI_FLAGS")) : 10;
  p.flags = epi_flags_env;
  // 256-bit accesses need 32-byte aligned rows
  if ((reinterpret_cast<uintptr_t>(ep.C) & 31) || (reinterpret_cast<uintptr_t>(ep.residual) & 31)) p.flags &= ~8;
  p.kb_per_split = ceil_div(p.kb_total, s);
  p.splits = ceil_div(p.kb_total, p.kb_per_split);

  // Hot call sites of the training step get a compile-time-specialised epilogue; anything else runs the generic one.
  const uint32_t f = epi_flags(ep);
#define VB_TC(AM, BM_, T, FL) \
  if (f == (FL)) return launch_tc<AM, BM_, BN, T, (FL)>(cg, ta, tb, p, ep, stream)
  if (a_major == 0 && b_major == 0) {
    if (out_f32) {
      VB_TC(0, 0, float, 0u);
      VB_TC(0, 0, float, kEpiBias | kEpiResidual);
      VB_TC(0, 0, float, kEpiBias | kEpiDrop | kEpiResidual);
      VB_TC(0, 0, float, kEpiBias | kEpiRowMap);
      return launch_tc<0, 0, BN, float, kEpiGeneric>(cg, ta, tb, p, ep, stream);
    }
    VB_TC(0, 0, bf16, 0u);
    VB_TC(0, 0, bf16, kEpiBias | kEpiRelu);
    VB_TC(0, 0, bf16, kEpiBias | kEpiRelu | kEpiBitsOut);
    return launch_tc<0, 0, BN, bf16, kEpiGeneric>(cg, ta, tb, p, ep, stream);
  } else if (a_major == 0 && b_major == 1) {
    if (out_f32) return launch_tc<0, 1, BN, float, kEpiGeneric>(cg, ta, tb, p, ep, stream);
    VB_TC(0, 1, bf16, 0u);
    VB_TC(0, 1, bf16, kEpiDot);
    VB_TC(0, 1, bf16, kEpiMask | kEpiColSum);
    VB_TC(0, 1, bf16, kEpiBitMask | kEpiColSum);
    return launch_tc<0, 1, BN, bf16, kEpiGeneric>(cg, ta, tb, p, ep, stream);
  } else if (a_major == 1 && b_major == 1) {
    if (out_f32) {
      VB_TC(1, 1, float, kEpiAccum);
      return launch_tc<1, 1, BN, float, kEpiGeneric>(cg, ta, tb, p, ep, stream);
    }
    return launch_tc<1, 1, BN, bf16, kEpiGeneric>(cg, ta, tb, p, ep, stream);
  }
#undef VB_TC
  set_error("gemm_tc: operand major combination a=%d b=%d not instantiated", a_major, b_major);
  return kErrUnsupported;
}

}  // namespace vb
