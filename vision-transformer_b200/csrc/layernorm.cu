// LayerNorm forward / backward as coalesced, vectorised bandwidth kernels (one warp per row, the row staged
// once in shared memory so HBM sees each element exactly once).
//   forward : x fp32 [M,E] -> y (fp32|bf16) [M,E], saves mean / rstd                       (transformer.py:72-73,77-78; vit.py:72)
//   backward: dx fused with the residual-gradient accumulate (g += dx), dgamma/dbeta reduced per CTA in
//             shared memory then atomically into fp32 grads, and optionally emits the dropout-masked,
//             down-cast gradient the next GEMM consumes (so no separate cast/mask pass exists).
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"
#include "ptx.cuh"

namespace vb {

namespace {

using namespace ptx;

constexpr int LN_WARPS = 8;
constexpr int LN_THREADS = LN_WARPS * 32;

template <typename OutT>
__global__ void __launch_bounds__(LN_THREADS)
ln_fwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
              OutT* __restrict__ y, float* __restrict__ mean_out, float* __restrict__ rstd_out, int M, int E,
              float eps) {
  extern __shared__ float ln_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  float* row = ln_smem + static_cast<size_t>(warp) * E;
  const bool vec = (E & 3) == 0;
  pdl_wait();
  pdl_trigger();
  for (int m = blockIdx.x * nw + warp; m < M; m += gridDim.x * nw) {
    const float* xr = x + static_cast<int64_t>(m) * E;
    float s = 0.f;
    if (vec) {
      for (int i = lane; i < E / 4; i += 32) {
        const float4 v = __ldcs(reinterpret_cast<const float4*>(xr) + i);
        reinterpret_cast<float4*>(row)[i] = v;
        s += (v.x + v.y) + (v.z + v.w);
      }
    } else {
      for (int i = lane; i < E; i += 32) { const float v = xr[i]; row[i] = v; s += v; }
    }
    const float mean = warp_sum(s) / static_cast<float>(E);
    __syncwarp();
    float sq = 0.f;
    for (int i = lane; i < E; i += 32) { const float d = row[i] - mean; sq = fmaf(d, d, sq); }
    const float rstd = rsqrtf(warp_sum(sq) / static_cast<float>(E) + eps);
    if (lane == 0) {
      if (mean_out) mean_out[m] = mean;
      if (rstd_out) rstd_out[m] = rstd;
    }
    OutT* yr = y + static_cast<int64_t>(m) * E;
    if (vec) {
      for (int i = lane; i < E / 4; i += 32) {
        const float4 v = reinterpret_cast<const float4*>(row)[i];
        const float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + i);
        const float4 b = __ldg(reinterpret_cast<const float4*>(beta) + i);
        const float o0 = (v.x - mean) * rstd * g.x + b.x, o1 = (v.y - mean) * rstd * g.y + b.y;
        const float o2 = (v.z - mean) * rstd * g.z + b.z, o3 = (v.w - mean) * rstd * g.w + b.w;
        if constexpr (sizeof(OutT) == 2) {
          uint2 o;
          o.x = pack_bf16x2(o0, o1);
          o.y = pack_bf16x2(o2, o3);
          reinterpret_cast<uint2*>(yr)[i] = o;
        } else {
          reinterpret_cast<float4*>(yr)[i] = make_float4(o0, o1, o2, o3);
        }
      }
    } else {
      for (int i = lane; i < E; i += 32)
        yr[i] = from_f32<OutT>((row[i] - mean) * rstd * __ldg(gamma + i) + __ldg(beta + i));
    }
    __syncwarp();
  }
}

// Register-resident forward for E % 4 == 0, E <= 128 * VPL: each lane owns float4 chunks lane + 32k of its warp's
// row, and the NEXT row's chunks are requested before the current row's statistics are reduced, so every warp keeps
// two rows of loads in flight and shared memory is not touched.  Same two-pass arithmetic as ln_fwd_kernel.
template <typename OutT, int VPL>
__global__ void __launch_bounds__(LN_THREADS)
ln_fwd_reg_kernel(const float* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
                  OutT* __restrict__ y, float* __restrict__ mean_out, float* __restrict__ rstd_out, int M, int E,
                  float eps) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const int nq = E >> 2;
  const int stride = gridDim.x * nw;
  pdl_wait();
  pdl_trigger();
  float4 gm[VPL], bt[VPL];
#pragma unroll
  for (int k = 0; k < VPL; ++k) {
    const int i = lane + 32 * k;
    gm[k] = i < nq ? __ldg(reinterpret_cast<const float4*>(gamma) + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    bt[k] = i < nq ? __ldg(reinterpret_cast<const float4*>(beta) + i) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  int m = blockIdx.x * nw + warp;
  float4 cur[VPL], nxt[VPL];
  if (m < M) {
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      const int i = lane + 32 * k;
      cur[k] = i < nq ? __ldcs(reinterpret_cast<const float4*>(x + static_cast<int64_t>(m) * E) + i)
                      : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  const float inv_e = 1.f / static_cast<float>(E);
  for (; m < M; m += stride) {
    const int mn = m + stride;
    if (mn < M) {
#pragma unroll
      for (int k = 0; k < VPL; ++k) {
        const int i = lane + 32 * k;
        nxt[k] = i < nq ? __ldcs(reinterpret_cast<const float4*>(x + static_cast<int64_t>(mn) * E) + i)
                        : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < VPL; ++k) s += (cur[k].x + cur[k].y) + (cur[k].z + cur[k].w);
    const float mean = warp_sum(s) / static_cast<float>(E);
    float sq = 0.f;
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      if (lane + 32 * k < nq) {
        const float d0 = cur[k].x - mean, d1 = cur[k].y - mean, d2 = cur[k].z - mean, d3 = cur[k].w - mean;
        sq = fmaf(d0, d0, sq); sq = fmaf(d1, d1, sq); sq = fmaf(d2, d2, sq); sq = fmaf(d3, d3, sq);
      }
    }
    const float rstd = rsqrtf(warp_sum(sq) * inv_e + eps);
    if (lane == 0) {
      if (mean_out) mean_out[m] = mean;
      if (rstd_out) rstd_out[m] = rstd;
    }
    OutT* yr = y + static_cast<int64_t>(m) * E;
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      const int i = lane + 32 * k;
      if (i < nq) {
        const float o0 = (cur[k].x - mean) * rstd * gm[k].x + bt[k].x, o1 = (cur[k].y - mean) * rstd * gm[k].y + bt[k].y;
        const float o2 = (cur[k].z - mean) * rstd * gm[k].z + bt[k].z, o3 = (cur[k].w - mean) * rstd * gm[k].w + bt[k].w;
        if constexpr (sizeof(OutT) == 2) {
          uint2 o;
          o.x = pack_bf16x2(o0, o1);
          o.y = pack_bf16x2(o2, o3);
          reinterpret_cast<uint2*>(yr)[i] = o;
        } else {
          reinterpret_cast<float4*>(yr)[i] = make_float4(o0, o1, o2, o3);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < VPL; ++k) cur[k] = nxt[k];
  }
}

// Forward, bulk-copy ring version (round 2; E % 4 == 0, E <= 128 * VPL): one persistent CTA per SM, LNF_WARPS warps,
// each warp owns a ring of `nst` row slots in shared memory that lane 0 fills with 1-D bulk copies nst-1 rows ahead.
// The bytes in flight (warps x (nst-1) x 4E, ~100 KB per SM at E = 768) no longer depend on registers, which is what
// limited the register-resident kernel above to ~4.1 TB/s of DRAM reads (two CTAs x 8 warps x one prefetched row).
constexpr int LNF_WARPS = 16;
constexpr int LNF_MAX_STAGES = 4;

template <typename OutT, int VPL, bool EXACT>   // EXACT: E == 128 * VPL, no per-chunk bounds tests (see ln_bwd_ring_kernel)
__global__ void __launch_bounds__(LNF_WARPS * 32, 1)
ln_fwd_ring_kernel(const float* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
                   OutT* __restrict__ y, float* __restrict__ mean_out, float* __restrict__ rstd_out, int M, int E,
                   float eps, int nst) {
  extern __shared__ __align__(128) uint8_t lnf_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  uint64_t* bars = reinterpret_cast<uint64_t*>(lnf_smem);
  uint64_t* bar = bars + warp * LNF_MAX_STAGES;
  const uint32_t row_bytes = static_cast<uint32_t>(E) * 4u;
  const uint32_t ring = smem_u32(bars + LNF_WARPS * LNF_MAX_STAGES) + static_cast<uint32_t>(warp * nst) * row_bytes;
  if (lane == 0) {
    for (int st = 0; st < nst; ++st) mbar_init(bar + st, 1);
    fence_barrier_init();
  }
  __syncwarp();
  pdl_wait();
  pdl_trigger();
  const int nq = E >> 2;
  constexpr bool GB_REG = VPL <= 8;   // wide rows (512 threads x 12 chunks): gamma / beta come from L1 instead of registers
  float4 gm[GB_REG ? VPL : 1], bt[GB_REG ? VPL : 1];
  if constexpr (GB_REG) {
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      const int i = lane + 32 * k;
      gm[k] = (EXACT || i < nq) ? __ldg(reinterpret_cast<const float4*>(gamma) + i) : make_float4(0.f, 0.f, 0.f, 0.f);
      bt[k] = (EXACT || i < nq) ? __ldg(reinterpret_cast<const float4*>(beta) + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  const int w0 = blockIdx.x * nw + warp, stride = gridDim.x * nw;
  const int nrows = w0 < M ? (M - w0 + stride - 1) / stride : 0;
  auto issue = [&](int it) {  // lane 0
    const int st = it % nst;
    mbar_arrive_expect_tx(bar + st, row_bytes);
    bulk_load_1d(ring + st * row_bytes, x + (static_cast<int64_t>(w0) + static_cast<int64_t>(it) * stride) * E, row_bytes,
                 bar + st);
  };
  if (lane == 0)
    for (int it = 0; it < nst - 1 && it < nrows; ++it) issue(it);
  const float inv_e = 1.f / static_cast<float>(E);
  for (int it = 0; it < nrows; ++it) {
    if (lane == 0 && it + nst - 1 < nrows) issue(it + nst - 1);   // refills the slot every lane left in iteration it-1
    const int st = it % nst;
    const int64_t m = static_cast<int64_t>(w0) + static_cast<int64_t>(it) * stride;
    mbar_wait(bar + st, (it / nst) & 1);
    const uint32_t sx = ring + st * row_bytes;
    float4 cur[VPL];
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      const int i = lane + 32 * k;
      cur[k] = (EXACT || i < nq) ? lds128(sx + 16 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
      s += (cur[k].x + cur[k].y) + (cur[k].z + cur[k].w);
    }
    const float mean = warp_sum(s) / static_cast<float>(E);
    float sq = 0.f;
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      if (EXACT || lane + 32 * k < nq) {
        const float d0 = cur[k].x - mean, d1 = cur[k].y - mean, d2 = cur[k].z - mean, d3 = cur[k].w - mean;
        sq = fmaf(d0, d0, sq); sq = fmaf(d1, d1, sq); sq = fmaf(d2, d2, sq); sq = fmaf(d3, d3, sq);
      }
    }
    const float rstd = rsqrtf(warp_sum(sq) * inv_e + eps);
    if (lane == 0) {
      if (mean_out) mean_out[m] = mean;
      if (rstd_out) rstd_out[m] = rstd;
    }
    OutT* yr = y + m * E;
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      const int i = lane + 32 * k;
      if (EXACT || i < nq) {
        float4 gk, bk;
        if constexpr (GB_REG) {
          gk = gm[k]; bk = bt[k];
        } else {
          gk = __ldg(reinterpret_cast<const float4*>(gamma) + i);
          bk = __ldg(reinterpret_cast<const float4*>(beta) + i);
        }
        const float o0 = (cur[k].x - mean) * rstd * gk.x + bk.x, o1 = (cur[k].y - mean) * rstd * gk.y + bk.y;
        const float o2 = (cur[k].z - mean) * rstd * gk.z + bk.z, o3 = (cur[k].w - mean) * rstd * gk.w + bk.w;
        if constexpr (sizeof(OutT) == 2) {
          uint2 o;
          o.x = pack_bf16x2(o0, o1);
          o.y = pack_bf16x2(o2, o3);
          reinterpret_cast<uint2*>(yr)[i] = o;
        } else {
          reinterpret_cast<float4*>(yr)[i] = make_float4(o0, o1, o2, o3);
        }
      }
    }
    __syncwarp();  // every lane has its chunks of this slot in registers before lane 0 refills it
  }
}

// dy: gradient w.r.t. the LN output (DyT).  g: fp32 residual-stream gradient, updated in place (g += dx) when
// add_to_g != 0, overwritten (g = dx) otherwise.  dyn: optional down-cast copy of the new g multiplied by the
// dropout factors of the layer that produced this LN's input.
// VPL > 0: E % 4 == 0 and E <= 128*VPL; each lane owns float4 chunks lane+32k and keeps its dgamma/dbeta
// partial sums in registers across all rows it visits.  VPL == 0: generic path (shared-memory atomics), used for
// odd E and for the very wide, very short head LayerNorm.
template <typename DyT, int VPL>
__global__ void __launch_bounds__(LN_THREADS)
ln_bwd_kernel(const DyT* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ mean_in,
              const float* __restrict__ rstd_in, const float* __restrict__ gamma, float* __restrict__ g,
              int add_to_g, float* __restrict__ dgamma, float* __restrict__ dbeta, DyT* __restrict__ dyn,
              const Dropout drop, int M, int E) {
  extern __shared__ float ln_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  float* acc_g = ln_smem;                 // [E] per-CTA dgamma partial
  float* acc_b = ln_smem + E;             // [E] per-CTA dbeta partial
  float* xh = ln_smem + 2 * static_cast<size_t>(E) + static_cast<size_t>(warp) * 2 * E;  // xhat row
  float* dyr = xh + E;                                                                   // dy row
  for (int i = threadIdx.x; i < 2 * E; i += blockDim.x) ln_smem[i] = 0.f;
  __syncthreads();
  pdl_wait();
  pdl_trigger();
  constexpr bool vec = VPL > 0;
  constexpr int NACC = VPL > 0 ? VPL * 4 : 1;
  float rg[NACC], rb[NACC];
#pragma unroll
  for (int j = 0; j < NACC; ++j) { rg[j] = 0.f; rb[j] = 0.f; }
  for (int m = blockIdx.x * nw + warp; m < M; m += gridDim.x * nw) {
    const int64_t base = static_cast<int64_t>(m) * E;
    const float mean = mean_in[m], rstd = rstd_in[m];
    float s1 = 0.f, s2 = 0.f;
    if constexpr (vec) {
#pragma unroll
      for (int k = 0; k < VPL; ++k) {
        const int i = lane + 32 * k;
        if (i >= E / 4) break;
        const float4 xv = __ldcs(reinterpret_cast<const float4*>(x + base) + i);
        float d[4];
        if constexpr (sizeof(DyT) == 2) {
          const uint2 raw = __ldcs(reinterpret_cast<const uint2*>(dy + base) + i);
          const __nv_bfloat162 p0 = *reinterpret_cast<const __nv_bfloat162*>(&raw.x);
          const __nv_bfloat162 p1 = *reinterpret_cast<const __nv_bfloat162*>(&raw.y);
          d[0] = __low2float(p0); d[1] = __high2float(p0); d[2] = __low2float(p1); d[3] = __high2float(p1);
        } else {
          const float4 dv = __ldcs(reinterpret_cast<const float4*>(dy + base) + i);
          d[0] = dv.x; d[1] = dv.y; d[2] = dv.z; d[3] = dv.w;
        }
        const float4 gm = __ldg(reinterpret_cast<const float4*>(gamma) + i);
        const float h[4] = {(xv.x - mean) * rstd, (xv.y - mean) * rstd, (xv.z - mean) * rstd, (xv.w - mean) * rstd};
        const float gg[4] = {gm.x, gm.y, gm.z, gm.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float t = d[j] * gg[j];
          s1 += t;
          s2 = fmaf(t, h[j], s2);
          xh[i * 4 + j] = h[j];
          dyr[i * 4 + j] = d[j];
        }
      }
    } else {
      for (int i = lane; i < E; i += 32) {
        const float h = (x[base + i] - mean) * rstd;
        const float d = to_f32(dy[base + i]);
        const float t = d * __ldg(gamma + i);
        s1 += t; s2 = fmaf(t, h, s2);
        xh[i] = h; dyr[i] = d;
      }
    }
    s1 = warp_sum(s1) / static_cast<float>(E);
    s2 = warp_sum(s2) / static_cast<float>(E);
    __syncwarp();
    if constexpr (vec) {
#pragma unroll
      for (int k = 0; k < VPL; ++k) {
        const int i = lane + 32 * k;
        if (i >= E / 4) break;
        const float4 gm = __ldg(reinterpret_cast<const float4*>(gamma) + i);
        const float gg[4] = {gm.x, gm.y, gm.z, gm.w};
        float o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float h = xh[i * 4 + j], d = dyr[i * 4 + j];
          o[j] = rstd * (d * gg[j] - s1 - h * s2);
          rg[k * 4 + j] = fmaf(d, h, rg[k * 4 + j]);
          rb[k * 4 + j] += d;
        }
        float4* gp = reinterpret_cast<float4*>(g + base) + i;
        if (add_to_g) {
          const float4 old = *gp;
          o[0] += old.x; o[1] += old.y; o[2] += old.z; o[3] += old.w;
        }
        *gp = make_float4(o[0], o[1], o[2], o[3]);
        if (dyn != nullptr) {
          if (drop.p > 0.f) {
            float f[4];
            dropout_factors4(drop, static_cast<uint64_t>(base) + 4 * i, f);
#pragma unroll
            for (int j = 0; j < 4; ++j) o[j] *= f[j];
          }
          if constexpr (sizeof(DyT) == 2) {
            uint2 pk;
            pk.x = pack_bf16x2(o[0], o[1]);
            pk.y = pack_bf16x2(o[2], o[3]);
            reinterpret_cast<uint2*>(dyn + base)[i] = pk;
          } else {
            reinterpret_cast<float4*>(dyn + base)[i] = make_float4(o[0], o[1], o[2], o[3]);
          }
        }
      }
    } else {
      for (int i = lane; i < E; i += 32) {
        const float h = xh[i], d = dyr[i];
        float o = rstd * (d * __ldg(gamma + i) - s1 - h * s2);
        atomicAdd(&acc_g[i], d * h);
        atomicAdd(&acc_b[i], d);
        if (add_to_g) o += g[base + i];
        g[base + i] = o;
        if (dyn != nullptr) {
          if (drop.p > 0.f) {
            const uint64_t lin = static_cast<uint64_t>(base) + i;
            o *= dropout_factor1(drop, lin);
          }
          dyn[base + i] = from_f32<DyT>(o);
        }
      }
    }
    __syncwarp();
  }
  if constexpr (vec) {
#pragma unroll
    for (int k = 0; k < VPL; ++k) {
      const int i = lane + 32 * k;
      if (i >= E / 4) break;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        atomicAdd(&acc_g[i * 4 + j], rg[k * 4 + j]);
        atomicAdd(&acc_b[i * 4 + j], rb[k * 4 + j]);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < E; i += blockDim.x) {
    if (dgamma) atomicAdd(dgamma + i, acc_g[i]);
    if (dbeta) atomicAdd(dbeta + i, acc_b[i]);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Backward, bulk-copy ring version (E % 8 == 0, E <= 1024): one persistent CTA per SM, 8 warps, each warp owns a ring
// of NST row slots in shared memory.  Lane 0 of a warp issues 1-D bulk copies (x row, old g row, dy row) NST-1 rows
// ahead of the row being processed, so ~2 rows per warp (~120 KB per SM) are in flight while the warp computes;
// the row statistics, dgamma / dbeta / bias-gradient partial sums stay in registers.  The dropout keep-bits of a
// row are generated once per 8 elements (lane l owns Philox blocks l, l+32, ...) and handed to the two lanes that
// need them with one shuffle per chunk.
// ---------------------------------------------------------------------------------------------------------------
constexpr int LNB_WARPS = 8;
constexpr int LNB_MAX_STAGES = 4;

// KEEP: xhat / dy of the row stay in registers between the two passes (CH <= 8); otherwise they are re-read from
// the shared-memory slot in the second pass so that wide rows (E up to 1536, CH = 10 / 12) do not spill.
// EXACT: E == 128 * CH (every lane owns CH whole chunks), so the per-chunk bounds tests and their divergence regions vanish.
template <typename DyT, int CH, bool KEEP, bool EXACT>
__global__ void __launch_bounds__(LNB_WARPS * 32, 1)
ln_bwd_ring_kernel(const DyT* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ mean_in,
                   const float* __restrict__ rstd_in, const float* __restrict__ gamma, float* __restrict__ g,
                   int add_to_g, float* __restrict__ dgamma, float* __restrict__ dbeta, DyT* __restrict__ dyn,
                   float* __restrict__ dyn_colsum, const Dropout drop, int M, int E, int nst) {
  extern __shared__ __align__(128) uint8_t lnb_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* s_gamma = reinterpret_cast<float*>(lnb_smem);
  float* s_acc = s_gamma + E;  // dgamma | dbeta | colsum(dyn)
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_acc + 3 * static_cast<size_t>(E));
  const uint32_t row_x = E * 4u, row_g = add_to_g ? E * 4u : 0u, row_dy = E * static_cast<uint32_t>(sizeof(DyT));
  const uint32_t stage_bytes = row_x + row_g + row_dy;
  uint64_t* bar = bars + warp * LNB_MAX_STAGES;
  const int nwarps = blockDim.x >> 5;   // 8, or 4 for wide rows (slot size)
  const uint32_t ring = smem_u32(bars + LNB_WARPS * LNB_MAX_STAGES) + warp * nst * stage_bytes;
  const uint32_t sgam = smem_u32(s_gamma);

  for (int i = threadIdx.x; i < 3 * E; i += blockDim.x) s_acc[i] = 0.f;
  if (lane == 0) {
    for (int st = 0; st < nst; ++st) mbar_init(bar + st, 1);
    fence_barrier_init();
  }
  pdl_wait();
  pdl_trigger();
  for (int i = threadIdx.x; i < E; i += blockDim.x) s_gamma[i] = __ldg(gamma + i);
  __syncthreads();

  const int E4 = E >> 2;
  const int w0 = blockIdx.x * nwarps + warp, stride = gridDim.x * nwarps;
  const int nrows = w0 < M ? (M - w0 + stride - 1) / stride : 0;
  auto issue = [&](int it) {  // lane 0
    const int st = it % nst;
    const int64_t off = (static_cast<int64_t>(w0) + static_cast<int64_t>(it) * stride) * E;
    const uint32_t dst = ring + st * stage_bytes;
    mbar_arrive_expect_tx(bar + st, stage_bytes);
    bulk_load_1d(dst, x + off, row_x, bar + st);
    if (add_to_g) bulk_load_1d(dst + row_x, g + off, row_g, bar + st);
    bulk_load_1d(dst + row_x + row_g, dy + off, row_dy, bar + st);
  };
  if (lane == 0)
    for (int it = 0; it < nst - 1 && it < nrows; ++it) issue(it);

  float rg[CH][4], rb[CH][4], rc[CH][4];
#pragma unroll
  for (int k = 0; k < CH; ++k)
#pragma unroll
    for (int j = 0; j < 4; ++j) { rg[k][j] = 0.f; rb[k][j] = 0.f; rc[k][j] = 0.f; }
  const bool masked = dyn != nullptr && drop.p > 0.f;
  const float inv_e = 1.f / static_cast<float>(E);
  float mean_l = 0.f, rstd_l = 0.f;

  for (int it = 0; it < nrows; ++it) {
    if ((it & 31) == 0) {  // statistics of this warp's next 32 rows, one per lane
      const int r = it + lane;
      if (r < nrows) {
        const int64_t m = static_cast<int64_t>(w0) + static_cast<int64_t>(r) * stride;
        mean_l = __ldg(mean_in + m);
        rstd_l = __ldg(rstd_in + m);
      }
    }
    if (lane == 0 && it + nst - 1 < nrows) issue(it + nst - 1);
    const float mean = __shfl_sync(0xffffffffu, mean_l, it & 31), rstd = __shfl_sync(0xffffffffu, rstd_l, it & 31);
    const int st = it % nst;
    const int64_t base = (static_cast<int64_t>(w0) + static_cast<int64_t>(it) * stride) * E;
    uint32_t keep[(CH + 1) / 2];
    if (masked) {  // independent of the loads: overlaps the wait below
#pragma unroll
      for (int c = 0; c < (CH + 1) / 2; ++c)
        keep[c] = dropout_keep8(drop, (static_cast<uint64_t>(base) >> 3) + lane + 32 * c);
    }
    mbar_wait(bar + st, (it / nst) & 1);
    const uint32_t sx = ring + st * stage_bytes, sg = sx + row_x, sdy = sg + row_g;

    constexpr int NK = KEEP ? CH : 1;
    float h[NK][4], d[NK][4];
    float s1 = 0.f, s2 = 0.f;
    // xhat and dy of chunk i of the current row, from the shared-memory slot
    auto row_chunk = [&](int i, float (&hh)[4], float (&dd)[4]) {
      const float4 xv = lds128(sx + 16 * i);
      if constexpr (sizeof(DyT) == 2) {
        const uint2 raw = lds64(sdy + 8 * i);
        dd[0] = __uint_as_float(raw.x << 16); dd[1] = __uint_as_float(raw.x & 0xffff0000u);
        dd[2] = __uint_as_float(raw.y << 16); dd[3] = __uint_as_float(raw.y & 0xffff0000u);
      } else {
        const float4 dv = lds128(sdy + 16 * i);
        dd[0] = dv.x; dd[1] = dv.y; dd[2] = dv.z; dd[3] = dv.w;
      }
      hh[0] = (xv.x - mean) * rstd; hh[1] = (xv.y - mean) * rstd;
      hh[2] = (xv.z - mean) * rstd; hh[3] = (xv.w - mean) * rstd;
    };
#pragma unroll
    for (int k = 0; k < CH; ++k) {
      const int i = lane + 32 * k;
      float hh[4] = {0.f, 0.f, 0.f, 0.f}, dd[4] = {0.f, 0.f, 0.f, 0.f};
      if (EXACT || i < E4) {
        row_chunk(i, hh, dd);
        const float4 gm = lds128(sgam + 16 * i);
        const float gg[4] = {gm.x, gm.y, gm.z, gm.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float t = dd[j] * gg[j];
          s1 += t;
          s2 = fmaf(t, hh[j], s2);
        }
      }
      if constexpr (KEEP) {
#pragma unroll
        for (int j = 0; j < 4; ++j) { h[k][j] = hh[j]; d[k][j] = dd[j]; }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      s1 += __shfl_xor_sync(0xffffffffu, s1, o);
      s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    }
    s1 *= inv_e;
    s2 *= inv_e;

#pragma unroll
    for (int k = 0; k < CH; ++k) {
      const int i = lane + 32 * k;
      uint32_t kb = 0xfu;
      if (masked) kb = __shfl_sync(0xffffffffu, keep[k >> 1], (k & 1) * 16 + (lane >> 1)) >> ((lane & 1) * 4);
      if (EXACT || i < E4) {
        const float4 gm = lds128(sgam + 16 * i);
        const float gg[4] = {gm.x, gm.y, gm.z, gm.w};
        float hh[4], dd[4];
        if constexpr (KEEP) {
#pragma unroll
          for (int j = 0; j < 4; ++j) { hh[j] = h[k][j]; dd[j] = d[k][j]; }
        } else {
          row_chunk(i, hh, dd);
        }
        float o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          o[j] = rstd * (dd[j] * gg[j] - s1 - hh[j] * s2);
          rg[k][j] = fmaf(dd[j], hh[j], rg[k][j]);
          rb[k][j] += dd[j];
        }
        if (add_to_g) {
          const float4 old = lds128(sg + 16 * i);
          o[0] += old.x; o[1] += old.y; o[2] += old.z; o[3] += old.w;
        }
        reinterpret_cast<float4*>(g + base)[i] = make_float4(o[0], o[1], o[2], o[3]);
        if (dyn != nullptr) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            o[j] = masked ? ((kb >> j) & 1u ? o[j] * drop.scale : 0.f) : o[j];
            rc[k][j] += o[j];
          }
          if constexpr (sizeof(DyT) == 2) {
            uint2 pk;
            pk.x = pack_bf16x2(o[0], o[1]);
            pk.y = pack_bf16x2(o[2], o[3]);
            reinterpret_cast<uint2*>(dyn + base)[i] = pk;
          } else {
            reinterpret_cast<float4*>(dyn + base)[i] = make_float4(o[0], o[1], o[2], o[3]);
          }
        }
      }
    }
    __syncwarp();  // every lane is done with this slot before lane 0 refills it in the next iteration
  }

#pragma unroll
  for (int k = 0; k < CH; ++k) {
    const int i = lane + 32 * k;
    if (EXACT || i < E4) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        atomicAdd(&s_acc[i * 4 + j], rg[k][j]);
        atomicAdd(&s_acc[E + i * 4 + j], rb[k][j]);
        if (dyn_colsum) atomicAdd(&s_acc[2 * E + i * 4 + j], rc[k][j]);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < E; i += blockDim.x) {
    if (dgamma) atomicAdd(dgamma + i, s_acc[i]);
    if (dbeta) atomicAdd(dbeta + i, s_acc[E + i]);
    if (dyn_colsum) atomicAdd(dyn_colsum + i, s_acc[2 * E + i]);
  }
}

// warps per CTA such that the per-warp row caches stay under ~96 KB of shared memory
int ln_warps(int E, int rows_per_warp, int fixed_rows) {
  int w = (96 * 1024 / (E * 4) - fixed_rows) / rows_per_warp;
  if (w > LN_WARPS) w = LN_WARPS;
  if (w < 1) w = 1;
  return w;
}
int ln_grid(int M, int nw, int per_sm) {
  const int blocks_needed = ceil_div(M, nw);
  const int cap = num_sms() * per_sm;
  return blocks_needed < cap ? blocks_needed : cap;
}

}  // namespace

int ln_forward(const float* x, const float* gamma, const float* beta, void* y, bool y_bf16, float* mean,
               float* rstd, int M, int E, float eps, cudaStream_t stream) {
  VB_REQUIRE(M > 0 && E > 0, "ln_forward: empty input");
  static const bool reg_path = !(getenv("VITB200_LN_REG") && atoi(getenv("VITB200_LN_REG")) == 0);
  if (reg_path && (E & 3) == 0 && E <= 128 * 12 && aligned16(x) && aligned16(y) && aligned16(gamma) &&
      aligned16(beta) && M >= 4 * num_sms()) {
    ProfScope prof(kProfLnFwd, static_cast<double>(M) * E * (4 + (y_bf16 ? 2 : 4)) + 8.0 * M, 1, stream);
    const int vpl = ceil_div(E, 128);
    // ---- bulk-copy ring kernel (VITB200_LN_FWD_RING=0 falls back to the register-resident kernel below)
    static const bool ring_path = !(getenv("VITB200_LN_FWD_RING") && atoi(getenv("VITB200_LN_FWD_RING")) == 0);
    const size_t lnf_fixed = LNF_WARPS * LNF_MAX_STAGES * 8;
    int nst = static_cast<int>((208 * 1024 - lnf_fixed) / (static_cast<size_t>(LNF_WARPS) * E * 4));
    if (nst > LNF_MAX_STAGES) nst = LNF_MAX_STAGES;
    if (ring_path && nst >= 2 && M >= 4 * LNF_WARPS) {
      const size_t smem = lnf_fixed + static_cast<size_t>(nst) * LNF_WARPS * E * 4;
      const int grid = std::min(num_sms(), ceil_div(M, LNF_WARPS));
#define VB_LNF_RING(T, V)                                                                                         \
  do {                                                                                                            \
    static bool attr_set = false;                                                                                 \
    if (!attr_set) {                                                                                              \
      VB_CHECK_CUDA(cudaFuncSetAttribute(ln_fwd_ring_kernel<T, V, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                         208 * 1024));                                                            \
      VB_CHECK_CUDA(cudaFuncSetAttribute(ln_fwd_ring_kernel<T, V, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                         208 * 1024));                                                            \
      attr_set = true;                                                                                            \
    }                                                                                                             \
    if (E == 128 * V)                                                                                             \
      VB_CHECK_CUDA(launch_pdl(ln_fwd_ring_kernel<T, V, true>, dim3(grid), dim3(LNF_WARPS * 32), smem, stream, 1, x, \
                               gamma, beta, static_cast<T*>(y), mean, rstd, M, E, eps, nst));                     \
    else                                                                                                          \
      VB_CHECK_CUDA(launch_pdl(ln_fwd_ring_kernel<T, V, false>, dim3(grid), dim3(LNF_WARPS * 32), smem, stream, 1, x, \
                               gamma, beta, static_cast<T*>(y), mean, rstd, M, E, eps, nst));                     \
  } while (0)
#define VB_LNF_RING_V(T)                                                                                          \
  do {                                                                                                            \
    if (vpl <= 3) VB_LNF_RING(T, 3); else if (vpl <= 6) VB_LNF_RING(T, 6); else if (vpl <= 8) VB_LNF_RING(T, 8);   \
    else VB_LNF_RING(T, 12);                                                                                      \
  } while (0)
      if (y_bf16) VB_LNF_RING_V(bf16); else VB_LNF_RING_V(float);
#undef VB_LNF_RING_V
#undef VB_LNF_RING
      VB_CHECK_LAUNCH();
      return kOk;
    }
    // register-resident rows, next row prefetched (the block LayerNorms; the [B, 4E] head LayerNorm stays below)
    const int per_sm = vpl <= 3 ? 4 : vpl <= 6 ? 2 : 1;  // CTAs resident per SM by register count
    const int grid = ln_grid(M, LN_WARPS, per_sm);
#define VB_LN_REG(V)                                                                                              \
  case V:                                                                                                         \
    if (y_bf16) VB_CHECK_CUDA(launch_pdl(ln_fwd_reg_kernel<bf16, V>, dim3(grid), dim3(LN_THREADS), 0, stream, 1, x, \
                                         gamma, beta, static_cast<bf16*>(y), mean, rstd, M, E, eps));             \
    else VB_CHECK_CUDA(launch_pdl(ln_fwd_reg_kernel<float, V>, dim3(grid), dim3(LN_THREADS), 0, stream, 1, x, gamma, \
                                  beta, static_cast<float*>(y), mean, rstd, M, E, eps));                          \
    break;
    switch (vpl <= 3 ? 3 : vpl <= 6 ? 6 : vpl <= 8 ? 8 : 12) {
      VB_LN_REG(3) VB_LN_REG(6) VB_LN_REG(8) VB_LN_REG(12)
    }
#undef VB_LN_REG
    VB_CHECK_LAUNCH();
    return kOk;
  }
  const int nw = ln_warps(E, 1, 0);
  const size_t smem = static_cast<size_t>(nw) * E * sizeof(float);
  VB_REQUIRE(smem <= 200 * 1024, "ln_forward: row too long (E=%d)", E);
  const int grid = ln_grid(M, nw, 4);
  ProfScope prof(kProfLnFwd, static_cast<double>(M) * E * (4 + (y_bf16 ? 2 : 4)) + 8.0 * M, 1, stream);
  if (y_bf16) {
    if (smem > 48 * 1024) VB_CHECK_CUDA(cudaFuncSetAttribute(ln_fwd_kernel<bf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    VB_CHECK_CUDA(launch_pdl(ln_fwd_kernel<bf16>, dim3(grid), dim3(nw * 32), smem, stream, 1, x, gamma, beta, static_cast<bf16*>(y), mean, rstd, M, E, eps));
  } else {
    if (smem > 48 * 1024) VB_CHECK_CUDA(cudaFuncSetAttribute(ln_fwd_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    VB_CHECK_CUDA(launch_pdl(ln_fwd_kernel<float>, dim3(grid), dim3(nw * 32), smem, stream, 1, x, gamma, beta, static_cast<float*>(y), mean, rstd, M, E, eps));
  }
  VB_CHECK_LAUNCH();
  return kOk;
}

int ln_backward(const void* dy, bool dy_bf16, const float* x, const float* mean, const float* rstd,
                const float* gamma, float* g, int add_to_g, float* dgamma, float* dbeta, void* dyn,
                float* dyn_colsum, const Dropout& drop, int M, int E, cudaStream_t stream) {
  VB_REQUIRE(M > 0 && E > 0, "ln_backward: empty input");
  VB_REQUIRE(dyn_colsum == nullptr || dyn != nullptr, "ln_backward: dyn_colsum needs dyn");
  const double es = dy_bf16 ? 2 : 4;  // bytes: x read, dy read, g write (+read when add_to_g), dyn write, stats
  ProfScope prof(kProfLnBwd, static_cast<double>(M) * E * (4 + es + 4 + (add_to_g ? 4 : 0) + (dyn ? es : 0)) + 8.0 * M,
                 1, stream);
  // ---- ring kernel
  if ((E & 7) == 0 && E <= 1536 && M >= 4 * LNB_WARPS) {
    static const int wide_warps = getenv("VITB200_LNB_WIDE_WARPS") ? atoi(getenv("VITB200_LNB_WIDE_WARPS")) : 6;
    const int nwarps = E <= 1024 ? LNB_WARPS : wide_warps;   // wide rows: fewer, deeper rings (6 warps: 4.15 -> 4.80 TB/s at E=1280 with dropout live)
    const size_t stage = static_cast<size_t>(E) * (4 + (add_to_g ? 4 : 0) + (dy_bf16 ? 2 : 4));
    const size_t fixed = static_cast<size_t>(E) * 16 + LNB_WARPS * LNB_MAX_STAGES * 8;
    int nst = static_cast<int>((216 * 1024 - fixed) / (nwarps * stage));
    if (nst > LNB_MAX_STAGES) nst = LNB_MAX_STAGES;
    if (nst >= 2) {
      const size_t smem = fixed + static_cast<size_t>(nst) * nwarps * stage;
      const int ch = ceil_div(E, 128);
      const int grid = std::min(num_sms(), ceil_div(M, nwarps));
      static const bool exact_on = !(getenv("VITB200_LNB_EXACT") && atoi(getenv("VITB200_LNB_EXACT")) == 0);
#define VB_LN_RING(T, C)                                                                                          \
  do {                                                                                                            \
    static bool attr_set = false; /* one flag per instantiation: the opt-in is the 216 KB bound, not this call's size */ \
    if (!attr_set) {                                                                                              \
      VB_CHECK_CUDA(cudaFuncSetAttribute(ln_bwd_ring_kernel<T, C, (C <= 8), false>,                               \
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024));               \
      VB_CHECK_CUDA(cudaFuncSetAttribute(ln_bwd_ring_kernel<T, C, (C <= 8), true>,                                \
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024));               \
      attr_set = true;                                                                                            \
    }                                                                                                             \
    if (exact_on && E == 128 * C)                                                                                 \
      VB_CHECK_CUDA(launch_pdl(ln_bwd_ring_kernel<T, C, (C <= 8), true>, dim3(grid), dim3(nwarps * 32), smem, stream, 1, \
                               static_cast<const T*>(dy), x, mean, rstd, gamma, g, add_to_g, dgamma, dbeta,       \
                               static_cast<T*>(dyn), dyn_colsum, drop, M, E, nst));                               \
    else                                                                                                          \
      VB_CHECK_CUDA(launch_pdl(ln_bwd_ring_kernel<T, C, (C <= 8), false>, dim3(grid), dim3(nwarps * 32), smem, stream, 1, \
                               static_cast<const T*>(dy), x, mean, rstd, gamma, g, add_to_g, dgamma, dbeta,       \
                               static_cast<T*>(dyn), dyn_colsum, drop, M, E, nst));                               \
  } while (0)
#define VB_LN_RING_CH(T)                                                                                          \
  do {                                                                                                            \
    if (ch <= 1) VB_LN_RING(T, 1); else if (ch <= 2) VB_LN_RING(T, 2); else if (ch <= 3) VB_LN_RING(T, 3);        \
    else if (ch <= 4) VB_LN_RING(T, 4); else if (ch <= 6) VB_LN_RING(T, 6); else if (ch <= 8) VB_LN_RING(T, 8);   \
    else if (ch <= 10) VB_LN_RING(T, 10); else VB_LN_RING(T, 12);                                                 \
  } while (0)
      if (dy_bf16) VB_LN_RING_CH(bf16); else VB_LN_RING_CH(float);
#undef VB_LN_RING_CH
#undef VB_LN_RING
      VB_CHECK_LAUNCH();
      return kOk;
    }
  }
  // ---- generic kernel (odd E, very wide rows, tiny M)
  const int nw = ln_warps(E, 2, 2);
  const size_t smem = (2 + 2 * static_cast<size_t>(nw)) * E * sizeof(float);
  VB_REQUIRE(smem <= 200 * 1024, "ln_backward: row too long (E=%d)", E);
  // fewer, fatter CTAs than forward: each CTA ends with 2*E global atomics
  const int grid = ln_grid(M, nw, 2);
  const int vpl = ((E & 3) == 0 && E <= 512) ? 4 : ((E & 3) == 0 && E <= 1024) ? 8 : ((E & 3) == 0 && E <= 1536) ? 12 : 0;
#define VB_LN_BWD(T, V)                                                                                          \
  do {                                                                                                           \
    if (smem > 48 * 1024)                                                                                        \
      VB_CHECK_CUDA(cudaFuncSetAttribute(ln_bwd_kernel<T, V>, cudaFuncAttributeMaxDynamicSharedMemorySize,       \
                                         static_cast<int>(smem)));                                               \
    VB_CHECK_CUDA(launch_pdl(ln_bwd_kernel<T, V>, dim3(grid), dim3(nw * 32), smem, stream, 1,                     \
                             static_cast<const T*>(dy), x, mean, rstd, gamma, g, add_to_g, dgamma, dbeta,        \
                             static_cast<T*>(dyn), drop, M, E));                                                 \
  } while (0)
  if (dy_bf16) {
    if (vpl == 4) VB_LN_BWD(bf16, 4); else if (vpl == 8) VB_LN_BWD(bf16, 8); else if (vpl == 12) VB_LN_BWD(bf16, 12); else VB_LN_BWD(bf16, 0);
  } else {
    if (vpl == 4) VB_LN_BWD(float, 4); else if (vpl == 8) VB_LN_BWD(float, 8); else if (vpl == 12) VB_LN_BWD(float, 12); else VB_LN_BWD(float, 0);
  }
#undef VB_LN_BWD
  VB_CHECK_LAUNCH();
  if (dyn_colsum) VB_PROPAGATE(colsum(dyn, dy_bf16, E, dyn_colsum, M, E, stream));
  return kOk;
}

}  // namespace vb
