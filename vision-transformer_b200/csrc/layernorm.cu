// LayerNorm forward / backward as coalesced, vectorised bandwidth kernels (one warp per row, the row staged
// once in shared memory so HBM sees each element exactly once).
//   forward : x fp32 [M,E] -> y (fp32|bf16) [M,E], saves mean / rstd                       (transformer.py:72-73,77-78; vit.py:72)
//   backward: dx fused with the residual-gradient accumulate (g += dx), dgamma/dbeta reduced per CTA in
//             shared memory then atomically into fp32 grads, and optionally emits the dropout-masked,
//             down-cast gradient the next GEMM consumes (so no separate cast/mask pass exists).
#include "common.cuh"

namespace vb {

namespace {

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
  const int nw = ln_warps(E, 1, 0);
  const size_t smem = static_cast<size_t>(nw) * E * sizeof(float);
  VB_REQUIRE(smem <= 200 * 1024, "ln_forward: row too long (E=%d)", E);
  const int grid = ln_grid(M, nw, 4);
  ProfScope prof(kProfLnFwd, static_cast<double>(M) * E * (4 + (y_bf16 ? 2 : 4)) + 8.0 * M, 1, stream);
  if (y_bf16) {
    if (smem > 48 * 1024) VB_CHECK_CUDA(cudaFuncSetAttribute(ln_fwd_kernel<bf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    ln_fwd_kernel<bf16><<<grid, nw * 32, smem, stream>>>(x, gamma, beta, static_cast<bf16*>(y), mean, rstd, M, E, eps);
  } else {
    if (smem > 48 * 1024) VB_CHECK_CUDA(cudaFuncSetAttribute(ln_fwd_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    ln_fwd_kernel<float><<<grid, nw * 32, smem, stream>>>(x, gamma, beta, static_cast<float*>(y), mean, rstd, M, E, eps);
  }
  VB_CHECK_LAUNCH();
  return kOk;
}

int ln_backward(const void* dy, bool dy_bf16, const float* x, const float* mean, const float* rstd,
                const float* gamma, float* g, int add_to_g, float* dgamma, float* dbeta, void* dyn,
                const Dropout& drop, int M, int E, cudaStream_t stream) {
  VB_REQUIRE(M > 0 && E > 0, "ln_backward: empty input");
  const int nw = ln_warps(E, 2, 2);
  const size_t smem = (2 + 2 * static_cast<size_t>(nw)) * E * sizeof(float);
  VB_REQUIRE(smem <= 200 * 1024, "ln_backward: row too long (E=%d)", E);
  // fewer, fatter CTAs than forward: each CTA ends with 2*E global atomics
  const int grid = ln_grid(M, nw, 2);
  const double es = dy_bf16 ? 2 : 4;  // bytes: x read, dy read, g write (+read when add_to_g), dyn write, stats
  ProfScope prof(kProfLnBwd, static_cast<double>(M) * E * (4 + es + 4 + (add_to_g ? 4 : 0) + (dyn ? es : 0)) + 8.0 * M,
                 1, stream);
  const int vpl = ((E & 3) == 0 && E <= 512) ? 4 : ((E & 3) == 0 && E <= 1024) ? 8 : ((E & 3) == 0 && E <= 1536) ? 12 : 0;
#define VB_LN_BWD(T, V)                                                                                          \
  do {                                                                                                           \
    if (smem > 48 * 1024)                                                                                        \
      VB_CHECK_CUDA(cudaFuncSetAttribute(ln_bwd_kernel<T, V>, cudaFuncAttributeMaxDynamicSharedMemorySize,       \
                                         static_cast<int>(smem)));                                               \
    ln_bwd_kernel<T, V><<<grid, nw * 32, smem, stream>>>(static_cast<const T*>(dy), x, mean, rstd, gamma, g,  \
                                                            add_to_g, dgamma, dbeta, static_cast<T*>(dyn), drop, \
                                                            M, E);                                               \
  } while (0)
  if (dy_bf16) {
    if (vpl == 4) VB_LN_BWD(bf16, 4); else if (vpl == 8) VB_LN_BWD(bf16, 8); else if (vpl == 12) VB_LN_BWD(bf16, 12); else VB_LN_BWD(bf16, 0);
  } else {
    if (vpl == 4) VB_LN_BWD(float, 4); else if (vpl == 8) VB_LN_BWD(float, 8); else if (vpl == 12) VB_LN_BWD(float, 12); else VB_LN_BWD(float, 0);
  }
#undef VB_LN_BWD
  VB_CHECK_LAUNCH();
  return kOk;
}

}  // namespace vb
