// fp32 self-attention (forward, backward, probability export) on CUDA cores, any head size / sequence length.
// This is the fp32-mode path (1e-4 parity gate, head_size 1 in the reference's default config) and the debug
// path that materialises `attention_probs`; the bf16 training path uses the tensor-core kernel in
// attention_bf16.cu.  Semantics follow reference transformer.py:20-31: logits = (q k^T) * scale with
// scale = +sqrt(head_size) (MULTIPLIED), softmax over keys, out = wei v; heads are column slices h*hs of the
// packed [M, 3E] qkv buffer (q | k | v), output written at column h*hs of [M, E] (the torch.cat at :45).
#include "common.cuh"
#include "kernels.h"

namespace vb {

namespace {

constexpr int AF_WARPS = 8;

// One warp per query row.  smem per warp: q[hs] + p[T].
template <bool WRITE_PROBS>
__global__ void __launch_bounds__(AF_WARPS * 32)
attn_f32_fwd_kernel(const float* __restrict__ qkv, float* __restrict__ out, float* __restrict__ lse,
                    float* __restrict__ probs, int B, int T, int H, int hs, float scale) {
  extern __shared__ float af_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int E = H * hs;
  const int bh = blockIdx.x, b = bh / H, h = bh % H;
  const int i = blockIdx.y * AF_WARPS + warp;
  if (i >= T) return;
  float* qs = af_smem + static_cast<size_t>(warp) * (hs + T);
  float* ps = qs + hs;
  const int64_t row3 = 3 * static_cast<int64_t>(E);
  const float* qrow = qkv + (static_cast<int64_t>(b) * T + i) * row3 + h * hs;
  const float* kbase = qkv + static_cast<int64_t>(b) * T * row3 + E + h * hs;
  const float* vbase = kbase + E;
  for (int d = lane; d < hs; d += 32) qs[d] = qrow[d];
  __syncwarp();
  float mx = -INFINITY;
  for (int j = lane; j < T; j += 32) {
    const float* kr = kbase + j * row3;
    float s = 0.f;
    for (int d = 0; d < hs; ++d) s = fmaf(qs[d], kr[d], s);
    s *= scale;
    ps[j] = s;
    mx = fmaxf(mx, s);
  }
  mx = warp_max(mx);
  float sum = 0.f;
  for (int j = lane; j < T; j += 32) {
    const float e = expf(ps[j] - mx);
    ps[j] = e;
    sum += e;
  }
  sum = warp_sum(sum);
  const float inv = 1.f / sum;
  __syncwarp();
  if (WRITE_PROBS) {
    float* pr = probs + (static_cast<int64_t>(bh) * T + i) * T;
    for (int j = lane; j < T; j += 32) pr[j] = ps[j] * inv;
  } else {
    if (lane == 0 && lse != nullptr) lse[static_cast<int64_t>(bh) * T + i] = mx + logf(sum);
    float* orow = out + (static_cast<int64_t>(b) * T + i) * E + h * hs;
    for (int d = lane; d < hs; d += 32) {
      float acc = 0.f;
      for (int j = 0; j < T; ++j) acc = fmaf(ps[j], vbase[j * row3 + d], acc);
      orow[d] = acc * inv;
    }
  }
}

// delta[bh, i] = sum_d dO[i, d] * O[i, d]
template <typename T_>
__global__ void attn_delta_kernel(const T_* __restrict__ o, const T_* __restrict__ d_o, float* __restrict__ delta,
                                  int B, int T, int H, int hs) {
  const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int total = B * H * T;
  if (warp_global >= total) return;
  const int i = warp_global % T, bh = warp_global / T, b = bh / H, h = bh % H;
  const int64_t off = (static_cast<int64_t>(b) * T + i) * (H * hs) + h * hs;
  float s = 0.f;
  for (int d = lane; d < hs; d += 32) s = fmaf(to_f32(o[off + d]), to_f32(d_o[off + d]), s);
  s = warp_sum(s);
  if (lane == 0) delta[static_cast<int64_t>(bh) * T + i] = s;
}

// Vectorised variant: one thread per 16-byte chunk, CPH consecutive lanes make up one (token, head) slice
// (CPH = hs * sizeof(T) / 16, a power of two), partial dot products combined with CPH-wide shuffles.  Fully
// coalesced: the kernel streams o and dO exactly once.
template <typename T_, int CPH>
__global__ void __launch_bounds__(256)
attn_delta_vec_kernel(const uint4* __restrict__ o, const uint4* __restrict__ d_o, float* __restrict__ delta,
                      int64_t total_chunks, int T, int H) {
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  pdl_wait();
  pdl_trigger();
  for (int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; c < total_chunks + (threadIdx.x & 31);
       c += stride) {  // whole warps stay in the loop (shuffles); blockDim and stride are multiples of CPH
    float s = 0.f;
    if (c < total_chunks) {
      const uint4 a = __ldcs(o + c), b = __ldcs(d_o + c);
      if constexpr (sizeof(T_) == 2) {
        const uint32_t av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          s = fmaf(__uint_as_float(av[j] << 16), __uint_as_float(bv[j] << 16), s);
          s = fmaf(__uint_as_float(av[j] & 0xffff0000u), __uint_as_float(bv[j] & 0xffff0000u), s);
        }
      } else {
        s = __uint_as_float(a.x) * __uint_as_float(b.x) + __uint_as_float(a.y) * __uint_as_float(b.y) +
            __uint_as_float(a.z) * __uint_as_float(b.z) + __uint_as_float(a.w) * __uint_as_float(b.w);
      }
    }
#pragma unroll
    for (int o2 = CPH / 2; o2 > 0; o2 >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o2);
    if (c < total_chunks && (c & (CPH - 1)) == 0) {
      const int64_t hr = c / CPH;                 // (token, head)
      const int64_t tok = hr / H;
      const int h = static_cast<int>(hr - tok * H);
      const int64_t b = tok / T;
      const int i = static_cast<int>(tok - b * T);
      delta[(b * H + h) * T + i] = s;
    }
  }
}

// Head slices that are a whole number of 16-byte chunks but not a power of two of them (head_size 80 / 96 in bf16):
// one thread owns one (token, head) slice, consecutive threads consecutive heads, so a warp still sweeps one
// contiguous span of o and dO.
template <typename T_>
__global__ void __launch_bounds__(256)
attn_delta_slice_kernel(const uint4* __restrict__ o, const uint4* __restrict__ d_o, float* __restrict__ delta,
                        int64_t slices, int cph, int T, int H) {
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  pdl_wait();
  pdl_trigger();
  for (int64_t hr = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; hr < slices; hr += stride) {
    const uint4* a4 = o + hr * cph;
    const uint4* b4 = d_o + hr * cph;
    float s0 = 0.f, s1 = 0.f;
#pragma unroll 2
    for (int j = 0; j < cph; ++j) {
      const uint4 a = __ldg(a4 + j), b = __ldg(b4 + j);
      if constexpr (sizeof(T_) == 2) {
        const uint32_t av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          s0 = fmaf(__uint_as_float(av[k] << 16), __uint_as_float(bv[k] << 16), s0);
          s1 = fmaf(__uint_as_float(av[k] & 0xffff0000u), __uint_as_float(bv[k] & 0xffff0000u), s1);
        }
      } else {
        s0 = fmaf(__uint_as_float(a.x), __uint_as_float(b.x), s0);
        s1 = fmaf(__uint_as_float(a.y), __uint_as_float(b.y), s1);
        s0 = fmaf(__uint_as_float(a.z), __uint_as_float(b.z), s0);
        s1 = fmaf(__uint_as_float(a.w), __uint_as_float(b.w), s1);
      }
    }
    const int64_t tok = hr / H;
    const int h = static_cast<int>(hr - tok * H);
    const int64_t b = tok / T;
    const int i = static_cast<int>(tok - b * T);
    delta[(b * H + h) * T + i] = s0 + s1;
  }
}

// dQ: one warp per query row i.  ds_ij = p_ij (dP_ij - delta_i) * scale ; dq_i = sum_j ds_ij k_j
__global__ void __launch_bounds__(AF_WARPS * 32)
attn_f32_bwd_dq_kernel(const float* __restrict__ qkv, const float* __restrict__ d_o, const float* __restrict__ lse,
                       const float* __restrict__ delta, float* __restrict__ dqkv, int B, int T, int H, int hs,
                       float scale) {
  extern __shared__ float af_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int E = H * hs;
  const int bh = blockIdx.x, b = bh / H, h = bh % H;
  const int i = blockIdx.y * AF_WARPS + warp;
  if (i >= T) return;
  float* qs = af_smem + static_cast<size_t>(warp) * (2 * hs + T);
  float* dos = qs + hs;
  float* ds = dos + hs;
  const int64_t row3 = 3 * static_cast<int64_t>(E);
  const int64_t tok = static_cast<int64_t>(b) * T + i;
  const float* kbase = qkv + static_cast<int64_t>(b) * T * row3 + E + h * hs;
  const float* vbase = kbase + E;
  for (int d = lane; d < hs; d += 32) {
    qs[d] = qkv[tok * row3 + h * hs + d];
    dos[d] = d_o[tok * E + h * hs + d];
  }
  __syncwarp();
  const float l = lse[static_cast<int64_t>(bh) * T + i], dl = delta[static_cast<int64_t>(bh) * T + i];
  for (int j = lane; j < T; j += 32) {
    const float* kr = kbase + j * row3;
    const float* vr = vbase + j * row3;
    float s = 0.f, dp = 0.f;
    for (int d = 0; d < hs; ++d) { s = fmaf(qs[d], kr[d], s); dp = fmaf(dos[d], vr[d], dp); }
    const float p = expf(s * scale - l);
    ds[j] = p * (dp - dl) * scale;
  }
  __syncwarp();
  for (int d = lane; d < hs; d += 32) {
    float acc = 0.f;
    for (int j = 0; j < T; ++j) acc = fmaf(ds[j], kbase[j * row3 + d], acc);
    dqkv[tok * row3 + h * hs + d] = acc;
  }
}

// dK, dV: one warp per key row j.  dv_j = sum_i p_ij dO_i ; dk_j = sum_i ds_ij q_i
__global__ void __launch_bounds__(AF_WARPS * 32)
attn_f32_bwd_dkv_kernel(const float* __restrict__ qkv, const float* __restrict__ d_o, const float* __restrict__ lse,
                        const float* __restrict__ delta, float* __restrict__ dqkv, int B, int T, int H, int hs,
                        float scale) {
  extern __shared__ float af_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int E = H * hs;
  const int bh = blockIdx.x, b = bh / H, h = bh % H;
  const int j = blockIdx.y * AF_WARPS + warp;
  if (j >= T) return;
  float* ks = af_smem + static_cast<size_t>(warp) * (2 * hs + 2 * T);
  float* vs = ks + hs;
  float* ps = vs + hs;
  float* ds = ps + T;
  const int64_t row3 = 3 * static_cast<int64_t>(E);
  const int64_t tok = static_cast<int64_t>(b) * T + j;
  const float* qbase = qkv + static_cast<int64_t>(b) * T * row3 + h * hs;
  const float* dobase = d_o + static_cast<int64_t>(b) * T * E + h * hs;
  for (int d = lane; d < hs; d += 32) {
    ks[d] = qkv[tok * row3 + E + h * hs + d];
    vs[d] = qkv[tok * row3 + 2 * E + h * hs + d];
  }
  __syncwarp();
  for (int i = lane; i < T; i += 32) {
    const float* qr = qbase + i * row3;
    const float* dor = dobase + static_cast<int64_t>(i) * E;
    float s = 0.f, dp = 0.f;
    for (int d = 0; d < hs; ++d) { s = fmaf(qr[d], ks[d], s); dp = fmaf(dor[d], vs[d], dp); }
    const float p = expf(s * scale - lse[static_cast<int64_t>(bh) * T + i]);
    ps[i] = p;
    ds[i] = p * (dp - delta[static_cast<int64_t>(bh) * T + i]) * scale;
  }
  __syncwarp();
  for (int d = lane; d < hs; d += 32) {
    float dk = 0.f, dv = 0.f;
    for (int i = 0; i < T; ++i) {
      dk = fmaf(ds[i], qbase[i * row3 + d], dk);
      dv = fmaf(ps[i], dobase[static_cast<int64_t>(i) * E + d], dv);
    }
    dqkv[tok * row3 + E + h * hs + d] = dk;
    dqkv[tok * row3 + 2 * E + h * hs + d] = dv;
  }
}

template <typename K>
int set_smem(K kern, size_t smem) {
  VB_REQUIRE(smem <= 220 * 1024, "attention_f32: sequence too long for the fp32 path (%zu B smem)", smem);
  if (smem > 48 * 1024) VB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  return kOk;
}

}  // namespace

int attn_delta(const void* o, const void* d_o, bool is_bf16, float* delta, int B, int T, int H, int hs,
               cudaStream_t stream) {
  ProfScope prof(kProfAttnBwd, 0.0, 1, stream);
  const int cph = hs * (is_bf16 ? 2 : 4) / 16;
  if (hs * (is_bf16 ? 2 : 4) % 16 == 0 && (cph & (cph - 1)) == 0 && cph >= 1 && cph <= 32 && aligned16(o) &&
      aligned16(d_o)) {
    const int64_t chunks = static_cast<int64_t>(B) * T * H * cph;
    const int grid_v = static_cast<int>(std::min<int64_t>((chunks + 255) / 256, static_cast<int64_t>(num_sms()) * 16));
    const uint4* o4 = static_cast<const uint4*>(o);
    const uint4* d4 = static_cast<const uint4*>(d_o);
#define VB_DELTA(T, C) \
  VB_CHECK_CUDA(launch_pdl(attn_delta_vec_kernel<T, C>, dim3(grid_v), dim3(256), 0, stream, 1, o4, d4, delta, chunks, T_, H))
    const int T_ = T;
    if (is_bf16) {
      switch (cph) {
        case 1: VB_DELTA(bf16, 1); break; case 2: VB_DELTA(bf16, 2); break; case 4: VB_DELTA(bf16, 4); break;
        case 8: VB_DELTA(bf16, 8); break; case 16: VB_DELTA(bf16, 16); break; default: VB_DELTA(bf16, 32); break;
      }
    } else {
      switch (cph) {
        case 1: VB_DELTA(float, 1); break; case 2: VB_DELTA(float, 2); break; case 4: VB_DELTA(float, 4); break;
        case 8: VB_DELTA(float, 8); break; case 16: VB_DELTA(float, 16); break; default: VB_DELTA(float, 32); break;
      }
    }
#undef VB_DELTA
    VB_CHECK_LAUNCH();
    return kOk;
  }
  if (hs * (is_bf16 ? 2 : 4) % 16 == 0 && cph >= 1 && aligned16(o) && aligned16(d_o)) {
    const int64_t slices = static_cast<int64_t>(B) * T * H;
    const int grid_s = static_cast<int>(std::min<int64_t>((slices + 255) / 256, static_cast<int64_t>(num_sms()) * 16));
    const uint4* o4 = static_cast<const uint4*>(o);
    const uint4* d4 = static_cast<const uint4*>(d_o);
    if (is_bf16)
      VB_CHECK_CUDA(launch_pdl(attn_delta_slice_kernel<bf16>, dim3(grid_s), dim3(256), 0, stream, 1, o4, d4, delta,
                               slices, cph, T, H));
    else
      VB_CHECK_CUDA(launch_pdl(attn_delta_slice_kernel<float>, dim3(grid_s), dim3(256), 0, stream, 1, o4, d4, delta,
                               slices, cph, T, H));
    VB_CHECK_LAUNCH();
    return kOk;
  }
  const int total_warps = B * H * T;
  const int threads = 256;
  const int grid = ceil_div(total_warps * 32, threads);
  if (is_bf16) attn_delta_kernel<bf16><<<grid, threads, 0, stream>>>(static_cast<const bf16*>(o), static_cast<const bf16*>(d_o), delta, B, T, H, hs);
  else attn_delta_kernel<float><<<grid, threads, 0, stream>>>(static_cast<const float*>(o), static_cast<const float*>(d_o), delta, B, T, H, hs);
  VB_CHECK_LAUNCH();
  return kOk;
}

int attn_f32_forward(const float* qkv, float* out, float* lse, int B, int T, int H, int hs, float scale,
                     cudaStream_t stream) {
  if (attn_f32_tiled_supported(T, hs, qkv, out)) {
    ProfScope prof(kProfAttnFwd, 4.0 * B * H * T * T * hs, 1, stream);
    return attn_f32_tiled_forward(qkv, out, lse, nullptr, B, T, H, hs, scale, stream);
  }
  const size_t smem = static_cast<size_t>(AF_WARPS) * (hs + T) * sizeof(float);
  VB_PROPAGATE(set_smem(attn_f32_fwd_kernel<false>, smem));
  dim3 grid(B * H, ceil_div(T, AF_WARPS));
  ProfScope prof(kProfAttnFwd, 4.0 * B * H * T * T * hs, 1, stream);
  attn_f32_fwd_kernel<false><<<grid, AF_WARPS * 32, smem, stream>>>(qkv, out, lse, nullptr, B, T, H, hs, scale);
  VB_CHECK_LAUNCH();
  return kOk;
}

int attn_f32_probs(const float* qkv, float* probs, int B, int T, int H, int hs, float scale, cudaStream_t stream) {
  if (attn_f32_tiled_supported(T, hs, qkv, probs)) {
    ProfScope prof(kProfAttnFwd, 2.0 * B * H * T * T * hs, 1, stream);
    return attn_f32_tiled_forward(qkv, nullptr, nullptr, probs, B, T, H, hs, scale, stream);
  }
  const size_t smem = static_cast<size_t>(AF_WARPS) * (hs + T) * sizeof(float);
  VB_PROPAGATE(set_smem(attn_f32_fwd_kernel<true>, smem));
  dim3 grid(B * H, ceil_div(T, AF_WARPS));
  ProfScope prof(kProfAttnFwd, 2.0 * B * H * T * T * hs, 1, stream);
  attn_f32_fwd_kernel<true><<<grid, AF_WARPS * 32, smem, stream>>>(qkv, nullptr, nullptr, probs, B, T, H, hs, scale);
  VB_CHECK_LAUNCH();
  return kOk;
}

int attn_f32_backward(const float* qkv, const float* out, const float* d_o, const float* lse, float* delta,
                      float* dqkv, int B, int T, int H, int hs, float scale, cudaStream_t stream) {
  VB_PROPAGATE(attn_delta(out, d_o, false, delta, B, T, H, hs, stream));
  if (attn_f32_tiled_supported(T, hs, qkv, d_o) && aligned16(dqkv)) {
    ProfScope prof(kProfAttnBwd, 8.0 * B * H * T * T * hs, 2, stream);
    return attn_f32_tiled_backward(qkv, d_o, lse, delta, dqkv, B, T, H, hs, scale, stream);
  }
  dim3 grid(B * H, ceil_div(T, AF_WARPS));
  ProfScope prof(kProfAttnBwd, 8.0 * B * H * T * T * hs, 2, stream);
  const size_t smem_q = static_cast<size_t>(AF_WARPS) * (2 * hs + T) * sizeof(float);
  VB_PROPAGATE(set_smem(attn_f32_bwd_dq_kernel, smem_q));
  attn_f32_bwd_dq_kernel<<<grid, AF_WARPS * 32, smem_q, stream>>>(qkv, d_o, lse, delta, dqkv, B, T, H, hs, scale);
  VB_CHECK_LAUNCH();
  const size_t smem_kv = static_cast<size_t>(AF_WARPS) * (2 * hs + 2 * T) * sizeof(float);
  VB_PROPAGATE(set_smem(attn_f32_bwd_dkv_kernel, smem_kv));
  attn_f32_bwd_dkv_kernel<<<grid, AF_WARPS * 32, smem_kv, stream>>>(qkv, d_o, lse, delta, dqkv, B, T, H, hs, scale);
  VB_CHECK_LAUNCH();
  return kOk;
}

}  // namespace vb
