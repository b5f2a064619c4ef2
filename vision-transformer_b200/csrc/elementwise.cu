// Bandwidth kernels around the contractions: patch im2col, CLS rows, embedding-gradient reductions, bias-gradient
// column sums, classifier-head pointwise ops, cross-entropy, casts and the dropout-mask export used by tests.
#include "common.cuh"

namespace vb {

namespace {

// ---------------------------------------------------------------------------------------------- im2col
// x fp32 [B,C,Hi,Wi] -> A [B*N, Kpad], row b*N + gy*gw + gx, column c*P*P + ph*P + pw (zero padded to Kpad).
// Replaces the gather implicit in Conv2d(C,E,P,stride=P) + Flatten + permute (reference vit.py:20-30,40).
template <typename OutT, bool VEC4>
__global__ void im2col_kernel(const float* __restrict__ x, OutT* __restrict__ A, int B, int C, int Hi, int Wi, int P,
                              int gh, int gw, int K, int Kpad) {
  const int kq = VEC4 ? Kpad / 4 : Kpad;
  const int64_t total = static_cast<int64_t>(B) * gh * gw * kq;
  for (int64_t idx = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int kk = static_cast<int>(idx % kq);
    const int64_t row = idx / kq;
    const int gx = static_cast<int>(row % gw);
    const int gy = static_cast<int>((row / gw) % gh);
    const int b = static_cast<int>(row / (static_cast<int64_t>(gw) * gh));
    if (VEC4) {
      const int k = kk * 4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (k < K) {
        const int c = k / (P * P), rem = k - c * P * P, ph = rem / P, pw = rem - ph * P;
        v = __ldg(reinterpret_cast<const float4*>(
            x + ((static_cast<int64_t>(b) * C + c) * Hi + gy * P + ph) * Wi + gx * P + pw));
      }
      if constexpr (sizeof(OutT) == 2) {
        uint2 o;
        o.x = pack_bf16x2(v.x, v.y);
        o.y = pack_bf16x2(v.z, v.w);
        *reinterpret_cast<uint2*>(A + row * Kpad + k) = o;
      } else {
        *reinterpret_cast<float4*>(A + row * Kpad + k) = v;
      }
    } else {
      float v = 0.f;
      if (kk < K) {
        const int c = kk / (P * P), rem = kk - c * P * P, ph = rem / P, pw = rem - ph * P;
        v = __ldg(x + ((static_cast<int64_t>(b) * C + c) * Hi + gy * P + ph) * Wi + gx * P + pw);
      }
      A[row * Kpad + kk] = from_f32<OutT>(v);
    }
  }
}

// uint8 pixels -> A with the ToTensor scaling fused in (reference train.py:151-155: ToTensor = uint8 / 255 in fp32,
// IEEE division).  NHWC = 0: x u8 [B,C,Hi,Wi] (what ToTensor's permute yields); NHWC = 1: x u8 [B,Hi,Wi,C] (decoder
// order).  VEC4 (NCHW only) moves 4 pixels of one patch row per thread.
template <typename OutT, bool NHWC, bool VEC4>
__global__ void im2col_u8_kernel(const uint8_t* __restrict__ x, OutT* __restrict__ A, int B, int C, int Hi, int Wi,
                                 int P, int gh, int gw, int K, int Kpad) {
  const int kq = VEC4 ? Kpad / 4 : Kpad;
  const int64_t total = static_cast<int64_t>(B) * gh * gw * kq;
  for (int64_t idx = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int kk = static_cast<int>(idx % kq);
    const int64_t row = idx / kq;
    const int gx = static_cast<int>(row % gw);
    const int gy = static_cast<int>((row / gw) % gh);
    const int b = static_cast<int>(row / (static_cast<int64_t>(gw) * gh));
    if (VEC4) {
      const int k = kk * 4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (k < K) {
        const int c = k / (P * P), rem = k - c * P * P, ph = rem / P, pw = rem - ph * P;
        const uchar4 u = __ldg(reinterpret_cast<const uchar4*>(
            x + ((static_cast<int64_t>(b) * C + c) * Hi + gy * P + ph) * Wi + gx * P + pw));
        v = make_float4(__fdiv_rn(static_cast<float>(u.x), 255.f), __fdiv_rn(static_cast<float>(u.y), 255.f),
                        __fdiv_rn(static_cast<float>(u.z), 255.f), __fdiv_rn(static_cast<float>(u.w), 255.f));
      }
      if constexpr (sizeof(OutT) == 2) {
        uint2 o;
        o.x = pack_bf16x2(v.x, v.y);
        o.y = pack_bf16x2(v.z, v.w);
        *reinterpret_cast<uint2*>(A + row * Kpad + k) = o;
      } else {
        *reinterpret_cast<float4*>(A + row * Kpad + k) = v;
      }
    } else {
      float v = 0.f;
      if (kk < K) {
        const int c = kk / (P * P), rem = kk - c * P * P, ph = rem / P, pw = rem - ph * P;
        const int64_t y = gy * P + ph, xx = gx * P + pw;
        const int64_t off = NHWC ? ((static_cast<int64_t>(b) * Hi + y) * Wi + xx) * C + c
                                 : ((static_cast<int64_t>(b) * C + c) * Hi + y) * Wi + xx;
        v = __fdiv_rn(static_cast<float>(__ldg(x + off)), 255.f);
      }
      A[row * Kpad + kk] = from_f32<OutT>(v);
    }
  }
}

// Patch-embedding weight rows whose length is not a multiple of 8 (P = 14: K = 588): bf16 copy with every row
// zero-padded to Kpad, so that the row pitch is TMA-legal (16 B) and the tensor-core GEMM can run on it; and the
// inverse for the weight gradient (fp32 [rows, Kpad] scratch accumulated into the dense [rows, K] gradient).
__global__ void cast_pad_rows_kernel(const float* __restrict__ w, bf16* __restrict__ out, int rows, int K, int Kpad) {
  const int64_t total = static_cast<int64_t>(rows) * Kpad;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int k = static_cast<int>(i % Kpad);
    const int64_t r = i / Kpad;
    out[i] = from_f32<bf16>(k < K ? w[r * K + k] : 0.f);
  }
}
__global__ void add_unpad_rows_kernel(float* __restrict__ dst, const float* __restrict__ src, int rows, int K, int Kpad) {
  const int64_t total = static_cast<int64_t>(rows) * K;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int k = static_cast<int>(i % K);
    const int64_t r = i / K;
    dst[i] += src[r * Kpad + k];
  }
}

// out[b, N, :] = cls[b, :] + pos[N, :]   (CLS appended LAST: reference vit.py:41-42)
__global__ void cls_rows_kernel(const float* __restrict__ cls, const float* __restrict__ pos, float* __restrict__ out,
                                int B, int T, int E) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * E) return;
  const int b = idx / E, e = idx - b * E;
  out[(static_cast<int64_t>(b) * T + (T - 1)) * E + e] = cls[idx] + pos[static_cast<int64_t>(T - 1) * E + e];
}

// Embedding backward reductions over the batch: g fp32 [B,T,E] ->
//   dpos[t,e] += sum_b g[b,t,e];  dcls[b,e] += g[b,T-1,e];  dY[b*N+t, e] = cast(g[b,t,e]) for t < N.
// Thread = 4 consecutive e of one token t (E % 4 == 0), blockIdx.y = a slice of the batch; dpos partial sums are
// combined with one atomic per thread, so the grid has B/bchunk times more loads in flight than a per-(t,e) loop.
template <typename OutT>
__global__ void embed_bwd_kernel(const float* __restrict__ g, float* __restrict__ dpos, float* __restrict__ dcls,
                                 OutT* __restrict__ dY, int B, int T, int E, int bchunk) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;      // float4 index inside one image
  const int E4 = E >> 2;
  if (idx >= T * E4) return;
  const int t = idx / E4, e = (idx - t * E4) * 4;
  const int N = T - 1;
  const int b0 = blockIdx.y * bchunk, b1 = min(B, b0 + bchunk);
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
  for (int b = b0; b < b1; ++b) {
    const float4 v = __ldcs(reinterpret_cast<const float4*>(g + (static_cast<int64_t>(b) * T + t) * E + e));
    s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    if (t < N) {
      OutT* o = dY + (static_cast<int64_t>(b) * N + t) * E + e;
      if constexpr (sizeof(OutT) == 2) {
        uint2 pk;
        pk.x = pack_bf16x2(v.x, v.y);
        pk.y = pack_bf16x2(v.z, v.w);
        *reinterpret_cast<uint2*>(o) = pk;
      } else {
        *reinterpret_cast<float4*>(o) = v;
      }
    } else {
      float* c = dcls + static_cast<int64_t>(b) * E + e;
      c[0] += v.x; c[1] += v.y; c[2] += v.z; c[3] += v.w;
    }
  }
  float* dp = dpos + static_cast<int64_t>(t) * E + e;
  atomicAdd(dp, s.x); atomicAdd(dp + 1, s.y); atomicAdd(dp + 2, s.z); atomicAdd(dp + 3, s.w);
}
// scalar fallback (E % 4 != 0)
template <typename OutT>
__global__ void embed_bwd_scalar_kernel(const float* __restrict__ g, float* __restrict__ dpos, float* __restrict__ dcls,
                                        OutT* __restrict__ dY, int B, int T, int E) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= T * E) return;
  const int t = idx / E, e = idx - t * E;
  const int N = T - 1;
  float s = 0.f;
  for (int b = 0; b < B; ++b) {
    const float v = g[(static_cast<int64_t>(b) * T + t) * E + e];
    s += v;
    if (t < N) dY[(static_cast<int64_t>(b) * N + t) * E + e] = from_f32<OutT>(v);
    else dcls[static_cast<int64_t>(b) * E + e] += v;
  }
  dpos[idx] += s;
}

// out[n] += sum_m in[m, n]   (bias gradients).  blockDim = (32, 8); each x-thread owns one column.
template <typename InT>
__global__ void colsum_kernel(const InT* __restrict__ in, int64_t ld, float* __restrict__ out, int M, int N) {
  __shared__ float red[8][33];
  const int n = blockIdx.x * 32 + threadIdx.x;
  float s = 0.f;
  if (n < N)
    for (int m = blockIdx.y * 8 + threadIdx.y; m < M; m += gridDim.y * 8) s += to_f32(in[static_cast<int64_t>(m) * ld + n]);
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && n < N) {
    float tsum = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) tsum += red[j][threadIdx.x];
    atomicAdd(out + n, tsum);
  }
}

// bf16 variant with 16-byte loads: each x-thread owns 8 consecutive columns (N % 8 == 0, ld % 8 == 0).
__global__ void colsum_bf16x8_kernel(const bf16* __restrict__ in, int64_t ld, float* __restrict__ out, int M, int N) {
  __shared__ float red[8][32][9];
  const int n0 = (blockIdx.x * 32 + threadIdx.x) * 8;
  float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (n0 < N) {
    for (int m = blockIdx.y * 8 + threadIdx.y; m < M; m += gridDim.y * 8) {
      const uint4 raw = __ldcs(reinterpret_cast<const uint4*>(in + static_cast<int64_t>(m) * ld + n0));
      const bf16* h = reinterpret_cast<const bf16*>(&raw);
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] += __bfloat162float(h[j]);
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[threadIdx.y][threadIdx.x][j] = s[j];
  __syncthreads();
  if (threadIdx.y == 0 && n0 < N) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float tsum = 0.f;
#pragma unroll
      for (int r = 0; r < 8; ++r) tsum += red[r][threadIdx.x][j];
      atomicAdd(out + n0 + j, tsum);
    }
  }
}

// x0[b, :] = x[b, 0, :]  (the classifier reads token 0, reference vit.py:80)
__global__ void gather_row0_kernel(const float* __restrict__ x, float* __restrict__ x0, int B, int T, int E) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * E) return;
  const int b = idx / E, e = idx - b * E;
  x0[idx] = x[static_cast<int64_t>(b) * T * E + e];
}

// g[b,0,:] = dx0[b,:], g[b,t>0,:] = 0 ; dyn = cast(g * dropout factors)   (top of the encoder backward)
template <typename OutT>
__global__ void head_scatter_kernel(const float* __restrict__ dx0, float* __restrict__ g, OutT* __restrict__ dyn,
                                    const Dropout drop, int B, int T, int E) {
  const int64_t total4 = static_cast<int64_t>(B) * T * E / 4;
  for (int64_t i4 = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i4 < total4;
       i4 += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t lin = i4 * 4;
    const int e = static_cast<int>(lin % E);
    const int64_t row = lin / E;
    const int t = static_cast<int>(row % T);
    const int b = static_cast<int>(row / T);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (t == 0) v = *reinterpret_cast<const float4*>(dx0 + static_cast<int64_t>(b) * E + e);
    *reinterpret_cast<float4*>(g + lin) = v;
    if (dyn != nullptr) {
      float o[4] = {v.x, v.y, v.z, v.w};
      if (drop.p > 0.f && t == 0) {
        float f[4];
        dropout_factors4(drop, static_cast<uint64_t>(lin), f);
#pragma unroll
        for (int j = 0; j < 4; ++j) o[j] *= f[j];
      }
      if constexpr (sizeof(OutT) == 2) {
        uint2 pk;
        pk.x = pack_bf16x2(o[0], o[1]);
        pk.y = pack_bf16x2(o[2], o[3]);
        *reinterpret_cast<uint2*>(dyn + lin) = pk;
      } else {
        *reinterpret_cast<float4*>(dyn + lin) = make_float4(o[0], o[1], o[2], o[3]);
      }
    }
  }
}
template <typename OutT>
__global__ void head_scatter_scalar_kernel(const float* __restrict__ dx0, float* __restrict__ g, OutT* __restrict__ dyn,
                                           const Dropout drop, int B, int T, int E) {
  const int64_t total = static_cast<int64_t>(B) * T * E;
  for (int64_t lin = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; lin < total;
       lin += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int e = static_cast<int>(lin % E);
    const int64_t row = lin / E;
    const int t = static_cast<int>(row % T);
    const int b = static_cast<int>(row / T);
    float v = t == 0 ? dx0[static_cast<int64_t>(b) * E + e] : 0.f;
    g[lin] = v;
    if (dyn != nullptr) {
      if (drop.p > 0.f) {
        v *= dropout_factor1(drop, static_cast<uint64_t>(lin));
      }
      dyn[lin] = from_f32<OutT>(v);
    }
  }
}

// exact-erf GELU (nn.GELU() default, reference vit.py:71) and its derivative
__global__ void gelu_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t n) {
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const float v = x[i];
    y[i] = 0.5f * v * (1.f + erff(v * 0.70710678118654752f));
  }
}
__global__ void gelu_bwd_kernel(const float* __restrict__ x, float* __restrict__ dy_inout, int64_t n) {
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const float v = x[i];
    const float cdf = 0.5f * (1.f + erff(v * 0.70710678118654752f));
    const float pdf = 0.3989422804014327f * expf(-0.5f * v * v);
    dy_inout[i] *= cdf + v * pdf;
  }
}

// Mean cross-entropy over B rows + its gradient (reference train.py:81,93).  One CTA; one warp per row.
__global__ void ce_kernel(const float* __restrict__ logits, const int64_t* __restrict__ labels, float* __restrict__ loss,
                          float* __restrict__ dlogits, int B, int C, float grad_scale) {
  __shared__ float warp_loss[32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  float acc = 0.f;
  const float gs = grad_scale / static_cast<float>(B);  // d(mean loss)/dlogits, times the caller's scale
  for (int b = warp; b < B; b += nwarps) {
    const float* row = logits + static_cast<int64_t>(b) * C;
    float mx = -INFINITY;
    for (int c = lane; c < C; c += 32) mx = fmaxf(mx, row[c]);
    mx = warp_max(mx);
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s += expf(row[c] - mx);
    s = warp_sum(s);
    const int64_t y64 = labels[b];
    const bool y_ok = y64 >= 0 && y64 < C;   // torch device-asserts on an out-of-range target; here the loss turns NaN
    const int y = y_ok ? static_cast<int>(y64) : 0;
    const float lse = mx + logf(s);
    if (lane == 0) acc += y_ok ? lse - row[y] : __int_as_float(0x7fc00000);
    if (dlogits != nullptr)
      for (int c = lane; c < C; c += 32)
        dlogits[static_cast<int64_t>(b) * C + c] = (expf(row[c] - lse) - (c == y ? 1.f : 0.f)) * gs;
  }
  if (lane == 0) warp_loss[warp] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < nwarps; ++w) t += warp_loss[w];
    *loss = t / static_cast<float>(B);
  }
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ in, bf16* __restrict__ out, int64_t n) {
  const int64_t n4 = n / 4;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const float4 v = *reinterpret_cast<const float4*>(in + i * 4);
    uint2 o;
    o.x = pack_bf16x2(v.x, v.y);
    o.y = pack_bf16x2(v.z, v.w);
    *reinterpret_cast<uint2*>(out + i * 4) = o;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (int64_t i = n4 * 4; i < n; ++i) out[i] = __float2bfloat16_rn(in[i]);
}

// Gradient exchange staging of the data-parallel trainer: out = bf16(g) (when out != null) and g = 0 in ONE pass —
// the bf16 copy is what the reduce-scatter sends, the cleared fp32 arena is what the next backward accumulates into
// (optimizer.zero_grad(), reference train.py:94).  out == null: clear only.
__global__ void __launch_bounds__(256)
grad_pack_kernel(float* __restrict__ g, bf16* __restrict__ out, int64_t n) {
  const int64_t n4 = n / 4;
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    if (out != nullptr) {
      const float4 v = __ldcs(reinterpret_cast<const float4*>(g + i * 4));
      uint2 o;
      o.x = pack_bf16x2(v.x, v.y);
      o.y = pack_bf16x2(v.z, v.w);
      *reinterpret_cast<uint2*>(out + i * 4) = o;
    }
    *reinterpret_cast<float4*>(g + i * 4) = z;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (int64_t i = n4 * 4; i < n; ++i) {
      if (out != nullptr) out[i] = __float2bfloat16_rn(g[i]);
      g[i] = 0.f;
    }
}

// Bilinear resize of uint8 image planes — the arithmetic of PIL's ImagingResample (8 bits per channel), which is what
// torchvision's transforms.Resize((H, W)) runs on the PIL images of the reference's input pipeline (train.py:151-155):
// a HORIZONTAL pass then a VERTICAL pass, each a normalised triangle filter (support widened by the scale factor when
// shrinking) with coefficients in 22-bit fixed point, each pass rounded and clipped to uint8 on its own.  The
// coefficient tables (bounds = first tap and tap count, k = fixed-point weights) are built on the host in double
// precision exactly as PIL builds them; here both passes are fused per output pixel (the intermediate uint8 of the
// horizontal pass is recomputed for each vertical tap), so the result is bit-identical to PIL.
__device__ __forceinline__ int resize_hpass(const uint8_t* __restrict__ row, int x, int x0, int nx,
                                            const int32_t* __restrict__ kx, bool skip) {
  if (skip) return row[x];                 // PIL skips a pass whose size does not change
  int acc = 1 << 21;
  for (int c = 0; c < nx; ++c) acc += static_cast<int>(row[x0 + c]) * kx[c];
  acc >>= 22;
  return acc < 0 ? 0 : (acc > 255 ? 255 : acc);
}
__global__ void __launch_bounds__(256)
resize_bilinear_u8_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int planes, int hin, int win,
                          int hout, int wout, const int32_t* __restrict__ xb, const int32_t* __restrict__ xk, int ksx,
                          const int32_t* __restrict__ yb, const int32_t* __restrict__ yk, int ksy) {
  const int64_t total = static_cast<int64_t>(planes) * hout * wout;
  const bool skip_h = win == wout, skip_v = hin == hout;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int x = static_cast<int>(i % wout);
    const int y = static_cast<int>((i / wout) % hout);
    const int64_t pl = i / (static_cast<int64_t>(wout) * hout);
    const uint8_t* sp = src + pl * hin * win;
    const int x0 = xb[2 * x], nx = xb[2 * x + 1];
    const int32_t* kx = xk + static_cast<int64_t>(x) * ksx;
    int v;
    if (skip_v) {
      v = resize_hpass(sp + static_cast<int64_t>(y) * win, x, x0, nx, kx, skip_h);
    } else {
      const int y0 = yb[2 * y], ny = yb[2 * y + 1];
      const int32_t* ky = yk + static_cast<int64_t>(y) * ksy;
      int acc = 1 << 21;
      for (int r = 0; r < ny; ++r)
        acc += resize_hpass(sp + static_cast<int64_t>(y0 + r) * win, x, x0, nx, kx, skip_h) * ky[r];
      v = acc >> 22;
      v = v < 0 ? 0 : (v > 255 ? 255 : v);
    }
    dst[i] = static_cast<uint8_t>(v);
  }
}

// keep[i] = 1 if element i of the (seed, stream) dropout stream is kept.  Test/debug export of the mask that the
// fused epilogues regenerate on the fly.
__global__ void dropout_mask_kernel(uint8_t* __restrict__ keep, int64_t n, const Dropout drop) {
  for (int64_t i8 = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i8 * 8 < n;
       i8 += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const uint32_t bits = dropout_keep8(drop, static_cast<uint64_t>(i8));
    for (int j = 0; j < 8 && i8 * 8 + j < n; ++j) keep[i8 * 8 + j] = (bits >> j) & 1u;
  }
}

// out = cast(g * dropout factors): the masked low-precision gradient a backward GEMM consumes, for call sites where
// no LayerNorm-backward kernel is upstream to fuse it into (stand-alone Block / FeedForward / MultiHeadAttention).
template <typename OutT>
__global__ void mask_cast_kernel(const float* __restrict__ g, OutT* __restrict__ out, const Dropout drop, int64_t n) {
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    float v = g[i];
    if (drop.p > 0.f) {
      v *= dropout_factor1(drop, static_cast<uint64_t>(i));
    }
    out[i] = from_f32<OutT>(v);
  }
}

inline int grid_for(int64_t work_items, int threads) {
  const int64_t blocks = ceil_div<int64_t>(work_items, threads);
  const int64_t cap = static_cast<int64_t>(num_sms()) * 16;
  return static_cast<int>(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

}  // namespace

int im2col(const float* x, void* A, bool out_bf16, int B, int C, int Hi, int Wi, int P, int Kpad, cudaStream_t s) {
  ProfScope prof(kProfEltwise, 0.0, 1, s);
  const int gh = Hi / P, gw = Wi / P, K = C * P * P;
  VB_REQUIRE(gh > 0 && gw > 0 && Kpad >= K, "im2col: bad geometry");
  const bool vec = (P % 4 == 0) && (Wi % 4 == 0) && (Kpad % 4 == 0) && aligned16(x) && aligned16(A);
  const int64_t items = static_cast<int64_t>(B) * gh * gw * (vec ? Kpad / 4 : Kpad);
  const int grid = grid_for(items, 256);
  if (out_bf16) {
    if (vec) im2col_kernel<bf16, true><<<grid, 256, 0, s>>>(x, static_cast<bf16*>(A), B, C, Hi, Wi, P, gh, gw, K, Kpad);
    else im2col_kernel<bf16, false><<<grid, 256, 0, s>>>(x, static_cast<bf16*>(A), B, C, Hi, Wi, P, gh, gw, K, Kpad);
  } else {
    if (vec) im2col_kernel<float, true><<<grid, 256, 0, s>>>(x, static_cast<float*>(A), B, C, Hi, Wi, P, gh, gw, K, Kpad);
    else im2col_kernel<float, false><<<grid, 256, 0, s>>>(x, static_cast<float*>(A), B, C, Hi, Wi, P, gh, gw, K, Kpad);
  }
  VB_CHECK_LAUNCH();
  return kOk;
}

int im2col_u8(const uint8_t* x, int nhwc, void* A, bool out_bf16, int B, int C, int Hi, int Wi, int P, int Kpad,
              cudaStream_t s) {
  ProfScope prof(kProfEltwise, 0.0, 1, s);
  const int gh = Hi / P, gw = Wi / P, K = C * P * P;
  VB_REQUIRE(gh > 0 && gw > 0 && Kpad >= K, "im2col_u8: bad geometry");
  VB_REQUIRE(nhwc == 0 || nhwc == 1, "im2col_u8: layout must be 0 (NCHW) or 1 (NHWC)");
  const bool vec = !nhwc && (P % 4 == 0) && (Wi % 4 == 0) && (Kpad % 4 == 0) &&
                   (reinterpret_cast<uintptr_t>(x) % 4 == 0) && aligned16(A);
  const int64_t items = static_cast<int64_t>(B) * gh * gw * (vec ? Kpad / 4 : Kpad);
  const int grid = grid_for(items, 256);
#define VB_U8_LAUNCH(OUT, NHWC, VEC) \
  im2col_u8_kernel<OUT, NHWC, VEC><<<grid, 256, 0, s>>>(x, static_cast<OUT*>(A), B, C, Hi, Wi, P, gh, gw, K, Kpad)
  if (out_bf16) {
    if (vec) VB_U8_LAUNCH(bf16, false, true);
    else if (nhwc) VB_U8_LAUNCH(bf16, true, false);
    else VB_U8_LAUNCH(bf16, false, false);
  } else {
    if (vec) VB_U8_LAUNCH(float, false, true);
    else if (nhwc) VB_U8_LAUNCH(float, true, false);
    else VB_U8_LAUNCH(float, false, false);
  }
#undef VB_U8_LAUNCH
  VB_CHECK_LAUNCH();
  return kOk;
}

int cls_rows(const float* cls, const float* pos, float* out, int B, int T, int E, cudaStream_t s) {
  ProfScope prof(kProfEltwise, 0.0, 1, s);
  cls_rows_kernel<<<ceil_div(B * E, 256), 256, 0, s>>>(cls, pos, out, B, T, E);
  VB_CHECK_LAUNCH();
  return kOk;
}

int embed_bwd_reduce(const float* g, float* dpos, float* dcls, void* dY, bool dy_bf16, int B, int T, int E,
                     cudaStream_t s) {
  ProfScope prof(kProfEltwise, 0.0, 1, s);
  if (E % 4 == 0 && aligned16(g) && aligned16(dY)) {
    const int bchunk = B >= 64 ? 16 : B;
    const dim3 grid(ceil_div(T * (E / 4), 128), ceil_div(B, bchunk));
    if (dy_bf16) embed_bwd_kernel<bf16><<<grid, 128, 0, s>>>(g, dpos, dcls, static_cast<bf16*>(dY), B, T, E, bchunk);
    else embed_bwd_kernel<float><<<grid, 128, 0, s>>>(g, dpos, dcls, static_cast<float*>(dY), B, T, E, bchunk);
  } else {
    const int grid = ceil_div(T * E, 128);
    if (dy_bf16) embed_bwd_scalar_kernel<bf16><<<grid, 128, 0, s>>>(g, dpos, dcls, static_cast<bf16*>(dY), B, T, E);
    else embed_bwd_scalar_kernel<float><<<grid, 128, 0, s>>>(g, dpos, dcls, static_cast<float*>(dY), B, T, E);
  }
  VB_CHECK_LAUNCH();
  return kOk;
}

int colsum(const void* in, bool in_bf16, int64_t ld, float* out, int M, int N, cudaStream_t s) {
  ProfScope prof(kProfEltwise, 0.0, 1, s);
  dim3 block(32, 8);
  if (in_bf16 && N % 8 == 0 && ld % 8 == 0 && aligned16(in)) {
    const int gx = ceil_div(N, 256);
    int gy = ceil_div(num_sms() * 4, gx);
    if (gy > ceil_div(M, 8)) gy = ceil_div(M, 8);
    colsum_bf16x8_kernel<<<dim3(gx, gy), block, 0, s>>>(static_cast<const bf16*>(in), ld, out, M, N);
  } else {
    const int gx = ceil_div(N, 32);
    int gy = ceil_div(num_sms() * 4, gx);
    if (gy > ceil_div(M, 8)) gy = ceil_div(M, 8);
    if (gy < 1) gy = 1;
    if (in_bf16) colsum_kernel<bf16><<<dim3(gx, gy), block, 0, s>>>(static_cast<const bf16*>(in), ld, out, M, N);
    else colsum_kernel<float><<<dim3(gx, gy), block, 0, s>>>(static_cast<const float*>(in), ld, out, M, N);
  }
  VB_CHECK_LAUNCH();
  return kOk;
}

int gather_row0(const float* x, float* x0, int B, int T, int E, cudaStream_t s) {
  ProfScope prof(kProfEltwise, 0.0, 1, s);
  gather_row0_kernel<<<ceil_div(B * E, 256), 256, 0, s>>>(x, x0, B, T, E);
  VB_CHECK_LAUNCH();
  return kOk;
}

int head_scatter(const float* dx0, float* g, void* dyn, bool dyn_bf16, const Dropout& drop, int B, int T, int E,
                 cudaStream_t s) {
  ProfScope prof(kProfEltwise, 0.0, 1, s);
  const int64_t total = static_cast<int64_t>(B) * T * E;
  if (E % 4 == 0) {
    const int grid = grid_for(total / 4, 256);
    if (dyn_bf16) head_scatter_kernel<bf16><<<grid, 256, 0, s>>>(dx0, g, static_cast<bf16*>(dyn), drop, B, T, E);
    else head_scatter_kernel<float><<<grid, 256, 0, s>>>(dx0, g, static_cast<float*>(dyn), drop, B, T, E);
  } else {
    const int grid = grid_for(total, 256);
    if (dyn_bf16) head_scatter_scalar_kernel<bf16><<<grid, 256, 0, s>>>(dx0, g, static_cast<bf16*>(dyn), drop, B, T, E);
    else head_scatter_scalar_kernel<float><<<grid, 256, 0, s>>>(dx0, g, static_cast<float*>(dyn), drop, B, T, E);
  }
  VB_CHECK_LAUNCH();
  return kOk;
}

int gelu_forward(const float* x, float* y, int64_t n, cudaStream_t s) {
  ProfScope prof(kProfEltwise, 0.0, 1, s);
  gelu_fwd_kernel<<<grid_for(n, 256), 256, 0, s>>>(x, y, n);
  VB_CHECK_LAUNCH();
  return kOk;
}
int gelu_backward(const float* x, float* dy_inout, int64_t n, cudaStream_t s) {
  ProfScope prof(kProfEltwise, 0.0, 1, s);
  gelu_bwd_kernel<<<grid_for(n, 256), 256, 0, s>>>(x, dy_inout, n);
  VB_CHECK_LAUNCH();
  return kOk;
}

int cross_entropy(const float* logits, const int64_t* labels, float* loss, float* dlogits, int B, int C,
                  float grad_scale, cudaStream_t s) {
  ProfScope prof(kProfEltwise, 0.0, 1, s);
  ce_kernel<<<1, 1024, 0, s>>>(logits, labels, loss, dlogits, B, C, grad_scale);
  VB_CHECK_LAUNCH();
  return kOk;
}

int cast_pad_rows(const float* w, bf16* out, int rows, int K, int Kpad, cudaStream_t s) {
  ProfScope prof(kProfEltwise, 0.0, 1, s);
  cast_pad_rows_kernel<<<grid_for(static_cast<int64_t>(rows) * Kpad, 256), 256, 0, s>>>(w, out, rows, K, Kpad);
  VB_CHECK_LAUNCH();
  return kOk;
}
int add_unpad_rows(float* dst, const float* src, int rows, int K, int Kpad, cudaStream_t s) {
  ProfScope prof(kProfEltwise, 0.0, 1, s);
  add_unpad_rows_kernel<<<grid_for(static_cast<int64_t>(rows) * K, 256), 256, 0, s>>>(dst, src, rows, K, Kpad);
  VB_CHECK_LAUNCH();
  return kOk;
}

int cast_f32_to_bf16(const float* in, bf16* out, int64_t n, cudaStream_t s) {
  ProfScope prof(kProfEltwise, 0.0, 1, s);
  VB_REQUIRE(aligned16(in) && (reinterpret_cast<uintptr_t>(out) & 7u) == 0, "cast: unaligned buffers");
  cast_f32_bf16_kernel<<<grid_for(n / 4 + 1, 256), 256, 0, s>>>(in, out, n);
  VB_CHECK_LAUNCH();
  return kOk;
}

int resize_bilinear_u8(const uint8_t* src, uint8_t* dst, int planes, int hin, int win, int hout, int wout,
                       const int32_t* xb, const int32_t* xk, int ksx, const int32_t* yb, const int32_t* yk, int ksy,
                       cudaStream_t s) {
  VB_REQUIRE(planes > 0 && hin > 0 && win > 0 && hout > 0 && wout > 0 && ksx > 0 && ksy > 0, "resize: bad geometry");
  ProfScope prof(kProfEltwise, 0.0, 1, s);
  const int64_t total = static_cast<int64_t>(planes) * hout * wout;
  resize_bilinear_u8_kernel<<<grid_for(total, 256), 256, 0, s>>>(src, dst, planes, hin, win, hout, wout, xb, xk, ksx,
                                                                  yb, yk, ksy);
  VB_CHECK_LAUNCH();
  return kOk;
}

int grad_pack(float* g, bf16* out, int64_t n, cudaStream_t s) {
  VB_REQUIRE(n > 0, "grad_pack: empty range");
  VB_REQUIRE(aligned16(g) && (reinterpret_cast<uintptr_t>(out) & 7u) == 0, "grad_pack: unaligned buffers");
  ProfScope prof(kProfEltwise, 0.0, 1, s);
  const int64_t blocks = ceil_div<int64_t>(n / 4 + 1, 256 * 4);
  const int64_t cap = static_cast<int64_t>(num_sms()) * 8;
  grad_pack_kernel<<<static_cast<int>(blocks < cap ? blocks : cap), 256, 0, s>>>(g, out, n);
  VB_CHECK_LAUNCH();
  return kOk;
}

int mask_cast(const float* g, void* out, bool out_bf16, const Dropout& drop, int64_t n, cudaStream_t s) {
  ProfScope prof(kProfEltwise, 0.0, 1, s);
  if (out_bf16) mask_cast_kernel<bf16><<<grid_for(n, 256), 256, 0, s>>>(g, static_cast<bf16*>(out), drop, n);
  else mask_cast_kernel<float><<<grid_for(n, 256), 256, 0, s>>>(g, static_cast<float*>(out), drop, n);
  VB_CHECK_LAUNCH();
  return kOk;
}

int dropout_mask(uint8_t* keep, int64_t n, const Dropout& drop, cudaStream_t s) {
  ProfScope prof(kProfEltwise, 0.0, 1, s);
  dropout_mask_kernel<<<grid_for(n / 8 + 1, 256), 256, 0, s>>>(keep, n, drop);
  VB_CHECK_LAUNCH();
  return kOk;
}

}  // namespace vb
