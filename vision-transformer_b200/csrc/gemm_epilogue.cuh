// Epilogue shared by the tcgen05 (bf16) and CUDA-core (fp32) contraction kernels.
// Order: +bias -> relu -> aux mask -> dropout -> +row_add -> +residual -> store / atomic add.
#pragma once
#include "common.cuh"

namespace vb {

struct GemmEpilogue {
  const float* bias;       // [N] or null
  const void* mask_aux;    // [M][ldc] (OutT) or null: keep v where aux > 0
  const float* row_add;    // [row_period][N] or null
  const float* residual;   // [rows][ldc] fp32 or null
  void* C;
  int64_t ldc;
  int relu;
  int row_period;          // 0 = identity
  int row_group_stride;
  int accumulate;          // atomic add into fp32 C
  Dropout drop;
};

__device__ __forceinline__ int64_t epi_out_row(const GemmEpilogue& ep, int m) {
  if (ep.row_period > 0) return static_cast<int64_t>(m / ep.row_period) * ep.row_group_stride + (m % ep.row_period);
  return m;
}

__device__ __forceinline__ void red_add_f32x4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

// Apply the epilogue to 8 consecutive columns [n, n+8) of logical row m and write them.
// Requirements: n % 8 == 0, N % 8 == 0, ldc % 8 == 0, all base pointers 16-byte aligned.
template <typename OutT>
__device__ __forceinline__ void epilogue_store8(const GemmEpilogue& ep, int m, int n, int N, float (&v)[8]) {
  if (ep.bias != nullptr) {
    const float4 b0 = __ldg(reinterpret_cast<const float4*>(ep.bias + n));
    const float4 b1 = __ldg(reinterpret_cast<const float4*>(ep.bias + n + 4));
    v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w;
    v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
  }
  if (ep.relu) {
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = fmaxf(v[i], 0.f);
  }
  const int64_t orow = epi_out_row(ep, m);
  const int64_t off = orow * ep.ldc + n;
  if (ep.mask_aux != nullptr) {
    const OutT* aux = reinterpret_cast<const OutT*>(ep.mask_aux) + off;
    if constexpr (sizeof(OutT) == 2) {
      const uint4 raw = *reinterpret_cast<const uint4*>(aux);
      const bf16* h = reinterpret_cast<const bf16*>(&raw);
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = __bfloat162float(h[i]) > 0.f ? v[i] : 0.f;
    } else {
      const float4 a0 = *reinterpret_cast<const float4*>(aux);
      const float4 a1 = *reinterpret_cast<const float4*>(aux + 4);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = a[i] > 0.f ? v[i] : 0.f;
    }
  }
  if (ep.drop.p > 0.f) {
    const uint64_t lin = static_cast<uint64_t>(m) * static_cast<uint64_t>(N) + static_cast<uint64_t>(n);
    float f0[4], f1[4];
    dropout_factors4(ep.drop, lin >> 2, f0);
    dropout_factors4(ep.drop, (lin >> 2) + 1, f1);
#pragma unroll
    for (int i = 0; i < 4; ++i) { v[i] *= f0[i]; v[4 + i] *= f1[i]; }
  }
  if (ep.row_add != nullptr) {
    const int pr = ep.row_period > 0 ? (m % ep.row_period) : m;
    const float* ra = ep.row_add + static_cast<int64_t>(pr) * N + n;
    const float4 a0 = __ldg(reinterpret_cast<const float4*>(ra));
    const float4 a1 = __ldg(reinterpret_cast<const float4*>(ra + 4));
    v[0] += a0.x; v[1] += a0.y; v[2] += a0.z; v[3] += a0.w;
    v[4] += a1.x; v[5] += a1.y; v[6] += a1.z; v[7] += a1.w;
  }
  if (ep.residual != nullptr) {
    const float4 r0 = *reinterpret_cast<const float4*>(ep.residual + off);
    const float4 r1 = *reinterpret_cast<const float4*>(ep.residual + off + 4);
    v[0] += r0.x; v[1] += r0.y; v[2] += r0.z; v[3] += r0.w;
    v[4] += r1.x; v[5] += r1.y; v[6] += r1.z; v[7] += r1.w;
  }
  if constexpr (sizeof(OutT) == 2) {
    uint4 o;
    o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]);
    o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
    *reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(ep.C) + off) = o;
  } else {
    float* c = reinterpret_cast<float*>(ep.C) + off;
    if (ep.accumulate) {
      red_add_f32x4(c, v[0], v[1], v[2], v[3]);
      red_add_f32x4(c + 4, v[4], v[5], v[6], v[7]);
    } else {
      *reinterpret_cast<float4*>(c) = make_float4(v[0], v[1], v[2], v[3]);
      *reinterpret_cast<float4*>(c + 4) = make_float4(v[4], v[5], v[6], v[7]);
    }
  }
}

}  // namespace vb
