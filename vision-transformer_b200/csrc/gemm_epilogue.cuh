// Epilogue shared by the tcgen05 (bf16) and CUDA-core (fp32) contraction kernels.
// Order: +bias -> relu (-> sign bits out) -> aux mask / bit mask -> dropout -> +row_add -> +residual -> store / atomic add.
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
  float* col_sum;          // [N] or null: col_sum[n] += sum over rows of the stored value (bias gradients)
  // 1 bit per element, word (n / 32) * bits_rows + m holds columns [32*(n/32), +32) of row m (bit = column % 32):
  uint32_t* bits_out;        // or null: bit = (stored value > 0)   — the ReLU mask, 1/16 the bytes of the bf16 tensor
  const uint32_t* mask_bits; // or null: keep v where the bit is set  (tensor-core kernel only, N % 32 == 0)
  int bits_rows;
  Dropout drop;
  // Fused attention-backward statistic (tensor-core kernel, 2-byte output, N % 64 == 0): dot_out[(b*dot_H + h)*dot_T + t] =
  // sum over the 64 columns of head h of stored(m, n) * dot_aux(m, n), m = b*dot_T + t  (delta = rowsum(dO o O), the dO GEMM)
  const void* dot_aux = nullptr;   // [M][ldc] (OutT), same layout as C
  float* dot_out = nullptr;
  int dot_T = 0, dot_H = 0;
};

// Compile-time epilogue specialisation: a kernel instantiated with EPI != kEpiGeneric evaluates each option as a
// constant (the hot call sites of the training step); kEpiGeneric keeps every option a run-time test so that any
// combination reachable through the C ABI still works.
enum : uint32_t {
  kEpiBias = 1u, kEpiRelu = 2u, kEpiMask = 4u, kEpiDrop = 8u, kEpiRowMap = 16u, kEpiResidual = 32u, kEpiAccum = 64u,
  kEpiColSum = 128u, kEpiBitsOut = 256u, kEpiBitMask = 512u, kEpiDot = 1024u, kEpiGeneric = 0x80000000u
};
inline uint32_t epi_flags(const GemmEpilogue& ep) {
  return (ep.bias ? kEpiBias : 0u) | (ep.relu ? kEpiRelu : 0u) | (ep.mask_aux ? kEpiMask : 0u) |
         (ep.drop.p > 0.f ? kEpiDrop : 0u) | ((ep.row_add || ep.row_period > 0) ? kEpiRowMap : 0u) |
         (ep.residual ? kEpiResidual : 0u) | (ep.accumulate ? kEpiAccum : 0u) | (ep.col_sum ? kEpiColSum : 0u) |
         (ep.bits_out ? kEpiBitsOut : 0u) | (ep.mask_bits ? kEpiBitMask : 0u) | (ep.dot_out ? kEpiDot : 0u);
}
#define VB_EPI(flag, cond) ((EPI == ::vb::kEpiGeneric) ? (cond) : ((EPI & (flag)) != 0u))

__device__ __forceinline__ int64_t epi_out_row(const GemmEpilogue& ep, int m) {
  if (ep.row_period > 0) return static_cast<int64_t>(m / ep.row_period) * ep.row_group_stride + (m % ep.row_period);
  return m;
}

__device__ __forceinline__ void red_add_f32x4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

// Read-side operands of the epilogue for 8 consecutive columns of one output row, fetched EARLY (before the
// accumulator is available) so that their latency overlaps the TMEM drain instead of serialising behind it.
struct EpiPre {
  float4 r0, r1;   // residual
  uint4 a0, a1;    // mask_aux: 8 x bf16 in a0, or 8 x fp32 in a0|a1
  uint32_t bits;   // mask_bits word of this row / 32-column group
};
__device__ __forceinline__ float4 ldg_f4_no_l1(const float* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
// 256-bit global accesses (sm_100: LDG/STG.E.256): the 8 consecutive fp32 columns of a lane in ONE request instead of two
// half-sector 16-byte ones.  The address must be 32-byte aligned.
__device__ __forceinline__ void ldg_f8(const float* p, float4& a, float4& b) {
  asm volatile("ld.global.nc.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p));
}
__device__ __forceinline__ void stg_f8(float* p, const float (&v)[8]) {
  asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]),
               "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]) : "memory");
}
template <typename OutT, uint32_t EPI>
__device__ __forceinline__ void epilogue_fetch8(const GemmEpilogue& ep, int64_t off, int m, int n, EpiPre& pre,
                                                int hints = 0) {
  if (VB_EPI(kEpiBitMask, ep.mask_bits != nullptr))
    pre.bits = __ldg(ep.mask_bits + static_cast<int64_t>(n >> 5) * ep.bits_rows + m);
  if (VB_EPI(kEpiResidual, ep.residual != nullptr)) {
    if (hints & 8) {
      ldg_f8(ep.residual + off, pre.r0, pre.r1);
    } else if (hints & 4) {   // read once, overwritten in place by this very thread: keep it out of L1
      pre.r0 = ldg_f4_no_l1(ep.residual + off);
      pre.r1 = ldg_f4_no_l1(ep.residual + off + 4);
    } else {
      pre.r0 = __ldg(reinterpret_cast<const float4*>(ep.residual + off));
      pre.r1 = __ldg(reinterpret_cast<const float4*>(ep.residual + off + 4));
    }
  }
  if (VB_EPI(kEpiMask, ep.mask_aux != nullptr)) {
    const OutT* aux = reinterpret_cast<const OutT*>(ep.mask_aux) + off;
    pre.a0 = __ldg(reinterpret_cast<const uint4*>(aux));
    if constexpr (sizeof(OutT) == 4) pre.a1 = __ldg(reinterpret_cast<const uint4*>(aux) + 1);
  }
  if constexpr (sizeof(OutT) == 2) {   // the dot operand shares the aux registers (never combined with the aux mask)
    if (VB_EPI(kEpiDot, ep.dot_out != nullptr))
      pre.a0 = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const OutT*>(ep.dot_aux) + off));
  }
}

// Apply the epilogue to 8 consecutive columns [n, n+8) of logical row m (output offset `off`) and write them.
// bias8: the 8 bias values of these columns (zeros when ep.bias is null).  v is left holding the stored values.
// Requirements: n % 8 == 0, N % 8 == 0, ldc % 8 == 0, all base pointers 16-byte aligned.
template <typename OutT, uint32_t EPI>
__device__ __forceinline__ void epilogue_finish8(const GemmEpilogue& ep, int m, int n, int N, int64_t off,
                                                 float (&v)[8], const EpiPre& pre, const float (&bias8)[8],
                                                 int hints = 0) {
  if (VB_EPI(kEpiBias, true)) {
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] += bias8[i];
  }
  if (VB_EPI(kEpiRelu, ep.relu)) {
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = fmaxf(v[i], 0.f);
  }
  if (VB_EPI(kEpiMask, ep.mask_aux != nullptr)) {
    if constexpr (sizeof(OutT) == 2) {
      const bf16* h = reinterpret_cast<const bf16*>(&pre.a0);
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = __bfloat162float(h[i]) > 0.f ? v[i] : 0.f;
    } else {
      const float a[8] = {__uint_as_float(pre.a0.x), __uint_as_float(pre.a0.y), __uint_as_float(pre.a0.z),
                          __uint_as_float(pre.a0.w), __uint_as_float(pre.a1.x), __uint_as_float(pre.a1.y),
                          __uint_as_float(pre.a1.z), __uint_as_float(pre.a1.w)};
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = a[i] > 0.f ? v[i] : 0.f;
    }
  }
  if (VB_EPI(kEpiBitMask, ep.mask_bits != nullptr)) {
    const uint32_t b = pre.bits >> (n & 31);   // n % 8 == 0: my 8 columns are one byte of the word
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = (b >> i) & 1u ? v[i] : 0.f;
  }
  if (VB_EPI(kEpiDrop, ep.drop.p > 0.f)) {
    const uint64_t lin = static_cast<uint64_t>(m) * static_cast<uint64_t>(N) + static_cast<uint64_t>(n);
    float f[8];
    dropout_factors8(ep.drop, lin >> 3, f);   // lin % 8 == 0 (n % 8 == 0 and N % 8 == 0)
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] *= f[i];
  }
  if (VB_EPI(kEpiRowMap, ep.row_add != nullptr) && ep.row_add != nullptr) {
    const int pr = ep.row_period > 0 ? (m % ep.row_period) : m;
    const float* ra = ep.row_add + static_cast<int64_t>(pr) * N + n;
    const float4 a0 = __ldg(reinterpret_cast<const float4*>(ra));
    const float4 a1 = __ldg(reinterpret_cast<const float4*>(ra + 4));
    v[0] += a0.x; v[1] += a0.y; v[2] += a0.z; v[3] += a0.w;
    v[4] += a1.x; v[5] += a1.y; v[6] += a1.z; v[7] += a1.w;
  }
  if (VB_EPI(kEpiResidual, ep.residual != nullptr)) {
    v[0] += pre.r0.x; v[1] += pre.r0.y; v[2] += pre.r0.z; v[3] += pre.r0.w;
    v[4] += pre.r1.x; v[5] += pre.r1.y; v[6] += pre.r1.z; v[7] += pre.r1.w;
  }
  if constexpr (sizeof(OutT) == 2) {
    uint4 o;
    o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]);
    o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
    *reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(ep.C) + off) = o;
  } else {
    float* c = reinterpret_cast<float*>(ep.C) + off;
    if (VB_EPI(kEpiAccum, ep.accumulate)) {
      red_add_f32x4(c, v[0], v[1], v[2], v[3]);
      red_add_f32x4(c + 4, v[4], v[5], v[6], v[7]);
    } else if (hints & 8) {
      stg_f8(c, v);
    } else {
      *reinterpret_cast<float4*>(c) = make_float4(v[0], v[1], v[2], v[3]);
      *reinterpret_cast<float4*>(c + 4) = make_float4(v[4], v[5], v[6], v[7]);
    }
  }
}

}  // namespace vb
