// Fused AdamW (torch.optim.AdamW semantics, reference train.py:66,96: decoupled weight decay, bias-corrected
// moments, eps added after the sqrt) as ONE launch over either a flat parameter arena or a table of tensors.
// Optionally folds the data-parallel 1/world gradient scale and emits the bf16 shadow copy the tensor-core
// GEMMs read, so no separate cast pass exists.  16 B/param read + 12 B/param written (+2 B shadow).
#include "kernels.h"

namespace vb {


namespace {

__device__ __forceinline__ float adamw_one(float p, float g, float& m, float& v, const AdamWHyper& h) {
  g *= h.grad_scale;
  p *= 1.f - h.lr * h.weight_decay;
  m = h.beta1 * m + (1.f - h.beta1) * g;
  v = h.beta2 * v + (1.f - h.beta2) * g * g;
  const float denom = sqrtf(v) * h.inv_sqrt_bc2 + h.eps;
  return p - (h.lr / h.bias_corr1) * (m / denom);
}

__device__ __forceinline__ void adamw_span(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                           float* __restrict__ v, bf16* __restrict__ shadow, int64_t begin,
                                           int64_t end, const AdamWHyper& h, int tid, int nthreads) {
  // vector body on 16-byte aligned quads, scalar head/tail
  const bool al = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                    reinterpret_cast<uintptr_t>(v)) & 15u) == 0 &&
                  (shadow == nullptr || (reinterpret_cast<uintptr_t>(shadow) & 7u) == 0) && (begin & 3) == 0;
  int64_t vec_end = al ? begin + ((end - begin) & ~int64_t(3)) : begin;
  for (int64_t i = begin + static_cast<int64_t>(tid) * 4; i < vec_end; i += static_cast<int64_t>(nthreads) * 4) {
    float4 pp = *reinterpret_cast<float4*>(p + i);
    const float4 gg = __ldcs(reinterpret_cast<const float4*>(g + i));
    float4 mm = *reinterpret_cast<float4*>(m + i);
    float4 vv = *reinterpret_cast<float4*>(v + i);
    pp.x = adamw_one(pp.x, gg.x, mm.x, vv.x, h);
    pp.y = adamw_one(pp.y, gg.y, mm.y, vv.y, h);
    pp.z = adamw_one(pp.z, gg.z, mm.z, vv.z, h);
    pp.w = adamw_one(pp.w, gg.w, mm.w, vv.w, h);
    *reinterpret_cast<float4*>(p + i) = pp;
    *reinterpret_cast<float4*>(m + i) = mm;
    *reinterpret_cast<float4*>(v + i) = vv;
    if (shadow != nullptr) {
      uint2 o;
      o.x = pack_bf16x2(pp.x, pp.y);
      o.y = pack_bf16x2(pp.z, pp.w);
      *reinterpret_cast<uint2*>(shadow + i) = o;
    }
  }
  for (int64_t i = vec_end + tid; i < end; i += nthreads) {
    float mm = m[i], vv = v[i];
    const float pp = adamw_one(p[i], g[i], mm, vv, h);
    p[i] = pp; m[i] = mm; v[i] = vv;
    if (shadow != nullptr) shadow[i] = __float2bfloat16_rn(pp);
  }
}

__global__ void __launch_bounds__(256)
adamw_flat_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                  bf16* __restrict__ shadow, int64_t n, const AdamWHyper h) {
  adamw_span(p, g, m, v, shadow, 0, n, h, blockIdx.x * blockDim.x + threadIdx.x, gridDim.x * blockDim.x);
}

// Range form for the data-parallel trainer: elements [0, n) of arenas already offset to the range start (a 16-byte
// aligned offset).  GB: the gradient is read from a bf16 exchange buffer (reduce-scattered bf16 gradients) instead of
// the fp32 arena.  ZERO: the fp32 gradient range is cleared in the same pass — the kernels of the next backward
// accumulate into it, so this is optimizer.zero_grad() (reference train.py:94) without a separate memset.
template <bool GB, bool ZERO>
__global__ void __launch_bounds__(256)
adamw_range_kernel(float* __restrict__ p, float* __restrict__ g, const bf16* __restrict__ gb, float* __restrict__ m,
                   float* __restrict__ v, bf16* __restrict__ shadow, int64_t n, const AdamWHyper h_in) {
  AdamWHyper h = h_in;
  if (h.corr_dev != nullptr) {          // captured CUDA graph: the step-dependent corrections come from device memory
    h.bias_corr1 = __ldg(h.corr_dev);
    h.inv_sqrt_bc2 = __ldg(h.corr_dev + 1);
  }
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x * 4;
  const int64_t n4 = n & ~int64_t(3);
  for (int64_t i = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 4; i < n4; i += stride) {
    float4 pp = *reinterpret_cast<float4*>(p + i);
    float4 gg;
    if (GB) {
      const uint2 raw = __ldcs(reinterpret_cast<const uint2*>(gb + i));
      gg.x = __uint_as_float(raw.x << 16); gg.y = __uint_as_float(raw.x & 0xFFFF0000u);
      gg.z = __uint_as_float(raw.y << 16); gg.w = __uint_as_float(raw.y & 0xFFFF0000u);
    } else {
      gg = __ldcs(reinterpret_cast<const float4*>(g + i));
    }
    float4 mm = *reinterpret_cast<float4*>(m + i);
    float4 vv = *reinterpret_cast<float4*>(v + i);
    pp.x = adamw_one(pp.x, gg.x, mm.x, vv.x, h);
    pp.y = adamw_one(pp.y, gg.y, mm.y, vv.y, h);
    pp.z = adamw_one(pp.z, gg.z, mm.z, vv.z, h);
    pp.w = adamw_one(pp.w, gg.w, mm.w, vv.w, h);
    *reinterpret_cast<float4*>(p + i) = pp;
    *reinterpret_cast<float4*>(m + i) = mm;
    *reinterpret_cast<float4*>(v + i) = vv;
    if (ZERO) *reinterpret_cast<float4*>(g + i) = make_float4(0.f, 0.f, 0.f, 0.f);
    if (shadow != nullptr) {
      uint2 o;
      o.x = pack_bf16x2(pp.x, pp.y);
      o.y = pack_bf16x2(pp.z, pp.w);
      *reinterpret_cast<uint2*>(shadow + i) = o;
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < (n - n4)) {
    const int64_t i = n4 + threadIdx.x;
    float mm = m[i], vv = v[i];
    const float gi = GB ? __bfloat162float(gb[i]) : g[i];
    const float pp = adamw_one(p[i], gi, mm, vv, h);
    p[i] = pp; m[i] = mm; v[i] = vv;
    if (ZERO) g[i] = 0.f;
    if (shadow != nullptr) shadow[i] = __float2bfloat16_rn(pp);
  }
}

// Per-step scalars of a CAPTURED step (CUDA graph: kernel arguments are frozen at capture, so whatever changes from
// step to step lives here).  Layout (32 bytes): uint64 seed_add | int32 step | float bias_corr1 | float inv_sqrt_bc2.
// One thread: step += 1, seed_add = (step - 1) * seed_stride (the per-step term of DPTrainer's dropout seed), corrections
// recomputed in double precision exactly as the host does (to_hyper in engine.cu).
struct StepState { unsigned long long seed_add; int step; float bias_corr1; float inv_sqrt_bc2; int pad[3]; };
__global__ void step_state_advance_kernel(StepState* st, float beta1, float beta2, unsigned long long seed_stride) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const int step = st->step + 1;
  st->step = step;
  st->seed_add = static_cast<unsigned long long>(step - 1) * seed_stride;   // the step being run, 0-based, as the host counts it
  st->bias_corr1 = static_cast<float>(1.0 - pow(static_cast<double>(beta1), static_cast<double>(step)));
  st->inv_sqrt_bc2 = static_cast<float>(1.0 / sqrt(1.0 - pow(static_cast<double>(beta2), static_cast<double>(step))));
}

// Multi-tensor form: block -> (tensor, chunk) through a device-resident table.
constexpr int64_t kChunk = 65536;
__global__ void __launch_bounds__(256)
adamw_multi_kernel(const uint64_t* __restrict__ p_ptrs, const uint64_t* __restrict__ g_ptrs,
                   const uint64_t* __restrict__ m_ptrs, const uint64_t* __restrict__ v_ptrs,
                   const uint64_t* __restrict__ s_ptrs, const int64_t* __restrict__ sizes,
                   const int32_t* __restrict__ blk_tensor, const int32_t* __restrict__ blk_chunk, const AdamWHyper h) {
  const int t = blk_tensor[blockIdx.x];
  const int64_t begin = static_cast<int64_t>(blk_chunk[blockIdx.x]) * kChunk;
  const int64_t end = begin + kChunk < sizes[t] ? begin + kChunk : sizes[t];
  adamw_span(reinterpret_cast<float*>(p_ptrs[t]), reinterpret_cast<const float*>(g_ptrs[t]),
             reinterpret_cast<float*>(m_ptrs[t]), reinterpret_cast<float*>(v_ptrs[t]),
             s_ptrs != nullptr ? reinterpret_cast<bf16*>(s_ptrs[t]) : nullptr, begin, end, h, threadIdx.x, blockDim.x);
}

}  // namespace

int adamw_flat(float* p, const float* g, float* m, float* v, bf16* shadow, int64_t n, const AdamWHyper& h,
               cudaStream_t s) {
  VB_REQUIRE(n > 0, "adamw: empty arena");
  const int64_t blocks = ceil_div<int64_t>(n, 256 * 4 * 4);
  const int64_t cap = static_cast<int64_t>(num_sms()) * 8;
  ProfScope prof(kProfAdamW, static_cast<double>(n) * (28 + (shadow ? 2 : 0)), 1, s);
  adamw_flat_kernel<<<static_cast<int>(blocks < cap ? blocks : cap), 256, 0, s>>>(p, g, m, v, shadow, n, h);
  VB_CHECK_LAUNCH();
  return kOk;
}

int adamw_range(float* p, float* g, const bf16* gb, float* m, float* v, bf16* shadow, int64_t n, bool zero_grad,
                const AdamWHyper& h, cudaStream_t s) {
  VB_REQUIRE(n > 0, "adamw: empty range");
  VB_REQUIRE(((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
               reinterpret_cast<uintptr_t>(v)) & 15u) == 0 && (reinterpret_cast<uintptr_t>(shadow) & 7u) == 0 &&
             (reinterpret_cast<uintptr_t>(gb) & 7u) == 0, "adamw_range: range start must be 16-byte aligned");
  const int64_t blocks = ceil_div<int64_t>(n, 256 * 4 * 4);
  const int64_t cap = static_cast<int64_t>(num_sms()) * 8;
  const int grid = static_cast<int>(blocks < cap ? blocks : cap);
  // bytes: p, m, v read + written (24), gradient read (4 fp32 / 2 bf16), gradient cleared (4), bf16 shadow written (2)
  ProfScope prof(kProfAdamW, static_cast<double>(n) * (24 + (gb ? 2 : 4) + (zero_grad ? 4 : 0) + (shadow ? 2 : 0)), 1, s);
  if (gb != nullptr) {
    if (zero_grad) adamw_range_kernel<true, true><<<grid, 256, 0, s>>>(p, g, gb, m, v, shadow, n, h);
    else adamw_range_kernel<true, false><<<grid, 256, 0, s>>>(p, g, gb, m, v, shadow, n, h);
  } else {
    if (zero_grad) adamw_range_kernel<false, true><<<grid, 256, 0, s>>>(p, g, gb, m, v, shadow, n, h);
    else adamw_range_kernel<false, false><<<grid, 256, 0, s>>>(p, g, gb, m, v, shadow, n, h);
  }
  VB_CHECK_LAUNCH();
  return kOk;
}

int step_state_advance(void* state, float beta1, float beta2, uint64_t seed_stride, cudaStream_t s) {
  static_assert(sizeof(StepState) == 32, "step state layout is part of the C ABI (VITB200_STEP_STATE_BYTES)");
  ProfScope prof(kProfAdamW, 0.0, 1, s);
  step_state_advance_kernel<<<1, 32, 0, s>>>(static_cast<StepState*>(state), beta1, beta2,
                                             static_cast<unsigned long long>(seed_stride));
  VB_CHECK_LAUNCH();
  return kOk;
}

int adamw_multi(const uint64_t* p_ptrs, const uint64_t* g_ptrs, const uint64_t* m_ptrs, const uint64_t* v_ptrs,
                const uint64_t* s_ptrs, const int64_t* sizes, const int32_t* blk_tensor, const int32_t* blk_chunk,
                int n_blocks, const AdamWHyper& h, cudaStream_t s) {
  VB_REQUIRE(n_blocks > 0, "adamw_multi: empty table");
  ProfScope prof(kProfAdamW, 0.0, 1, s);
  adamw_multi_kernel<<<n_blocks, 256, 0, s>>>(p_ptrs, g_ptrs, m_ptrs, v_ptrs, s_ptrs, sizes, blk_tensor, blk_chunk, h);
  VB_CHECK_LAUNCH();
  return kOk;
}

}  // namespace vb
