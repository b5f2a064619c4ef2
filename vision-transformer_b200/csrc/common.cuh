// Shared host/device utilities for libvitb200: error reporting across the C ABI, launch checks, small math.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

namespace vb {

// Thread-local error string returned by vitb200_last_error().  Nothing throws or exits across the ABI.
void set_error(const char* fmt, ...);
const char* last_error();

constexpr int kOk = 0;
constexpr int kErrInvalid = -1;   // bad argument (shape / alignment / dtype)
constexpr int kErrCuda = -2;      // CUDA runtime / driver error
constexpr int kErrUnsupported = -3;

#define VB_CHECK_CUDA(expr)                                                                  \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess) {                                                                 \
      ::vb::set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__, cudaGetErrorString(_e)); \
      return ::vb::kErrCuda;                                                                 \
    }                                                                                        \
  } while (0)

#define VB_CHECK_LAUNCH() VB_CHECK_CUDA(cudaGetLastError())

// ---------------------------------------------------------------------------------------------------
// Programmatic dependent launch.  A kernel launched through launch_pdl() may begin (block scheduling + its prologue:
// barrier init, tensor-memory allocation, descriptor prefetch) while the previous kernel of the stream is still
// draining; it MUST execute pdl_wait() before its first access to global memory, which blocks until the previous grid
// has completed and flushed.  pdl_trigger() lets the next kernel of the stream start launching.  VITB200_PDL=0 turns
// the launch attribute off (the device-side instructions are then no-ops).
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
bool pdl_enabled();  // api.cu
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              int cluster_x, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  unsigned n = 0;
  if (pdl_enabled()) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  if (cluster_x > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = static_cast<unsigned>(cluster_x);
    attr[n].val.clusterDim.y = 1;
    attr[n].val.clusterDim.z = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

#define VB_REQUIRE(cond, ...)          \
  do {                                 \
    if (!(cond)) {                     \
      ::vb::set_error(__VA_ARGS__);    \
      return ::vb::kErrInvalid;        \
    }                                  \
  } while (0)

#define VB_PROPAGATE(expr)        \
  do {                            \
    int _rc = (expr);             \
    if (_rc != ::vb::kOk) return _rc; \
  } while (0)

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

template <typename T>
__host__ __device__ constexpr T ceil_div(T a, T b) {
  return (a + b - 1) / b;
}

int num_sms();  // SM count of the current device (cached)

// ---------------------------------------------------------------------------------------------------
// Launch accounting + optional live profiling (vitb200_prof_begin / vitb200_prof_end): every host-side kernel
// wrapper opens a ProfScope naming its kernel class, its algorithmic work (FLOPs for contractions / attention,
// bytes for bandwidth kernels) and how many kernels it launches.  Disabled: one relaxed atomic add.  Enabled:
// a cudaEvent pair recorded on the launching stream around the launches.
// ---------------------------------------------------------------------------------------------------
enum ProfClass {
  kProfGemmTc = 0, kProfGemmF32, kProfAttnFwd, kProfAttnBwd, kProfLnFwd, kProfLnBwd, kProfEltwise, kProfAdamW,
  kProfCount
};
struct ProfScope {
  int slot;
  cudaStream_t stream;
  ProfScope(int cls, double work, int launches, cudaStream_t s);
  ~ProfScope();
};

// ---------------------------------------------------------------------------------------------------
// Counter-based dropout RNG (Philox4x32-10).  One call yields 4 uniform 32-bit words for a
// (seed, stream, index/4) triple, so forward and backward regenerate identical masks without storing them.
// ---------------------------------------------------------------------------------------------------
struct Philox4 {
  uint32_t x, y, z, w;
};
__host__ __device__ __forceinline__ uint32_t mulhi32(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
  return __umulhi(a, b);
#else
  return static_cast<uint32_t>((static_cast<uint64_t>(a) * b) >> 32);
#endif
}
// Philox4x32 with a template round count.  Dropout uses 7 rounds (the fewest that pass BigCrush, Salmon et al.
// SC'11) and 16 random bits per element, so ONE call decides 8 consecutive elements: that keeps the mask cheap enough
// to regenerate inside a GEMM epilogue instead of storing it.
template <int ROUNDS>
__host__ __device__ __forceinline__ Philox4 philox4x32(uint64_t seed, uint64_t stream, uint64_t ctr) {
  uint32_t c0 = static_cast<uint32_t>(ctr), c1 = static_cast<uint32_t>(ctr >> 32);
  uint32_t c2 = static_cast<uint32_t>(stream), c3 = static_cast<uint32_t>(stream >> 32);
  uint32_t k0 = static_cast<uint32_t>(seed), k1 = static_cast<uint32_t>(seed >> 32);
#pragma unroll
  for (int r = 0; r < ROUNDS; ++r) {
    const uint32_t hi0 = mulhi32(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = mulhi32(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return Philox4{c0, c1, c2, c3};
}
// keep-probability threshold on a 16-bit lane: keep iff lane < thresh (thresh = round(keep * 2^16), saturating).
// p = 0.2 -> 52429 / 65536 = 0.800003 (the 4e-6 excess over 0.8 is far below the sampling noise of any mask).
__host__ __device__ __forceinline__ uint32_t keep_threshold(float p_drop) {
  const double keep = 1.0 - static_cast<double>(p_drop);
  const double t = keep * 65536.0 + 0.5;
  return t >= 65536.0 ? 65536u : static_cast<uint32_t>(t);
}

// Dropout descriptor passed by value to kernels.  p == 0 disables (eval mode / parity mode).
struct Dropout {
  float p;            // drop probability
  float scale;        // 1 / (1 - p)
  uint32_t thresh;    // keep iff 16-bit lane < thresh
  uint64_t seed;
  uint64_t stream;    // distinguishes call sites (layer id * 2 + {proj, fc2})
  const uint64_t* seed_dev;   // optional DEVICE word added to `seed` when the mask is drawn: lets a captured CUDA graph
                              // draw fresh masks on every replay (the per-step part of the seed lives in device memory)
};
inline Dropout make_dropout(float p, uint64_t seed, uint64_t stream, const uint64_t* seed_dev = nullptr) {
  Dropout d;
  d.p = p;
  d.scale = p > 0.f ? 1.f / (1.f - p) : 1.f;
  d.thresh = keep_threshold(p);
  d.seed = seed;
  d.stream = stream;
  d.seed_dev = seed_dev;
  return d;
}
// Element `lin` of a (seed, stream) mask lives in lane lin & 7 of Philox counter lin >> 3; lanes are the 16-bit halves
// of the four output words in order (x.lo, x.hi, y.lo, y.hi, z.lo, z.hi, w.lo, w.hi).
__host__ __device__ __forceinline__ uint32_t dropout_keep8(const Dropout& d, uint64_t idx8) {
  uint64_t seed = d.seed;
#ifdef __CUDA_ARCH__
  if (d.seed_dev != nullptr) seed += __ldg(reinterpret_cast<const unsigned long long*>(d.seed_dev));
#endif
  const Philox4 r = philox4x32<7>(seed, d.stream, idx8);
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
  uint32_t bits = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    bits |= ((w[i] & 0xFFFFu) < d.thresh ? 1u : 0u) << (2 * i);
    bits |= ((w[i] >> 16) < d.thresh ? 1u : 0u) << (2 * i + 1);
  }
  return bits;
}
// multiplicative factors (0 or scale) of 8 consecutive elements starting at linear index idx8 * 8
__device__ __forceinline__ void dropout_factors8(const Dropout& d, uint64_t idx8, float (&f)[8]) {
  const uint32_t bits = dropout_keep8(d, idx8);
#pragma unroll
  for (int i = 0; i < 8; ++i) f[i] = (bits >> i) & 1u ? d.scale : 0.f;
}
// 4 consecutive elements starting at linear index lin4 (lin4 % 4 == 0)
__device__ __forceinline__ void dropout_factors4(const Dropout& d, uint64_t lin4, float (&f)[4]) {
  const uint32_t bits = dropout_keep8(d, lin4 >> 3) >> (lin4 & 4);
#pragma unroll
  for (int i = 0; i < 4; ++i) f[i] = (bits >> i) & 1u ? d.scale : 0.f;
}
__device__ __forceinline__ float dropout_factor1(const Dropout& d, uint64_t lin) {
  return (dropout_keep8(d, lin >> 3) >> (lin & 7)) & 1u ? d.scale : 0.f;
}

// ---------------------------------------------------------------------------------------------------
// dtype helpers
// ---------------------------------------------------------------------------------------------------
using bf16 = __nv_bfloat16;
__device__ __forceinline__ float to_f32(float v) { return v; }
__device__ __forceinline__ float to_f32(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f32<bf16>(float v) { return __float2bfloat16_rn(v); }

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

}  // namespace vb
