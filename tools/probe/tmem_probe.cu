// Micro-probe: tensor-memory read / write bandwidth per SM, MUFU.EX2 and bf16x2 pack throughput (sm_100a).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/probe/tmem_probe tools/probe/tmem_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_bf16.h>

__device__ __forceinline__ void ld32(uint32_t a, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(a) : "memory");
}
__device__ __forceinline__ void st16(uint32_t a, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(a),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}

// mode 0: TMEM ld x32 stream; 1: ld + wait each; 2: st x16 stream; 3: ex2 stream; 4: cvt bf16x2 stream; 5: ld+ex2+cvt
__global__ void probe(int mode, int iters, long long* out, float* sink) {
  __shared__ uint32_t tptr;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&tptr)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t base = tptr + ((uint32_t)((warp & 3) * 32) << 16);
  uint32_t r[32];
  float acc = threadIdx.x * 1e-3f;
  uint32_t pk[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) pk[i] = i + threadIdx.x;
  // initialise TMEM so loads return defined data
  for (int c = 0; c < 512; c += 16) st16(base + c, pk);
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  __syncthreads();
  const long long t0 = clock64();
  if (mode == 0) {
    for (int i = 0; i < iters; ++i) {
      ld32(base + ((i & 7) * 32), r);
      if ((i & 3) == 3) { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); acc += __uint_as_float(r[i & 31]); }
    }
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  } else if (mode == 1) {
    for (int i = 0; i < iters; ++i) {
      ld32(base + ((i & 7) * 32), r);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      acc += __uint_as_float(r[i & 31]);
    }
  } else if (mode == 2) {
    for (int i = 0; i < iters; ++i) { pk[i & 15] += i; st16(base + ((i & 15) * 16), pk); }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  } else if (mode == 3) {
    float v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = acc + k;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int k = 0; k < 8; ++k) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(v[k]));
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) acc += v[k];
  } else if (mode == 4) {
    float v[8];
    uint32_t o = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = acc + k;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int k = 0; k < 8; k += 2) {
        uint32_t t;
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(t) : "f"(v[k + 1]), "f"(v[k]));
        o ^= t; v[k] += 1.f;
      }
    }
    acc += __uint_as_float(o);
  } else {
    for (int i = 0; i < iters; ++i) {
      ld32(base + ((i & 7) * 32), r);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        float e0 = __uint_as_float(r[2 * j]) * 1.1f - acc, e1 = __uint_as_float(r[2 * j + 1]) * 1.1f - acc;
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(e0));
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(e1));
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(pk[j]) : "f"(e1), "f"(e0));
      }
      st16(base + 256 + ((i & 7) * 16), pk);
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  const long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc + __uint_as_float(pk[3]);
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tptr));
}

int main() {
  long long* out; float* sink;
  cudaMalloc(&out, 8 * 256); cudaMalloc(&sink, 4 * 1024 * 256);
  const char* names[] = {"tmem ld x32 (wait every 4)", "tmem ld x32 + wait", "tmem st x16", "ex2 x8", "cvt.bf16x2 x4", "ld+ex2+cvt+st"};
  const int iters = 4096;
  for (int mode = 0; mode < 6; ++mode)
    for (int warps : {1, 4, 8, 16}) {
      probe<<<1, warps * 32, 0>>>(mode, iters, out, sink);
      long long h = 0;
      cudaError_t e = cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
      const double cyc = (double)h / iters;
      double per = 0; const char* unit = "";
      if (mode <= 1) { per = warps * 32 * 32 * 4 / cyc; unit = "B/clk/SM read"; }
      else if (mode == 2) { per = warps * 32 * 16 * 4 / cyc; unit = "B/clk/SM written"; }
      else if (mode == 3) { per = warps * 32 * 8 / cyc; unit = "ex2/clk/SM"; }
      else if (mode == 4) { per = warps * 32 * 4 / cyc; unit = "cvt/clk/SM"; }
      else { per = warps * 32 * 32 / cyc; unit = "elements/clk/SM"; }
      printf("%-28s warps=%2d  %.1f clk/iter  %.1f %s\n", names[mode], warps, cyc, per, unit);
    }
  return 0;
}
