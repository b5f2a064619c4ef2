// Micro-probe: tcgen05.mma issue rate per SM for the shapes the attention kernels use (sm_100a).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I vision-transformer_b200/csrc -o tools/probe/umma_probe tools/probe/umma_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace vb::ptx;

// kind 0: SS, A K-major, B K-major, N = n.  kind 1: TS (A from TMEM), B MN-major, N = n.  kind 2: SS with B MN-major.
__global__ void probe(int kind, int n, int count, int reps, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tptr;
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (threadIdx.x < 32) { tmem_alloc(&tptr, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  fence_proxy_async();
  const uint32_t tm = tptr;
  if (kind >= 10) {
    // two issuing threads (warps 0 and 2), each count/2 MMAs of N = n into its own accumulator
    __shared__ uint64_t bar2[2];
    if (threadIdx.x == 0) { mbar_init(&bar2[0], 1); mbar_init(&bar2[1], 1); fence_barrier_init(); }
    __syncthreads();
    const int w = threadIdx.x >> 5;
    if ((w == 0 || w == 2) && (threadIdx.x & 31) == 0) {
      const int me = w >> 1;
      const uint32_t sa = smem_u32(smem) + me * 65536, sb = sa + 32768;
      const uint32_t idesc = make_idesc_bf16_f32(128, n, 0, kind == 10 ? 0 : 1);
      long long best = 1ll << 60;
      for (int r = 0; r < reps; ++r) {
        const long long t0 = clock64();
        for (int i = 0; i < count / 2; ++i) {
          const int k = i & 3;
          if (kind == 10)
            umma_f16(tm + me * 256, make_smem_desc_sw128(sa + k * 32, 16, 1024), make_smem_desc_sw128(sb + k * 32, 16, 1024), idesc, i > 0);
          else
            umma_f16_ts(tm + me * 256 + 128, tm + me * 256 + k * 8, make_smem_desc_sw128(sb + k * 2048, 64 * 128, 1024), idesc, i > 0);
        }
        umma_commit(&bar2[me]);
        mbar_wait(&bar2[me], r & 1);
        const long long t1 = clock64();
        if (t1 - t0 < best) best = t1 - t0;
      }
      out[me] = best;
    }
  } else if (kind >= 20) {
    if (threadIdx.x == 0) {
      const bool ts = kind >= 30;
      const int nacc = kind % 10;
      const uint32_t sa = smem_u32(smem), sb = sa + 32768;
      const uint32_t idesc = make_idesc_bf16_f32(128, n, 0, ts ? 1 : 0);
      long long best = 1ll << 60;
      for (int r = 0; r < reps; ++r) {
        const long long t0 = clock64();
        for (int i = 0; i < count; ++i) {
          const int k = (i / nacc) & 3, a = i % nacc;
          if (!ts)
            umma_f16(tm + a * 256, make_smem_desc_sw128(sa + k * 32, 16, 1024), make_smem_desc_sw128(sb + k * 32, 16, 1024), idesc, i >= nacc);
          else
            umma_f16_ts(tm + 256 + a * 64, tm + a * 32 + k * 8, make_smem_desc_sw128(sb + k * 2048, 64 * 128, 1024), idesc, i >= nacc);
        }
        umma_commit(&bar);
        mbar_wait(&bar, r & 1);
        const long long t1 = clock64();
        if (t1 - t0 < best) best = t1 - t0;
      }
      out[0] = best;
    }
  } else if (threadIdx.x == 0) {
    const uint32_t sa = smem_u32(smem), sb = sa + 32768;
    const uint32_t idesc = make_idesc_bf16_f32(128, n, 0, kind == 0 ? 0 : 1);
    long long best = 1ll << 60;
    for (int r = 0; r < reps; ++r) {
      const long long t0 = clock64();
      for (int i = 0; i < count; ++i) {
        const int k = i & 3;
        if (kind == 0)
          umma_f16(tm, make_smem_desc_sw128(sa + k * 32, 16, 1024), make_smem_desc_sw128(sb + k * 32, 16, 1024), idesc, i > 0);
        else if (kind == 1)
          umma_f16_ts(tm + 256, tm + k * 8, make_smem_desc_sw128(sb + k * 2048, 64 * 128, 1024), idesc, i > 0);
        else
          umma_f16(tm, make_smem_desc_sw128(sa + k * 32, 16, 1024), make_smem_desc_sw128(sb + k * 2048, 64 * 128, 1024), idesc, i > 0);
      }
      umma_commit(&bar);
      mbar_wait(&bar, r & 1);
      const long long t1 = clock64();
      if (t1 - t0 < best) best = t1 - t0;
    }
    out[0] = best;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tm, 512);
}

int main() {
  long long* out;
  cudaMalloc(&out, 64);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
  struct C { int kind, n; } cases[] = {{0, 256}, {0, 208}, {0, 128}, {0, 64},
                                       {1, 64}, {1, 128}, {1, 256}, {2, 64}, {2, 256},
                                       {10, 64}, {10, 256}, {11, 64}, {22, 64}, {22, 112}, {32, 64}, {34, 64}};
  for (auto c : cases)
    for (int count : {1, 8, 64}) {
      probe<<<1, 128, 160 * 1024>>>(c.kind, c.n, count, 5, out);
      long long h = 0;
      cudaError_t e = cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
      const double ideal = 128.0 * c.n * 16 * 2 / 8192.0;
      printf("kind=%d N=%3d count=%2d: %6lld clk total, %.1f clk/MMA (ideal %.1f at 8192 flop/clk)\n", c.kind, c.n, count, h,
             (double)h / count, ideal);
    }
  return 0;
}
