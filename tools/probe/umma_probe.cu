// Micro-probe: tcgen05.mma issue rate per SM for the shapes the attention kernels use (sm_100a).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I vision-transformer_b200/csrc -o tools/probe/umma_probe tools/probe/umma_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace vb::ptx;

// kind 0: SS, A K-major, B K-major, N = n.  kind 1: TS (A from TMEM), B MN-major, N = n.  kind 2: SS with B MN-major.
__global__ void probe(int kind, int n, int count, int reps, long long* out, int off2 = 256) {
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
  if (kind >= 60 && kind < 70) {
    // leanest issue loop: descriptors precomputed, 8 MMAs unrolled, nacc = 4 (kind 60) or 2 (61) accumulators, N = n;
    // kind 62: TS form with 2 accumulators
    if (threadIdx.x == 0) {
      const uint32_t sa = smem_u32(smem), sb = sa + 32768;
      const uint32_t idesc = make_idesc_bf16_f32(128, n, 0, kind == 62 ? 1 : 0);
      uint64_t da[4], db[4];
      for (int k = 0; k < 4; ++k) {
        da[k] = make_smem_desc_sw128(sa + k * 32, 16, 1024);
        db[k] = kind == 62 ? make_smem_desc_sw128(sb + k * 2048, 64 * 128, 1024) : make_smem_desc_sw128(sb + k * 32, 16, 1024);
      }
      const uint32_t t0a = tm, t1a = tm + 128, t2a = tm + 256, t3a = tm + 384;
      long long best = 1ll << 60;
      for (int r = 0; r < reps; ++r) {
        const long long t0 = clock64();
        for (int i = 0; i < count; i += 8) {
          const uint32_t acc = i > 0;
          if (kind == 60) {
            umma_f16(t0a, da[0], db[0], idesc, acc); umma_f16(t1a, da[0], db[0], idesc, acc);
            umma_f16(t2a, da[0], db[0], idesc, acc); umma_f16(t3a, da[0], db[0], idesc, acc);
            umma_f16(t0a, da[1], db[1], idesc, 1); umma_f16(t1a, da[1], db[1], idesc, 1);
            umma_f16(t2a, da[1], db[1], idesc, 1); umma_f16(t3a, da[1], db[1], idesc, 1);
          } else if (kind == 61) {
            umma_f16(t0a, da[0], db[0], idesc, acc); umma_f16(t2a, da[0], db[0], idesc, acc);
            umma_f16(t0a, da[1], db[1], idesc, 1); umma_f16(t2a, da[1], db[1], idesc, 1);
            umma_f16(t0a, da[2], db[2], idesc, 1); umma_f16(t2a, da[2], db[2], idesc, 1);
            umma_f16(t0a, da[3], db[3], idesc, 1); umma_f16(t2a, da[3], db[3], idesc, 1);
          } else {
            umma_f16_ts(t2a, t0a, db[0], idesc, acc); umma_f16_ts(t3a, t1a, db[0], idesc, acc);
            umma_f16_ts(t2a, t0a + 8, db[1], idesc, 1); umma_f16_ts(t3a, t1a + 8, db[1], idesc, 1);
            umma_f16_ts(t2a, t0a + 16, db[2], idesc, 1); umma_f16_ts(t3a, t1a + 16, db[2], idesc, 1);
            umma_f16_ts(t2a, t0a + 24, db[3], idesc, 1); umma_f16_ts(t3a, t1a + 24, db[3], idesc, 1);
          }
        }
        umma_commit(&bar);
        mbar_wait(&bar, r & 1);
        const long long t1 = clock64();
        if (t1 - t0 < best) best = t1 - t0;
      }
      out[0] = best;
    }
  } else if (kind >= 50 && kind < 60) {
    // two SS chains at columns 0 and off2; kind 51 restarts the accumulation every 4 MMAs per chain (like S / dP)
    if (threadIdx.x == 0) {
      const uint32_t sa = smem_u32(smem), sb = sa + 32768;
      const uint32_t idesc = make_idesc_bf16_f32(128, n, 0, 0);
      long long best = 1ll << 60;
      for (int r = 0; r < reps; ++r) {
        const long long t0 = clock64();
        for (int i = 0; i < count; ++i) {
          const int k = (i >> 1) & 3, a = i & 1;
          umma_f16(tm + a * off2, make_smem_desc_sw128(sa + k * 32, 16, 1024), make_smem_desc_sw128(sb + k * 32, 16, 1024), idesc,
                   kind == 51 ? (k > 0) : (i >= 2));
        }
        umma_commit(&bar);
        mbar_wait(&bar, r & 1);
        const long long t1 = clock64();
        if (t1 - t0 < best) best = t1 - t0;
      }
      out[0] = best;
    }
  } else if (kind >= 40 && kind < 50) {
    // the attention-backward issue pattern: per round 8 SS MMAs (S | dP interleaved, N = n, distinct operands), commit,
    // then 10 TS MMAs (two accumulators), commit.  kind 41: no intermediate commits.  42: SS part only.  43: TS part only.
    // 44: like 40 but every MMA reuses the same smem operands.
    __shared__ uint64_t bars[3];
    if (threadIdx.x == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_init(&bars[2], 1); fence_barrier_init(); }
    __syncthreads();
    if (threadIdx.x == 0) {
      const uint32_t s0 = smem_u32(smem);
      const uint32_t a1 = s0, b1 = s0 + 32768, a2 = s0 + 65536, b2 = s0 + 98304, bo1 = s0 + 131072, bo2 = s0 + 131072 + 12288;
      const uint32_t idesc_s = make_idesc_bf16_f32(128, n, 0, 0), idesc_o = make_idesc_bf16_f32(128, 64, 0, 1);
      long long best = 1ll << 60;
      for (int r = 0; r < reps; ++r) {
        const long long t0 = clock64();
        for (int i = 0; i < count; ++i) {
          const int slot = i & 1;
          const uint32_t tS = tm + slot * 192, tP = tS + 96;
          if (kind != 43) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const int kk = kind == 44 ? 0 : k;
              umma_f16(tS, make_smem_desc_sw128(a1 + kk * 32, 16, 1024), make_smem_desc_sw128((kind == 44 ? b1 : b1 + slot * 10240) + kk * 32, 16, 1024), idesc_s, k > 0);
              umma_f16(tP, make_smem_desc_sw128((kind == 44 ? a1 : a2) + kk * 32, 16, 1024), make_smem_desc_sw128((kind == 44 ? b1 : b2 + slot * 10240) + kk * 32, 16, 1024), idesc_s, k > 0);
            }
            if (kind != 41) umma_commit(&bars[0]);
          }
          if (kind != 42) {
            const uint32_t oS = tm + (slot ^ 1) * 192, oP = oS + 96;
            for (int k = 0; k < 5; ++k) {
              umma_f16_ts(tm + 384, oS + k * 8, make_smem_desc_sw128(bo1 + k * 2048, 64 * 128, 1024), idesc_o, (i | k) > 0);
              umma_f16_ts(tm + 448, oP + k * 8, make_smem_desc_sw128(bo2 + k * 2048, 64 * 128, 1024), idesc_o, (i | k) > 0);
            }
            if (kind != 41) umma_commit(&bars[1]);
          }
        }
        umma_commit(&bars[2]);
        mbar_wait(&bars[2], r & 1);
        const long long t1 = clock64();
        if (t1 - t0 < best) best = t1 - t0;
      }
      out[0] = best;
    }
  } else if (kind >= 10 && kind < 20) {
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
                                       {10, 64}, {10, 256}, {11, 64}, {22, 64}, {22, 112}, {32, 64}, {34, 64},
                                       {40, 80}, {41, 80}, {42, 80}, {43, 80}, {44, 80}};
  for (auto c : cases)
    for (int count : {1, 8, 64}) {
      probe<<<1, 128, 160 * 1024>>>(c.kind, c.n, count, 5, out);
      long long h = 0;
      cudaError_t e = cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
      const double ideal = 128.0 * c.n * 16 * 2 / 8192.0;
      if (c.kind >= 40) { printf("kind=%d N=%3d rounds=%2d: %6lld clk total, %.1f clk/round\n", c.kind, c.n, count, h, (double)h / count); continue; }
      printf("kind=%d N=%3d count=%2d: %6lld clk total, %.1f clk/MMA (ideal %.1f at 8192 flop/clk)\n", c.kind, c.n, count, h,
             (double)h / count, ideal);
    }
  for (int kind : {60, 61, 62})
    for (int n : {64, 96, 128}) {
      if (kind == 62 && n != 64) continue;
      probe<<<1, 128, 160 * 1024>>>(kind, n, 256, 5, out, 0);
      long long h = 0;
      cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
      printf("kind=%d N=%3d: %.1f clk/MMA\n", kind, n, (double)h / 256);
    }
  for (int kind : {50, 51})
    for (int n : {64, 80, 96, 128})
      for (int off : {64, 96, 128, 256}) {
        if (off < n) continue;
        probe<<<1, 128, 160 * 1024>>>(kind, n, 64, 5, out, off);
        long long h = 0;
        cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
        printf("kind=%d N=%3d off2=%3d: %.1f clk/MMA\n", kind, n, off, (double)h / 64);
      }
  return 0;
}
