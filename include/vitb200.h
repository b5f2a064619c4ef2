/* libvitb200 — C ABI of the B200-native ViT training hot path.
 *
 * The reference (SiddhantSKarki/Vision-Transformer) has no native extension: its hot path is stock
 * PyTorch ops called from src/VisionTransformer/{vit,transformer}.py.  Each entry point below names the
 * reference call site (file:line under the reference root) whose arithmetic it replaces.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer owned by the caller (PyTorch), contiguous, 16-byte aligned;
 *  - all work is enqueued on `stream` (a cudaStream_t passed as void*); nothing synchronises;
 *  - return 0 on success, negative on failure; vitb200_last_error() returns a thread-local message;
 *  - dtype enums: VITB200_F32 = 0, VITB200_BF16 = 1.
 */
#ifndef VITB200_H_
#define VITB200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VITB200_F32 0
#define VITB200_BF16 1

#define VITB200_MAJOR_K 0  /* operand stored with the reduction index contiguous */
#define VITB200_MAJOR_MN 1 /* operand stored with the M (or N) index contiguous  */

const char* vitb200_last_error(void);
int vitb200_version(void);
/* 1 if the current device is sm_100 (B200); the library refuses to run anywhere else. */
int vitb200_device_ok(void);

/* Dropout descriptor: p == 0 disables.  Masks are regenerated from (seed, stream, element index) with
 * Philox4x32-10, so backward never stores them.  Replaces nn.Dropout(0.2) at transformer.py:40,47,59. */
typedef struct {
  float p;
  uint64_t seed;
  uint64_t stream;
} vitb200_dropout;

/* General contraction  C[m,n] = epilogue( sum_k A(m,k) * B(n,k) ).
 *   a_major == MAJOR_K : A stored row-major [M][lda];   MAJOR_MN: A stored [K][lda] (transposed view)
 *   b_major == MAJOR_K : B stored row-major [N][ldb];   MAJOR_MN: B stored [K][ldb]
 * Epilogue, applied in this order to v = acc:
 *   +bias[n]  ->  relu  ->  (mask_aux[m,n] > 0 ? v : 0)  ->  dropout  ->  +row_add[(m % row_period), n]
 *   -> +residual[orow, n]  ->  store / atomic-add into C[orow, n]
 *   with orow = (m / row_period) * row_group_stride + m % row_period when row_period > 0, else m.
 * in_dtype BF16 runs the tcgen05/TMEM/TMA tensor-core kernel (fp32 accumulation); in_dtype F32 runs the
 * fp32 CUDA-core kernel used for the fp32 parity gate.  `splits` > 1 partitions K and accumulates with
 * fp32 atomics into a pre-zeroed fp32 C (used for weight gradients); 0 lets the library choose.
 * Replaces nn.Linear / Conv2d-as-GEMM forward and their autograd backward:
 *   transformer.py:12-17,22-28 (q/k/v), :38,46 (proj), :56-58 (MLP), vit.py:21-28 (patch conv). */
typedef struct {
  int M, N, K;
  int a_major, b_major;
  int in_dtype;  /* dtype of A and B */
  int out_dtype; /* dtype of C (and mask_aux) */
  const void* A; int64_t lda;
  const void* B; int64_t ldb;
  void* C; int64_t ldc;
  const float* bias;     /* [N] or NULL */
  int relu;
  const void* mask_aux;  /* [M][ldc] of out_dtype or NULL */
  vitb200_dropout dropout;
  const float* row_add;  /* [row_period][N] fp32 or NULL */
  int row_period;        /* 0 = identity row map */
  int row_group_stride;
  const float* residual; /* [..][ldc] fp32 or NULL (requires out_dtype F32) */
  int splits;            /* split-K factor (atomic accumulate), 0 = auto only when accumulate != 0 */
  int accumulate;        /* 1: C += result via fp32 atomics (C must be fp32) */
} vitb200_gemm_args;

int vitb200_gemm(const vitb200_gemm_args* args, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VITB200_H_ */
