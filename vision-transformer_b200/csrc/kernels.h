// Internal (C++) entry points of the individual kernels; the C ABI in include/vitb200.h wraps these.
#pragma once
#include "common.cuh"
#include "gemm_epilogue.cuh"

namespace vb {

struct AdamWHyper {
  float lr, beta1, beta2, eps, weight_decay;
  float bias_corr1;      // 1 - beta1^step
  float inv_sqrt_bc2;    // 1 / sqrt(1 - beta2^step)
  float grad_scale;      // multiply gradients (1/world_size for the data-parallel mean)
  const float* corr_dev; // optional DEVICE pair {bias_corr1, inv_sqrt_bc2} overriding the two fields above (CUDA graphs)
};

// contractions
int gemm_tc_auto_splits(int tiles, int kb_total, int walkers);
int gemm_tc_bf16(int M, int N, int K, int a_major, int b_major, const void* A, int64_t lda, const void* B,
                 int64_t ldb, bool out_f32, int splits, const GemmEpilogue& ep, cudaStream_t stream);
int gemm_f32(int M, int N, int K, int a_major, int b_major, const float* A, int64_t lda, const float* B,
             int64_t ldb, int splits, const GemmEpilogue& ep, cudaStream_t stream);

// 2-D bf16 TMA descriptor over a row-major matrix [rows][ld] of logical width `cols`; box = box_cols x box_rows,
// 128-byte swizzle, out-of-bounds reads return zero (defined in gemm_tc.cu; opaque 128-byte CUtensorMap)
int make_tmap_bf16(void* map, const void* base, int64_t rows, int64_t cols, int64_t ld, int box_cols, int box_rows);
int make_tmap_bf16_heads(void* map, const void* base, int64_t rows, int sections, int hs, int64_t ld, int box_rows);

// layer norm
int ln_forward(const float* x, const float* gamma, const float* beta, void* y, bool y_bf16, float* mean,
               float* rstd, int M, int E, float eps, cudaStream_t stream);
int ln_backward(const void* dy, bool dy_bf16, const float* x, const float* mean, const float* rstd,
                const float* gamma, float* g, int add_to_g, float* dgamma, float* dbeta, void* dyn,
                float* dyn_colsum, const Dropout& drop, int M, int E, cudaStream_t stream);

// attention
// tiled flash-style fp32 attention (attention_f32_tiled.cu): head sizes that are multiples of 4 up to 128
bool attn_f32_tiled_supported(int T, int hs, const void* a, const void* b);
int attn_f32_tiled_forward(const float* qkv, float* out, float* lse, float* probs, int B, int T, int H, int hs,
                           float scale, cudaStream_t stream);
int attn_f32_tiled_backward(const float* qkv, const float* d_o, const float* lse, const float* delta, float* dqkv, int B,
                            int T, int H, int hs, float scale, cudaStream_t stream);
int attn_delta(const void* o, const void* d_o, bool is_bf16, float* delta, int B, int T, int H, int hs,
               cudaStream_t stream);
int attn_f32_forward(const float* qkv, float* out, float* lse, int B, int T, int H, int hs, float scale,
                     cudaStream_t stream);
int attn_f32_probs(const float* qkv, float* probs, int B, int T, int H, int hs, float scale, cudaStream_t stream);
int attn_f32_backward(const float* qkv, const float* out, const float* d_o, const float* lse, float* delta,
                      float* dqkv, int B, int T, int H, int hs, float scale, cudaStream_t stream);
int attn_bf16_forward(const bf16* qkv, bf16* out, float* lse, int B, int T, int H, int hs, float scale,
                      cudaStream_t stream);
int attn_bf16_backward(const bf16* qkv, const bf16* out, const bf16* d_o, const float* lse, float* delta, bf16* dqkv,
                       int B, int T, int H, int hs, float scale, cudaStream_t stream, bool delta_ready = false);

// tcgen05 attention (head_size 64, T <= 256); returns kErrUnsupported outside that envelope
int attn_tc_forward(const bf16* qkv, bf16* out, float* lse, int B, int T, int H, int hs, float scale,
                    cudaStream_t stream);
bool attn_tc_supported(int T, int hs);
// tcgen05 forward for any T and head sizes 16..128 (multiple of 16): keys in chunks of 128, two-pass softmax
bool attn_tc_gen_supported(int T, int hs);
int attn_tc_gen_forward(const bf16* qkv, bf16* out, float* lse, int B, int T, int H, int hs, float scale,
                        cudaStream_t stream);
int attn_tc_gen_backward_dkv(const bf16* qkv, const bf16* d_o, const float* lse, const float* delta, bf16* dqkv, int B,
                             int T, int H, int hs, float scale, cudaStream_t stream);
int attn_tc_gen_backward_dq(const bf16* qkv, const bf16* d_o, const float* lse, const float* delta, bf16* dqkv, int B,
                            int T, int H, int hs, float scale, cudaStream_t stream);
int attn_tc_backward(const bf16* qkv, const bf16* d_o, const float* lse, const float* delta, bf16* dqkv, int B, int T,
                     int H, int hs, float scale, cudaStream_t stream);

// bandwidth kernels
int im2col(const float* x, void* A, bool out_bf16, int B, int C, int Hi, int Wi, int P, int Kpad, cudaStream_t s);
int im2col_u8(const uint8_t* x, int nhwc, void* A, bool out_bf16, int B, int C, int Hi, int Wi, int P, int Kpad,
              cudaStream_t s);
int cls_rows(const float* cls, const float* pos, float* out, int B, int T, int E, cudaStream_t s);
int embed_bwd_reduce(const float* g, float* dpos, float* dcls, void* dY, bool dy_bf16, int B, int T, int E,
                     cudaStream_t s);
int colsum(const void* in, bool in_bf16, int64_t ld, float* out, int M, int N, cudaStream_t s);
int gather_row0(const float* x, float* x0, int B, int T, int E, cudaStream_t s);
int head_scatter(const float* dx0, float* g, void* dyn, bool dyn_bf16, const Dropout& drop, int B, int T, int E,
                 cudaStream_t s);
int gelu_forward(const float* x, float* y, int64_t n, cudaStream_t s);
int gelu_backward(const float* x, float* dy_inout, int64_t n, cudaStream_t s);
int cross_entropy(const float* logits, const int64_t* labels, float* loss, float* dlogits, int B, int C,
                  float grad_scale, cudaStream_t s);
int cast_f32_to_bf16(const float* in, bf16* out, int64_t n, cudaStream_t s);
int cast_pad_rows(const float* w, bf16* out, int rows, int K, int Kpad, cudaStream_t s);
int add_unpad_rows(float* dst, const float* src, int rows, int K, int Kpad, cudaStream_t s);
int mask_cast(const float* g, void* out, bool out_bf16, const Dropout& drop, int64_t n, cudaStream_t s);
int dropout_mask(uint8_t* keep, int64_t n, const Dropout& drop, cudaStream_t s);

// optimiser
int adamw_flat(float* p, const float* g, float* m, float* v, bf16* shadow, int64_t n, const AdamWHyper& h,
               cudaStream_t s);
int grad_pack(float* g, bf16* out, int64_t n, cudaStream_t s);
int step_state_advance(void* state, float beta1, float beta2, uint64_t seed_stride, cudaStream_t s);
int resize_bilinear_u8(const uint8_t* src, uint8_t* dst, int planes, int hin, int win, int hout, int wout,
                       const int32_t* xb, const int32_t* xk, int ksx, const int32_t* yb, const int32_t* yk, int ksy,
                       cudaStream_t s);
int adamw_range(float* p, float* g, const bf16* gb, float* m, float* v, bf16* shadow, int64_t n, bool zero_grad,
                const AdamWHyper& h, cudaStream_t s);
int adamw_multi(const uint64_t* p_ptrs, const uint64_t* g_ptrs, const uint64_t* m_ptrs, const uint64_t* v_ptrs,
                const uint64_t* s_ptrs, const int64_t* sizes, const int32_t* blk_tensor, const int32_t* blk_chunk,
                int n_blocks, const AdamWHyper& h, cudaStream_t s);

}  // namespace vb
