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

/* Launch accounting and live profiling (bench.py's roofline / gpu_launches; no reference counterpart).
 * vitb200_launch_count: kernels launched by this library so far (process-wide).
 * vitb200_prof_begin .. vitb200_prof_end: every kernel launch in between is bracketed by a cudaEvent pair on its
 * launching stream; prof_end synchronises the device and fills out[VITB200_PROF_CLASSES] with, per kernel class,
 * the summed device time, the summed algorithmic work (FLOPs for gemm/attn classes, bytes for the others) and the
 * launch count. */
#define VITB200_PROF_CLASSES 8
typedef struct { double ms; double work; int64_t launches; } vitb200_prof_entry;
int64_t vitb200_launch_count(void);
const char* vitb200_prof_class_name(int cls);
int vitb200_prof_begin(void);
int vitb200_prof_end(vitb200_prof_entry* out);

/* Dropout descriptor: p == 0 disables.  Masks are regenerated from (seed, stream, element index) with
 * Philox4x32-7 (16 random bits per element, one call per 8 elements), so backward never stores them.  Replaces nn.Dropout(0.2) at transformer.py:40,47,59. */
typedef struct {
  float p;
  uint64_t seed;
  uint64_t stream;
} vitb200_dropout;

/* General contraction  C[m,n] = epilogue( sum_k A(m,k) * B(n,k) ).
 *   a_major == MAJOR_K : A stored row-major [M][lda];   MAJOR_MN: A stored [K][lda] (transposed view)
 *   b_major == MAJOR_K : B stored row-major [N][ldb];   MAJOR_MN: B stored [K][ldb]
 * Epilogue, applied in this order to v = acc:
 *   +bias[n]  ->  relu  ->  (mask_aux[m,n] > 0 ? v : 0)  ->  (mask_bits bit set ? v : 0)  ->  dropout
 *   ->  +row_add[(m % row_period), n]  ->  +residual[orow, n]  ->  store / atomic-add into C[orow, n]
 *   (-> column sums into col_sum; -> sign bits of the stored values into bits_out)
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
  int splits;            /* split-K factor (atomic accumulate), 0 = auto only when accumulate != 0; -1 (fp32 inputs):
                          * also allow an atomic split-K into a zeroed C when accumulate == 0 (not bit-reproducible) */
  int accumulate;        /* 1: C += result via fp32 atomics (C must be fp32) */
  float* col_sum;        /* [N] fp32 or NULL: col_sum[n] += sum_m (value stored at C[m,n]) — fused bias gradient */
  /* Bit masks (BF16 inputs only, N % 32 == 0): word (n / 32) * bits_rows + m covers columns [32*(n/32), +32) of row m.
   * bits_out: bit = stored value > 0 (the ReLU mask at 1/16 of the activation's bytes, transformer.py:57);
   * mask_bits: keep v where the bit is set (ReLU backward). */
  uint32_t* bits_out;
  const uint32_t* mask_bits;
  int bits_rows;         /* >= M */
  /* Fused per-head row dot (BF16 inputs, BF16 output, N == 64 * dot_H, M % dot_T == 0, no mask_aux / row map):
   * dot_out[(b * dot_H + h) * dot_T + t] = sum over the 64 columns of head h of (value stored at C[m,n]) * dot_aux[m,n],
   * m = b * dot_T + t.  With C = dO (the gradient of the concatenated head outputs, reference transformer.py:45-46) and
   * dot_aux = O this is delta = rowsum(dO o O), the softmax-backward statistic of transformer.py:26-27's autograd. */
  const void* dot_aux;   /* [M][ldc], C's dtype and layout, or NULL */
  float* dot_out;        /* [M / dot_T][dot_H][dot_T] fp32, or NULL */
  int dot_T;
  int dot_H;
} vitb200_gemm_args;

int vitb200_gemm(const vitb200_gemm_args* args, void* stream);

/* Host arithmetic only (no GPU needed): the split-K count vitb200_gemm picks for an atomically accumulated BF16 contraction
 * (accumulate = 1, splits = 0) of M x N x K on a device with `sms` SMs — the weight-gradient GEMMs dW = dY^T X behind
 * nn.Linear's backward (reference transformer.py:22-28,46,56-58).  Units of equal depth are walked statically by sms / 2 CTA
 * pairs; the rule minimises rounds x (k-blocks per split + hand-over).  -1 on bad arguments. */
int vitb200_gemm_auto_splits(int M, int N, int K, int sms);


/* ------------------------------------------------------------------------------------------------------
 * Single-op entry points (unit-tested one by one against the oracle / torch fp32).
 * ---------------------------------------------------------------------------------------------------- */

/* nn.LayerNorm(E), eps, affine (transformer.py:72-73; vit.py:72).  x fp32 [M,E] -> y [M,E] of y_dtype. */
int vitb200_ln_fwd(const float* x, const float* gamma, const float* beta, void* y, int y_dtype, float* mean,
                   float* rstd, int M, int E, float eps, void* stream);
/* LayerNorm backward.  g (fp32 [M,E]) receives dx (add_to_g: g += dx).  dgamma/dbeta are ACCUMULATED.
 * dyn (optional, dy_dtype) = cast(g_new * dropout factors) — the operand of the next backward GEMM;
 * dyn_colsum (optional, fp32 [E], needs dyn) += column sums of dyn = the bias gradient of the Linear whose
 * dropout-ed output fed this LayerNorm's residual input (transformer.py:46,58). */
int vitb200_ln_bwd(const void* dy, int dy_dtype, const float* x, const float* mean, const float* rstd,
                   const float* gamma, float* g, int add_to_g, float* dgamma, float* dbeta, void* dyn,
                   float* dyn_colsum, vitb200_dropout dyn_dropout, int M, int E, void* stream);

/* Multi-head self-attention core (transformer.py:20-31,44-45): qkv [B*T, 3E] packed (q | k | v, head h at
 * columns h*hs), logits * scale (the reference MULTIPLIES by sqrt(hs)), softmax, @v, heads concatenated into
 * out [B*T, E]; lse [B,H,T] fp32 saved for backward. */
int vitb200_attn_fwd(const void* qkv, void* out, float* lse, int dtype, int B, int T, int H, int hs, float scale,
                     void* stream);
/* dqkv [B*T,3E] from d_out; delta is fp32 scratch [B,H,T]. */
int vitb200_attn_bwd(const void* qkv, const void* out, const void* d_out, const float* lse, float* delta,
                     void* dqkv, int dtype, int B, int T, int H, int hs, float scale, void* stream);
/* NO COUNTERPART IN THE REFERENCE.  Allocation-free variant of
 * vitb200_attn_bwd for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it): same inputs and outputs, but the fp32 scratch `delta` [B,H,T] lives in a file-scope
 * `static __device__` array of VITB200_NOALLOC_MAX_DELTA floats (4 MiB, held by every GPU that loads the library
 * until exit).  Fails with a clear error when B*H*T exceeds that limit.  Calls no cudaMalloc / cudaFree / allocating
 * library routine.  All calls share the one array (it keeps the last call's contents; every value a call reads is
 * written by that call first, so nothing needs clearing): calls that OVERLAP on the GPU (different streams) race —
 * serialise them.  The calling thread's current device must be the device of `stream` and of the buffers. */
#define VITB200_NOALLOC_MAX_DELTA (1 << 20)
int vitb200_attn_bwd_noalloc(const void* qkv, const void* out, const void* d_out, const float* lse, void* dqkv,
                             int dtype, int B, int T, int H, int hs, float scale, void* stream);
/* Debug/export: probs fp32 [B,H,T,T] = MultiHeadAttention.attention_probs (transformer.py:48); fp32 qkv only. */
int vitb200_attn_probs(const float* qkv, float* probs, int B, int T, int H, int hs, float scale, void* stream);

/* Patch gather of Conv2d(C,E,P,stride=P) (vit.py:20-30): x fp32 [B,C,Hi,Wi] -> A [B*N, Kpad] of out_dtype,
 * row b*N + gy*(Wi/P) + gx, column c*P*P + ph*P + pw. */
int vitb200_im2col(const float* x, void* A, int out_dtype, int B, int C, int Hi, int Wi, int P, int Kpad,
                   void* stream);
/* The same gather from uint8 pixels with the ToTensor scaling (uint8 / 255 in fp32, train.py:151-155) fused in, so
 * the host ships 1 byte per pixel instead of 4.  layout 0: x u8 [B,C,Hi,Wi]; layout 1: x u8 [B,Hi,Wi,C]. */
#define VITB200_U8_NCHW 0
#define VITB200_U8_NHWC 1
int vitb200_im2col_u8(const uint8_t* x, int layout, void* A, int out_dtype, int B, int C, int Hi, int Wi, int P,
                      int Kpad, void* stream);
/* out[n] += sum_m in[m,n] (bias gradients). */
int vitb200_colsum(const void* in, int in_dtype, int64_t ld, float* out, int M, int N, void* stream);
/* nn.CrossEntropyLoss() mean + gradient (train.py:81,93): dlogits = grad_scale * d(mean loss)/dlogits, may be NULL. */
int vitb200_cross_entropy(const float* logits, const int64_t* labels, float* loss, float* dlogits, int B, int C,
                          float grad_scale, void* stream);
int vitb200_cast_f32_bf16(const float* in, void* out_bf16, int64_t n, void* stream);
/* out = cast(g * dropout factors of (seed, stream)) — the masked gradient operand of a backward GEMM. */
int vitb200_mask_cast(const float* g, void* out, int out_dtype, vitb200_dropout d, int64_t n, void* stream);
/* keep[i] in {0,1}: the mask the fused dropout epilogues regenerate for (seed, stream). */
int vitb200_dropout_mask(uint8_t* keep, int64_t n, vitb200_dropout d, void* stream);

/* torch.optim.AdamW step (train.py:66,96) over a flat fp32 arena in ONE launch; shadow_bf16 (optional) receives
 * the bf16 copy of the updated parameters; grad_scale folds the data-parallel mean. */
typedef struct {
  float lr, beta1, beta2, eps, weight_decay;
  int step;          /* 1-based */
  float grad_scale;
} vitb200_adamw_hyper;
int vitb200_adamw_flat(float* p, const float* g, float* m, float* v, void* shadow_bf16, int64_t n,
                       vitb200_adamw_hyper h, void* stream);
/* Device-side input pipeline: transforms.Resize((H, W)) of the reference's loader (train.py:151-155)
 * on uint8 planes [planes = B*C, hin, win] -> [planes, hout, wout], bit-identical to PIL's bilinear ImagingResample
 * (horizontal then vertical pass, 22-bit fixed-point triangle filter, each pass rounded to uint8).  *_bounds: int32
 * [out, 2] = (first tap, tap count); *_coeffs: int32 [out, taps] fixed-point weights, built on the host exactly as
 * PIL does (VisionTransformer/input_pipeline.py).  Feed the result to vitb200_embed_fwd_u8 (the /255 is fused there). */
int vitb200_resize_bilinear_u8(const uint8_t* src, uint8_t* dst, int planes, int hin, int win, int hout, int wout,
                               const int32_t* x_bounds, const int32_t* x_coeffs, int x_taps, const int32_t* y_bounds,
                               const int32_t* y_coeffs, int y_taps, void* stream);
/* Gradient-exchange staging (data-parallel trainer): out_bf16 = bf16(g) when out_bf16 != NULL, and g = 0, in one
 * pass over n elements — the bf16 copy feeds the reduce-scatter, the cleared arena is optimizer.zero_grad()
 * (train.py:94).  out_bf16 == NULL: clear only. */
int vitb200_grad_pack(float* g, void* out_bf16, int64_t n, void* stream);
/* Range form (data-parallel trainer, ZeRO-1 shards): AdamW over n elements of arenas already offset to a 16-byte
 * aligned range start.  g_bf16 != NULL: gradients are read from that bf16 buffer (reduce-scattered bf16 gradient
 * exchange) instead of the fp32 g.  zero_grad != 0: the fp32 gradient range is cleared in the same pass
 * (optimizer.zero_grad(), train.py:94 — the backward kernels accumulate into it). */
int vitb200_adamw_range(float* p, float* g, const void* g_bf16, float* m, float* v, void* shadow_bf16, int64_t n,
                        int zero_grad, vitb200_adamw_hyper h, void* stream);
/* Per-step scalars for a training step captured into a CUDA graph (the reference has no counterpart):
 * a 32-byte device struct { uint64 seed_add; int32 step; float bias_corr1; float inv_sqrt_bc2; pad }.
 * vitb200_step_state_advance (one tiny kernel, capturable): step += 1, seed_add = (step - 1) * seed_stride, Adam bias
 * corrections of the new step (double precision, as the host computes them).  Point vitb200_dims.seed_dev at the
 * struct and pass it to vitb200_adamw_range_dev, which is vitb200_adamw_range with the two corrections read from it. */
#define VITB200_STEP_STATE_BYTES 32
int vitb200_step_state_advance(void* step_state, float beta1, float beta2, uint64_t seed_stride, void* stream);
int vitb200_adamw_range_dev(float* p, float* g, const void* g_bf16, float* m, float* v, void* shadow_bf16, int64_t n,
                            int zero_grad, vitb200_adamw_hyper h, const void* step_state, void* stream);
/* Multi-tensor form: device arrays of tensor addresses / sizes and a (block -> tensor, chunk of 65536) table. */
int vitb200_adamw_multi(const uint64_t* p_ptrs, const uint64_t* g_ptrs, const uint64_t* m_ptrs,
                        const uint64_t* v_ptrs, const uint64_t* shadow_ptrs, const int64_t* sizes,
                        const int32_t* blk_tensor, const int32_t* blk_chunk, int n_blocks, vitb200_adamw_hyper h,
                        void* stream);

/* ------------------------------------------------------------------------------------------------------
 * Stage-level entry points: the model forward/backward is embed -> L x block -> head, each stage one call.
 * `dtype` is the activation / GEMM-operand dtype (BF16: tcgen05 path; F32: fp32 parity path).  The residual
 * stream, LayerNorm statistics, biases, LayerNorm parameters, the classifier head and ALL gradients are fp32.
 * ---------------------------------------------------------------------------------------------------- */
typedef struct {
  int B, T, E, H;       /* batch, tokens (= patches + 1), embedding, heads */
  int C, P, Hi, Wi;     /* image channels, patch size, image height / width */
  int classes;
  int dtype;
  float dropout_p;      /* 0 in eval mode; the reference hard-codes 0.2 when training (transformer.py:35,53) */
  uint64_t seed;        /* dropout seed for this step */
  float ln_eps;
  const uint64_t* seed_dev;  /* optional DEVICE word added to `seed` when a mask is drawn (NULL: none).  Kernel arguments
                              * are frozen when a step is captured into a CUDA graph; the per-step part of the seed then
                              * lives in a vitb200 step state (below) that the graph itself advances. */
} vitb200_dims;

/* PatchEmbedding (vit.py:9-43).  w_lp is the dtype copy of w_f32 ([E, C*P*P]); used when dtype == BF16 and
 * C*P*P % 8 == 0, otherwise the fp32 kernel runs on w_f32. */
typedef struct {
  const float* w_f32; const void* w_lp; const float* bias;  /* Conv2d weight (E,C,P,P) and bias (E) */
  const float* cls;   /* (B,1,E) one CLS vector per batch slot, appended LAST */
  const float* pos;   /* (1,T,E) */
} vitb200_embed_params;
typedef struct { float *w, *bias, *cls, *pos; } vitb200_embed_grads;
int64_t vitb200_embed_act_bytes(const vitb200_dims* d);
int vitb200_embed_fwd(const vitb200_dims* d, const vitb200_embed_params* p, const float* images, float* x_out,
                      void* acts, void* stream);
/* Same stage fed with uint8 pixels (the input pipeline of train.py:91,151-162 moved onto the device):
 * identical results to vitb200_embed_fwd on images_u8 / 255. */
int vitb200_embed_fwd_u8(const vitb200_dims* d, const vitb200_embed_params* p, const uint8_t* images_u8, int layout,
                         float* x_out, void* acts, void* stream);
/* g: fp32 [B,T,E] gradient of the embedding output; scratch >= vitb200_embed_act_bytes. Grads are ACCUMULATED. */
int vitb200_embed_bwd(const vitb200_dims* d, const vitb200_embed_grads* gr, const float* g, const void* acts,
                      void* scratch, void* stream);

/* Block (transformer.py:66-79): x + MHA(LN1 x), then x + MLP(LN2 x).  Weights in `dtype`, rest fp32.
 * w_qkv packs the 3H per-head matrices: row s*E + h*hs + j = heads[h].{query,key,value}[s].weight[j]. */
typedef struct {
  const float *ln1_w, *ln1_b, *ln2_w, *ln2_b;
  const void* w_qkv;                      /* [3E, E] */
  const void* w_proj; const float* b_proj; /* [E, E]  */
  const void* w_fc1;  const float* b_fc1;  /* [4E, E] */
  const void* w_fc2;  const float* b_fc2;  /* [E, 4E] */
} vitb200_block_params;
/* b_fc2 == NULL: the producer of dy_in already accumulated this layer's fc2 bias gradient (see b_fc2_prev).
 * b_fc2_prev (optional): fc2 bias gradient of layer-1; the last kernel of this call adds colsum(dy_out) to it. */
typedef struct {
  float *ln1_w, *ln1_b, *ln2_w, *ln2_b, *w_qkv, *w_proj, *b_proj, *w_fc1, *b_fc1, *w_fc2, *b_fc2, *b_fc2_prev;
} vitb200_block_grads;
int64_t vitb200_block_act_bytes(const vitb200_dims* d);
int64_t vitb200_block_scratch_bytes(const vitb200_dims* d);
/* Byte offset, inside a block's `acts` arena, of one saved tensor (-1: unknown id).  Lets the host export
 * MultiHeadAttention.attention_probs (transformer.py:48) from the qkv a whole-model forward already stored. */
enum { VITB200_ACT_QKV = 0, VITB200_ACT_ATTN = 1, VITB200_ACT_LSE = 2, VITB200_ACT_XMID = 3 };
int64_t vitb200_block_act_offset(const vitb200_dims* d, int which);
/* x_in, x_out fp32 [B*T, E] (may not alias).  acts: saved activations for backward (or scratch in eval). */
int vitb200_block_fwd(const vitb200_dims* d, int layer, const vitb200_block_params* p, const float* x_in,
                      float* x_out, void* acts, void* stream);
/* g: fp32 [B*T,E], in: dL/dx_out, out: dL/dx_in.  dy_in: `dtype` [B*T,E] = cast(dL/dx_out * fc2 dropout mask of
 * this layer).  dy_out (optional): same quantity for layer-1, produced for the next call.  Grads ACCUMULATED. */
int vitb200_block_bwd(const vitb200_dims* d, int layer, const vitb200_block_params* p, const vitb200_block_grads* gr,
                      const float* x_in, const void* acts, float* g, const void* dy_in, void* dy_out, void* scratch,
                      void* stream);

/* Last-block variants: the classifier reads ONLY token 0 of the encoder output (vit.py:80 `x[:, 0, :]`), so in the LAST
 * block the output projection, LayerNorm-2 and the MLP (transformer.py:46-47,56-59,78) are needed for the B token-0
 * rows only; LayerNorm-1, q/k/v and attention still run over all tokens.  Logits and every gradient equal those of
 * vitb200_block_fwd / vitb200_block_bwd (the other rows are dead values with exactly zero gradient).
 *   fwd_row0: x0_out fp32 [B, E] = token 0 of each image.
 *   bwd_row0: on entry g fp32 [B, E] and dy0_in `dtype` [B, E] are the compact token-0 rows (as written by
 *             vitb200_head_bwd_row0); on exit g is the full [B*T, E] gradient w.r.t. x_in, dy_out as in block_bwd.
 * g must be the usual [B*T, E] buffer in both cases. */
int vitb200_block_fwd_row0(const vitb200_dims* d, int layer, const vitb200_block_params* p, const float* x_in,
                           float* x0_out, void* acts, void* stream);
int vitb200_block_bwd_row0(const vitb200_dims* d, int layer, const vitb200_block_params* p,
                           const vitb200_block_grads* gr, const float* x_in, const void* acts, float* g,
                           const void* dy0_in, void* dy_out, void* scratch, void* stream);

/* Classifier head on token 0 (vit.py:69-74,80): Linear(E,4E) -> GELU(erf) -> LayerNorm(4E) -> Linear(4E,cls); fp32. */
/* w1_lp (optional, bf16 copy of w1): in bf16 mode the two large head products (x0 W1^T and its backward pair) run
 * on the tensor-core path, as they do under the reference's autocast; NULL keeps the whole head in fp32. */
typedef struct { const float *w1, *b1, *ln_w, *ln_b, *w3, *b3; const void* w1_lp; } vitb200_head_params;
typedef struct { float *w1, *b1, *ln_w, *ln_b, *w3, *b3; } vitb200_head_grads;
int64_t vitb200_head_act_bytes(const vitb200_dims* d);
int vitb200_head_fwd(const vitb200_dims* d, const vitb200_head_params* p, const float* x_last, float* logits,
                     void* acts, void* stream);
/* Produces g (fp32 [B*T,E], zero except token 0) and dy_out = cast(g * fc2-dropout mask of layer n_layers-1). */
int vitb200_head_bwd(const vitb200_dims* d, int n_layers, const vitb200_head_params* p, const vitb200_head_grads* gr,
                     const void* acts, const float* dlogits, float* g, void* dy_out, void* scratch, void* stream);
/* Head fed by / feeding vitb200_block_{fwd,bwd}_row0: x0 fp32 [B, E] compact; g0 fp32 [B, E] and dy0 `dtype` [B, E]
 * compact (dy0 masked with the fc2 dropout of layer n_layers-1 at compact element indices, as block_fwd_row0 drew it). */
int vitb200_head_fwd_row0(const vitb200_dims* d, const vitb200_head_params* p, const float* x0, float* logits,
                          void* acts, void* stream);
int vitb200_head_bwd_row0(const vitb200_dims* d, int n_layers, const vitb200_head_params* p,
                          const vitb200_head_grads* gr, const void* acts, const float* dlogits, float* g0, void* dy0,
                          void* scratch, void* stream);

/* sizeof() of the ABI structs as this library was compiled, for binding self-checks.
 * which: 0 dims, 1 dropout, 2 gemm_args, 3 adamw_hyper, 4 block_params, 5 block_grads, 6 head_params, 7 prof_entry;
 * -1 for anything else. */
int64_t vitb200_abi_sizeof(int which);

#ifdef __cplusplus
}
#endif
#endif /* VITB200_H_ */
