// Stage-level host orchestration: embed -> L x block -> head, forward and backward, each stage one C-ABI call that
// enqueues its kernels on the caller's stream.  No allocation, no synchronisation: every buffer (parameters,
// gradients, saved activations, scratch) is carved from caller-owned memory with the layouts defined here.
#include <cmath>
#include <mutex>

#include "../../include/vitb200.h"
#include "kernels.h"

namespace vb {

namespace {

inline size_t align256(size_t n) { return (n + 255) & ~size_t(255); }

struct Carver {
  uint8_t* base;
  size_t off = 0;
  explicit Carver(void* p) : base(static_cast<uint8_t*>(p)) {}
  template <typename T = void>
  T* take(size_t bytes) {
    T* r = reinterpret_cast<T*>(base + off);
    off += align256(bytes);
    return r;
  }
};

inline size_t esize(int dtype) { return dtype == VITB200_BF16 ? 2 : 4; }
inline bool is_bf16(const vitb200_dims* d) { return d->dtype == VITB200_BF16; }

int check_dims(const vitb200_dims* d) {
  VB_REQUIRE(d != nullptr, "null dims");
  VB_REQUIRE(d->B > 0 && d->T > 1 && d->E > 0 && d->H > 0 && d->E % d->H == 0, "bad dims B=%d T=%d E=%d H=%d", d->B,
             d->T, d->E, d->H);
  VB_REQUIRE(d->dtype == VITB200_F32 || d->dtype == VITB200_BF16, "bad dtype %d", d->dtype);
  if (d->dtype == VITB200_BF16)
    VB_REQUIRE(d->E % 8 == 0, "bf16 path needs embedding_size %% 8 == 0 (got %d); use precision=float32", d->E);
  VB_REQUIRE(static_cast<int64_t>(d->B) * d->H <= 65535, "B*H = %lld exceeds the attention grid limit 65535",
             static_cast<long long>(d->B) * d->H);
  return kOk;
}

// One linear layer call: C = epilogue(A @ W^T) etc., dispatched on dtype.
struct Lin {
  int M, N, K;
  int a_major = 0, b_major = 0;
  const void* A; int64_t lda;
  const void* B; int64_t ldb;
  GemmEpilogue ep;
  bool out_f32;
  int splits = 0;
};
GemmEpilogue make_ep(void* C, int64_t ldc) {
  GemmEpilogue ep;
  ep.bias = nullptr; ep.mask_aux = nullptr; ep.row_add = nullptr; ep.residual = nullptr;
  ep.C = C; ep.ldc = ldc; ep.relu = 0; ep.row_period = 0; ep.row_group_stride = 0; ep.accumulate = 0; ep.col_sum = nullptr;
  ep.bits_out = nullptr; ep.mask_bits = nullptr; ep.bits_rows = 0;
  ep.drop = make_dropout(0.f, 0, 0);
  return ep;
}
int run_lin(bool bf, const Lin& l, cudaStream_t s) {
  if (bf) return gemm_tc_bf16(l.M, l.N, l.K, l.a_major, l.b_major, l.A, l.lda, l.B, l.ldb, l.out_f32, l.splits, l.ep, s);
  return gemm_f32(l.M, l.N, l.K, l.a_major, l.b_major, static_cast<const float*>(l.A), l.lda,
                  static_cast<const float*>(l.B), l.ldb, l.splits, l.ep, s);
}

inline uint64_t drop_stream(int layer, int which) { return static_cast<uint64_t>(layer) * 2 + which; }  // 0 proj, 1 fc2

// ------------------------------------------------------------------------------------------------ embed
struct EmbedGeom {
  int N, K, Kpad;
  bool lp;   // patch GEMM runs on the bf16 tensor-core path
  bool pad;  // K = C*P*P is not a multiple of 8 (P = 14): operands are zero-padded to Kpad columns, the bf16 weight is
             // re-padded from the fp32 master in the stage's own buffer, dW goes through a padded fp32 scratch
};
EmbedGeom embed_geom(const vitb200_dims* d) {
  EmbedGeom g;
  g.N = (d->Hi / d->P) * (d->Wi / d->P);
  g.K = d->C * d->P * d->P;
  g.lp = is_bf16(d);
  g.pad = g.lp && (g.K % 8 != 0);
  g.Kpad = g.pad ? (g.K + 7) / 8 * 8 : g.K;
  return g;
}
struct EmbedBufs { void* A; void* dY; bf16* w_pad; float* dw_pad; };
EmbedBufs carve_embed(const vitb200_dims* d, const EmbedGeom& g, void* base, size_t* total = nullptr) {
  Carver c(base);
  EmbedBufs b;
  const size_t rows = static_cast<size_t>(d->B) * g.N, es = g.lp ? 2 : 4;
  b.A = c.take(rows * g.Kpad * es);        // im2col
  b.dY = c.take(rows * d->E * es);         // backward: compacted gradient of the patch rows
  b.w_pad = g.pad ? c.take<bf16>(static_cast<size_t>(d->E) * g.Kpad * 2) : nullptr;
  b.dw_pad = g.pad ? c.take<float>(static_cast<size_t>(d->E) * g.Kpad * 4) : nullptr;
  if (total) *total = c.off;
  return b;
}

// ------------------------------------------------------------------------------------------------ head
struct HeadActs {
  float *x0, *h1, *a, *n, *mean, *rstd;
  bf16* x0_lp;   // [B,E] bf16 copy of token 0 (tensor-core path of the first Linear)
};
inline bool head_lp(const vitb200_dims* d, const vitb200_head_params* p) {
  return is_bf16(d) && p->w1_lp != nullptr && d->E % 8 == 0;
}
HeadActs carve_head(const vitb200_dims* d, void* acts, size_t* total = nullptr) {
  Carver c(acts);
  HeadActs h;
  const size_t B = d->B, E = d->E;
  h.x0 = c.take<float>(B * E * 4);
  h.h1 = c.take<float>(B * 4 * E * 4);
  h.a = c.take<float>(B * 4 * E * 4);
  h.n = c.take<float>(B * 4 * E * 4);
  h.mean = c.take<float>(B * 4);
  h.rstd = c.take<float>(B * 4);
  h.x0_lp = c.take<bf16>(B * E * 2);
  if (total) *total = c.off;
  return h;
}

// ------------------------------------------------------------------------------------------------ block
struct BlockActs {
  void* ln1;     // [M,E]   dtype
  float *mean1, *rstd1;
  void* qkv;     // [M,3E]  dtype
  void* attn;    // [M,E]   dtype
  float* lse;    // [B,H,T]
  float* x_mid;  // [M,E]   fp32 residual after attention
  void* ln2;     // [M,E]   dtype
  float *mean2, *rstd2;
  void* h;       // [M,4E]  dtype (post-ReLU)
  uint32_t* hbits;  // [4E/32][M] sign bits of h (bf16 mode, 4E % 32 == 0; else null): the ReLU mask of the backward
};
inline bool use_hbits(const vitb200_dims* d) { return is_bf16(d) && (4 * d->E) % 32 == 0; }
BlockActs carve_block(const vitb200_dims* d, void* acts, size_t* total = nullptr) {
  Carver c(acts);
  BlockActs a;
  const size_t M = static_cast<size_t>(d->B) * d->T, E = d->E, es = esize(d->dtype);
  a.ln1 = c.take(M * E * es);
  a.mean1 = c.take<float>(M * 4);
  a.rstd1 = c.take<float>(M * 4);
  a.qkv = c.take(M * 3 * E * es);
  a.attn = c.take(M * E * es);
  a.lse = c.take<float>(static_cast<size_t>(d->B) * d->H * d->T * 4);
  a.x_mid = c.take<float>(M * E * 4);
  a.ln2 = c.take(M * E * es);
  a.mean2 = c.take<float>(M * 4);
  a.rstd2 = c.take<float>(M * 4);
  a.h = c.take(M * 4 * E * es);
  a.hbits = use_hbits(d) ? c.take<uint32_t>(M * (4 * E / 32) * 4) : nullptr;
  if (total) *total = c.off;
  return a;
}
struct BlockScratch {
  void* s1;      // [M,4E] dtype : dh, then dqkv
  void* s2;      // [M,E]  dtype : d(ln2 out), dO, d(ln1 out)
  void* s3;      // [M,E]  dtype : masked dy for proj
  float* delta;  // [B,H,T]
};
BlockScratch carve_scratch(const vitb200_dims* d, void* scratch, size_t* total = nullptr) {
  Carver c(scratch);
  BlockScratch s;
  const size_t M = static_cast<size_t>(d->B) * d->T, E = d->E, es = esize(d->dtype);
  s.s1 = c.take(M * 4 * E * es);
  s.s2 = c.take(M * E * es);
  s.s3 = c.take(M * E * es);
  s.delta = c.take<float>(static_cast<size_t>(d->B) * d->H * d->T * 4);
  if (total) *total = c.off;
  return s;
}

}  // namespace
}  // namespace vb

using namespace vb;

extern "C" {

// ================================================================================================ embed
int64_t vitb200_embed_act_bytes(const vitb200_dims* d) {
  const EmbedGeom g = embed_geom(d);
  size_t total = 0;
  carve_embed(d, g, nullptr, &total);
  return static_cast<int64_t>(total);
}

// images_f32 XOR images_u8 (u8_layout: 0 NCHW, 1 NHWC; the ToTensor /255 is fused into the gather)
static int embed_forward(const vitb200_dims* d, const vitb200_embed_params* p, const float* images,
                         const uint8_t* images_u8, int u8_layout, float* x_out, void* acts, void* stream) {
  VB_PROPAGATE(check_dims(d));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const EmbedGeom g = embed_geom(d);
  VB_REQUIRE(g.N + 1 == d->T, "num_patches (%d) != (H/P)*(W/P) (%d): pos_embd add would fail (reference vit.py:42)",
             d->T - 1, g.N);
  VB_REQUIRE(!g.lp || g.pad || p->w_lp != nullptr, "embed: bf16 weight copy missing");
  const EmbedBufs eb = carve_embed(d, g, acts);
  void* A = eb.A;
  if (g.pad) VB_PROPAGATE(cast_pad_rows(p->w_f32, eb.w_pad, d->E, g.K, g.Kpad, s));
  if (images_u8) VB_PROPAGATE(im2col_u8(images_u8, u8_layout, A, g.lp, d->B, d->C, d->Hi, d->Wi, d->P, g.Kpad, s));
  else VB_PROPAGATE(im2col(images, A, g.lp, d->B, d->C, d->Hi, d->Wi, d->P, g.Kpad, s));
  Lin l;
  l.M = d->B * g.N; l.N = d->E; l.K = g.Kpad;   // pad columns are zero in both operands
  l.A = A; l.lda = g.Kpad;
  l.B = g.pad ? static_cast<const void*>(eb.w_pad) : g.lp ? p->w_lp : static_cast<const void*>(p->w_f32);
  l.ldb = g.Kpad;
  l.ep = make_ep(x_out, d->E);
  l.ep.bias = p->bias;
  l.ep.row_add = p->pos;           // + pos[token]
  l.ep.row_period = g.N;           // token = m % N, image = m / N
  l.ep.row_group_stride = d->T;    // output row = image*T + token (row T-1 is the CLS slot)
  l.out_f32 = true;
  VB_PROPAGATE(run_lin(g.lp, l, s));
  return cls_rows(p->cls, p->pos, x_out, d->B, d->T, d->E, s);
}

int vitb200_embed_fwd(const vitb200_dims* d, const vitb200_embed_params* p, const float* images, float* x_out,
                      void* acts, void* stream) {
  VB_REQUIRE(images != nullptr, "embed_fwd: images is NULL");
  return embed_forward(d, p, images, nullptr, 0, x_out, acts, stream);
}

int vitb200_embed_fwd_u8(const vitb200_dims* d, const vitb200_embed_params* p, const uint8_t* images_u8, int layout,
                         float* x_out, void* acts, void* stream) {
  VB_REQUIRE(images_u8 != nullptr, "embed_fwd_u8: images is NULL");
  return embed_forward(d, p, nullptr, images_u8, layout, x_out, acts, stream);
}

int vitb200_embed_bwd(const vitb200_dims* d, const vitb200_embed_grads* gr, const float* g_in, const void* acts,
                      void* scratch, void* stream) {
  VB_PROPAGATE(check_dims(d));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const EmbedGeom g = embed_geom(d);
  const size_t rows = static_cast<size_t>(d->B) * g.N;
  const void* A = acts;
  // dY (and the padded dW scratch) live after A in the same carve order as forward
  const EmbedBufs eb = carve_embed(d, g, scratch);
  void* dY = eb.dY;
  VB_PROPAGATE(embed_bwd_reduce(g_in, gr->pos, gr->cls, dY, g.lp, d->B, d->T, d->E, s));
  VB_PROPAGATE(colsum(dY, g.lp, d->E, gr->bias, static_cast<int>(rows), d->E, s));
  Lin l;  // dW[E, K] += dY^T[E, rows] @ A[rows, K]
  l.M = d->E; l.N = g.Kpad; l.K = static_cast<int>(rows);
  l.a_major = 1; l.b_major = 1;
  l.A = dY; l.lda = d->E;
  l.B = A; l.ldb = g.Kpad;
  l.ep = make_ep(g.pad ? eb.dw_pad : gr->w, g.Kpad);
  l.ep.accumulate = 1;
  l.out_f32 = true;
  if (g.pad) VB_CHECK_CUDA(cudaMemsetAsync(eb.dw_pad, 0, static_cast<size_t>(d->E) * g.Kpad * 4, s));
  VB_PROPAGATE(run_lin(g.lp, l, s));
  if (g.pad) VB_PROPAGATE(add_unpad_rows(gr->w, eb.dw_pad, d->E, g.K, g.Kpad, s));
  return kOk;
}

// ================================================================================================ head
int64_t vitb200_head_act_bytes(const vitb200_dims* d) {
  size_t total = 0;
  carve_head(d, nullptr, &total);
  // backward scratch (dn [B,4E] + dx0 [B,E]) is carved from a separate buffer of the same size bound
  return static_cast<int64_t>(total);
}

// Tx: tokens per image in x_last / g (d->T for the full residual stream, 1 when the last block ran in row-0 mode and
// handed over only token 0 of every image)
static int head_forward(const vitb200_dims* d, const vitb200_head_params* p, const float* x_last, float* logits,
                        void* acts, void* stream, int Tx) {
  VB_PROPAGATE(check_dims(d));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const HeadActs a = carve_head(d, acts);
  const int B = d->B, E = d->E, F = 4 * d->E;
  VB_PROPAGATE(gather_row0(x_last, a.x0, B, Tx, E, s));
  GemmEpilogue ep = make_ep(a.h1, F);
  ep.bias = p->b1;
  if (head_lp(d, p)) {
    VB_PROPAGATE(cast_f32_to_bf16(a.x0, a.x0_lp, static_cast<int64_t>(B) * E, s));
    VB_PROPAGATE(gemm_tc_bf16(B, F, E, 0, 0, a.x0_lp, E, p->w1_lp, E, true, 1, ep, s));
  } else {
    VB_PROPAGATE(gemm_f32(B, F, E, 0, 0, a.x0, E, p->w1, E, 1, ep, s));
  }
  VB_PROPAGATE(gelu_forward(a.h1, a.a, static_cast<int64_t>(B) * F, s));
  VB_PROPAGATE(ln_forward(a.a, p->ln_w, p->ln_b, a.n, false, a.mean, a.rstd, B, F, d->ln_eps, s));
  GemmEpilogue ep3 = make_ep(logits, d->classes);
  ep3.bias = p->b3;
  return gemm_f32(B, d->classes, F, 0, 0, a.n, F, p->w3, F, 1, ep3, s);
}

int vitb200_head_fwd(const vitb200_dims* d, const vitb200_head_params* p, const float* x_last, float* logits,
                     void* acts, void* stream) {
  return head_forward(d, p, x_last, logits, acts, stream, d ? d->T : 0);
}
int vitb200_head_fwd_row0(const vitb200_dims* d, const vitb200_head_params* p, const float* x0, float* logits,
                          void* acts, void* stream) {
  return head_forward(d, p, x0, logits, acts, stream, 1);
}

static int head_backward(const vitb200_dims* d, int n_layers, const vitb200_head_params* p, const vitb200_head_grads* gr,
                         const void* acts, const float* dlogits, float* g, void* dy_out, void* scratch, void* stream,
                         int Tx) {
  VB_PROPAGATE(check_dims(d));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const HeadActs a = carve_head(d, const_cast<void*>(acts));
  const int B = d->B, E = d->E, F = 4 * d->E, Cc = d->classes;
  Carver c(scratch);
  float* dn = c.take<float>(static_cast<size_t>(B) * F * 4);
  float* dx0 = c.take<float>(static_cast<size_t>(B) * E * 4);
  // classifier: dW3 += dlogits^T n ; db3 += colsum ; dn = dlogits W3
  GemmEpilogue e = make_ep(gr->w3, F);
  e.accumulate = 1;
  VB_PROPAGATE(gemm_f32(Cc, F, B, 1, 1, dlogits, Cc, a.n, F, 1, e, s));
  VB_PROPAGATE(colsum(dlogits, false, Cc, gr->b3, B, Cc, s));
  e = make_ep(dn, F);
  VB_PROPAGATE(gemm_f32(B, F, Cc, 0, 1, dlogits, Cc, p->w3, F, 1, e, s));
  // LayerNorm(4E) backward in place (dn := d a), then GELU'
  VB_PROPAGATE(ln_backward(dn, false, a.a, a.mean, a.rstd, p->ln_w, dn, 0, gr->ln_w, gr->ln_b, nullptr, nullptr,
                           make_dropout(0.f, 0, 0), B, F, s));
  VB_PROPAGATE(gelu_backward(a.h1, dn, static_cast<int64_t>(B) * F, s));
  // first linear: dW1 += dh1^T x0 ; db1 ; dx0 = dh1 W1
  VB_PROPAGATE(colsum(dn, false, F, gr->b1, B, F, s));
  if (head_lp(d, p)) {
    bf16* dn_lp = c.take<bf16>(static_cast<size_t>(B) * F * 2);
    VB_PROPAGATE(cast_f32_to_bf16(dn, dn_lp, static_cast<int64_t>(B) * F, s));
    e = make_ep(gr->w1, E);
    e.accumulate = 1;
    VB_PROPAGATE(gemm_tc_bf16(F, E, B, 1, 1, dn_lp, F, a.x0_lp, E, true, 0, e, s));
    e = make_ep(dx0, E);
    VB_PROPAGATE(gemm_tc_bf16(B, E, F, 0, 1, dn_lp, F, p->w1_lp, E, true, 1, e, s));
  } else {
    e = make_ep(gr->w1, E);
    e.accumulate = 1;
    VB_PROPAGATE(gemm_f32(F, E, B, 1, 1, dn, F, a.x0, E, 1, e, s));
    e = make_ep(dx0, E);
    VB_PROPAGATE(gemm_f32(B, E, F, 0, 1, dn, F, p->w1, E, -1, e, s));  // few tiles, K = 4E: atomic split-K allowed
  }
  // scatter into the residual-stream gradient (token 0 only) + masked low-precision copy for the last fc2
  const Dropout dr = make_dropout(d->dropout_p, d->seed, drop_stream(n_layers - 1, 1), d->seed_dev);
  return head_scatter(dx0, g, dy_out, is_bf16(d), dr, B, Tx, E, s);
}

int vitb200_head_bwd(const vitb200_dims* d, int n_layers, const vitb200_head_params* p, const vitb200_head_grads* gr,
                     const void* acts, const float* dlogits, float* g, void* dy_out, void* scratch, void* stream) {
  return head_backward(d, n_layers, p, gr, acts, dlogits, g, dy_out, scratch, stream, d ? d->T : 0);
}
int vitb200_head_bwd_row0(const vitb200_dims* d, int n_layers, const vitb200_head_params* p,
                          const vitb200_head_grads* gr, const void* acts, const float* dlogits, float* g0, void* dy0,
                          void* scratch, void* stream) {
  return head_backward(d, n_layers, p, gr, acts, dlogits, g0, dy0, scratch, stream, 1);
}

}  // extern "C"

// ================================================================================================ block
extern "C" {

int64_t vitb200_block_act_bytes(const vitb200_dims* d) {
  size_t total = 0;
  carve_block(d, nullptr, &total);
  return static_cast<int64_t>(total);
}
int64_t vitb200_block_scratch_bytes(const vitb200_dims* d) {
  size_t total = 0;
  carve_scratch(d, nullptr, &total);
  return static_cast<int64_t>(total);
}
int64_t vitb200_block_act_offset(const vitb200_dims* d, int which) {
  const BlockActs a = carve_block(d, nullptr);
  const void* ptr = nullptr;
  switch (which) {
    case VITB200_ACT_QKV: ptr = a.qkv; break;
    case VITB200_ACT_ATTN: ptr = a.attn; break;
    case VITB200_ACT_LSE: ptr = a.lse; break;
    case VITB200_ACT_XMID: ptr = a.x_mid; break;
    default: return -1;
  }
  return static_cast<int64_t>(reinterpret_cast<uintptr_t>(ptr));
}

}  // extern "C"

// row0: ONLY token 0 of every image leaves this block (the classifier reads x[:, 0, :] of the LAST block, reference
// vit.py:80).  LayerNorm-1, the q/k/v projection and attention still run on all B*T rows (token 0 attends to every
// key), but the output projection, LayerNorm-2 and the MLP run on the B token-0 rows only; every stored tensor after
// attention is then a compact [B, ...] matrix at the front of its usual buffer.  Logits and all gradients are those
// of the full computation (the other rows are dead: nothing reads them, their gradient is exactly zero).
static int block_forward(const vitb200_dims* d, int layer, const vitb200_block_params* p, const float* x_in,
                         float* x_out, void* acts, void* stream, bool row0) {
  VB_PROPAGATE(check_dims(d));
  VB_REQUIRE(x_in != x_out, "block_fwd: x_in and x_out may not alias");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const bool bf = is_bf16(d);
  const int M = d->B * d->T, E = d->E, hs = E / d->H;
  const int Mt = row0 ? d->B : M;                        // rows of the block tail (proj .. fc2)
  const int64_t attn_ld = row0 ? static_cast<int64_t>(d->T) * E : E;   // token-0 rows of attn, strided view
  const BlockActs a = carve_block(d, acts);
  const float scale = sqrtf(static_cast<float>(hs));  // reference MULTIPLIES by sqrt(head_size): transformer.py:24

  // x_mid = x_in + dropout(proj(attn(LN1 x_in)))
  VB_PROPAGATE(ln_forward(x_in, p->ln1_w, p->ln1_b, a.ln1, bf, a.mean1, a.rstd1, M, E, d->ln_eps, s));
  Lin l;
  l.M = M; l.N = 3 * E; l.K = E; l.A = a.ln1; l.lda = E; l.B = p->w_qkv; l.ldb = E;
  l.ep = make_ep(a.qkv, 3 * E); l.out_f32 = !bf;
  VB_PROPAGATE(run_lin(bf, l, s));
  if (bf) VB_PROPAGATE(attn_bf16_forward(static_cast<const bf16*>(a.qkv), static_cast<bf16*>(a.attn), a.lse, d->B, d->T, d->H, hs, scale, s));
  else VB_PROPAGATE(attn_f32_forward(static_cast<const float*>(a.qkv), static_cast<float*>(a.attn), a.lse, d->B, d->T, d->H, hs, scale, s));
  const float* res_in = x_in;
  if (row0) {   // compact copy of the token-0 rows of x_in behind x_mid's own B rows (the residual operand shares ldc)
    float* x_in0 = a.x_mid + static_cast<size_t>(d->B) * E;
    VB_CHECK_CUDA(cudaMemcpy2DAsync(x_in0, static_cast<size_t>(E) * 4, x_in, static_cast<size_t>(d->T) * E * 4,
                                    static_cast<size_t>(E) * 4, d->B, cudaMemcpyDeviceToDevice, s));
    res_in = x_in0;
  }
  l = Lin();
  l.M = Mt; l.N = E; l.K = E; l.A = a.attn; l.lda = attn_ld; l.B = p->w_proj; l.ldb = E;
  l.ep = make_ep(a.x_mid, E); l.ep.bias = p->b_proj; l.ep.residual = res_in;
  l.ep.drop = make_dropout(d->dropout_p, d->seed, drop_stream(layer, 0), d->seed_dev);
  l.out_f32 = true;
  VB_PROPAGATE(run_lin(bf, l, s));

  // x_out = x_mid + dropout(fc2(relu(fc1(LN2 x_mid))))
  VB_PROPAGATE(ln_forward(a.x_mid, p->ln2_w, p->ln2_b, a.ln2, bf, a.mean2, a.rstd2, Mt, E, d->ln_eps, s));
  l = Lin();
  l.M = Mt; l.N = 4 * E; l.K = E; l.A = a.ln2; l.lda = E; l.B = p->w_fc1; l.ldb = E;
  l.ep = make_ep(a.h, 4 * E); l.ep.bias = p->b_fc1; l.ep.relu = 1; l.out_f32 = !bf;
  if (a.hbits) { l.ep.bits_out = a.hbits; l.ep.bits_rows = Mt; }
  VB_PROPAGATE(run_lin(bf, l, s));
  l = Lin();
  l.M = Mt; l.N = E; l.K = 4 * E; l.A = a.h; l.lda = 4 * E; l.B = p->w_fc2; l.ldb = 4 * E;
  l.ep = make_ep(x_out, E); l.ep.bias = p->b_fc2; l.ep.residual = a.x_mid;
  l.ep.drop = make_dropout(d->dropout_p, d->seed, drop_stream(layer, 1), d->seed_dev);
  l.out_f32 = true;
  return run_lin(bf, l, s);
}

extern "C" {

int vitb200_block_fwd(const vitb200_dims* d, int layer, const vitb200_block_params* p, const float* x_in,
                      float* x_out, void* acts, void* stream) {
  return block_forward(d, layer, p, x_in, x_out, acts, stream, false);
}
int vitb200_block_fwd_row0(const vitb200_dims* d, int layer, const vitb200_block_params* p, const float* x_in,
                           float* x0_out, void* acts, void* stream) {
  return block_forward(d, layer, p, x_in, x0_out, acts, stream, true);
}

}  // extern "C"

static int block_backward(const vitb200_dims* d, int layer, const vitb200_block_params* p, const vitb200_block_grads* gr,
                      const float* x_in, const void* acts, float* g, const void* dy_in, void* dy_out, void* scratch,
                      void* stream, bool row0) {
  VB_PROPAGATE(check_dims(d));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const bool bf = is_bf16(d);
  const int M_all = d->B * d->T, E = d->E, F = 4 * d->E, hs = E / d->H;
  // row0 (see block_forward): on entry g and dy_in hold the B token-0 rows only, compact; the MLP / LayerNorm-2 /
  // projection backward runs on those B rows, then the gradient is spread back over the full [B*T, E] layout (zero
  // in every other row) for the attention / q-k-v / LayerNorm-1 backward, which see all tokens.
  int M = row0 ? d->B : M_all;
  const int64_t attn_ld = row0 ? static_cast<int64_t>(d->T) * E : E;
  const BlockActs a = carve_block(d, const_cast<void*>(acts));
  const BlockScratch sc = carve_scratch(d, scratch);
  const float scale = sqrtf(static_cast<float>(hs));
  Lin l;

  // ---- MLP branch.  dy_in = d(fc2 dropout output) already masked & cast.
  // dW2[E,4E] += dy^T h ; db2 += colsum(dy)
  l = Lin(); l.M = E; l.N = F; l.K = M; l.a_major = 1; l.b_major = 1; l.A = dy_in; l.lda = E; l.B = a.h; l.ldb = F;
  l.ep = make_ep(gr->w_fc2, F); l.ep.accumulate = 1; l.out_f32 = true;
  VB_PROPAGATE(run_lin(bf, l, s));
  if (gr->b_fc2) VB_PROPAGATE(colsum(dy_in, bf, E, gr->b_fc2, M, E, s));
  // dh[M,4E] = (dy W2) * (h > 0)
  l = Lin(); l.M = M; l.N = F; l.K = E; l.b_major = 1; l.A = dy_in; l.lda = E; l.B = p->w_fc2; l.ldb = F;
  l.ep = make_ep(sc.s1, F); l.ep.col_sum = gr->b_fc1; l.out_f32 = !bf;  // db1 += colsum(dh), fused
  if (a.hbits) { l.ep.mask_bits = a.hbits; l.ep.bits_rows = M; }   // 1 bit per element instead of re-reading h
  else l.ep.mask_aux = a.h;
  VB_PROPAGATE(run_lin(bf, l, s));
  // dW1[4E,E] += dh^T ln2
  l = Lin(); l.M = F; l.N = E; l.K = M; l.a_major = 1; l.b_major = 1; l.A = sc.s1; l.lda = F; l.B = a.ln2; l.ldb = E;
  l.ep = make_ep(gr->w_fc1, E); l.ep.accumulate = 1; l.out_f32 = true;
  VB_PROPAGATE(run_lin(bf, l, s));
  // d(ln2 out)[M,E] = dh W1
  l = Lin(); l.M = M; l.N = E; l.K = F; l.b_major = 1; l.A = sc.s1; l.lda = F; l.B = p->w_fc1; l.ldb = E;
  l.ep = make_ep(sc.s2, E); l.out_f32 = !bf;
  VB_PROPAGATE(run_lin(bf, l, s));
  // LN2 backward: g += dx ; emits s3 = cast(g * proj-dropout mask)
  VB_PROPAGATE(ln_backward(sc.s2, bf, a.x_mid, a.mean2, a.rstd2, p->ln2_w, g, 1, gr->ln2_w, gr->ln2_b, sc.s3,
                           gr->b_proj, make_dropout(d->dropout_p, d->seed, drop_stream(layer, 0), d->seed_dev), M, E, s));

  // ---- attention branch
  // dWproj[E,E] += s3^T attn  (dbproj += colsum(s3) came out of the LN2 backward above)
  l = Lin(); l.M = E; l.N = E; l.K = M; l.a_major = 1; l.b_major = 1; l.A = sc.s3; l.lda = E; l.B = a.attn; l.ldb = attn_ld;
  l.ep = make_ep(gr->w_proj, E); l.ep.accumulate = 1; l.out_f32 = true;
  VB_PROPAGATE(run_lin(bf, l, s));
  if (row0) {
    // dO is zero except in the token-0 rows (written by the GEMM below through ldc = T*E); the residual-stream
    // gradient g goes from compact [B,E] to full [B*T,E] with zeros elsewhere (staged through the free s1 buffer)
    const size_t esz = bf ? 2 : 4;
    VB_CHECK_CUDA(cudaMemsetAsync(sc.s2, 0, static_cast<size_t>(M_all) * E * esz, s));
    VB_CHECK_CUDA(cudaMemcpyAsync(sc.s1, g, static_cast<size_t>(d->B) * E * 4, cudaMemcpyDeviceToDevice, s));
    VB_CHECK_CUDA(cudaMemsetAsync(g, 0, static_cast<size_t>(M_all) * E * 4, s));
    VB_CHECK_CUDA(cudaMemcpy2DAsync(g, static_cast<size_t>(d->T) * E * 4, sc.s1, static_cast<size_t>(E) * 4,
                                    static_cast<size_t>(E) * 4, d->B, cudaMemcpyDeviceToDevice, s));
  }
  // dO[M,E] = s3 Wproj
  l = Lin(); l.M = M; l.N = E; l.K = E; l.b_major = 1; l.A = sc.s3; l.lda = E; l.B = p->w_proj; l.ldb = E;
  l.ep = make_ep(sc.s2, attn_ld); l.out_f32 = !bf;
  // VITB200_FUSED_DELTA=1: delta = rowsum(dO o O) per (image, head, token) straight from this GEMM's epilogue (64-wide
  // heads, every row computed; bit-identical to the stand-alone kernel).  Off by default: measured NEUTRAL — the K = 768
  // GEMM is epilogue-bound and grows 47.7 -> 68.1 us, the delta kernel it replaces takes ~29 us; step 32.65 vs 32.63 ms
  // (profiles/r02_fused_delta_ab.txt).
  static const bool fuse_delta_on = getenv("VITB200_FUSED_DELTA") && atoi(getenv("VITB200_FUSED_DELTA")) != 0;
  const bool fuse_delta = fuse_delta_on && bf && !row0 && hs == 64 && M == M_all;
  if (fuse_delta) { l.ep.dot_aux = a.attn; l.ep.dot_out = sc.delta; l.ep.dot_T = d->T; l.ep.dot_H = d->H; }
  VB_PROPAGATE(run_lin(bf, l, s));
  M = M_all;   // from here on every token takes part
  // dqkv[M,3E] (into s1)
  if (bf) VB_PROPAGATE(attn_bf16_backward(static_cast<const bf16*>(a.qkv), static_cast<const bf16*>(a.attn), static_cast<const bf16*>(sc.s2), a.lse, sc.delta, static_cast<bf16*>(sc.s1), d->B, d->T, d->H, hs, scale, s, fuse_delta));
  else VB_PROPAGATE(attn_f32_backward(static_cast<const float*>(a.qkv), static_cast<const float*>(a.attn), static_cast<const float*>(sc.s2), a.lse, sc.delta, static_cast<float*>(sc.s1), d->B, d->T, d->H, hs, scale, s));
  // dWqkv[3E,E] += dqkv^T ln1
  l = Lin(); l.M = 3 * E; l.N = E; l.K = M; l.a_major = 1; l.b_major = 1; l.A = sc.s1; l.lda = 3 * E; l.B = a.ln1; l.ldb = E;
  l.ep = make_ep(gr->w_qkv, E); l.ep.accumulate = 1; l.out_f32 = true;
  VB_PROPAGATE(run_lin(bf, l, s));
  // d(ln1 out)[M,E] = dqkv Wqkv
  l = Lin(); l.M = M; l.N = E; l.K = 3 * E; l.b_major = 1; l.A = sc.s1; l.lda = 3 * E; l.B = p->w_qkv; l.ldb = E;
  l.ep = make_ep(sc.s2, E); l.out_f32 = !bf;
  VB_PROPAGATE(run_lin(bf, l, s));
  // LN1 backward: g += dx ; emits dy_out = cast(g * fc2-dropout mask of layer-1) for the previous block
  const Dropout prev = make_dropout(layer > 0 ? d->dropout_p : 0.f, d->seed, drop_stream(layer - 1, 1), d->seed_dev);
  return ln_backward(sc.s2, bf, x_in, a.mean1, a.rstd1, p->ln1_w, g, 1, gr->ln1_w, gr->ln1_b,
                     layer > 0 ? dy_out : nullptr, layer > 0 ? gr->b_fc2_prev : nullptr, prev, M, E, s);
}

extern "C" {

int vitb200_block_bwd(const vitb200_dims* d, int layer, const vitb200_block_params* p, const vitb200_block_grads* gr,
                      const float* x_in, const void* acts, float* g, const void* dy_in, void* dy_out, void* scratch,
                      void* stream) {
  return block_backward(d, layer, p, gr, x_in, acts, g, dy_in, dy_out, scratch, stream, false);
}
int vitb200_block_bwd_row0(const vitb200_dims* d, int layer, const vitb200_block_params* p,
                           const vitb200_block_grads* gr, const float* x_in, const void* acts, float* g,
                           const void* dy0_in, void* dy_out, void* scratch, void* stream) {
  return block_backward(d, layer, p, gr, x_in, acts, g, dy0_in, dy_out, scratch, stream, true);
}

// ================================================================================================ single ops
int vitb200_ln_fwd(const float* x, const float* gamma, const float* beta, void* y, int y_dtype, float* mean,
                   float* rstd, int M, int E, float eps, void* stream) {
  return ln_forward(x, gamma, beta, y, y_dtype == VITB200_BF16, mean, rstd, M, E, eps, static_cast<cudaStream_t>(stream));
}
int vitb200_ln_bwd(const void* dy, int dy_dtype, const float* x, const float* mean, const float* rstd,
                   const float* gamma, float* g, int add_to_g, float* dgamma, float* dbeta, void* dyn,
                   float* dyn_colsum, vitb200_dropout dd, int M, int E, void* stream) {
  return ln_backward(dy, dy_dtype == VITB200_BF16, x, mean, rstd, gamma, g, add_to_g, dgamma, dbeta, dyn, dyn_colsum,
                     make_dropout(dd.p, dd.seed, dd.stream), M, E, static_cast<cudaStream_t>(stream));
}
int vitb200_attn_fwd(const void* qkv, void* out, float* lse, int dtype, int B, int T, int H, int hs, float scale,
                     void* stream) {
  VB_REQUIRE(static_cast<int64_t>(B) * H <= 65535, "attention: B*H exceeds 65535");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (dtype == VITB200_BF16)
    return attn_bf16_forward(static_cast<const bf16*>(qkv), static_cast<bf16*>(out), lse, B, T, H, hs, scale, s);
  return attn_f32_forward(static_cast<const float*>(qkv), static_cast<float*>(out), lse, B, T, H, hs, scale, s);
}
int vitb200_attn_bwd(const void* qkv, const void* out, const void* d_out, const float* lse, float* delta,
                     void* dqkv, int dtype, int B, int T, int H, int hs, float scale, void* stream) {
  VB_REQUIRE(static_cast<int64_t>(B) * H <= 65535, "attention: B*H exceeds 65535");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (dtype == VITB200_BF16)
    return attn_bf16_backward(static_cast<const bf16*>(qkv), static_cast<const bf16*>(out), static_cast<const bf16*>(d_out),
                              lse, delta, static_cast<bf16*>(dqkv), B, T, H, hs, scale, s);
  return attn_f32_backward(static_cast<const float*>(qkv), static_cast<const float*>(out), static_cast<const float*>(d_out),
                           lse, delta, static_cast<float*>(dqkv), B, T, H, hs, scale, s);
}
// ---- allocation-free variant of vitb200_attn_bwd (see include/vitb200.h) --
// The backward attention kernels need one fp32 scratch value per (image, head, query): delta = rowsum(dO o O).  The
// regular entry point takes that buffer from the caller; this one keeps it in a file-scope device array, so callers
// that may not allocate (CUDA graph capture, serving loops) pass inputs and outputs only.  In HOST code the array's
// name is only a host-side stand-in: its device address is fetched once per device with cudaGetSymbolAddress (with
// that device current) and that pointer is what the kernels receive.
static __device__ __align__(256) float g_noalloc_delta[VITB200_NOALLOC_MAX_DELTA];

static int noalloc_delta_ptr(float** out) {
  constexpr int kMaxDev = 64;
  static float* ptrs[kMaxDev] = {};
  static std::mutex mu;
  int dev = -1;
  VB_CHECK_CUDA(cudaGetDevice(&dev));
  VB_REQUIRE(dev >= 0 && dev < kMaxDev, "attn_bwd_noalloc: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    void* p = nullptr;
    VB_CHECK_CUDA(cudaGetSymbolAddress(&p, g_noalloc_delta));   // set-up, once per device, that device current
    VB_REQUIRE(p != nullptr, "attn_bwd_noalloc: cudaGetSymbolAddress returned a null address");
    ptrs[dev] = static_cast<float*>(p);
  }
  *out = ptrs[dev];
  return kOk;
}

int vitb200_attn_bwd_noalloc(const void* qkv, const void* out, const void* d_out, const float* lse, void* dqkv,
                             int dtype, int B, int T, int H, int hs, float scale, void* stream) {
  VB_REQUIRE(B > 0 && T > 0 && H > 0 && hs > 0, "attn_bwd_noalloc: empty problem B=%d T=%d H=%d hs=%d", B, T, H, hs);
  const int64_t need = static_cast<int64_t>(B) * H * T;
  VB_REQUIRE(need <= static_cast<int64_t>(VITB200_NOALLOC_MAX_DELTA),
             "attn_bwd_noalloc: B*H*T = %lld scratch values exceed the static limit VITB200_NOALLOC_MAX_DELTA = %d; "
             "use vitb200_attn_bwd with a caller-provided delta buffer", static_cast<long long>(need),
             VITB200_NOALLOC_MAX_DELTA);
  float* delta = nullptr;
  VB_PROPAGATE(noalloc_delta_ptr(&delta));
  // every one of the B*H*T values is written by the delta kernel before the backward kernels read it: nothing in the
  // shared array has to start at zero, and stale contents of an earlier (larger) call are never read
  return vitb200_attn_bwd(qkv, out, d_out, lse, delta, dqkv, dtype, B, T, H, hs, scale, stream);
}
int vitb200_attn_probs(const float* qkv, float* probs, int B, int T, int H, int hs, float scale, void* stream) {
  VB_REQUIRE(static_cast<int64_t>(B) * H <= 65535, "attention: B*H exceeds 65535");
  return attn_f32_probs(qkv, probs, B, T, H, hs, scale, static_cast<cudaStream_t>(stream));
}
int vitb200_im2col(const float* x, void* A, int out_dtype, int B, int C, int Hi, int Wi, int P, int Kpad,
                   void* stream) {
  return im2col(x, A, out_dtype == VITB200_BF16, B, C, Hi, Wi, P, Kpad, static_cast<cudaStream_t>(stream));
}
int vitb200_im2col_u8(const uint8_t* x, int layout, void* A, int out_dtype, int B, int C, int Hi, int Wi, int P,
                      int Kpad, void* stream) {
  return im2col_u8(x, layout, A, out_dtype == VITB200_BF16, B, C, Hi, Wi, P, Kpad, static_cast<cudaStream_t>(stream));
}
int vitb200_colsum(const void* in, int in_dtype, int64_t ld, float* out, int M, int N, void* stream) {
  return colsum(in, in_dtype == VITB200_BF16, ld, out, M, N, static_cast<cudaStream_t>(stream));
}
int vitb200_cross_entropy(const float* logits, const int64_t* labels, float* loss, float* dlogits, int B, int C,
                          float grad_scale, void* stream) {
  return cross_entropy(logits, labels, loss, dlogits, B, C, grad_scale, static_cast<cudaStream_t>(stream));
}
int vitb200_cast_f32_bf16(const float* in, void* out_bf16, int64_t n, void* stream) {
  return cast_f32_to_bf16(in, static_cast<bf16*>(out_bf16), n, static_cast<cudaStream_t>(stream));
}
int vitb200_mask_cast(const float* g, void* out, int out_dtype, vitb200_dropout dd, int64_t n, void* stream) {
  return mask_cast(g, out, out_dtype == VITB200_BF16, make_dropout(dd.p, dd.seed, dd.stream), n, static_cast<cudaStream_t>(stream));
}
int vitb200_dropout_mask(uint8_t* keep, int64_t n, vitb200_dropout dd, void* stream) {
  return dropout_mask(keep, n, make_dropout(dd.p, dd.seed, dd.stream), static_cast<cudaStream_t>(stream));
}

static AdamWHyper to_hyper(const vitb200_adamw_hyper& h) {
  AdamWHyper k;
  k.lr = h.lr; k.beta1 = h.beta1; k.beta2 = h.beta2; k.eps = h.eps; k.weight_decay = h.weight_decay;
  k.bias_corr1 = static_cast<float>(1.0 - std::pow(static_cast<double>(h.beta1), h.step));
  k.inv_sqrt_bc2 = static_cast<float>(1.0 / std::sqrt(1.0 - std::pow(static_cast<double>(h.beta2), h.step)));
  k.grad_scale = h.grad_scale;
  k.corr_dev = nullptr;
  return k;
}
int vitb200_adamw_flat(float* p, const float* g, float* m, float* v, void* shadow_bf16, int64_t n,
                       vitb200_adamw_hyper h, void* stream) {
  VB_REQUIRE(h.step >= 1, "adamw: step must be >= 1");
  return adamw_flat(p, g, m, v, static_cast<bf16*>(shadow_bf16), n, to_hyper(h), static_cast<cudaStream_t>(stream));
}
int vitb200_resize_bilinear_u8(const uint8_t* src, uint8_t* dst, int planes, int hin, int win, int hout, int wout,
                               const int32_t* x_bounds, const int32_t* x_coeffs, int x_taps, const int32_t* y_bounds,
                               const int32_t* y_coeffs, int y_taps, void* stream) {
  return resize_bilinear_u8(src, dst, planes, hin, win, hout, wout, x_bounds, x_coeffs, x_taps, y_bounds, y_coeffs,
                            y_taps, static_cast<cudaStream_t>(stream));
}
int vitb200_grad_pack(float* g, void* out_bf16, int64_t n, void* stream) {
  return grad_pack(g, static_cast<bf16*>(out_bf16), n, static_cast<cudaStream_t>(stream));
}
int vitb200_adamw_range(float* p, float* g, const void* g_bf16, float* m, float* v, void* shadow_bf16, int64_t n,
                        int zero_grad, vitb200_adamw_hyper h, void* stream) {
  VB_REQUIRE(h.step >= 1, "adamw: step must be >= 1");
  return adamw_range(p, g, static_cast<const bf16*>(g_bf16), m, v, static_cast<bf16*>(shadow_bf16), n, zero_grad != 0,
                     to_hyper(h), static_cast<cudaStream_t>(stream));
}
int vitb200_adamw_range_dev(float* p, float* g, const void* g_bf16, float* m, float* v, void* shadow_bf16, int64_t n,
                            int zero_grad, vitb200_adamw_hyper h, const void* step_state, void* stream) {
  VB_REQUIRE(step_state != nullptr, "adamw_range_dev: null step state");
  AdamWHyper k = to_hyper(vitb200_adamw_hyper{h.lr, h.beta1, h.beta2, h.eps, h.weight_decay, 1, h.grad_scale});
  k.corr_dev = reinterpret_cast<const float*>(static_cast<const uint8_t*>(step_state) + 12);   // {bias_corr1, inv_sqrt_bc2}
  return adamw_range(p, g, static_cast<const bf16*>(g_bf16), m, v, static_cast<bf16*>(shadow_bf16), n, zero_grad != 0, k,
                     static_cast<cudaStream_t>(stream));
}
int vitb200_step_state_advance(void* step_state, float beta1, float beta2, uint64_t seed_stride, void* stream) {
  VB_REQUIRE(step_state != nullptr, "step_state_advance: null step state");
  return step_state_advance(step_state, beta1, beta2, seed_stride, static_cast<cudaStream_t>(stream));
}
int vitb200_adamw_multi(const uint64_t* p_ptrs, const uint64_t* g_ptrs, const uint64_t* m_ptrs,
                        const uint64_t* v_ptrs, const uint64_t* shadow_ptrs, const int64_t* sizes,
                        const int32_t* blk_tensor, const int32_t* blk_chunk, int n_blocks, vitb200_adamw_hyper h,
                        void* stream) {
  VB_REQUIRE(h.step >= 1, "adamw: step must be >= 1");
  return adamw_multi(p_ptrs, g_ptrs, m_ptrs, v_ptrs, shadow_ptrs, sizes, blk_tensor, blk_chunk, n_blocks, to_hyper(h),
                     static_cast<cudaStream_t>(stream));
}

}  // extern "C"
