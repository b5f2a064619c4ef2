// C-ABI plumbing: error string, device check, GEMM dispatch (see include/vitb200.h).
#include <atomic>
#include <cstdarg>
#include <cstring>
#include <mutex>
#include <vector>

#include "../../include/vitb200.h"
#include <cstdlib>
#include "common.cuh"
#include "kernels.h"

namespace vb {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* last_error() { return g_err; }

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

// ---- launch accounting / live profiling --------------------------------------------------------------
namespace {
constexpr int kProfMaxScopes = 16384;
std::atomic<long long> g_launches{0};
std::atomic<bool> g_prof_on{false};
std::mutex g_prof_mu;
struct ProfRec { cudaEvent_t e0, e1; int cls; double work; int launches; };
std::vector<ProfRec> g_prof_recs;      // scopes recorded since prof_begin
std::vector<cudaEvent_t> g_prof_pool;  // reusable events
cudaEvent_t prof_take_event() {
  if (!g_prof_pool.empty()) { cudaEvent_t e = g_prof_pool.back(); g_prof_pool.pop_back(); return e; }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}
}  // namespace

ProfScope::ProfScope(int cls, double work, int launches, cudaStream_t s) : slot(-1), stream(s) {
  g_launches.fetch_add(launches, std::memory_order_relaxed);
  if (!g_prof_on.load(std::memory_order_relaxed)) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (static_cast<int>(g_prof_recs.size()) >= kProfMaxScopes) return;
  ProfRec r;
  r.e0 = prof_take_event(); r.e1 = prof_take_event();
  r.cls = cls; r.work = work; r.launches = launches;
  if (r.e0 == nullptr || r.e1 == nullptr) return;
  cudaEventRecord(r.e0, s);
  slot = static_cast<int>(g_prof_recs.size());
  g_prof_recs.push_back(r);
}
ProfScope::~ProfScope() {
  if (slot < 0) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (slot < static_cast<int>(g_prof_recs.size())) cudaEventRecord(g_prof_recs[slot].e1, stream);
}

}  // namespace vb

namespace vb {
bool pdl_enabled() {
  static const bool on = !(getenv("VITB200_PDL") && atoi(getenv("VITB200_PDL")) == 0);
  return on;
}
}  // namespace vb

extern "C" {

int64_t vitb200_launch_count(void) { return vb::g_launches.load(); }

const char* vitb200_prof_class_name(int cls) {
  static const char* names[vb::kProfCount] = {"gemm_tcgen05", "gemm_f32", "attn_fwd", "attn_bwd", "ln_fwd", "ln_bwd",
                                              "elementwise", "adamw"};
  return cls >= 0 && cls < vb::kProfCount ? names[cls] : "?";
}

int vitb200_prof_begin(void) {
  std::lock_guard<std::mutex> lk(vb::g_prof_mu);
  for (auto& r : vb::g_prof_recs) { vb::g_prof_pool.push_back(r.e0); vb::g_prof_pool.push_back(r.e1); }
  vb::g_prof_recs.clear();
  vb::g_prof_on.store(true);
  return 0;
}

int vitb200_prof_end(vitb200_prof_entry* out) {
  using namespace vb;
  g_prof_on.store(false);
  VB_CHECK_CUDA(cudaDeviceSynchronize());
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (int c = 0; c < kProfCount; ++c) { out[c].ms = 0; out[c].work = 0; out[c].launches = 0; }
  for (auto& r : g_prof_recs) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r.e0, r.e1) == cudaSuccess) {
      out[r.cls].ms += ms; out[r.cls].work += r.work; out[r.cls].launches += r.launches;
    }
    g_prof_pool.push_back(r.e0); g_prof_pool.push_back(r.e1);
  }
  g_prof_recs.clear();
  (void)cudaGetLastError();
  return kOk;
}

const char* vitb200_last_error(void) { return vb::last_error(); }
int vitb200_version(void) { return 100; }

int vitb200_device_ok(void) {
  int dev = 0, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return 0;
  return major == 10 ? 1 : 0;
}

int vitb200_gemm(const vitb200_gemm_args* a, void* stream) {
  using namespace vb;
  VB_REQUIRE(a != nullptr, "vitb200_gemm: null args");
  VB_REQUIRE(a->A && a->B && a->C, "vitb200_gemm: null operand");
  GemmEpilogue ep;
  ep.bias = a->bias;
  ep.mask_aux = a->mask_aux;
  ep.row_add = a->row_add;
  ep.residual = a->residual;
  ep.C = a->C;
  ep.ldc = a->ldc;
  ep.relu = a->relu;
  ep.row_period = a->row_period;
  ep.row_group_stride = a->row_group_stride;
  ep.accumulate = a->accumulate;
  ep.col_sum = a->col_sum;
  ep.bits_out = a->bits_out;
  ep.mask_bits = a->mask_bits;
  ep.bits_rows = a->bits_rows;
  ep.dot_aux = a->dot_aux;
  ep.dot_out = a->dot_out;
  ep.dot_T = a->dot_T;
  ep.dot_H = a->dot_H;
  VB_REQUIRE(a->dot_out == nullptr || a->in_dtype == VITB200_BF16, "vitb200_gemm: the fused head dot needs BF16 inputs");
  ep.drop = make_dropout(a->dropout.p, a->dropout.seed, a->dropout.stream);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (a->in_dtype == VITB200_BF16) {
    return gemm_tc_bf16(a->M, a->N, a->K, a->a_major, a->b_major, a->A, a->lda, a->B, a->ldb,
                        a->out_dtype == VITB200_F32, a->splits, ep, s);
  }
  VB_REQUIRE(a->in_dtype == VITB200_F32 && a->out_dtype == VITB200_F32,
             "vitb200_gemm: fp32 inputs require an fp32 output");
  return gemm_f32(a->M, a->N, a->K, a->a_major, a->b_major, static_cast<const float*>(a->A), a->lda,
                  static_cast<const float*>(a->B), a->ldb, a->splits, ep, s);
}

int vitb200_gemm_auto_splits(int M, int N, int K, int sms) {
  using namespace vb;
  if (M <= 0 || N <= 0 || K <= 0 || sms <= 0) return -1;
  const int cg = M > 128 ? 2 : 1;                       // as gemm_tc_bf16: CTA pairs own 256-row tiles
  const int tiles = ((M + 128 * cg - 1) / (128 * cg)) * ((N + 255) / 256);
  const int walkers = sms / cg > 0 ? sms / cg : 1;
  return gemm_tc_auto_splits(tiles, (K + 63) / 64, walkers);
}

int64_t vitb200_abi_sizeof(int which) {
  switch (which) {
    case 0: return sizeof(vitb200_dims);
    case 1: return sizeof(vitb200_dropout);
    case 2: return sizeof(vitb200_gemm_args);
    case 3: return sizeof(vitb200_adamw_hyper);
    case 4: return sizeof(vitb200_block_params);
    case 5: return sizeof(vitb200_block_grads);
    case 6: return sizeof(vitb200_head_params);
    case 7: return sizeof(vitb200_prof_entry);
    default: return -1;
  }
}

}  // extern "C"
