// C-ABI plumbing: error string, device check, GEMM dispatch (see include/vitb200.h).
#include <cstdarg>
#include <cstring>

#include "../../include/vitb200.h"
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

}  // namespace vb

extern "C" {

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

}  // extern "C"
