// Library-level entry points: version, per-thread error string, device check, and the
// acai_linear_fwd dispatcher (tcgen05 tensor-core GEMM when the problem qualifies, SIMT otherwise).
#include <stdarg.h>

#include <cstdlib>

#include "common.cuh"

namespace acai {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

bool pdl_enabled() {
  static const bool on = [] { const char* e = getenv("ACAI_PDL"); return !(e && atoi(e) == 0); }();
  return on;
}

int num_sms() {
  static int cached = 0;
  if (cached == 0) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
      cached = n;
    else
      return 148;
  }
  return cached;
}

int linear_simt(const void* a, int a_dt, int64_t lda, const void* w, int w_dt, int64_t ldw, const float* bias,
                const void* res, int res_dt, int64_t ldr, void* out, int out_dt, int64_t ldo, int M, int N, int K,
                int act, cudaStream_t st);
// returns ACAI_ERR_UNSUPPORTED when the problem does not qualify for the tensor-core path
int linear_gemv_bf16(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias, const void* res, int res_dt,
                     int64_t ldr, void* out, int out_dt, int64_t ldo, int M, int N, int K, int act, cudaStream_t st);
int linear_tcgen05(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias, const void* res,
                   int res_dt, int64_t ldr, void* out, int out_dt, int64_t ldo, int M, int N, int K, int act,
                   cudaStream_t st);
size_t linear_splitk_workspace_bytes(int M, int N, int K);
int linear_splitk_tcgen05(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias, const void* res, int res_dt,
                          int64_t ldr, const float* gamma, const float* beta, float eps, void* out, int out_dt, int64_t ldo,
                          void* out2, int out2_dt, int64_t ldo2, int M,
                          int N, int K, int act, void* ws, size_t ws_bytes, cudaStream_t st);
int linear_dropout_tcgen05(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias, const void* res, int res_dt,
                           int64_t ldr, void* out, int64_t ldo, int M, int N, int K, float dropout_p, uint64_t seed, cudaStream_t st);
int linear_gelu_train_tcgen05(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias, void* z, int64_t ldz,
                              void* h, int64_t ldh, int M, int N, int K, float dropout_p, uint64_t seed, cudaStream_t st);
int linear_dgrad_tcgen05(const void* dz, int64_t ldz, const void* w, int64_t ldw, const void* res, int64_t ldr, const void* z,
                         int64_t ldzz, void* out, int64_t ldo, int M, int N, int K, float dropout_p, uint64_t seed, cudaStream_t st);
int linear_gelu_bwd_tcgen05(const void* dy, int64_t ldd, const void* wt, int64_t ldw, const void* z, int64_t ldz, void* dz,
                            int64_t ldo, int M, int N, int K, float dropout_p, uint64_t seed, cudaStream_t st);

}  // namespace acai

using namespace acai;

extern "C" int acai_version(void) { return 100; }
extern "C" const char* acai_last_error(void) { return g_err; }

extern "C" int acai_device_ok(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { set_error("no CUDA device"); cudaGetLastError(); return 0; }
  int dev = 0, major = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  if (major != 10) { set_error("device compute capability %d.x is not sm_100", major); return 0; }
  return 1;
}

extern "C" int acai_linear_fwd(const void* a, int32_t a_dtype, int64_t lda, const void* w, int32_t w_dtype,
                               int64_t ldw, const float* bias, const void* residual, int32_t res_dtype, int64_t ldr,
                               void* out, int32_t out_dtype, int64_t ldo, int32_t M, int32_t N, int32_t K,
                               int32_t act, void* stream) {
  ACAI_CHECK_ARG(a && w && out, "acai_linear_fwd: null pointer");
  ACAI_CHECK_ARG(M >= 0 && N > 0 && K > 0 && lda >= K && ldw >= K && ldo >= N, "acai_linear_fwd: bad sizes M=%d N=%d K=%d", M, N, K);
  ACAI_CHECK_ARG(dtype_valid(a_dtype) && dtype_valid(w_dtype) && dtype_valid(out_dtype) &&
                 (!residual || (dtype_valid(res_dtype) && ldr >= N)), "acai_linear_fwd: bad dtype/ldr");
  ACAI_CHECK_ARG(act == ACAI_ACT_NONE || act == ACAI_ACT_GELU, "acai_linear_fwd: bad activation %d", act);
  if (M == 0) return ACAI_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (a_dtype == ACAI_BF16 && w_dtype == ACAI_BF16) {
    // decode steps at very small batch: stream the weights once through all SMs instead of one 128-row tile column
    // (measured cross-over against the tcgen05 path: M = 4; ACAI_GEMV_MAX_M overrides it for benchmarking)
    static const int gemv_max_m = [] { const char* e = getenv("ACAI_GEMV_MAX_M"); return e ? atoi(e) : 4; }();
    if (M <= gemv_max_m) {
      int rc = linear_gemv_bf16(a, lda, w, ldw, bias, residual, res_dtype, ldr, out, out_dtype, ldo, M, N, K, act, st);
      if (rc != ACAI_ERR_UNSUPPORTED) return rc;
    }
    int rc = linear_tcgen05(a, lda, w, ldw, bias, residual, res_dtype, ldr, out, out_dtype, ldo, M, N, K, act, st);
    if (rc != ACAI_ERR_UNSUPPORTED) return rc;
  }
  return linear_simt(a, a_dtype, lda, w, w_dtype, ldw, bias, residual, res_dtype, ldr, out, out_dtype, ldo, M, N, K, act, st);
}

// Small-M (decode step) projection with everything that follows it in the layer: split-K tcgen05 GEMM + one reduce kernel
// that adds bias / residual, applies the activation and, with ln_gamma, the LayerNorm.
extern "C" size_t acai_linear_splitk_workspace_bytes(int32_t M, int32_t N, int32_t K) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  return acai::linear_splitk_workspace_bytes(M, N, K);
}

extern "C" int acai_linear_splitk(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias, const void* residual,
                                  int32_t res_dtype, int64_t ldr, const float* ln_gamma, const float* ln_beta, float ln_eps,
                                  void* out, int32_t out_dtype, int64_t ldo, void* out2, int32_t out2_dtype, int64_t ldo2,
                                  int32_t M, int32_t N, int32_t K, int32_t act,
                                  void* workspace, size_t workspace_bytes, void* stream) {
  ACAI_CHECK_ARG(a && w && out && workspace, "acai_linear_splitk: null pointer");
  ACAI_CHECK_ARG(M >= 0 && N > 0 && K > 0 && lda >= K && ldw >= K && ldo >= N, "acai_linear_splitk: bad sizes M=%d N=%d K=%d", M, N, K);
  ACAI_CHECK_ARG(dtype_valid(out_dtype) && (!residual || (dtype_valid(res_dtype) && ldr >= N)), "acai_linear_splitk: bad dtype/ldr");
  ACAI_CHECK_ARG(act == ACAI_ACT_NONE || act == ACAI_ACT_GELU, "acai_linear_splitk: bad activation %d", act);
  ACAI_CHECK_ARG(!ln_gamma == !ln_beta, "acai_linear_splitk: ln_gamma and ln_beta come together");
  ACAI_CHECK_ARG(!out2 || (dtype_valid(out2_dtype) && ldo2 >= N), "acai_linear_splitk: bad out2 dtype / ldo2");
  if (M == 0) return ACAI_OK;
  return acai::linear_splitk_tcgen05(a, lda, w, ldw, bias, residual, res_dtype, ldr, ln_gamma, ln_beta, ln_eps, out, out_dtype, ldo,
                                     out2, out2_dtype, ldo2, M, N, K, act, workspace, workspace_bytes, (cudaStream_t)stream);
}

// out = residual + dropout(a @ w^T + bias): the sub-layer output projection of a post-norm block with its nn.Dropout and
// residual add in the GEMM epilogue (bf16 operands and output, dense [M, N] output).  ACAI_ERR_UNSUPPORTED, nothing written,
// when the problem does not qualify - the caller then runs acai_linear_fwd + acai_dropout (same mask).
extern "C" int acai_linear_dropout(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias, const void* residual,
                                   int32_t res_dtype, int64_t ldr, void* out, int64_t ldo, int32_t M, int32_t N, int32_t K,
                                   float dropout_p, uint64_t dropout_seed, void* stream) {
  ACAI_CHECK_ARG(a && w && out, "acai_linear_dropout: null pointer");
  ACAI_CHECK_ARG(M >= 0 && N > 0 && K > 0 && lda >= K && ldw >= K && ldo >= N && (!residual || ldr >= N), "acai_linear_dropout: bad sizes");
  ACAI_CHECK_ARG(dropout_p > 0.0f && dropout_p < 1.0f, "acai_linear_dropout: dropout_p must be in (0, 1)");
  ACAI_CHECK_ARG(!residual || res_dtype == ACAI_BF16 || res_dtype == ACAI_F32, "acai_linear_dropout: bad residual dtype");
  if (M == 0) return ACAI_OK;
  return acai::linear_dropout_tcgen05(a, lda, w, ldw, bias, residual, res_dtype, ldr, out, ldo, M, N, K, dropout_p, dropout_seed,
                                      (cudaStream_t)stream);
}

// Feed-forward up-projection for training: z = a @ w^T + bias and h = dropout(gelu(z)), both bf16, one GEMM.
extern "C" int acai_linear_gelu_train(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias, void* z,
                                      int64_t ldz, void* h, int64_t ldh, int32_t M, int32_t N, int32_t K, float dropout_p,
                                      uint64_t dropout_seed, void* stream) {
  ACAI_CHECK_ARG(a && w && z && h, "acai_linear_gelu_train: null pointer");
  ACAI_CHECK_ARG(M >= 0 && N > 0 && K > 0 && lda >= K && ldw >= K && ldz >= N && ldh >= N, "acai_linear_gelu_train: bad sizes");
  ACAI_CHECK_ARG(dropout_p >= 0.0f && dropout_p < 1.0f, "acai_linear_gelu_train: dropout_p must be in [0, 1)");
  if (M == 0) return ACAI_OK;
  return acai::linear_gelu_train_tcgen05(a, lda, w, ldw, bias, z, ldz, h, ldh, M, N, K, dropout_p, dropout_seed, (cudaStream_t)stream);
}

// Its backward through the activation: dz = (dy @ wt^T) * gelu'(z) * mask, bf16, one GEMM (wt = the [N, K] transpose
// of the down-projection weight).
extern "C" int acai_linear_dgrad(const void* dz, int64_t ldz, const void* w, int64_t ldw, const void* residual, int64_t ldr,
                                const void* z, int64_t ldzz, void* dx, int64_t ldo, int32_t M, int32_t N, int32_t K, float dropout_p,
                                uint64_t dropout_seed, void* stream) {
  ACAI_CHECK_ARG(dz && w && dx, "acai_linear_dgrad: null pointer");
  ACAI_CHECK_ARG(M >= 0 && N > 0 && K > 0 && ldz >= K && ldw >= N && ldo >= N && (!residual || ldr >= N) && (!z || ldzz >= N),
                 "acai_linear_dgrad: bad sizes M=%d N=%d K=%d", M, N, K);
  ACAI_CHECK_ARG(!(z && residual), "acai_linear_dgrad: z (GELU backward) and residual are exclusive");
  ACAI_CHECK_ARG(dropout_p >= 0.0f && dropout_p < 1.0f, "acai_linear_dgrad: dropout_p must be in [0, 1)");
  ACAI_CHECK_ARG(!z || ldzz == N || dropout_p == 0.0f, "acai_linear_dgrad: the dropout mask is indexed over a dense [M, N] tensor");
  if (M == 0) return ACAI_OK;
  int rc = acai::linear_dgrad_tcgen05(dz, ldz, w, ldw, residual, ldr, z, ldzz, dx, ldo, M, N, K, dropout_p, dropout_seed, (cudaStream_t)stream);
  if (rc == ACAI_ERR_UNSUPPORTED) acai::set_error("acai_linear_dgrad: operands are not TMA-addressable (bf16, 16-byte aligned rows, N >= 128, K %% 64 == 0)");
  return rc;
}

extern "C" int acai_linear_gelu_bwd(const void* dy, int64_t ldd, const void* wt, int64_t ldw, const void* z, int64_t ldz, void* dz,
                                    int64_t ldo, int32_t M, int32_t N, int32_t K, float dropout_p, uint64_t dropout_seed,
                                    void* stream) {
  ACAI_CHECK_ARG(dy && wt && z && dz, "acai_linear_gelu_bwd: null pointer");
  ACAI_CHECK_ARG(M >= 0 && N > 0 && K > 0 && ldd >= K && ldw >= K && ldz >= N && ldo >= N, "acai_linear_gelu_bwd: bad sizes");
  ACAI_CHECK_ARG(dropout_p >= 0.0f && dropout_p < 1.0f, "acai_linear_gelu_bwd: dropout_p must be in [0, 1)");
  ACAI_CHECK_ARG(ldz == N || dropout_p == 0.0f, "acai_linear_gelu_bwd: the dropout mask is indexed over a dense [M, N] tensor");
  if (M == 0) return ACAI_OK;
  return acai::linear_gelu_bwd_tcgen05(dy, ldd, wt, ldw, z, ldz, dz, ldo, M, N, K, dropout_p, dropout_seed, (cudaStream_t)stream);
}
