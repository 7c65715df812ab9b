// C-ABI of the vega_b200 CUDA library (see include/vega_b200.h).
#include <cstdarg>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/vega_b200.h"
#include "vg_common.cuh"
#include "vg_gemm.cuh"

static thread_local char g_error[1024] = "";

void vg_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

extern "C" {

const char* vg_last_error(void) { return g_error; }

int vg_version(void) { return 100; }

int vg_dgemm_tn(const double* A, const double* B, double* C, int M, int N, int K,
                long long lda, long long ldb, long long ldc, double alpha, double beta,
                int splits, double* workspace, void* stream) {
  vg::GemmArgs g{};
  g.A = A; g.B = B; g.M = M; g.N = N; g.K = K;
  g.lda = lda; g.ldb = ldb; g.ldc = ldc;
  g.batch = 1; g.splits = splits < 1 ? 1 : splits;
  g.alpha = alpha; g.beta = beta;
  cudaStream_t st = (cudaStream_t)stream;
  if (g.splits == 1) {
    g.C = C;
    return vg::launch_dgemm(g, vg::GEMM_STORE, st);
  }
  if (!workspace || beta != 0.0) { vg_set_error("split-K needs a workspace and beta == 0"); return VG_ERR_ARG; }
  g.C = workspace; g.split_stride = (long long)M * ldc;
  int rc = vg::launch_dgemm(g, vg::GEMM_STORE, st);
  if (rc) return rc;
  long long n = (long long)M * ldc;
  vg::reduce_splits_kernel<<<vg_cdiv(n, 256), 256, 0, st>>>(workspace, C, n, g.splits, g.split_stride);
  VG_CUDA_CHECK(cudaGetLastError());
  return VG_OK;
}

}  // extern "C"
