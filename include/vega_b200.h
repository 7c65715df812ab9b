/* vega_b200 C ABI -- placeholder, filled as the library grows */
#ifndef VEGA_B200_H
#define VEGA_B200_H
#ifdef __cplusplus
extern "C" {
#endif
const char* vg_last_error(void);
int vg_version(void);
int vg_dgemm_tn(const double* A, const double* B, double* C, int M, int N, int K,
                long long lda, long long ldb, long long ldc, double alpha, double beta,
                int splits, double* workspace, void* stream);
#ifdef __cplusplus
}
#endif
#endif
