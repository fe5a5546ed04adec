/*
 * oracle/sgemm_fallback.c - plain-C cblas_sgemm used for the reference build
 * only when no OpenBLAS shared object can be found (test infrastructure).
 * Row-major only, alpha/beta general, the three transpose combinations the
 * reference uses (/root/reference/voxtral_kernels.c:56,73,91,451).
 */
#include "cblas.h"
#include <stddef.h>

#ifndef VOX_REF_SCIPY_OPENBLAS
void cblas_sgemm(enum CBLAS_ORDER order, enum CBLAS_TRANSPOSE ta,
                 enum CBLAS_TRANSPOSE tb, int M, int N, int K, float alpha,
                 const float *A, int lda, const float *B, int ldb, float beta,
                 float *C, int ldc) {
    (void)order;
    #pragma omp parallel for schedule(static)
    for (int m = 0; m < M; m++) {
        float *c = C + (size_t)m * ldc;
        for (int n = 0; n < N; n++) c[n] = (beta == 0.0f) ? 0.0f : beta * c[n];
        if (tb == CblasNoTrans) {
            for (int k = 0; k < K; k++) {
                float a = alpha * (ta == CblasNoTrans ? A[(size_t)m * lda + k]
                                                      : A[(size_t)k * lda + m]);
                const float *b = B + (size_t)k * ldb;
                for (int n = 0; n < N; n++) c[n] += a * b[n];
            }
        } else {
            for (int n = 0; n < N; n++) {
                const float *b = B + (size_t)n * ldb;
                float s = 0.0f;
                if (ta == CblasNoTrans) {
                    const float *a = A + (size_t)m * lda;
                    for (int k = 0; k < K; k++) s += a[k] * b[k];
                } else {
                    for (int k = 0; k < K; k++) s += A[(size_t)k * lda + m] * b[k];
                }
                c[n] += alpha * s;
            }
        }
    }
}
#endif
