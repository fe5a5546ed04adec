/*
 * oracle/cblas.h - minimal CBLAS declaration used ONLY to compile the reference
 * sources (test infrastructure, never linked into the product).
 *
 * The reference calls exactly one BLAS symbol, cblas_sgemm
 * (/root/reference/voxtral_kernels.c:56,73,91,451).  No libopenblas-dev is
 * installed in this image, so oracle/build_ref.sh maps the symbol onto the
 * OpenBLAS bundled in the scipy wheel (scipy_cblas_sgemm) or, when that is
 * absent, onto oracle/sgemm_fallback.c.
 */
#ifndef VOX_ORACLE_CBLAS_H
#define VOX_ORACLE_CBLAS_H

enum CBLAS_ORDER { CblasRowMajor = 101, CblasColMajor = 102 };
enum CBLAS_TRANSPOSE { CblasNoTrans = 111, CblasTrans = 112, CblasConjTrans = 113 };

#ifdef VOX_REF_SCIPY_OPENBLAS
#define cblas_sgemm scipy_cblas_sgemm
#endif

void cblas_sgemm(enum CBLAS_ORDER order, enum CBLAS_TRANSPOSE ta,
                 enum CBLAS_TRANSPOSE tb, int M, int N, int K, float alpha,
                 const float *A, int lda, const float *B, int ldb, float beta,
                 float *C, int ldc);

#endif
