/* Minimal <cblas.h> for the -DUSE_CUDA build of the reference's host sources: the one BLAS symbol they
 * reference (voxtral_kernels.c:56,73,91,451).  It resolves to the abort() stub in vox_cuda_dropin.c. */
#ifndef VOX_DROPIN_CBLAS_H
#define VOX_DROPIN_CBLAS_H
enum { CblasRowMajor = 101, CblasColMajor = 102 };
enum { CblasNoTrans = 111, CblasTrans = 112 };
void cblas_sgemm(int order, int ta, int tb, int M, int N, int K, float alpha, const float *A, int lda,
                 const float *B, int ldb, float beta, float *C, int ldc);
#endif
