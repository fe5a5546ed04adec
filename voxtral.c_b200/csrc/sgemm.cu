// sgemm.cu - fp32-activation x bf16-weight GEMM on the CUDA cores.
//
//   C[M,N] = epi( A[M,K] (fp32, row stride lda) x W[N,K]^T (bf16) + bias )
//
// This is the "exact" matmul of the backend: bf16 -> fp32 widening is exact and
// accumulation is fp32, i.e. the arithmetic of the reference's M>1 path
// (bf16_to_f32_buf + cblas_sgemm, /root/reference/voxtral_kernels.c:197-264)
// up to summation order.  It serves decoder prefill (M=38), the precise
// encoder mode, the adapter, the conv stem (im2col-free: overlapping A rows via
// lda) and the vox_cuda_sgemm_bf16 op hook.  The throughput path of the
// encoder is the tcgen05 kernel in gemm_tc.cu.
#include "vox_internal.h"

constexpr int SG_BM = 64, SG_BN = 64, SG_BK = 32, SG_T = 256;

__global__ void __launch_bounds__(SG_T) sgemm_bf16w_kernel(
    int M, int N, int K, const float *__restrict__ A, long lda,
    const uint16_t *__restrict__ W, float *__restrict__ C, long ldc,
    const float *__restrict__ bias, int epi) {
    __shared__ float As[2][SG_BK][SG_BM + 4];
    __shared__ float Ws[2][SG_BK][SG_BN + 4];
    const int t = threadIdx.x;
    const int m0 = blockIdx.y * SG_BM, n0 = blockIdx.x * SG_BN;
    // loader mapping: tile 64 rows x 32 k -> thread loads row (t%64), k-chunk (t/64)*8 .. +8
    // (a warp writes 32 consecutive rows of one k -> conflict-free transposed stores)
    const int lr = t & 63, lk = (t >> 6) * 8;
    const bool a_ok = (m0 + lr) < M;
    const float *ap = A + (size_t)(m0 + lr) * lda + lk;
    const uint16_t *wp = W + (size_t)(n0 + lr) * K + lk;
    // compute mapping: 16x16 threads, each 4x4 outputs
    const int ty = t >> 4, tx = t & 15;
    float acc[4][4];
    #pragma unroll
    for (int i = 0; i < 4; i++)
        #pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = 0.f;

    float4 ra0, ra1; uint4 rw;
    auto gload = [&](int k0) {
        if (a_ok) {
            ra0 = *reinterpret_cast<const float4 *>(ap + k0);
            ra1 = *reinterpret_cast<const float4 *>(ap + k0 + 4);
        } else { ra0 = make_float4(0, 0, 0, 0); ra1 = ra0; }
        rw = *reinterpret_cast<const uint4 *>(wp + k0);
    };
    auto sstore = [&](int b) {
        As[b][lk + 0][lr] = ra0.x; As[b][lk + 1][lr] = ra0.y; As[b][lk + 2][lr] = ra0.z; As[b][lk + 3][lr] = ra0.w;
        As[b][lk + 4][lr] = ra1.x; As[b][lk + 5][lr] = ra1.y; As[b][lk + 6][lr] = ra1.z; As[b][lk + 7][lr] = ra1.w;
        Ws[b][lk + 0][lr] = bf16_lo(rw.x); Ws[b][lk + 1][lr] = bf16_hi(rw.x);
        Ws[b][lk + 2][lr] = bf16_lo(rw.y); Ws[b][lk + 3][lr] = bf16_hi(rw.y);
        Ws[b][lk + 4][lr] = bf16_lo(rw.z); Ws[b][lk + 5][lr] = bf16_hi(rw.z);
        Ws[b][lk + 6][lr] = bf16_lo(rw.w); Ws[b][lk + 7][lr] = bf16_hi(rw.w);
    };
    gload(0);
    sstore(0);
    __syncthreads();
    const int nk = K / SG_BK;
    for (int kb = 0; kb < nk; kb++) {
        const int b = kb & 1;
        if (kb + 1 < nk) gload((kb + 1) * SG_BK);
        #pragma unroll
        for (int k = 0; k < SG_BK; k++) {
            const float4 a = *reinterpret_cast<const float4 *>(&As[b][k][ty * 4]);
            const float4 w = *reinterpret_cast<const float4 *>(&Ws[b][k][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, wv[4] = {w.x, w.y, w.z, w.w};
            #pragma unroll
            for (int i = 0; i < 4; i++)
                #pragma unroll
                for (int j = 0; j < 4; j++) acc[i][j] = fmaf(av[i], wv[j], acc[i][j]);
        }
        if (kb + 1 < nk) sstore(b ^ 1);
        __syncthreads();
    }
    #pragma unroll
    for (int i = 0; i < 4; i++) {
        const int m = m0 + ty * 4 + i;
        if (m >= M) continue;
        #pragma unroll
        for (int j = 0; j < 4; j++) {
            const int n = n0 + tx * 4 + j;
            float v = acc[i][j];
            if (epi & EPI_BIAS) v += bias[n];
            if (epi & EPI_GELU) v = gelu_tanh(v);
            float *c = C + (size_t)m * ldc + n;
            if (epi & EPI_RESIDUAL) v += *c;
            *c = v;
        }
    }
}

int sgemm_bf16w(cudaStream_t st, int M, int N, int K, const float *A, long lda,
                const uint16_t *W, float *C, long ldc, const float *bias, int epi) {
    if (M <= 0) return 0;
    if (N % SG_BN || K % SG_BK) {
        vox_set_error("sgemm_bf16w: N=%d must be a multiple of %d and K=%d of %d", N, SG_BN, K, SG_BK);
        return -1;
    }
    dim3 grid(N / SG_BN, (M + SG_BM - 1) / SG_BM);
    sgemm_bf16w_kernel<<<grid, SG_T, 0, st>>>(M, N, K, A, lda, W, C, ldc, bias, epi);
    LAUNCHED();
    CUDA_TRY(cudaGetLastError());
    return 0;
}
