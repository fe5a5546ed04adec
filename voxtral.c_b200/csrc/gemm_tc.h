// gemm_tc.h - host interface of the tcgen05 GEMM (gemm_tc.cu)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

enum {
    TC_BIAS = 1,        // + bias[n]
    TC_GELU = 2,        // tanh-GELU (voxtral_kernels.c:505-513)
    TC_RESIDUAL = 4,    // out_f32[m][n] += value
    TC_SWIGLU = 8,      // columns are (gate, up) pairs: out[m][n/2] = silu(g) * u
    TC_OUT_F32 = 16,    // write fp32 to out_f32 (ldo_f32)
    TC_OUT_BF16 = 32,   // write bf16 to out_bf16 (ldo_bf16)
    TC_OUT_SPLIT = 64,  // with TC_OUT_BF16: also write the bf16 rounding residual at column offset split_off
};

struct TcEpilogue {
    float *out_f32;
    long ldo_f32;
    uint16_t *out_bf16;
    long ldo_bf16;
    const float *bias;
    int flags;
    long split_off;     // TC_OUT_SPLIT: lo part goes to out_bf16[row][split_off + col]
    // Q8 weights: W holds the int8 values as bf16 and the product is multiplied by col_scale[n] (the row scale of the
    // quantised matrix, voxtral_kernels.c:277-310) before bias / activation; null = plain bf16 weights
    const float *col_scale = nullptr;
};

// A: device bf16 [M rows, K] with row stride lda (elements; rows may overlap: lda < K is allowed,
// lda*2 must be a multiple of 16 bytes).  W: device bf16 [N, K] row-major.
// N % 128 == 0, K % 64 == 0.  Returns 0 on success.
// Split activations ("bf16x2"): when KA == 2*K the A rows hold [hi(K) | lo(K)] with
// a = hi + lo to ~16 mantissa bits; the kernel walks KA and wraps the W coordinate
// modulo K, i.e. C = hi W^T + lo W^T, which restores fp32-class accuracy at twice
// the (cheap) tensor work.  KA == K is the plain single-bf16 product.
int gemm_tc(cudaStream_t st, int M, int N, int K, int KA, const uint16_t *A, long lda, const uint16_t *W,
            const TcEpilogue &epi);
// fp32 -> bf16 rows; split != 0 writes [hi | lo] rows of 2*cols
int f32_to_bf16_rows(cudaStream_t st, const float *in, long ldi, uint16_t *out, long ldo, int rows, int cols, int split);
// Same product with the split activations as two planes (A_hi, A_lo: identical shape and row stride); rows may
// overlap (lda < K) - the causal convolutions read their [3 x channels] windows this way, no im2col copy.
int gemm_tc_planes(cudaStream_t st, int M, int N, int K, const uint16_t *A_hi, const uint16_t *A_lo, long lda,
                   const uint16_t *W, const TcEpilogue &epi);
// fp32 -> two bf16 planes (hi = rn(x), lo = rn(x - hi)); n even
int f32_to_bf16_planes(cudaStream_t st, const float *in, uint16_t *hi, uint16_t *lo, size_t n);
