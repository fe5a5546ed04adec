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
};

struct TcEpilogue {
    float *out_f32;
    long ldo_f32;
    uint16_t *out_bf16;
    long ldo_bf16;
    const float *bias;
    int flags;
};

// A: device bf16 [M rows, K] with row stride lda (elements; rows may overlap: lda < K is allowed,
// lda*2 must be a multiple of 16 bytes).  W: device bf16 [N, K] row-major.
// N % 128 == 0, K % 64 == 0.  Returns 0 on success.
int gemm_tc(cudaStream_t st, int M, int N, int K, const uint16_t *A, long lda, const uint16_t *W,
            const TcEpilogue &epi);
