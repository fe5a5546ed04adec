// gemm_wt.h - host interface of the skinny multi-stream GEMM (gemm_wt.cu)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

enum { WT_OUT_F32 = 0, WT_OUT_SWIGLU_BF16X2 = 1 };

// One prepared product out[b][n] = sum_k act[b][k] W[n][k] (tensor maps built once, launched every step).
struct WtOp {
    alignas(64) unsigned char map_w[128];   // CUtensorMap of W   [N, K]
    alignas(64) unsigned char map_a[128];   // CUtensorMap of act [NB, K or 2K]
    int NB, B, N, K, split, ksplit, mode;
    // WT_OUT_F32: out_f32[s * split_stride + b * ldo + n] for K-range s (the consumer adds the ranges in order)
    float *out_f32; long ldo; long split_stride;
    // WT_OUT_SWIGLU_BF16X2 (ksplit == 1, W rows (2i, 2i+1) = (gate, up)):
    // h = silu(g) * u as split bf16: out_bf16[b * ldb + i] = hi, out_bf16[b * ldb + lo_off + i] = lo
    uint16_t *out_bf16; long ldb; long lo_off;
};

// act: device bf16 [NB rows, K] (split == 0) or [NB rows, hi(K) | lo(K)] (split == 1), row stride lda elements;
// rows >= B must be allocated (their products are discarded).  NB in {32, 64, 128}, N % 128 == 0, K % 64 == 0.
int gemm_wt_prepare(WtOp *op, int NB, int B, int N, int K, int split, const uint16_t *act, long lda,
                    const uint16_t *W, int ksplit);
int gemm_wt_launch(cudaStream_t st, const WtOp *op);
