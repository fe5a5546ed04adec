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
    // Optional tiled operands (both or neither).  w_tiled: wt_tile_weights() image of W - tile (n/128, k/64) is a
    // contiguous 16 KB block [128 rows][128 B] with the 16-byte chunk index XORed by (row & 7) (the UMMA
    // SWIZZLE_128B K-major image), so one 1-D bulk copy replaces 128 row requests.  act_tiled: per k-block a
    // contiguous [hi: NB rows x 128 B | lo: NB rows x 128 B] block in the same swizzle (wt_act_offset()).
    const uint16_t *w_tiled, *act_tiled;
    // Q8 weights: W holds the int8 values as bf16, the product (each K-range partial) is multiplied by row_scale[n];
    // null = plain bf16 weights
    const float *row_scale;
};

// byte offset of element (row b, column c) of the hi half inside a tiled activation buffer; lo = + NB * 128
static inline size_t wt_act_offset(int NB, int b, int c) {
    return (size_t)(c >> 6) * (2 * NB * 128) + (size_t)b * 128 + ((((c & 63) >> 3) ^ (b & 7)) << 4) + (c & 7) * 2;
}
// device repack W [N,K] row-major bf16 -> tiled image (N % 128 == 0, K % 64 == 0); dst holds N*K elements
int wt_tile_weights(cudaStream_t st, const uint16_t *W, uint16_t *dst, int N, int K);

// act: device bf16 [NB rows, K] (split == 0) or [NB rows, hi(K) | lo(K)] (split == 1), row stride lda elements;
// rows >= B must be allocated (their products are discarded).  NB in {32, 64, 128}, N % 128 == 0, K % 64 == 0.
int gemm_wt_prepare(WtOp *op, int NB, int B, int N, int K, int split, const uint16_t *act, long lda,
                    const uint16_t *W, int ksplit);
int gemm_wt_launch(cudaStream_t st, const WtOp *op);
