// rows.h - host launchers of the row-wise fp32 kernels (rows.cu)
#pragma once
#include <cuda_runtime.h>

int rows_rmsnorm(cudaStream_t st, const float *x, long ldx, const float *w, const float *ada,
                 float *out, long ldo, int rows, int dim);
int rows_rope_kv(cudaStream_t st, float *qkv, int rows, int q_dim, int kv_dim, int head_dim,
                 const float *inv_freq, int pos0, int ring_mask, float *kc, float *vc);
// pos_dev (optional): device int that overrides pos0 (CUDA-graph replays)
int rows_rope_kv_dev(cudaStream_t st, float *qkv, int rows, int q_dim, int kv_dim, int head_dim,
                     const float *inv_freq, int pos0, int ring_mask, float *kc, float *vc, const int *pos_dev);
int rows_attention(cudaStream_t st, int head_dim, const float *q, long ldq, const float *kc,
                   const float *vc, int kv_dim, int ring_mask, int n_heads, int heads_per_kv,
                   int rows, int q_pos0, int first_valid, int window, float *out, long ldo);
int rows_attention_split128(cudaStream_t st, const float *q, long ldq, const float *kc, const float *vc, int kv_dim,
                            int ring_mask, int n_heads, int heads_per_kv, int rows, int q_pos0, int first_valid, int window,
                            uint16_t *out_b, long ldb, long lo_off);
int rows_swiglu(cudaStream_t st, const float *gu, float *h, int rows, int hidden);
int rows_rmsnorm_bf16(cudaStream_t st, const float *x, long ldx, const float *w, const float *ada, uint16_t *out,
                      long ldo, int rows, int dim, int split);
