// vox_common.cuh - shared device/host helpers for the sm_100a backend.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/vox_cuda.h"

// ---- model dims (mirror /root/reference/voxtral.h:19-50) -------------------
constexpr int ENC_DIM = VOX_CUDA_ENC_DIM;          // 1280
constexpr int ENC_HEADS = VOX_CUDA_ENC_HEADS;      // 32
constexpr int ENC_HD = VOX_CUDA_ENC_HEAD_DIM;      // 64
constexpr int ENC_QKV = ENC_HEADS * ENC_HD;        // 2048
constexpr int ENC_HIDDEN = VOX_CUDA_ENC_HIDDEN;    // 5120
constexpr int ENC_WINDOW = VOX_CUDA_ENC_WINDOW;    // 750
constexpr int DEC_DIM = VOX_CUDA_DEC_DIM;          // 3072
constexpr int DEC_HEADS = VOX_CUDA_DEC_HEADS;      // 32
constexpr int DEC_KVH = VOX_CUDA_DEC_KV_HEADS;     // 8
constexpr int DEC_HD = VOX_CUDA_DEC_HEAD_DIM;      // 128
constexpr int DEC_QDIM = DEC_HEADS * DEC_HD;       // 4096
constexpr int DEC_KVDIM = DEC_KVH * DEC_HD;        // 1024
constexpr int DEC_QKV = DEC_QDIM + 2 * DEC_KVDIM;  // 6144
constexpr int DEC_HIDDEN = VOX_CUDA_DEC_HIDDEN;    // 9216
constexpr int DEC_WINDOW = VOX_CUDA_DEC_WINDOW;    // 8192
constexpr int VOCAB = VOX_CUDA_VOCAB;              // 131072
constexpr float NORM_EPS = 1e-5f;
constexpr float ROPE_THETA = 1000000.0f;

// ---- error handling ----------------------------------------------------------
void vox_set_error(const char *fmt, ...);
extern unsigned long long g_launch_count;

#define CUDA_TRY(expr)                                                         \
    do {                                                                       \
        cudaError_t _e = (expr);                                               \
        if (_e != cudaSuccess) {                                               \
            vox_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,        \
                          cudaGetErrorString(_e));                             \
            return -1;                                                         \
        }                                                                      \
    } while (0)

#define CUDA_TRY_PTR(expr)                                                     \
    do {                                                                       \
        cudaError_t _e = (expr);                                               \
        if (_e != cudaSuccess) {                                               \
            vox_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,        \
                          cudaGetErrorString(_e));                             \
            return nullptr;                                                    \
        }                                                                      \
    } while (0)

#define LAUNCHED() (++g_launch_count)

// ---- device helpers ------------------------------------------------------------
__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xFFFF0000u); }

// 16-byte streaming load that does not pollute L1 (weights are read once).
__device__ __forceinline__ uint4 ldg_stream(const void *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

__device__ __forceinline__ float warp_sum(float v) {
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// dot of 8 bf16 weights (one uint4) with 8 fp32 activations
__device__ __forceinline__ float dot8(const uint4 &w, const float4 &a, const float4 &b, float acc) {
    acc = fmaf(bf16_lo(w.x), a.x, acc); acc = fmaf(bf16_hi(w.x), a.y, acc);
    acc = fmaf(bf16_lo(w.y), a.z, acc); acc = fmaf(bf16_hi(w.y), a.w, acc);
    acc = fmaf(bf16_lo(w.z), b.x, acc); acc = fmaf(bf16_hi(w.z), b.y, acc);
    acc = fmaf(bf16_lo(w.w), b.z, acc); acc = fmaf(bf16_hi(w.w), b.w, acc);
    return acc;
}

// tanh-approximation GELU, /root/reference/voxtral_kernels.c:505-513
__device__ __forceinline__ float gelu_tanh(float v) {
    float x3 = v * v * v;
    float inner = 0.7978845608028654f * (v + 0.044715f * x3);
    return 0.5f * v * (1.0f + tanhf(inner));
}
// SiLU, /root/reference/voxtral_kernels.c:498-503
__device__ __forceinline__ float silu(float v) { return v / (1.0f + expf(-v)); }
