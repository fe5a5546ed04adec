// attn_tc.cu - sliding-window causal attention of the encoder on tensor cores.
//
// vox_causal_attention as the encoder calls it (/root/reference/voxtral_kernels.c:541-611,
// voxtral_encoder.c:622-624): 32 heads, head_dim 64, query at logical position g attends
// to keys [max(first, g - 749), g] of the KV ring.  Flash-style: a CTA owns 64 queries of
// one head (4 warps x 16 rows), walks the <= 13 key tiles of 64 keys, keeps the running
// max / sum / output in registers.  S = Q K^T and O = P V run on mma.sync.m16n8k16 bf16
// with fp32 accumulation; fp32 operands are split into two bf16 terms (x = hi + lo) and
// the three significant products hi*hi + hi*lo + lo*hi are accumulated (TERMS = 3), which
// keeps fp32-class accuracy; TERMS = 1 is plain bf16.  K/V tiles are converted from the
// fp32 ring on their way into shared memory.
//
// (The projections around it use tcgen05/TMEM, gemm_tc.cu; moving this kernel to
// tcgen05 with P kept in TMEM is the follow-up - attention is < 10 % of encoder FLOPs.)
#include "vox_internal.h"
#include "enc_state.h"
#include "rows.h"
#include "gemm_tc.h"
#include "tc_ptx.cuh"
#include <stdlib.h>

namespace at {

constexpr int BQ = 64, BKEY = 64, HD = ENC_HD, LDS = 72;    // smem row stride in bf16 (144 B: conflict-free ldmatrix)
constexpr int NT = 128;
constexpr int TILE = BQ * LDS;                               // elements per smem tile
constexpr int SMEM_BYTES = 6 * TILE * 2;                     // Q, K, V x (hi, lo)

__device__ __forceinline__ void ldsm4(uint32_t (&r)[4], const void *p) {
    const uint32_t a = (uint32_t)__cvta_generic_to_shared(p);
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
}
__device__ __forceinline__ void ldsm4t(uint32_t (&r)[4], const void *p) {
    const uint32_t a = (uint32_t)__cvta_generic_to_shared(p);
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
}
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack2(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&v);
}
// split two floats into packed bf16 hi and lo
__device__ __forceinline__ void split2(float a, float b, uint32_t &hi, uint32_t &lo) {
    hi = pack2(a, b);
    lo = pack2(a - bf16_lo(hi), b - bf16_hi(hi));
}

// Load a 64 x 64 fp32 tile (row r at src(r), or zeros if src(r) == nullptr), scale, split, store to smem.
template <typename RowPtr>
__device__ __forceinline__ void load_tile(uint16_t *s_hi, uint16_t *s_lo, float scale, RowPtr rowptr) {
    const int c4 = threadIdx.x & 15, r0 = threadIdx.x >> 4;
    #pragma unroll
    for (int pass = 0; pass < 8; pass++) {
        const int r = pass * 8 + r0;
        const float *src = rowptr(r);
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (src) v = *reinterpret_cast<const float4 *>(src + c4 * 4);
        uint32_t h0, l0, h1, l1;
        split2(v.x * scale, v.y * scale, h0, l0);
        split2(v.z * scale, v.w * scale, h1, l1);
        *reinterpret_cast<uint2 *>(s_hi + r * LDS + c4 * 4) = make_uint2(h0, h1);
        *reinterpret_cast<uint2 *>(s_lo + r * LDS + c4 * 4) = make_uint2(l0, l1);
    }
}

template <int TERMS>
__global__ void __launch_bounds__(NT) enc_attn_kernel(
    const float *__restrict__ qkv, long ldq, const float *__restrict__ kc, const float *__restrict__ vc,
    int M, int pos0, int first_valid, int window, int ring_mask, uint16_t *__restrict__ out, long ldo, int split,
    const int *__restrict__ pos_dev) {
    extern __shared__ __align__(16) uint16_t sm[];
    tcx::pdl_launch_dependents();
    tcx::pdl_wait();
    if (pos_dev) { pos0 = pos_dev[0]; first_valid = pos_dev[1]; }   // replayed CUDA graphs keep {position, oldest valid key} in device memory
    uint16_t *Qh = sm, *Ql = sm + TILE, *Kh = sm + 2 * TILE, *Kl = sm + 3 * TILE, *Vh = sm + 4 * TILE, *Vl = sm + 5 * TILE;
    const int head = blockIdx.y, q0 = blockIdx.x * BQ;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;

    // Q tile, pre-scaled by 1/sqrt(64) (exact power of two)
    load_tile(Qh, Ql, 0.125f, [&](int r) -> const float * {
        return (q0 + r < M) ? qkv + (size_t)(q0 + r) * ldq + head * HD : nullptr;
    });
    __syncthreads();
    uint32_t qh[4][4], ql[4][4];
    {
        const int mi = lane >> 3, r = lane & 7;
        #pragma unroll
        for (int kk = 0; kk < 4; kk++) {
            const int off = (warp * 16 + 8 * (mi & 1) + r) * LDS + kk * 16 + 8 * (mi >> 1);
            ldsm4(qh[kk], Qh + off);
            if (TERMS > 1) ldsm4(ql[kk], Ql + off);
        }
    }
    const int gq0 = pos0 + q0 + warp * 16 + g, gq1 = gq0 + 8;      // logical positions of this thread's two rows
    const int gmin = pos0 + q0, gmax = pos0 + min(q0 + BQ - 1, M - 1);
    int k_lo = gmin - (window - 1);
    if (k_lo < first_valid) k_lo = first_valid;

    float o[8][4];
    #pragma unroll
    for (int n = 0; n < 8; n++) { o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f; }
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;

    for (int kt = k_lo; kt <= gmax; kt += BKEY) {
        __syncthreads();                                   // previous tile fully consumed
        load_tile(Kh, Kl, 1.0f, [&](int r) -> const float * {
            const int j = kt + r;
            return (j <= gmax) ? kc + (size_t)(j & ring_mask) * ENC_QKV + head * HD : nullptr;
        });
        load_tile(Vh, Vl, 1.0f, [&](int r) -> const float * {
            const int j = kt + r;
            return (j <= gmax) ? vc + (size_t)(j & ring_mask) * ENC_QKV + head * HD : nullptr;
        });
        __syncthreads();

        // ---- S = Q K^T (16 x 64 per warp) ----
        float s[8][4];
        #pragma unroll
        for (int n = 0; n < 8; n++) { s[n][0] = s[n][1] = s[n][2] = s[n][3] = 0.f; }
        {
            const int mi = lane >> 3, r = lane & 7;
            #pragma unroll
            for (int n = 0; n < 8; n++) {
                #pragma unroll
                for (int kp = 0; kp < 2; kp++) {
                    const int off = (n * 8 + r) * LDS + kp * 32 + 8 * mi;
                    uint32_t bh[4], bl[4];
                    ldsm4(bh, Kh + off);
                    if (TERMS > 1) {
                        ldsm4(bl, Kl + off);
                        mma16816(s[n], ql[2 * kp], bh[0], bh[1]);
                        mma16816(s[n], ql[2 * kp + 1], bh[2], bh[3]);
                        mma16816(s[n], qh[2 * kp], bl[0], bl[1]);
                        mma16816(s[n], qh[2 * kp + 1], bl[2], bl[3]);
                    }
                    mma16816(s[n], qh[2 * kp], bh[0], bh[1]);
                    mma16816(s[n], qh[2 * kp + 1], bh[2], bh[3]);
                }
            }
        }
        // ---- mask + online softmax ----
        float mx0 = -INFINITY, mx1 = -INFINITY;
        #pragma unroll
        for (int n = 0; n < 8; n++) {
            #pragma unroll
            for (int e = 0; e < 2; e++) {
                const int jk = kt + n * 8 + 2 * t + e;
                const bool v0 = jk <= gq0 && jk > gq0 - window && jk >= first_valid;
                const bool v1 = jk <= gq1 && jk > gq1 - window && jk >= first_valid;
                s[n][e] = v0 ? s[n][e] : -INFINITY;
                s[n][2 + e] = v1 ? s[n][2 + e] : -INFINITY;
                mx0 = fmaxf(mx0, s[n][e]);
                mx1 = fmaxf(mx1, s[n][2 + e]);
            }
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
        const float mu0 = (mn0 == -INFINITY) ? 0.f : mn0, mu1 = (mn1 == -INFINITY) ? 0.f : mn1;
        const float c0 = (m0 == -INFINITY) ? 0.f : expf(m0 - mu0), c1 = (m1 == -INFINITY) ? 0.f : expf(m1 - mu1);
        m0 = mn0; m1 = mn1;
        l0 *= c0; l1 *= c1;
        #pragma unroll
        for (int n = 0; n < 8; n++) { o[n][0] *= c0; o[n][1] *= c0; o[n][2] *= c1; o[n][3] *= c1; }
        uint32_t ph[4][4], pl[4][4];
        #pragma unroll
        for (int n = 0; n < 8; n++) {
            const float p0 = expf(s[n][0] - mu0), p1 = expf(s[n][1] - mu0);
            const float p2 = expf(s[n][2] - mu1), p3 = expf(s[n][3] - mu1);
            l0 += p0 + p1; l1 += p2 + p3;
            // C fragment of n-tiles (2k, 2k+1) = A fragment of key step k
            split2(p0, p1, ph[n >> 1][(n & 1) * 2], pl[n >> 1][(n & 1) * 2]);
            split2(p2, p3, ph[n >> 1][(n & 1) * 2 + 1], pl[n >> 1][(n & 1) * 2 + 1]);
        }
        // ---- O += P V ----
        {
            const int mi = lane >> 3, r = lane & 7;
            #pragma unroll
            for (int kk = 0; kk < 4; kk++) {
                #pragma unroll
                for (int dp = 0; dp < 4; dp++) {
                    const int off = (kk * 16 + 8 * (mi & 1) + r) * LDS + dp * 16 + 8 * (mi >> 1);
                    uint32_t bh[4], bl[4];
                    ldsm4t(bh, Vh + off);
                    if (TERMS > 1) {
                        ldsm4t(bl, Vl + off);
                        mma16816(o[2 * dp], pl[kk], bh[0], bh[1]);
                        mma16816(o[2 * dp + 1], pl[kk], bh[2], bh[3]);
                        mma16816(o[2 * dp], ph[kk], bl[0], bl[1]);
                        mma16816(o[2 * dp + 1], ph[kk], bl[2], bl[3]);
                    }
                    mma16816(o[2 * dp], ph[kk], bh[0], bh[1]);
                    mma16816(o[2 * dp + 1], ph[kk], bh[2], bh[3]);
                }
            }
        }
    }
    // ---- normalise and store [hi | lo] ----
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float i0 = l0 > 0.f ? 1.0f / l0 : 0.f, i1 = l1 > 0.f ? 1.0f / l1 : 0.f;
    const int row0 = q0 + warp * 16 + g, row1 = row0 + 8;
    #pragma unroll
    for (int n = 0; n < 8; n++) {
        const int col = head * HD + n * 8 + 2 * t;
        uint32_t h, l;
        if (row0 < M) {
            split2(o[n][0] * i0, o[n][1] * i0, h, l);
            *reinterpret_cast<uint32_t *>(out + (size_t)row0 * ldo + col) = h;
            if (split) *reinterpret_cast<uint32_t *>(out + (size_t)row0 * ldo + ENC_QKV + col) = l;
        }
        if (row1 < M) {
            split2(o[n][2] * i1, o[n][3] * i1, h, l);
            *reinterpret_cast<uint32_t *>(out + (size_t)row1 * ldo + col) = h;
            if (split) *reinterpret_cast<uint32_t *>(out + (size_t)row1 * ldo + ENC_QKV + col) = l;
        }
    }
}

}  // namespace at

static int enc_attention_launch(vox_cuda_state *s, const float *qkv, const float *kc, const float *vc, int M, int pos0,
                                uint16_t *out_b, long ldo, int split, int terms, const int *pos_dev) {
    EncState *e = (EncState *)s->enc_priv;
    cudaStream_t st = s->stream;
    dim3 grid((M + at::BQ - 1) / at::BQ, ENC_HEADS);
    if (terms == 3)
        CUDA_TRY(launch_pdl(at::enc_attn_kernel<3>, grid, at::NT, at::SMEM_BYTES, st, qkv, 3L * ENC_QKV, kc, vc, M, pos0, e->pos_min, ENC_WINDOW,
                            ENC_RING - 1, out_b, ldo, split, pos_dev));
    else
        CUDA_TRY(launch_pdl(at::enc_attn_kernel<1>, grid, at::NT, at::SMEM_BYTES, st, qkv, 3L * ENC_QKV, kc, vc, M, pos0, e->pos_min, ENC_WINDOW,
                            ENC_RING - 1, out_b, ldo, split, pos_dev));
    LAUNCHED();
    return 0;
}

// qkv: fp32 [M][3*2048] (q rotated in place by rows_rope_kv); kc / vc: this layer's fp32 ring.
int enc_attention_dev(vox_cuda_state *s, const float *qkv, const float *kc, const float *vc, int M, int pos0,
                      uint16_t *out_b, long ldo, int split, const int *pos_dev);
int enc_attention(vox_cuda_state *s, const float *qkv, const float *kc, const float *vc, int M, int pos0,
                  uint16_t *out_b, long ldo, int split) {
    return enc_attention_dev(s, qkv, kc, vc, M, pos0, out_b, ldo, split, nullptr);
}
// pos_dev (optional): device int holding pos0 (tensor-core path only)
int enc_attention_dev(vox_cuda_state *s, const float *qkv, const float *kc, const float *vc, int M, int pos0,
                      uint16_t *out_b, long ldo, int split, const int *pos_dev) {
    EncState *e = (EncState *)s->enc_priv;
    cudaStream_t st = s->stream;
    static int mode = -1;     // 0: tensor cores, 1: fp32 CUDA-core kernel + conversion (A/B debugging)
    if (mode < 0) {
        const char *env = getenv("VOX_CUDA_ENC_ATTN_F32");
        mode = (env && atoi(env)) ? 1 : 0;
        cudaFuncSetAttribute(at::enc_attn_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, at::SMEM_BYTES);
        cudaFuncSetAttribute(at::enc_attn_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, at::SMEM_BYTES);
    }
    if (mode == 1) {
        if (pos_dev) { vox_set_error("enc_attention: device position needs the tensor-core path"); return -1; }
        if (rows_attention(st, ENC_HD, qkv, 3 * ENC_QKV, kc, vc, ENC_QKV, ENC_RING - 1, ENC_HEADS, 1, M, pos0, e->pos_min,
                           ENC_WINDOW, e->att, ENC_QKV)) return -1;
        return f32_to_bf16_rows(st, e->att, ENC_QKV, out_b, ldo, M, ENC_QKV, split);
    }
    return enc_attention_launch(s, qkv, kc, vc, M, pos0, out_b, ldo, split, split ? 3 : 1, pos_dev);
}

int enc_attention_terms(vox_cuda_state *s, const float *qkv, const float *kc, const float *vc, int M, int pos0,
                        uint16_t *out_b, long ldo, int split, int terms) {
    static int init = 0;
    if (!init) {
        cudaFuncSetAttribute(at::enc_attn_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, at::SMEM_BYTES);
        cudaFuncSetAttribute(at::enc_attn_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, at::SMEM_BYTES);
        init = 1;
    }
    return enc_attention_launch(s, qkv, kc, vc, M, pos0, out_b, ldo, split, terms, nullptr);
}
