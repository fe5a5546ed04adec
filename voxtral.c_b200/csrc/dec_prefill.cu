// dec_prefill.cu - prompt prefill (vox_decoder_prefill, /root/reference/voxtral_decoder.c:447-612) for blocks of
// <= 64 rows on the skinny swap-AB tcgen05 GEMM (gemm_wt.cu) instead of the large-tile one.
//
// With M = 38 prompt rows the large-tile GEMM (gemm_tc.cu) has N / 128 = 24 .. 144 tiles: wqkv, wo and w2 stream their
// 25 - 57 MB through 24 - 48 SMs, and the whole prefill took 5.05 ms for 6.05 GB of weights (0.92 ms at the HBM peak).
// Here every product is cut into ~144 (weight tile, K range) work items as the stream groups do (api_group.cu), the
// K-range partials are summed by the small row kernels below - which also apply RoPE / append K and V, add the residual
// and produce the next product's split-bf16 rows - and everything is chained with programmatic dependent launch.
// Arithmetic: split-bf16 activations x bf16 weights, fp32 accumulation, like the large-tile path; Q8 models multiply by
// the row scale in the GEMM epilogue (WtOp::row_scale).
#include "vox_internal.h"
#include "gemm_tc.h"
#include "gemm_wt.h"
#include "rows.h"
#include "tc_ptx.cuh"
#include <stdlib.h>

namespace {

constexpr int PW_NB = 64;                                    // rows per block = narrow UMMA operand width
constexpr int KS_QKV = 3, KS_WO = 6, KS_W2 = 6;              // K ranges: 48 x 3, 24 x 6, 24 x 6 work items (w1|w3: 144 x 1)
constexpr int NT_ROW = DEC_DIM / 4;                          // 768 threads: one float4 of a [3072] row each

struct PfWt {
    uint16_t *xn_t, *att_t, *hh_t;                           // split-bf16 rows [64][2*3072], [64][2*4096], [64][2*9216]
    float *part;                                             // K-range partials
    float *qkv;                                              // [64][6144] fp32: q (rotated) for the attention; k / v go to the ring
    WtOp *ops;                                               // [n_dec][4]
};

__device__ __forceinline__ void store_split4(uint16_t *row, int ld_lo, int col, float y0, float y1, float y2, float y3) {
    const __nv_bfloat162 h0 = __floats2bfloat162_rn(y0, y1), h1 = __floats2bfloat162_rn(y2, y3);
    const __nv_bfloat162 l0 = __floats2bfloat162_rn(y0 - __bfloat162float(h0.x), y1 - __bfloat162float(h0.y));
    const __nv_bfloat162 l1 = __floats2bfloat162_rn(y2 - __bfloat162float(h1.x), y3 - __bfloat162float(h1.y));
    uint2 hv, lv;
    hv.x = *reinterpret_cast<const uint32_t *>(&h0); hv.y = *reinterpret_cast<const uint32_t *>(&h1);
    lv.x = *reinterpret_cast<const uint32_t *>(&l0); lv.y = *reinterpret_cast<const uint32_t *>(&l1);
    *reinterpret_cast<uint2 *>(row + col) = hv;
    *reinterpret_cast<uint2 *>(row + ld_lo + col) = lv;
}

// x[m] += sum_s part[s][m] (KS > 0: wo / w2 + residual, voxtral_decoder.c:556-558,590-592); then, if nw, RMSNorm(x[m]) * nw
// (* (1 + ada), :564-570) as a split-bf16 row for the next product.  One CTA per row.
template <int KS>
__global__ void __launch_bounds__(NT_ROW) pf_resid_norm_kernel(float *__restrict__ x, const float *__restrict__ part, long stride,
                                                               const float *__restrict__ nw, const float *__restrict__ ada,
                                                               uint16_t *__restrict__ out_b) {
    tcx::pdl_launch_dependents();
    tcx::pdl_wait();
    const int m = blockIdx.x, col = threadIdx.x * 4;
    float *xr = x + (size_t)m * DEC_DIM;
    float4 a = *reinterpret_cast<const float4 *>(xr + col);
    if (KS > 0) {
        const float *p = part + (size_t)m * DEC_DIM + col;
        float4 v[KS > 0 ? KS : 1];
        #pragma unroll
        for (int s = 0; s < KS; s++) v[s] = *reinterpret_cast<const float4 *>(p + (size_t)s * stride);   // all loads in flight
        float4 acc = v[0];
        #pragma unroll
        for (int s = 1; s < KS; s++) { acc.x += v[s].x; acc.y += v[s].y; acc.z += v[s].z; acc.w += v[s].w; }
        a.x += acc.x; a.y += acc.y; a.z += acc.z; a.w += acc.w;
        *reinterpret_cast<float4 *>(xr + col) = a;
    }
    if (!nw) return;
    __shared__ float red[NT_ROW / 32];
    const float ss = warp_sum(a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
    #pragma unroll
    for (int i = 0; i < NT_ROW / 32; i++) tot += red[i];
    const float rinv = 1.0f / sqrtf(tot / (float)DEC_DIM + NORM_EPS);
    const float4 wv = *reinterpret_cast<const float4 *>(nw + col);
    float y0 = a.x * rinv * wv.x, y1 = a.y * rinv * wv.y, y2 = a.z * rinv * wv.z, y3 = a.w * rinv * wv.w;
    if (ada) {
        const float4 av = *reinterpret_cast<const float4 *>(ada + col);
        y0 *= (1.0f + av.x); y1 *= (1.0f + av.y); y2 *= (1.0f + av.z); y3 *= (1.0f + av.w);
    }
    store_split4(out_b + (size_t)m * 2 * DEC_DIM, DEC_DIM, col, y0, y1, y2, y3);
}

// q | k | v row = sum_s part[s]; RoPE on q and k at the row's logical position (interleaved pairs, voxtral_kernels.c:617-655),
// k and v appended to the ring slot (pos & mask), q left in qkv for the attention.  grid (rows, 6): 1024 columns per CTA.
__global__ void __launch_bounds__(256) pf_qkv_rope_kernel(const float *__restrict__ part, int ks, long stride, float *__restrict__ qkv,
                                                          const float *__restrict__ inv_freq, int pos0, int ring_mask,
                                                          float *__restrict__ kc, float *__restrict__ vc) {
    tcx::pdl_launch_dependents();
    tcx::pdl_wait();
    const int row = blockIdx.x, pos = pos0 + row;
    const int col = blockIdx.y * 1024 + threadIdx.x * 4;
    const size_t slot = (size_t)(pos & ring_mask) * DEC_KVDIM;
    const float *p = part + (size_t)row * DEC_QKV + col;
    float4 a = *reinterpret_cast<const float4 *>(p);
    for (int s = 1; s < ks; s++) {
        const float4 v = *reinterpret_cast<const float4 *>(p + (size_t)s * stride);
        a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }
    if (col >= DEC_QDIM + DEC_KVDIM) {
        *reinterpret_cast<float4 *>(vc + slot + col - DEC_QDIM - DEC_KVDIM) = a;
        return;
    }
    const int d = (col & (DEC_HD - 1)) >> 1;                  // pair indices d, d + 1 inside the head
    const float a0 = (float)pos * inv_freq[d], a1 = (float)pos * inv_freq[d + 1];
    const float c0 = cosf(a0), s0 = sinf(a0), c1 = cosf(a1), s1 = sinf(a1);
    const float4 y = make_float4(a.x * c0 - a.y * s0, a.x * s0 + a.y * c0, a.z * c1 - a.w * s1, a.z * s1 + a.w * c1);
    if (col < DEC_QDIM) *reinterpret_cast<float4 *>(qkv + (size_t)row * DEC_QKV + col) = y;
    else *reinterpret_cast<float4 *>(kc + slot + col - DEC_QDIM) = y;
}

PfWt *pf_get(vox_cuda_state *s) {
    if (s->pf_wt) return (PfWt *)s->pf_wt;
    vox_cuda_model *m = s->m;
    PfWt *q = (PfWt *)calloc(1, sizeof(PfWt));
    if (!q) return nullptr;
    const size_t part_f = (size_t)KS_QKV * PW_NB * DEC_QKV > (size_t)KS_W2 * PW_NB * DEC_DIM ? (size_t)KS_QKV * PW_NB * DEC_QKV
                                                                                           : (size_t)KS_W2 * PW_NB * DEC_DIM;
    const size_t bytes[5] = {(size_t)PW_NB * 2 * DEC_DIM * 2, (size_t)PW_NB * 2 * DEC_QDIM * 2, (size_t)PW_NB * 2 * DEC_HIDDEN * 2,
                             part_f * 4, (size_t)PW_NB * DEC_QKV * 4};
    void **ptrs[5] = {(void **)&q->xn_t, (void **)&q->att_t, (void **)&q->hh_t, (void **)&q->part, (void **)&q->qkv};
    bool ok = true;
    for (int i = 0; i < 5 && ok; i++)
        ok = cudaMalloc(ptrs[i], bytes[i]) == cudaSuccess && cudaMemsetAsync(*ptrs[i], 0, bytes[i], s->stream) == cudaSuccess;
    q->ops = ok ? (WtOp *)aligned_alloc(64, sizeof(WtOp) * 4 * m->n_dec) : nullptr;
    ok = ok && q->ops;
    for (int l = 0; l < m->n_dec && ok; l++) {
        const DecLayerDev &L = m->dec[l];
        WtOp *o = q->ops + 4 * l;
        ok = !gemm_wt_prepare(&o[0], PW_NB, 1, DEC_QKV, DEC_DIM, 1, q->xn_t, 2 * DEC_DIM, L.wqkv, KS_QKV) &&
             !gemm_wt_prepare(&o[1], PW_NB, 1, DEC_DIM, DEC_QDIM, 1, q->att_t, 2 * DEC_QDIM, L.wo, KS_WO) &&
             !gemm_wt_prepare(&o[2], PW_NB, 1, 2 * DEC_HIDDEN, DEC_DIM, 1, q->xn_t, 2 * DEC_DIM, L.w13, 1) &&
             !gemm_wt_prepare(&o[3], PW_NB, 1, DEC_DIM, DEC_HIDDEN, 1, q->hh_t, 2 * DEC_HIDDEN, L.w2, KS_W2);
        if (!ok) break;
        o[0].out_f32 = q->part; o[0].ldo = DEC_QKV;
        o[1].out_f32 = q->part; o[1].ldo = DEC_DIM;
        o[2].mode = WT_OUT_SWIGLU_BF16X2; o[2].out_bf16 = q->hh_t; o[2].ldb = 2 * DEC_HIDDEN; o[2].lo_off = DEC_HIDDEN;
        o[3].out_f32 = q->part; o[3].ldo = DEC_DIM;
        o[0].row_scale = L.s_wqkv; o[1].row_scale = L.s_wo; o[2].row_scale = L.s_w13; o[3].row_scale = L.s_w2;   // Q8 model, else null
    }
    if (!ok) {
        vox_set_error("prefill (skinny GEMM path): setup failed");
        cudaFree(q->xn_t); cudaFree(q->att_t); cudaFree(q->hh_t); cudaFree(q->part); cudaFree(q->qkv);
        free(q->ops); free(q);
        return nullptr;
    }
    s->pf_wt = q;
    return q;
}

}  // namespace

void dec_prefill_wt_free(vox_cuda_state *s) {
    PfWt *q = (PfWt *)s->pf_wt;
    if (!q) return;
    cudaFree(q->xn_t); cudaFree(q->att_t); cudaFree(q->hh_t); cudaFree(q->part); cudaFree(q->qkv);
    free(q->ops); free(q);
    s->pf_wt = nullptr;
}

// One block of M <= 64 rows x (the residual stream, updated in place); K/V rows go to the ring kbase / vbase as in
// dec_prefill_into.  Enqueues only.
int dec_prefill_block_wt(vox_cuda_state *s, float *x, int M, float *kbase, float *vbase, size_t kv_layer, int mask, int pos0,
                         int first_valid) {
    PfWt *q = pf_get(s);
    if (!q || M < 1 || M > PW_NB) return -1;
    vox_cuda_model *m = s->m;
    cudaStream_t st = s->stream;
    CUDA_TRY(launch_pdl(pf_resid_norm_kernel<0>, M, NT_ROW, 0, st, x, (const float *)nullptr, 0L, m->dec[0].attn_norm,
                             (const float *)nullptr, q->xn_t));
    LAUNCHED();
    for (int l = 0; l < m->n_dec; l++) {
        const DecLayerDev &L = m->dec[l];
        float *kc = kbase + l * kv_layer, *vc = vbase + l * kv_layer;
        WtOp *o = q->ops + 4 * l;
        for (int i = 0; i < 4; i++) o[i].B = M;
        o[0].split_stride = (long)M * DEC_QKV; o[1].split_stride = (long)M * DEC_DIM; o[3].split_stride = (long)M * DEC_DIM;
        if (gemm_wt_launch(st, &o[0])) return -1;
        CUDA_TRY(launch_pdl(pf_qkv_rope_kernel, dim3(M, DEC_QKV / 1024), 256, 0, st, (const float *)q->part, KS_QKV,
                                 (long)M * DEC_QKV, q->qkv, (const float *)m->inv_freq_dec, pos0, mask, kc, vc));
        LAUNCHED();
        if (rows_attention_split128(st, q->qkv, DEC_QKV, kc, vc, DEC_KVDIM, mask, DEC_HEADS, DEC_HEADS / DEC_KVH, M, pos0,
                                    first_valid, DEC_WINDOW, q->att_t, 2 * DEC_QDIM, DEC_QDIM)) return -1;
        if (gemm_wt_launch(st, &o[1])) return -1;
        CUDA_TRY(launch_pdl(pf_resid_norm_kernel<KS_WO>, M, NT_ROW, 0, st, x, (const float *)q->part, (long)M * DEC_DIM, L.ffn_norm,
                                 (const float *)(m->ada_scale + (size_t)l * DEC_DIM), q->xn_t));
        LAUNCHED();
        if (gemm_wt_launch(st, &o[2])) return -1;
        if (gemm_wt_launch(st, &o[3])) return -1;
        const bool last = l == m->n_dec - 1;                 // no logits in the prefill (voxtral_decoder.c:447-612): no final norm
        CUDA_TRY(launch_pdl(pf_resid_norm_kernel<KS_W2>, M, NT_ROW, 0, st, x, (const float *)q->part, (long)M * DEC_DIM,
                                 last ? (const float *)nullptr : m->dec[l + 1].attn_norm, (const float *)nullptr, q->xn_t));
        LAUNCHED();
    }
    return 0;
}
