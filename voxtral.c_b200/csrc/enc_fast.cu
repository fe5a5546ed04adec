// enc_fast.cu - throughput mode of the encoder layers: every linear layer of
// vox_encoder_forward_incremental (/root/reference/voxtral_encoder.c:562-686)
// runs on the tcgen05 GEMM (gemm_tc.cu) with its element-wise neighbours fused
// into the epilogue:
//     rmsnorm -> bf16            (rows.cu)
//     QKV  + bias                (gemm_tc, fp32 out)        voxtral_encoder.c:591-593
//     RoPE + KV-ring append      (rows.cu)                  :598-607
//     sliding-window attention   (rows.cu / attn_tc.cu)     :622-624
//     wo   + bias + residual     (gemm_tc, x += ...)        :644-646
//     rmsnorm -> bf16
//     w1|w3 + SwiGLU -> bf16     (gemm_tc epilogue)         :670-676
//     w2   + bias + residual     (gemm_tc, x += ...)        :677-679
// The residual stream x stays fp32; only GEMM A operands are rounded to bf16.
// VOX_CUDA_ENC_PRECISE=1 selects the all-fp32 CUDA-core path (api_enc.cu) instead.
#include "vox_internal.h"
#include "enc_state.h"
#include "rows.h"
#include "gemm_tc.h"
#include <stdlib.h>

// attn_tc.cu: sliding-window attention on tensor cores, output [hi | lo] bf16 rows
int enc_attention(vox_cuda_state *s, const float *qkv, const float *kc, const float *vc, int M, int pos0,
                  uint16_t *out_b, long ldo, int split);
int enc_attention_dev(vox_cuda_state *s, const float *qkv, const float *kc, const float *vc, int M, int pos0,
                      uint16_t *out_b, long ldo, int split, const int *pos_dev);
// terms = 3: fp32-class products (hi*hi + hi*lo + lo*hi), 1: plain bf16; split = output rows as [hi | lo]
int enc_attention_terms(vox_cuda_state *s, const float *qkv, const float *kc, const float *vc, int M, int pos0,
                        uint16_t *out_b, long ldo, int split, int terms);

void vox_mem_account(long delta);

#ifndef VOX_ENC_X1_DEFAULT
#define VOX_ENC_X1_DEFAULT 0
#endif
// Activations feeding the tensor cores are stored as TWO bf16 terms, [hi | lo] per
// row ("bf16x2", a = hi + lo to ~16 mantissa bits); the GEMM walks both halves against
// the same weights.  That keeps the encoder at fp32-class accuracy (max |d| ~1e-4
// instead of ~3e-2 with a single bf16 term, measured) for twice the - cheap - tensor
// work.  VOX_CUDA_ENC_BF16X1=1 selects single-term operands.
static int split_mode(void) {
    static int sp = -1;
    if (sp < 0) {
        const char *env = getenv("VOX_CUDA_ENC_BF16X1");
        sp = (env && atoi(env)) ? 0 : 1;
    }
    return sp;
}
// Which products take SINGLE-bf16 activations (bit set) instead of the hi/lo pair: 1 = wq|wk|wv, 2 = wo,
// 4 = w1|w3, 8 = w2, 16 = adapter, 32 = attention (QK^T and PV with one term instead of three).
// VOX_CUDA_ENC_BF16X1=1 is the same as mask 63.  The default is set from the per-product accuracy study in
// profiles/ (each cleared bit doubles that product's tensor work).
enum { X1_QKV = 1, X1_WO = 2, X1_W13 = 4, X1_W2 = 8, X1_ADAPTER = 16, X1_ATTN = 32 };
static int x1_mask(void) {
    static int m = -1;
    if (m < 0) {
        const char *env = getenv("VOX_CUDA_ENC_X1_MASK");
        m = env ? atoi(env) & 63 : VOX_ENC_X1_DEFAULT;
        if (!split_mode()) m = 63;
    }
    return m;
}

struct EncFast {
    uint16_t *xn_b;     // [R][2*1280]
    uint16_t *att_b;    // [R][2*2048]
    uint16_t *hh_b;     // [R][2*5120]
    uint16_t *enc_b;    // [R/4][2*5120] final-normed rows regrouped for the adapter
    uint16_t *mid_b;    // [R/4 + 1][2*3072]
};

static EncFast *fast_get(vox_cuda_state *s) {
    EncState *e = (EncState *)s->enc_priv;
    if (e->fast) return (EncFast *)e->fast;
    EncFast *f = (EncFast *)calloc(1, sizeof(EncFast));
    const size_t R = ENC_CHUNK + 8;
    const size_t bytes[5] = {R * ENC_DIM * 4, R * ENC_QKV * 4, R * ENC_HIDDEN * 4, R * ENC_DIM * 4, (R / 4 + 2) * DEC_DIM * 4};
    uint16_t **ptrs[5] = {&f->xn_b, &f->att_b, &f->hh_b, &f->enc_b, &f->mid_b};
    for (int i = 0; i < 5; i++) {
        if (cudaMalloc((void **)ptrs[i], bytes[i]) != cudaSuccess) {
            vox_set_error("enc_fast: cudaMalloc(%zu) failed", bytes[i]);
            return nullptr;
        }
        vox_mem_account((long)bytes[i]);
    }
    e->fast = f;
    return f;
}

void enc_fast_free(vox_cuda_state *s) {
    EncState *e = (EncState *)s->enc_priv;
    if (!e || !e->fast) return;
    EncFast *f = (EncFast *)e->fast;
    cudaFree(f->xn_b); cudaFree(f->att_b); cudaFree(f->hh_b); cudaFree(f->enc_b); cudaFree(f->mid_b);
    free(f);
    e->fast = nullptr;
}

static TcEpilogue epi(float *of, long ldf, uint16_t *ob, long ldb, const float *bias, int flags, long split_off = 0,
                      const float *col_scale = nullptr) {
    TcEpilogue t;
    t.out_f32 = of; t.ldo_f32 = ldf; t.out_bf16 = ob; t.ldo_bf16 = ldb; t.bias = bias; t.flags = flags;
    t.split_off = split_off;
    t.col_scale = col_scale;             // Q8 model: row scales of the quantised matrix, else null
    return t;
}

int enc_layers_fast(vox_cuda_state *s, float *x, int M, float *out) {
    EncState *e = (EncState *)s->enc_priv;
    EncFast *f = fast_get(s);
    if (!f) return -1;
    vox_cuda_model *m = s->m;
    cudaStream_t st = s->stream;
    const size_t kv_layer = (size_t)ENC_RING * ENC_QKV;
    const int pos0 = e->pos;
    const int mask = x1_mask();
    // per product: split = 1 -> A operand is [hi | lo] (row length 2K), 0 -> one bf16 term
    const int s_qkv = !(mask & X1_QKV), s_wo = !(mask & X1_WO), s_w13 = !(mask & X1_W13), s_w2 = !(mask & X1_W2);
    const int s_att = !(mask & X1_ATTN);
    for (int l = 0; l < m->n_enc; l++) {
        const EncLayerDev &L = m->enc[l];
        float *kc = e->kv_k + l * kv_layer, *vc = e->kv_v + l * kv_layer;
        if (rows_rmsnorm_bf16(st, x, ENC_DIM, L.attn_norm, nullptr, f->xn_b, (1 + s_qkv) * ENC_DIM, M, ENC_DIM, s_qkv)) return -1;
        if (gemm_tc(st, M, 3 * ENC_QKV, ENC_DIM, (1 + s_qkv) * ENC_DIM, f->xn_b, (1 + s_qkv) * ENC_DIM, L.wqkv,
                    epi(e->qkv, 3 * ENC_QKV, nullptr, 0, L.bqkv, TC_BIAS | TC_OUT_F32, 0, L.s_wqkv))) return -1;
        if (rows_rope_kv(st, e->qkv, M, ENC_QKV, ENC_QKV, ENC_HD, m->inv_freq_enc, pos0, ENC_RING - 1, kc, vc)) return -1;
        if (enc_attention_terms(s, e->qkv, kc, vc, M, pos0, f->att_b, (1 + s_wo) * ENC_QKV, s_wo, s_att ? 3 : 1)) return -1;
        if (gemm_tc(st, M, ENC_DIM, ENC_QKV, (1 + s_wo) * ENC_QKV, f->att_b, (1 + s_wo) * ENC_QKV, L.wo,
                    epi(x, ENC_DIM, nullptr, 0, L.bo, TC_BIAS | TC_RESIDUAL, 0, L.s_wo))) return -1;
        if (rows_rmsnorm_bf16(st, x, ENC_DIM, L.ffn_norm, nullptr, f->xn_b, (1 + s_w13) * ENC_DIM, M, ENC_DIM, s_w13)) return -1;
        if (gemm_tc(st, M, 2 * ENC_HIDDEN, ENC_DIM, (1 + s_w13) * ENC_DIM, f->xn_b, (1 + s_w13) * ENC_DIM, L.w13,
                    epi(nullptr, 0, f->hh_b, (1 + s_w2) * ENC_HIDDEN, nullptr, TC_SWIGLU | TC_OUT_BF16 | (s_w2 ? TC_OUT_SPLIT : 0), ENC_HIDDEN, L.s_w13))) return -1;
        if (gemm_tc(st, M, ENC_DIM, ENC_HIDDEN, (1 + s_w2) * ENC_HIDDEN, f->hh_b, (1 + s_w2) * ENC_HIDDEN, L.w2,
                    epi(x, ENC_DIM, nullptr, 0, L.b2, TC_BIAS | TC_RESIDUAL, 0, L.s_w2))) return -1;
    }
    if (rows_rmsnorm(st, x, ENC_DIM, m->enc_norm, nullptr, out, ENC_DIM, M, ENC_DIM)) return -1;
    e->pos += M;
    return 0;
}

// adapter (vox_adapter_forward, voxtral_encoder.c:699-737) on the tensor cores:
// [k, 5120] -> Linear 5120->3072 -> GELU -> Linear 3072->3072
int adapter_fast(vox_cuda_state *s, const float *enc_rows, int k_tokens, float *out) {
    EncFast *f = fast_get(s);
    if (!f) return -1;
    vox_cuda_model *m = s->m;
    cudaStream_t st = s->stream;
    const int sp = !(x1_mask() & X1_ADAPTER), mul = sp ? 2 : 1;
    const int K0 = 4 * ENC_DIM;
    if (f32_to_bf16_rows(st, enc_rows, K0, f->enc_b, mul * K0, k_tokens, K0, sp)) return -1;
    if (gemm_tc(st, k_tokens, DEC_DIM, K0, mul * K0, f->enc_b, mul * K0, m->adapter_w0,
                epi(nullptr, 0, f->mid_b, mul * DEC_DIM, nullptr, TC_GELU | TC_OUT_BF16 | (sp ? TC_OUT_SPLIT : 0), DEC_DIM, m->s_adapter0))) return -1;
    return gemm_tc(st, k_tokens, DEC_DIM, DEC_DIM, mul * DEC_DIM, f->mid_b, mul * DEC_DIM, m->adapter_w1,
                   epi(out, DEC_DIM, nullptr, 0, nullptr, TC_OUT_F32, 0, m->s_adapter1));
}

// ---- small-M path (streaming chunks: M <= 128 new positions per call) -----------------------------
// With M <= 128 rows the GEMMs above are weight-streaming problems (1.93 GB of encoder weights per
// call) and a 128 x 128 tile grid gives only N/128 = 10..80 CTAs, each pulling its slab at one SM's
// ingress rate: 5.9 ms per call, measured, against a 0.30 ms HBM floor.  Here the layers run on the
// skinny swap-AB GEMM of the stream groups (gemm_wt.cu: weight tile = UMMA A operand, the M rows =
// narrow B operand, K-range splits so that ~140 CTAs stream weights); K-range partials are summed
// (+ bias, + residual) by the small kernels below, which also produce the next GEMM's split-bf16 rows.
#include "gemm_wt.h"
#include "tc_ptx.cuh"

namespace {

constexpr int SM_KS_QKV = 3, SM_KS_WO = 14, SM_KS_W2 = 14, SM_KS_MAX = 14;

// out[m][n] = sum_s part[s][m][n] + bias[n]           (QKV + bias, voxtral_encoder.c:591-593)
__global__ void __launch_bounds__(256) enc_sum_bias_kernel(const float *__restrict__ part, int ks, long stride,
                                                           const float *__restrict__ bias, float *__restrict__ out, int N) {
    tcx::pdl_launch_dependents();
    tcx::pdl_wait();
    const int m = blockIdx.x, col = (blockIdx.y * 256 + threadIdx.x) * 4;
    if (col >= N) return;
    float4 a = *reinterpret_cast<const float4 *>(bias + col);
    const float *p = part + (size_t)m * N + col;
    for (int s = 0; s < ks; s++) {
        const float4 v = *reinterpret_cast<const float4 *>(p + (size_t)s * stride);
        a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }
    *reinterpret_cast<float4 *>(out + (size_t)m * N + col) = a;
}

// The two kernels that followed the QKV product in one: q / k / v row = sum_s part[s] + bias (as enc_sum_bias_kernel), RoPE on
// q and k at the row's logical position, k and v appended to the ring, q left in qkv (as rows_rope_kv_kernel, rows.cu; same
// float expressions in the same order).
__global__ void __launch_bounds__(256) enc_sum_bias_rope_kernel(const float *__restrict__ part, int ks, long stride,
                                                                const float *__restrict__ bias, float *__restrict__ qkv,
                                                                const float *__restrict__ inv_freq, int pos0,
                                                                const int *__restrict__ pos_dev, int ring_mask,
                                                                float *__restrict__ kc, float *__restrict__ vc) {
    tcx::pdl_launch_dependents();
    tcx::pdl_wait();
    if (pos_dev) pos0 = *pos_dev;
    constexpr int LD = 3 * ENC_QKV;
    // grid (rows, 6): 1024 columns per CTA - y 0,1 = q, 2,3 = k, 4,5 = v; four columns (two RoPE pairs) per thread
    const int row = blockIdx.x, pos = pos0 + row;
    const int col = blockIdx.y * 1024 + threadIdx.x * 4;
    const size_t slot = (size_t)(pos & ring_mask) * ENC_QKV;
    const float *p = part + (size_t)row * LD + col;
    float4 a = *reinterpret_cast<const float4 *>(bias + col);
    for (int s = 0; s < ks; s++) {
        const float4 v = *reinterpret_cast<const float4 *>(p + (size_t)s * stride);
        a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }
    if (col >= 2 * ENC_QKV) {                                 // v: appended as it is
        *reinterpret_cast<float4 *>(vc + slot + col - 2 * ENC_QKV) = a;
        return;
    }
    const int d = (col & (ENC_HD - 1)) >> 1;                  // pair indices d, d + 1 inside the head
    const float a0 = (float)pos * inv_freq[d], a1 = (float)pos * inv_freq[d + 1];
    const float c0 = cosf(a0), s0 = sinf(a0), c1 = cosf(a1), s1 = sinf(a1);
    const float4 y = make_float4(a.x * c0 - a.y * s0, a.x * s0 + a.y * c0, a.z * c1 - a.w * s1, a.z * s1 + a.w * c1);
    if (col < ENC_QKV) *reinterpret_cast<float4 *>(qkv + (size_t)row * LD + col) = y;
    else *reinterpret_cast<float4 *>(kc + slot + col - ENC_QKV) = y;
}

// x[m] += sum_s part[s][m] + bias; then RMSNorm(x[m]) * w -> split bf16 row (out_b) or fp32 row (out_f)
// (wo / w2 + bias + residual, voxtral_encoder.c:644-646,677-679, followed by the next norm :562-566,652-656)
template <int KS>
__global__ void __launch_bounds__(320) enc_resid_norm_kernel(float *__restrict__ x, const float *__restrict__ part, long stride,
                                                             const float *__restrict__ bias, const float *__restrict__ w,
                                                             uint16_t *__restrict__ out_b, float *__restrict__ out_f) {
    tcx::pdl_launch_dependents();
    tcx::pdl_wait();
    const int m = blockIdx.x, col = threadIdx.x * 4;
    float *xr = x + (size_t)m * ENC_DIM;
    float4 a = *reinterpret_cast<const float4 *>(xr + col);
    const float4 b = *reinterpret_cast<const float4 *>(bias + col);
    float4 acc = b;
    const float *p = part + (size_t)m * ENC_DIM + col;
    float4 v[KS];
    #pragma unroll
    for (int s = 0; s < KS; s++) v[s] = *reinterpret_cast<const float4 *>(p + (size_t)s * stride);   // all loads in flight
    #pragma unroll
    for (int s = 0; s < KS; s++) { acc.x += v[s].x; acc.y += v[s].y; acc.z += v[s].z; acc.w += v[s].w; }
    a.x += acc.x; a.y += acc.y; a.z += acc.z; a.w += acc.w;
    *reinterpret_cast<float4 *>(xr + col) = a;
    __shared__ float red[10];
    float ss = warp_sum(a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
    #pragma unroll
    for (int i = 0; i < 10; i++) tot += red[i];
    const float rinv = 1.0f / sqrtf(tot / (float)ENC_DIM + NORM_EPS);
    const float4 wv = *reinterpret_cast<const float4 *>(w + col);
    const float y0 = a.x * rinv * wv.x, y1 = a.y * rinv * wv.y, y2 = a.z * rinv * wv.z, y3 = a.w * rinv * wv.w;
    if (out_f) *reinterpret_cast<float4 *>(out_f + (size_t)m * ENC_DIM + col) = make_float4(y0, y1, y2, y3);
    if (out_b) {
        const __nv_bfloat162 h0 = __floats2bfloat162_rn(y0, y1), h1 = __floats2bfloat162_rn(y2, y3);
        const __nv_bfloat162 l0 = __floats2bfloat162_rn(y0 - __bfloat162float(h0.x), y1 - __bfloat162float(h0.y));
        const __nv_bfloat162 l1 = __floats2bfloat162_rn(y2 - __bfloat162float(h1.x), y3 - __bfloat162float(h1.y));
        uint2 hv, lv;
        hv.x = *reinterpret_cast<const uint32_t *>(&h0); hv.y = *reinterpret_cast<const uint32_t *>(&h1);
        lv.x = *reinterpret_cast<const uint32_t *>(&l0); lv.y = *reinterpret_cast<const uint32_t *>(&l1);
        uint16_t *orow = out_b + (size_t)m * 2 * ENC_DIM;
        *reinterpret_cast<uint2 *>(orow + col) = hv;
        *reinterpret_cast<uint2 *>(orow + ENC_DIM + col) = lv;
    }
}

constexpr int SM_GRAPHS = 16;
struct EncSmall {
    uint16_t *xn_t, *att_t, *hh_t;      // [128][2*1280], [128][2*2048], [128][2*5120] split bf16 rows
    float *part;                        // K-range partials, max(14*128*1280, 3*128*6144) floats
    WtOp *ops[3];                       // per NB in {32, 64, 128}: [n_enc][4] prepared products (lazily)
    // The call is launch bound on the host (~290 launches); a chunk shape seen twice is captured into a CUDA
    // graph (position in device memory) and replayed.  Key = (M, x, out).
    int *d_pos, *h_pos;                 // device {position, oldest valid key} + pinned staging ring
    unsigned h_pos_i;
    struct { int M, seen; const float *x, *out; cudaGraphExec_t exec; } graphs[SM_GRAPHS];
};

}  // namespace

void enc_small_free(void *p) {
    EncSmall *q = (EncSmall *)p;
    if (!q) return;
    cudaFree(q->xn_t); cudaFree(q->att_t); cudaFree(q->hh_t); cudaFree(q->part); cudaFree(q->d_pos);
    if (q->h_pos) cudaFreeHost(q->h_pos);
    for (int i = 0; i < SM_GRAPHS; i++) if (q->graphs[i].exec) cudaGraphExecDestroy(q->graphs[i].exec);
    for (int i = 0; i < 3; i++) free(q->ops[i]);
    free(q);
}

static EncSmall *small_get(vox_cuda_state *s, int NB) {
    EncState *e = (EncState *)s->enc_priv;
    EncSmall *q = (EncSmall *)e->small;
    vox_cuda_model *m = s->m;
    if (!q) {
        q = (EncSmall *)calloc(1, sizeof(EncSmall));
        if (!q) return nullptr;
        const size_t part_f = (size_t)SM_KS_QKV * 128 * 3 * ENC_QKV > (size_t)SM_KS_MAX * 128 * ENC_DIM
                                  ? (size_t)SM_KS_QKV * 128 * 3 * ENC_QKV : (size_t)SM_KS_MAX * 128 * ENC_DIM;
        const size_t bytes[4] = {(size_t)128 * 2 * ENC_DIM * 2, (size_t)128 * 2 * ENC_QKV * 2, (size_t)128 * 2 * ENC_HIDDEN * 2, part_f * 4};
        void **ptrs[4] = {(void **)&q->xn_t, (void **)&q->att_t, (void **)&q->hh_t, (void **)&q->part};
        for (int i = 0; i < 4; i++) {
            if (cudaMalloc(ptrs[i], bytes[i]) != cudaSuccess || cudaMemset(*ptrs[i], 0, bytes[i]) != cudaSuccess) {
                vox_set_error("enc_small: cudaMalloc(%zu) failed", bytes[i]);
                enc_small_free(q);
                return nullptr;
            }
        }
        if (cudaMalloc((void **)&q->d_pos, 8) != cudaSuccess || cudaMallocHost((void **)&q->h_pos, 64 * 8) != cudaSuccess) {
            vox_set_error("enc_small: position buffers alloc failed");
            enc_small_free(q);
            return nullptr;
        }
        e->small = q;
    }
    const int idx = NB == 32 ? 0 : (NB == 64 ? 1 : 2);
    if (!q->ops[idx]) {
        WtOp *ops = (WtOp *)aligned_alloc(64, sizeof(WtOp) * 4 * m->n_enc);
        if (!ops) return nullptr;
        for (int l = 0; l < m->n_enc; l++) {
            const EncLayerDev &L = m->enc[l];
            WtOp *o = ops + 4 * l;
            if (gemm_wt_prepare(&o[0], NB, 1, 3 * ENC_QKV, ENC_DIM, 1, q->xn_t, 2 * ENC_DIM, L.wqkv, SM_KS_QKV) ||
                gemm_wt_prepare(&o[1], NB, 1, ENC_DIM, ENC_QKV, 1, q->att_t, 2 * ENC_QKV, L.wo, SM_KS_WO) ||
                gemm_wt_prepare(&o[2], NB, 1, 2 * ENC_HIDDEN, ENC_DIM, 1, q->xn_t, 2 * ENC_DIM, L.w13, 1) ||
                gemm_wt_prepare(&o[3], NB, 1, ENC_DIM, ENC_HIDDEN, 1, q->hh_t, 2 * ENC_HIDDEN, L.w2, SM_KS_W2)) {
                free(ops);
                return nullptr;
            }
            o[0].out_f32 = q->part; o[0].ldo = 3 * ENC_QKV;
            o[1].out_f32 = q->part; o[1].ldo = ENC_DIM;
            o[2].mode = WT_OUT_SWIGLU_BF16X2; o[2].out_bf16 = q->hh_t; o[2].ldb = 2 * ENC_HIDDEN; o[2].lo_off = ENC_HIDDEN;
            o[3].out_f32 = q->part; o[3].ldo = ENC_DIM;
            o[0].row_scale = L.s_wqkv; o[1].row_scale = L.s_wo; o[2].row_scale = L.s_w13; o[3].row_scale = L.s_w2;   // Q8 model, else null
        }
        q->ops[idx] = ops;
    }
    return q;
}

// VOX_CUDA_ENC_FUSED_ROPE=0 keeps the two-kernel form (sum + bias, then RoPE / ring append) for A/B runs
static int small_fused_rope(void) {
    static int v = -1;
    if (v < 0) { const char *e = getenv("VOX_CUDA_ENC_FUSED_ROPE"); v = (e && !atoi(e)) ? 0 : 1; }
    return v;
}

// the 32 layers for M <= 128 rows; pos_dev: device position (graph capture) or null (pos0 by value)
static int small_enqueue(vox_cuda_state *s, EncSmall *q, WtOp *ops, float *x, int M, float *out, int pos0, const int *pos_dev) {
    EncState *e = (EncState *)s->enc_priv;
    vox_cuda_model *m = s->m;
    cudaStream_t st = s->stream;
    const size_t kv_layer = (size_t)ENC_RING * ENC_QKV;
    if (rows_rmsnorm_bf16(st, x, ENC_DIM, m->enc[0].attn_norm, nullptr, q->xn_t, 2 * ENC_DIM, M, ENC_DIM, 1)) return -1;
    for (int l = 0; l < m->n_enc; l++) {
        const EncLayerDev &L = m->enc[l];
        float *kc = e->kv_k + l * kv_layer, *vc = e->kv_v + l * kv_layer;
        WtOp *o = ops + 4 * l;
        for (int i = 0; i < 4; i++) o[i].B = M;
        o[0].split_stride = (long)M * 3 * ENC_QKV; o[1].split_stride = (long)M * ENC_DIM; o[3].split_stride = (long)M * ENC_DIM;
        if (gemm_wt_launch(st, &o[0])) return -1;
        if (small_fused_rope()) {
            CUDA_TRY(launch_pdl(enc_sum_bias_rope_kernel, dim3(M, 3 * ENC_QKV / 1024), 256, 0, st, (const float *)q->part, SM_KS_QKV, (long)M * 3 * ENC_QKV,
                                L.bqkv, e->qkv, (const float *)m->inv_freq_enc, pos0, pos_dev, ENC_RING - 1, kc, vc));
            LAUNCHED();
        } else {
            CUDA_TRY(launch_pdl(enc_sum_bias_kernel, dim3(M, (3 * ENC_QKV + 1023) / 1024), 256, 0, st, (const float *)q->part, SM_KS_QKV,
                                (long)M * 3 * ENC_QKV, L.bqkv, e->qkv, 3 * ENC_QKV));
            LAUNCHED();
            if (rows_rope_kv_dev(st, e->qkv, M, ENC_QKV, ENC_QKV, ENC_HD, m->inv_freq_enc, pos0, ENC_RING - 1, kc, vc, pos_dev)) return -1;
        }
        if (enc_attention_dev(s, e->qkv, kc, vc, M, pos0, q->att_t, 2 * ENC_QKV, 1, pos_dev)) return -1;
        if (gemm_wt_launch(st, &o[1])) return -1;
        CUDA_TRY(launch_pdl(enc_resid_norm_kernel<SM_KS_WO>, M, 320, 0, st, x, (const float *)q->part, (long)M * ENC_DIM, L.bo, L.ffn_norm,
                            q->xn_t, (float *)nullptr));
        LAUNCHED();
        if (gemm_wt_launch(st, &o[2])) return -1;
        if (gemm_wt_launch(st, &o[3])) return -1;
        const bool last = l == m->n_enc - 1;
        CUDA_TRY(launch_pdl(enc_resid_norm_kernel<SM_KS_W2>, M, 320, 0, st, x, (const float *)q->part, (long)M * ENC_DIM, L.b2,
                            last ? m->enc_norm : m->enc[l + 1].attn_norm, last ? (uint16_t *)nullptr : q->xn_t, last ? out : (float *)nullptr));
        LAUNCHED();
    }
    return 0;
}

static int small_graph_mode(void) {
    static int v = -1;
    if (v < 0) { const char *e = getenv("VOX_CUDA_ENC_GRAPH"); v = (e && !atoi(e)) ? 0 : 1; }
    return v;
}

// M <= 128 new positions: same layer arithmetic as enc_layers_fast (split-bf16 operands, fp32 accumulation)
int enc_layers_small(vox_cuda_state *s, float *x, int M, float *out) {
    EncState *e = (EncState *)s->enc_priv;
    if (x1_mask()) return enc_layers_fast(s, x, M, out);       // single-bf16 products live on the large-tile path only
    if (!fast_get(s)) return -1;
    const int NB = M <= 32 ? 32 : (M <= 64 ? 64 : 128);
    EncSmall *q = small_get(s, NB);
    if (!q) return -1;
    WtOp *ops = q->ops[NB == 32 ? 0 : (NB == 64 ? 1 : 2)];
    cudaStream_t st = s->stream;
    const int pos0 = e->pos;
    int slot = -1;
    if (small_graph_mode()) {
        for (int i = 0; i < SM_GRAPHS; i++)
            if (q->graphs[i].seen && q->graphs[i].M == M && q->graphs[i].x == x && q->graphs[i].out == out) { slot = i; break; }
        if (slot < 0)
            for (int i = 0; i < SM_GRAPHS; i++)
                if (!q->graphs[i].seen) { q->graphs[i].M = M; q->graphs[i].x = x; q->graphs[i].out = out; q->graphs[i].seen = 1; break; }
        // first sighting of a shape runs eagerly (slot stays -1); the second one is captured
    }
    if (slot >= 0) {
        int *hp = q->h_pos + 2 * (q->h_pos_i++ & 63);
        hp[0] = pos0; hp[1] = e->pos_min;
        CUDA_TRY(cudaMemcpyAsync(q->d_pos, hp, 8, cudaMemcpyHostToDevice, st));
        if (!q->graphs[slot].exec) {
            cudaGraph_t graph = nullptr;
            CUDA_TRY(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
            const int rc = small_enqueue(s, q, ops, x, M, out, 0, q->d_pos);
            const cudaError_t ce = cudaStreamEndCapture(st, &graph);
            if (rc || ce != cudaSuccess || !graph) {
                if (graph) cudaGraphDestroy(graph);
                vox_set_error("enc_small: graph capture failed (%s)", cudaGetErrorString(ce));
                return -1;
            }
            const cudaError_t ie = cudaGraphInstantiate(&q->graphs[slot].exec, graph, 0);
            cudaGraphDestroy(graph);
            if (ie != cudaSuccess) { vox_set_error("enc_small: graph instantiate failed (%s)", cudaGetErrorString(ie)); return -1; }
        }
        CUDA_TRY(cudaGraphLaunch(q->graphs[slot].exec, st));
        LAUNCHED();
    } else if (small_enqueue(s, q, ops, x, M, out, pos0, nullptr)) return -1;
    e->pos += M;
    return 0;
}
