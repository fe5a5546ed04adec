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

void vox_mem_account(long delta);

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

static TcEpilogue epi(float *of, long ldf, uint16_t *ob, long ldb, const float *bias, int flags, long split_off = 0) {
    TcEpilogue t;
    t.out_f32 = of; t.ldo_f32 = ldf; t.out_bf16 = ob; t.ldo_bf16 = ldb; t.bias = bias; t.flags = flags;
    t.split_off = split_off;
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
    const int sp = split_mode(), mul = sp ? 2 : 1;
    const int osp = sp ? TC_OUT_SPLIT : 0;
    for (int l = 0; l < m->n_enc; l++) {
        const EncLayerDev &L = m->enc[l];
        float *kc = e->kv_k + l * kv_layer, *vc = e->kv_v + l * kv_layer;
        if (rows_rmsnorm_bf16(st, x, ENC_DIM, L.attn_norm, nullptr, f->xn_b, mul * ENC_DIM, M, ENC_DIM, sp)) return -1;
        if (gemm_tc(st, M, 3 * ENC_QKV, ENC_DIM, mul * ENC_DIM, f->xn_b, mul * ENC_DIM, L.wqkv,
                    epi(e->qkv, 3 * ENC_QKV, nullptr, 0, L.bqkv, TC_BIAS | TC_OUT_F32))) return -1;
        if (rows_rope_kv(st, e->qkv, M, ENC_QKV, ENC_QKV, ENC_HD, m->inv_freq_enc, pos0, ENC_RING - 1, kc, vc)) return -1;
        if (enc_attention(s, e->qkv, kc, vc, M, pos0, f->att_b, mul * ENC_QKV, sp)) return -1;
        if (gemm_tc(st, M, ENC_DIM, ENC_QKV, mul * ENC_QKV, f->att_b, mul * ENC_QKV, L.wo,
                    epi(x, ENC_DIM, nullptr, 0, L.bo, TC_BIAS | TC_RESIDUAL))) return -1;
        if (rows_rmsnorm_bf16(st, x, ENC_DIM, L.ffn_norm, nullptr, f->xn_b, mul * ENC_DIM, M, ENC_DIM, sp)) return -1;
        if (gemm_tc(st, M, 2 * ENC_HIDDEN, ENC_DIM, mul * ENC_DIM, f->xn_b, mul * ENC_DIM, L.w13,
                    epi(nullptr, 0, f->hh_b, mul * ENC_HIDDEN, nullptr, TC_SWIGLU | TC_OUT_BF16 | osp, ENC_HIDDEN))) return -1;
        if (gemm_tc(st, M, ENC_DIM, ENC_HIDDEN, mul * ENC_HIDDEN, f->hh_b, mul * ENC_HIDDEN, L.w2,
                    epi(x, ENC_DIM, nullptr, 0, L.b2, TC_BIAS | TC_RESIDUAL))) return -1;
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
    const int sp = split_mode(), mul = sp ? 2 : 1;
    const int K0 = 4 * ENC_DIM;
    if (f32_to_bf16_rows(st, enc_rows, K0, f->enc_b, mul * K0, k_tokens, K0, sp)) return -1;
    if (gemm_tc(st, k_tokens, DEC_DIM, K0, mul * K0, f->enc_b, mul * K0, m->adapter_w0,
                epi(nullptr, 0, f->mid_b, mul * DEC_DIM, nullptr, TC_GELU | TC_OUT_BF16 | (sp ? TC_OUT_SPLIT : 0), DEC_DIM))) return -1;
    return gemm_tc(st, k_tokens, DEC_DIM, DEC_DIM, mul * DEC_DIM, f->mid_b, mul * DEC_DIM, m->adapter_w1,
                   epi(out, DEC_DIM, nullptr, 0, nullptr, TC_OUT_F32));
}
