/*
 * vox_cuda_dropin.c - glue between the reference's host (built with -DUSE_CUDA from
 * integration/use_cuda.patch) and libvoxcuda.so.  Compiled together with the patched reference
 * sources by `make -C voxtral.c_b200 dropin`; includes the reference's own voxtral.h from
 * /root/reference (never copied into this repository).
 *
 * One vox_ctx_t <-> one (model, state) pair, kept in a small table keyed by the ctx pointer - the
 * reference is single-threaded and holds one stream per ctx (voxtral.h:188-239).
 */
#include "voxtral.h"
#include "vox_cuda_dropin.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

extern int vox_verbose;

typedef struct {
    vox_ctx_t *ctx;
    vox_cuda_model_t *model;
    vox_cuda_state_t *state;
} slot_t;
static slot_t g_slots[8];

static slot_t *find(vox_ctx_t *ctx) {
    for (int i = 0; i < 8; i++) if (g_slots[i].ctx == ctx && ctx) return &g_slots[i];
    return NULL;
}

static void complain(const char *what) {
    fprintf(stderr, "voxtral (CUDA): %s failed: %s\n", what, vox_cuda_last_error());
}

/* the pointer fields of vox_ctx_t, as the loaders left them (voxtral_encoder.c:58-146,
 * voxtral_decoder.c:57-145, voxtral.c:102-125) */
static void fill_weights(const vox_ctx_t *c, vox_cuda_weights_t *w) {
    memset(w, 0, sizeof(*w));
    w->n_enc_layers = VOX_ENC_LAYERS;
    w->n_dec_layers = VOX_DEC_LAYERS;
    w->conv0_weight = c->encoder.conv0_weight; w->conv0_bias = c->encoder.conv0_bias;
    w->conv1_weight = c->encoder.conv1_weight; w->conv1_bias = c->encoder.conv1_bias;
    for (int i = 0; i < VOX_ENC_LAYERS; i++) {
        const vox_enc_layer_t *l = &c->encoder.layers[i];
        vox_cuda_enc_layer_w_t *d = &w->enc[i];
        d->wq = l->wq_weight_bf16; d->wk = l->wk_weight_bf16; d->wv = l->wv_weight_bf16; d->wo = l->wo_weight_bf16;
        d->w1 = l->w1_weight_bf16; d->w2 = l->w2_weight_bf16; d->w3 = l->w3_weight_bf16;
        d->wq_bias = l->wq_bias; d->wv_bias = l->wv_bias; d->wo_bias = l->wo_bias; d->w2_bias = l->w2_bias;
        d->attention_norm = l->attention_norm; d->ffn_norm = l->ffn_norm;
    }
    w->enc_norm = c->encoder.norm;
    w->adapter_w0 = c->adapter.linear0_weight_bf16;
    w->adapter_w1 = c->adapter.linear1_weight_bf16;
    w->tok_embeddings = c->decoder.tok_embeddings_bf16;
    for (int i = 0; i < VOX_DEC_LAYERS; i++) {
        const vox_dec_layer_t *l = &c->decoder.layers[i];
        vox_cuda_dec_layer_w_t *d = &w->dec[i];
        d->wq = l->wq_weight_bf16; d->wk = l->wk_weight_bf16; d->wv = l->wv_weight_bf16; d->wo = l->wo_weight_bf16;
        d->w1 = l->w1_weight_bf16; d->w2 = l->w2_weight_bf16; d->w3 = l->w3_weight_bf16;
        d->attention_norm = l->attention_norm; d->ffn_norm = l->ffn_norm;
    }
    w->dec_norm = c->decoder.norm;
    w->ada_scale = c->ada_scale;
}

/* Q8 checkpoint (ctx->use_q8, voxtral.c:175-181): the *_weight_q8 / *_scale_q8 pairs the loaders took straight from
 * the mmap (voxtral_encoder.c:94-110, voxtral_decoder.c:109-125, voxtral.c:113-120) */
static void fill_weights_q8(const vox_ctx_t *c, vox_cuda_weights_q8_t *q) {
    memset(q, 0, sizeof(*q));
#define Q8BIND(dst, src) do { (dst).q = (src##_weight_q8); (dst).scale = (src##_scale_q8); } while (0)
    for (int i = 0; i < VOX_ENC_LAYERS; i++) {
        const vox_enc_layer_t *l = &c->encoder.layers[i];
        vox_cuda_layer_q8_t *d = &q->enc[i];
        Q8BIND(d->wq, l->wq); Q8BIND(d->wk, l->wk); Q8BIND(d->wv, l->wv); Q8BIND(d->wo, l->wo);
        Q8BIND(d->w1, l->w1); Q8BIND(d->w2, l->w2); Q8BIND(d->w3, l->w3);
    }
    Q8BIND(q->adapter_w0, c->adapter.linear0); Q8BIND(q->adapter_w1, c->adapter.linear1);
    q->tok_embeddings.q = c->decoder.tok_embeddings_q8;
    q->tok_embeddings.scale = c->decoder.tok_embeddings_scale_q8;
    for (int i = 0; i < VOX_DEC_LAYERS; i++) {
        const vox_dec_layer_t *l = &c->decoder.layers[i];
        vox_cuda_layer_q8_t *d = &q->dec[i];
        Q8BIND(d->wq, l->wq); Q8BIND(d->wk, l->wk); Q8BIND(d->wv, l->wv); Q8BIND(d->wo, l->wo);
        Q8BIND(d->w1, l->w1); Q8BIND(d->w2, l->w2); Q8BIND(d->w3, l->w3);
    }
#undef Q8BIND
}

int vox_cuda_dropin_load(vox_ctx_t *ctx) {
    if (!vox_cuda_init()) { complain("vox_cuda_init"); return -1; }
    slot_t *s = find(NULL);
    for (int i = 0; i < 8 && !s; i++) if (!g_slots[i].ctx) s = &g_slots[i];
    if (!s) { fprintf(stderr, "voxtral (CUDA): too many contexts\n"); return -1; }
    vox_cuda_weights_t *w = (vox_cuda_weights_t *)malloc(sizeof(*w));
    if (!w) return -1;
    fill_weights(ctx, w);
    if (ctx->use_q8) {
        vox_cuda_weights_q8_t *q = (vox_cuda_weights_q8_t *)malloc(sizeof(*q));
        if (!q) { free(w); return -1; }
        fill_weights_q8(ctx, q);
        s->model = vox_cuda_model_create_q8(w, q);
        free(q);
    } else {
        s->model = vox_cuda_model_create(w);
    }
    free(w);
    if (!s->model) { complain(ctx->use_q8 ? "vox_cuda_model_create_q8" : "vox_cuda_model_create"); return -1; }
    s->state = vox_cuda_state_create(s->model);
    if (!s->state) { complain("vox_cuda_state_create"); vox_cuda_model_free(s->model); s->model = NULL; return -1; }
    s->ctx = ctx;
    if (vox_verbose >= 1)
        fprintf(stderr, "CUDA GPU: %.1f MB\n", vox_cuda_memory_used() / (1024.0 * 1024.0));
    return 0;
}

void vox_cuda_dropin_free(vox_ctx_t *ctx) {
    slot_t *s = find(ctx);
    if (!s) return;
    vox_cuda_state_free(s->state);
    vox_cuda_model_free(s->model);
    memset(s, 0, sizeof(*s));
}

void vox_cuda_dropin_set_delay(vox_ctx_t *ctx) {
    slot_t *s = find(ctx);
    if (s && vox_cuda_model_set_ada_scale(s->model, ctx->ada_scale)) complain("vox_cuda_model_set_ada_scale");
}

/* the host's integers are the source of truth: whatever it reset or compacted since the last call
 * (voxtral.c:781-782,819-820,1050-1051; voxtral_decoder.c:354-385) reaches the device as a position */
static int sync_decoder(slot_t *s, const vox_ctx_t *ctx) {
    const int logical = ctx->kv_pos_offset + ctx->kv_cache_len;
    return vox_cuda_decoder_set_position(s->state, logical, ctx->kv_cache_len);
}

int vox_cuda_dropin_decoder_forward(vox_ctx_t *ctx, const float *embed, float *logits) {
    slot_t *s = find(ctx);
    if (!s || sync_decoder(s, ctx)) { complain("decoder position"); return 2; }
    const int tok = vox_cuda_decoder_step(s->state, embed, logits);
    if (tok < 0) { complain("vox_cuda_decoder_step"); return 2; }
    ctx->kv_cache_len++;
    return tok;
}

void vox_cuda_dropin_decoder_prefill(vox_ctx_t *ctx, const float *x, int seq_len) {
    slot_t *s = find(ctx);
    if (!s || sync_decoder(s, ctx) || vox_cuda_decoder_prefill(s->state, x, seq_len)) {
        complain("vox_cuda_decoder_prefill");
        return;
    }
    ctx->kv_cache_len += seq_len;
}

int vox_cuda_dropin_encoder_step(vox_ctx_t *ctx, float *x, int new_len, int cache_len) {
    slot_t *s = find(ctx);
    if (!s) return -1;
    if (vox_cuda_encoder_set_position(s->state, ctx->enc_kv_pos_offset + cache_len) ||
        vox_cuda_encoder_step(s->state, x, new_len)) {
        complain("vox_cuda_encoder_step");
        return -1;
    }
    return 0;
}

float *vox_cuda_dropin_adapter(vox_ctx_t *ctx, const float *enc_out, int enc_len, int *out_len) {
    slot_t *s = find(ctx);
    const int k = enc_len / VOX_DOWNSAMPLE;
    *out_len = 0;
    if (!s || k <= 0) return NULL;
    float *out = (float *)malloc((size_t)k * VOX_DEC_DIM * sizeof(float));
    if (!out) return NULL;
    if (vox_cuda_adapter(s->state, enc_out, enc_len, out)) { complain("vox_cuda_adapter"); free(out); return NULL; }
    *out_len = k;
    return out;
}

float *vox_cuda_dropin_conv_stem(vox_ctx_t *ctx, int first, const float *mel, int n_mel, int *out_len) {
    slot_t *s = find(ctx);
    *out_len = 0;
    if (!s || n_mel <= 0) return NULL;
    if (first) vox_cuda_state_reset_stem(s->state);
    float *out = (float *)malloc((size_t)(n_mel / 2 + 2) * VOX_ENC_DIM * sizeof(float));
    if (!out) return NULL;
    int n = 0;
    if (vox_cuda_conv_stem(s->state, mel, n_mel, out, &n)) { complain("vox_cuda_conv_stem"); free(out); return NULL; }
    if (n <= 0) { free(out); return NULL; }
    *out_len = n;
    return out;
}

int vox_cuda_dropin_sgemm_bf16(int M, int N, int K, const float *A, const uint16_t *B, float *C) {
    vox_cuda_model_t *m = NULL;
    for (int i = 0; i < 8; i++) if (g_slots[i].ctx) { m = g_slots[i].model; break; }
    if (vox_cuda_sgemm_bf16(m, M, N, K, A, B, C)) { complain("vox_cuda_sgemm_bf16"); return 0; }
    return 1;
}

/* The reference calls cblas_sgemm unconditionally from vox_causal_conv1d (voxtral_kernels.c:451-457) and
 * from its CPU matmuls.  None of them is on a path this build leaves on the host; reaching one means a hook
 * is missing, and the CUDA build has no CPU fallback - say so and stop. */
void cblas_sgemm(int order, int ta, int tb, int M, int N, int K, float alpha, const float *A, int lda,
                 const float *B, int ldb, float beta, float *C, int ldc) {
    (void)order; (void)ta; (void)tb; (void)alpha; (void)A; (void)lda; (void)B; (void)ldb; (void)beta; (void)C; (void)ldc;
    fprintf(stderr, "voxtral (CUDA): CPU sgemm (%d x %d x %d) reached - this build has no CPU fallback\n", M, N, K);
    abort();
}
