/*
 * oracle/ref_shim.c - accessors compiled INTO the reference build
 * (oracle/_ref/libvoxref.so).  TEST INFRASTRUCTURE ONLY: nothing under
 * voxtral.c_b200/ may link or load this.
 *
 * It includes the reference's own voxtral.h from /root/reference (never
 * copied into this repo) so that Python/ctypes harnesses can read and write
 * the public vox_ctx_t fields (voxtral.h:188-239) without knowing the struct
 * layout, and it contains the weight-table binding a maintainer would add to
 * vox_load() to hand the mmap'd pointers to the CUDA backend (see
 * INTEGRATION.md, "vox_load hook").
 */
#include "voxtral.h"
#include "voxtral_audio.h"
#include "vox_cuda.h"
#include <string.h>
#include <stdlib.h>

extern int vox_verbose;

void refshim_set_verbose(int v) { vox_verbose = v; }

int refshim_sizeof_ctx(void) { return (int)sizeof(vox_ctx_t); }

/* ---- integer fields ---------------------------------------------------- */
int refshim_get_int(vox_ctx_t *c, const char *f) {
    if (!strcmp(f, "kv_cache_len")) return c->kv_cache_len;
    if (!strcmp(f, "kv_cache_max")) return c->kv_cache_max;
    if (!strcmp(f, "kv_pos_offset")) return c->kv_pos_offset;
    if (!strcmp(f, "enc_kv_cache_len")) return c->enc_kv_cache_len;
    if (!strcmp(f, "enc_kv_cache_max")) return c->enc_kv_cache_max;
    if (!strcmp(f, "enc_kv_pos_offset")) return c->enc_kv_pos_offset;
    if (!strcmp(f, "delay_tokens")) return c->delay_tokens;
    if (!strcmp(f, "use_bf16")) return c->use_bf16;
    return -999999;
}

int refshim_set_int(vox_ctx_t *c, const char *f, int v) {
    if (!strcmp(f, "kv_cache_len")) { c->kv_cache_len = v; return 0; }
    if (!strcmp(f, "kv_pos_offset")) { c->kv_pos_offset = v; return 0; }
    if (!strcmp(f, "enc_kv_cache_len")) { c->enc_kv_cache_len = v; return 0; }
    if (!strcmp(f, "enc_kv_pos_offset")) { c->enc_kv_pos_offset = v; return 0; }
    return -1;
}

/* ---- pointer fields ---------------------------------------------------- */
float *refshim_kv_k(vox_ctx_t *c) { return c->kv_cache_k; }
float *refshim_kv_v(vox_ctx_t *c) { return c->kv_cache_v; }
float *refshim_enc_kv_k(vox_ctx_t *c) { return c->enc_kv_cache_k; }
float *refshim_enc_kv_v(vox_ctx_t *c) { return c->enc_kv_cache_v; }
const float *refshim_ada_scale(vox_ctx_t *c) { return c->ada_scale; }
const uint16_t *refshim_tok_embeddings(vox_ctx_t *c) { return c->decoder.tok_embeddings_bf16; }

/* Limit the number of layers the reference walks is impossible (compile-time
 * constants, voxtral.h:28,41), so reduced-depth parity cases are served by the
 * restatement in oracle/vox_oracle.c instead. */

/* ---- the vox_load hook: reference ctx -> CUDA weight table ------------- */
void refshim_fill_weights(const vox_ctx_t *c, vox_cuda_weights_t *w) {
    memset(w, 0, sizeof(*w));
    w->n_enc_layers = VOX_ENC_LAYERS;
    w->n_dec_layers = VOX_DEC_LAYERS;
    w->conv0_weight = c->encoder.conv0_weight;
    w->conv0_bias = c->encoder.conv0_bias;
    w->conv1_weight = c->encoder.conv1_weight;
    w->conv1_bias = c->encoder.conv1_bias;
    for (int i = 0; i < VOX_ENC_LAYERS; i++) {
        const vox_enc_layer_t *l = &c->encoder.layers[i];
        vox_cuda_enc_layer_w_t *d = &w->enc[i];
        d->wq = l->wq_weight_bf16; d->wk = l->wk_weight_bf16;
        d->wv = l->wv_weight_bf16; d->wo = l->wo_weight_bf16;
        d->w1 = l->w1_weight_bf16; d->w2 = l->w2_weight_bf16;
        d->w3 = l->w3_weight_bf16;
        d->wq_bias = l->wq_bias; d->wv_bias = l->wv_bias;
        d->wo_bias = l->wo_bias; d->w2_bias = l->w2_bias;
        d->attention_norm = l->attention_norm; d->ffn_norm = l->ffn_norm;
    }
    w->enc_norm = c->encoder.norm;
    w->adapter_w0 = c->adapter.linear0_weight_bf16;
    w->adapter_w1 = c->adapter.linear1_weight_bf16;
    w->tok_embeddings = c->decoder.tok_embeddings_bf16;
    for (int i = 0; i < VOX_DEC_LAYERS; i++) {
        const vox_dec_layer_t *l = &c->decoder.layers[i];
        vox_cuda_dec_layer_w_t *d = &w->dec[i];
        d->wq = l->wq_weight_bf16; d->wk = l->wk_weight_bf16;
        d->wv = l->wv_weight_bf16; d->wo = l->wo_weight_bf16;
        d->w1 = l->w1_weight_bf16; d->w2 = l->w2_weight_bf16;
        d->w3 = l->w3_weight_bf16;
        d->attention_norm = l->attention_norm; d->ffn_norm = l->ffn_norm;
    }
    w->dec_norm = c->decoder.norm;
    w->ada_scale = c->ada_scale;
}

/* Q8 checkpoint (ctx->use_q8): the *_weight_q8 / *_scale_q8 pairs of the reference's tables (voxtral.h:56-182)
 * -> vox_cuda_weights_q8_t.  Returns ctx->use_q8. */
int refshim_fill_weights_q8(const vox_ctx_t *c, vox_cuda_weights_q8_t *q) {
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
    return c->use_q8;
}
const int8_t *refshim_tok_embeddings_q8(vox_ctx_t *c) { return c->decoder.tok_embeddings_q8; }
const float *refshim_tok_embeddings_scale_q8(vox_ctx_t *c) { return c->decoder.tok_embeddings_scale_q8; }

int refshim_sizeof_weights(void) { return (int)sizeof(vox_cuda_weights_t); }
int refshim_sizeof_weights_q8(void) { return (int)sizeof(vox_cuda_weights_q8_t); }

/* ---- mel helpers (the incremental ctx is opaque in voxtral_audio.h) ---- */
/* Runs the reference streaming mel exactly as vox_stream_* drives it
 * (voxtral.c:1255,1291,1645-1658,1306): left pad 32*1280, feed, flush
 * padding, finish.  Returns malloc'd [frames,128]. */
float *refshim_stream_mel(const float *samples, int n, int delay_tokens,
                          int chunk, int *out_frames) {
    vox_mel_ctx_t *m = vox_mel_ctx_init(32 * 1280);
    if (!m) return NULL;
    if (chunk <= 0) chunk = n > 0 ? n : 1;
    for (int off = 0; off < n; off += chunk) {
        int c = n - off < chunk ? n - off : chunk;
        vox_mel_feed(m, samples + off, c);
    }
    int align_pad = (1280 - (n % 1280)) % 1280;
    int right_pad = align_pad + (delay_tokens + 1 + 10) * 1280;
    float *z = (float *)calloc((size_t)right_pad + 1, sizeof(float));
    vox_mel_feed(m, z, right_pad);
    free(z);
    vox_mel_finish(m, 0);
    int nf = 0;
    float *d = vox_mel_data(m, &nf);
    float *out = (float *)malloc((size_t)(nf > 0 ? nf : 1) * 128 * sizeof(float));
    memcpy(out, d, (size_t)nf * 128 * sizeof(float));
    vox_mel_free(m);
    *out_frames = nf;
    return out;
}

void refshim_free(void *p) { free(p); }

/* ---- call log of the reference's OWN stream path ------------------------
 * libvoxref.so is linked with -Wl,--wrap=vox_decoder_forward
 * -Wl,--wrap=vox_encoder_forward_incremental (oracle/build_ref.sh), so the calls
 * that stream_run_decoder / stream_run_encoder (voxtral.c:1012-1113, 826-951) make
 * are routed through the two recorders below.  They call the real function and
 * keep what it returned: the greedy token id with the top-2 logit margin, and the
 * row count of every encoder call.  The arithmetic is untouched; the harness
 * calling vox_decoder_forward directly still reaches the real symbol. */
int __real_vox_decoder_forward(vox_ctx_t *ctx, const float *input_embeds, float *logits);
float *__real_vox_encoder_forward_incremental(vox_ctx_t *ctx, const float *x_new, int new_len, int *out_len);

#define REFSHIM_LOG_MAX (1 << 16)
static int g_log_tok[REFSHIM_LOG_MAX];
static float g_log_margin[REFSHIM_LOG_MAX];
static int g_log_n;
static float g_log_top[REFSHIM_LOG_MAX][5];   /* logit of the returned id, then the 4 best logits among ids >= 1000 other than it */
static int g_enc_rows[REFSHIM_LOG_MAX];
static int g_enc_n;

int __wrap_vox_decoder_forward(vox_ctx_t *ctx, const float *input_embeds, float *logits) {
    int tok = __real_vox_decoder_forward(ctx, input_embeds, logits);
    if (g_log_n < REFSHIM_LOG_MAX) {
        float m = -1.0f;
        if (logits && tok >= 0) {
            float second = -3.0e38f;
            for (int i = 0; i < VOX_VOCAB_SIZE; i++)
                if (i != tok && logits[i] > second) second = logits[i];
            m = logits[tok] - second;
        }
        float *top = g_log_top[g_log_n];
        top[0] = top[1] = top[2] = top[3] = top[4] = -3.0e38f;
        if (logits && tok >= 0) {
            top[0] = logits[tok];
            for (int i = 1000; i < VOX_VOCAB_SIZE; i++) {
                const float v = logits[i];
                if (i == tok || !(v > top[4])) continue;
                int k = 4;
                while (k > 1 && v > top[k - 1]) { top[k] = top[k - 1]; k--; }
                top[k] = v;
            }
        }
        g_log_tok[g_log_n] = tok;
        g_log_margin[g_log_n] = m;
        g_log_n++;
    }
    return tok;
}

float *__wrap_vox_encoder_forward_incremental(vox_ctx_t *ctx, const float *x_new, int new_len, int *out_len) {
    if (g_enc_n < REFSHIM_LOG_MAX) g_enc_rows[g_enc_n++] = new_len;
    return __real_vox_encoder_forward_incremental(ctx, x_new, new_len, out_len);
}

void refshim_log_reset(void) { g_log_n = 0; g_enc_n = 0; }
int refshim_log_count(void) { return g_log_n; }
const int *refshim_log_tokens(void) { return g_log_tok; }
const float *refshim_log_margins(void) { return g_log_margin; }
const float *refshim_log_top(void) { return &g_log_top[0][0]; }
int refshim_enc_call_count(void) { return g_enc_n; }
const int *refshim_enc_call_rows(void) { return g_enc_rows; }
