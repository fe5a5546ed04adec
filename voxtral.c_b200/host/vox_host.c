#define _GNU_SOURCE
/*
 * vox_host.c - C host over libvoxcuda.so.  Mirrors, for the hot path only, what the reference's
 * voxtral.c does around its backends: vox_load (voxtral.c:131-290: mmap the safetensors file,
 * point at the bf16 matrices, widen the small tensors, precompute ada_scale), the vox_stream_*
 * state machine (voxtral.c:826-1145, 1242-1317, 1640-1667: mel -> encoder chunk -> adapter ->
 * prefill -> greedy steps) and vox_transcribe_audio (voxtral.c:1390-1396).  Written from the
 * behaviour, not from the source: all arithmetic lives in the CUDA backend, this file only
 * moves pointers, pads audio and keeps counters.
 */
#include "vox_host.h"
#include "vox_cuda.h"

#include <fcntl.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/time.h>
#include <unistd.h>

#define HOP 160
#define WIN 400
#define TOKEN_SAMPLES 1280             /* 80 ms of audio per adapter token */
#define LEFT_PAD_TOKENS 32
#define ENC_PREFIX "mm_streams_embeddings.embedding_module.whisper_encoder"
#define EMB_PREFIX "mm_streams_embeddings.embedding_module"

struct vox_b200_ctx {
    int fd;
    uint8_t *map;
    size_t map_bytes;
    size_t map_total;       /* length passed to mmap (map_bytes is later reduced to the data section) */
    const char *json;                  /* safetensors header (not NUL terminated) */
    size_t json_len;
    const uint8_t *data;               /* first byte after the header */
    vox_cuda_weights_t w;
    float *small[1024];                /* widened small tensors, freed in vox_b200_free */
    int n_small;
    float *ada_scale;
    int delay_tokens;
    vox_cuda_model_t *model;
    vox_cuda_state_t *state;           /* one stream per context, like the reference */
};

static double now_ms(void) {
    struct timeval t;
    gettimeofday(&t, NULL);
    return t.tv_sec * 1e3 + t.tv_usec * 1e-3;
}

/* ---- safetensors: u64 header length, JSON header, raw tensors ---------------------------- */
/* Finds "name":{... "data_offsets":[a,b] ...}; the header is flat, so a scan is enough. */
static int st_find(const vox_b200_ctx *c, const char *name, size_t *a, size_t *b) {
    char key[320];
    const int klen = snprintf(key, sizeof(key), "\"%s\"", name);
    if (klen <= 0 || (size_t)klen >= sizeof(key)) return -1;
    const char *p = c->json, *end = c->json + c->json_len;
    while (p + klen < end) {
        p = (const char *)memmem(p, (size_t)(end - p), key, (size_t)klen);
        if (!p) return -1;
        const char *q = p + klen;
        while (q < end && (*q == ' ' || *q == '\t')) q++;
        if (q < end && *q == ':') {
            static const char off[] = "\"data_offsets\"";
            const char *o = (const char *)memmem(q, (size_t)(end - q), off, sizeof(off) - 1);
            const char *close = (const char *)memchr(q, '}', (size_t)(end - q));
            if (!o || !close || o > close) return -1;
            o = (const char *)memchr(o, '[', (size_t)(end - o));
            if (!o) return -1;
            char *e1 = NULL, *e2 = NULL;
            *a = (size_t)strtoull(o + 1, &e1, 10);
            if (!e1 || *e1 != ',') return -1;
            *b = (size_t)strtoull(e1 + 1, &e2, 10);
            return (*b >= *a && *b <= c->map_bytes) ? 0 : -1;
        }
        p += klen;
    }
    return -1;
}

static const uint16_t *bf16_ptr(vox_b200_ctx *c, const char *name, size_t expect_elems) {
    size_t a, b;
    if (st_find(c, name, &a, &b)) { fprintf(stderr, "vox_b200_load: tensor %s not found\n", name); return NULL; }
    if (expect_elems && (b - a) != expect_elems * 2) {
        fprintf(stderr, "vox_b200_load: tensor %s has %zu bytes, expected %zu\n", name, b - a, expect_elems * 2);
        return NULL;
    }
    return (const uint16_t *)(c->data + a);
}

/* small tensors are stored bf16 and used as f32 (voxtral_encoder.c:32-39) */
static const float *f32_copy(vox_b200_ctx *c, const char *name, size_t elems) {
    const uint16_t *src = bf16_ptr(c, name, elems);
    if (!src || c->n_small >= (int)(sizeof(c->small) / sizeof(c->small[0]))) return NULL;
    float *dst = (float *)malloc(elems * sizeof(float));
    if (!dst) return NULL;
    for (size_t i = 0; i < elems; i++) {
        const uint32_t u = (uint32_t)src[i] << 16;
        memcpy(&dst[i], &u, 4);
    }
    c->small[c->n_small++] = dst;
    return dst;
}

static float gelu_tanh(float x) {
    const float inner = 0.7978845608028654f * (x + 0.044715f * x * x * x);
    return 0.5f * x * (1.0f + tanhf(inner));
}

/* ada_scale[l] = up_l . gelu(down_l . t_cond), t_cond = [cos, sin](delay * 10000^(-i/1536))
 * (what vox_update_time_conditioning produces, voxtral.c:31-80) */
static int compute_ada_scale(vox_b200_ctx *c) {
    const int dim = VOX_CUDA_DEC_DIM, half = dim / 2, r = 32, n = c->w.n_dec_layers;
    float *t_cond = (float *)malloc(sizeof(float) * dim);
    if (!c->ada_scale) c->ada_scale = (float *)malloc(sizeof(float) * (size_t)n * dim);
    if (!t_cond || !c->ada_scale) { free(t_cond); return -1; }
    const float log_theta = logf(10000.0f);
    for (int i = 0; i < half; i++) {
        const float e = (float)c->delay_tokens * expf(-log_theta * (float)i / (float)half);
        t_cond[i] = cosf(e);
        t_cond[i + half] = sinf(e);
    }
    for (int l = 0; l < n; l++) {
        char name[128];
        snprintf(name, sizeof(name), "layers.%d.ada_rms_norm_t_cond.0.weight", l);
        const uint16_t *down = bf16_ptr(c, name, (size_t)r * dim);
        snprintf(name, sizeof(name), "layers.%d.ada_rms_norm_t_cond.2.weight", l);
        const uint16_t *up = bf16_ptr(c, name, (size_t)dim * r);
        if (!down || !up) { free(t_cond); return -1; }
        float hidden[32];
        for (int i = 0; i < r; i++) {
            float s = 0.f;
            for (int j = 0; j < dim; j++) {
                const uint32_t u = (uint32_t)down[(size_t)i * dim + j] << 16;
                float wv;
                memcpy(&wv, &u, 4);
                s += wv * t_cond[j];
            }
            hidden[i] = gelu_tanh(s);
        }
        for (int i = 0; i < dim; i++) {
            float s = 0.f;
            for (int j = 0; j < r; j++) {
                const uint32_t u = (uint32_t)up[(size_t)i * r + j] << 16;
                float wv;
                memcpy(&wv, &u, 4);
                s += wv * hidden[j];
            }
            c->ada_scale[(size_t)l * dim + i] = s;
        }
    }
    free(t_cond);
    c->w.ada_scale = c->ada_scale;
    return 0;
}

static int count_layers(const vox_b200_ctx *c, const char *fmt, int max) {
    int n = 0;
    for (; n < max; n++) {
        char name[256];
        size_t a, b;
        snprintf(name, sizeof(name), fmt, n);
        if (st_find(c, name, &a, &b)) break;
    }
    return n;
}

#define NEED(x) do { if (!(x)) goto fail; } while (0)

vox_b200_ctx *vox_b200_load(const char *model_dir) {
    if (!vox_cuda_init()) { fprintf(stderr, "vox_b200_load: %s\n", vox_cuda_last_error()); return NULL; }
    vox_b200_ctx *c = (vox_b200_ctx *)calloc(1, sizeof(*c));
    if (!c) return NULL;
    c->fd = -1;
    c->delay_tokens = 6;
    char path[1024];
    snprintf(path, sizeof(path), "%s/consolidated.safetensors", model_dir);
    c->fd = open(path, O_RDONLY);
    struct stat st;
    if (c->fd < 0 || fstat(c->fd, &st)) { fprintf(stderr, "vox_b200_load: cannot open %s\n", path); goto fail; }
    c->map_bytes = c->map_total = (size_t)st.st_size;
    c->map = (uint8_t *)mmap(NULL, c->map_bytes, PROT_READ, MAP_PRIVATE, c->fd, 0);
    if (c->map == MAP_FAILED) { c->map = NULL; fprintf(stderr, "vox_b200_load: mmap failed\n"); goto fail; }
    uint64_t hlen;
    memcpy(&hlen, c->map, 8);
    if (hlen + 8 > c->map_bytes) { fprintf(stderr, "vox_b200_load: bad safetensors header\n"); goto fail; }
    c->json = (const char *)c->map + 8;
    c->json_len = (size_t)hlen;
    c->data = c->map + 8 + hlen;
    c->map_bytes -= 8 + (size_t)hlen;              /* offsets are relative to the data section */

    if (memmem(c->json, c->json_len, "\"dtype\":\"Q8\"", 12)) {
        /* Q8 containers reach the backend through vox_cuda_model_create_q8 - bound by the drop-in host
         * (dropin/vox_cuda_dropin.c, the reference's own loader) and by hostmodel.py; this small host reads bf16 only */
        fprintf(stderr, "vox_b200_load: %s is a Q8 container - use voxtral_cuda (make dropin) for Q8 checkpoints\n", path);
        goto fail;
    }
    vox_cuda_weights_t *w = &c->w;
    w->n_enc_layers = count_layers(c, ENC_PREFIX ".transformer.layers.%d.attention.wq.weight", VOX_CUDA_ENC_LAYERS);
    w->n_dec_layers = count_layers(c, "layers.%d.attention.wq.weight", VOX_CUDA_DEC_LAYERS);
    if (w->n_enc_layers < 1 || w->n_dec_layers < 1) { fprintf(stderr, "vox_b200_load: no layers found\n"); goto fail; }
    NEED(w->conv0_weight = f32_copy(c, ENC_PREFIX ".conv_layers.0.conv.weight", 1280u * 128 * 3));
    NEED(w->conv0_bias = f32_copy(c, ENC_PREFIX ".conv_layers.0.conv.bias", 1280));
    NEED(w->conv1_weight = f32_copy(c, ENC_PREFIX ".conv_layers.1.conv.weight", 1280u * 1280 * 3));
    NEED(w->conv1_bias = f32_copy(c, ENC_PREFIX ".conv_layers.1.conv.bias", 1280));
    for (int i = 0; i < w->n_enc_layers; i++) {
        char n[256];
        vox_cuda_enc_layer_w_t *d = &w->enc[i];
#define ENC(field, suffix, elems) do { snprintf(n, sizeof(n), ENC_PREFIX ".transformer.layers.%d." suffix, i); \
                                       NEED(d->field = bf16_ptr(c, n, (elems))); } while (0)
#define ENCF(field, suffix, elems) do { snprintf(n, sizeof(n), ENC_PREFIX ".transformer.layers.%d." suffix, i); \
                                        NEED(d->field = f32_copy(c, n, (elems))); } while (0)
        ENC(wq, "attention.wq.weight", 2048u * 1280); ENC(wk, "attention.wk.weight", 2048u * 1280);
        ENC(wv, "attention.wv.weight", 2048u * 1280); ENC(wo, "attention.wo.weight", 1280u * 2048);
        ENC(w1, "feed_forward.w1.weight", 5120u * 1280); ENC(w2, "feed_forward.w2.weight", 1280u * 5120);
        ENC(w3, "feed_forward.w3.weight", 5120u * 1280);
        ENCF(wq_bias, "attention.wq.bias", 2048); ENCF(wv_bias, "attention.wv.bias", 2048);
        ENCF(wo_bias, "attention.wo.bias", 1280); ENCF(w2_bias, "feed_forward.w2.bias", 1280);
        ENCF(attention_norm, "attention_norm.weight", 1280); ENCF(ffn_norm, "ffn_norm.weight", 1280);
    }
    NEED(w->enc_norm = f32_copy(c, ENC_PREFIX ".transformer.norm.weight", 1280));
    NEED(w->adapter_w0 = bf16_ptr(c, EMB_PREFIX ".audio_language_projection.0.weight", 3072u * 5120));
    NEED(w->adapter_w1 = bf16_ptr(c, EMB_PREFIX ".audio_language_projection.2.weight", 3072u * 3072));
    NEED(w->tok_embeddings = bf16_ptr(c, EMB_PREFIX ".tok_embeddings.weight", (size_t)VOX_CUDA_VOCAB * 3072));
    for (int i = 0; i < w->n_dec_layers; i++) {
        char n[256];
        vox_cuda_dec_layer_w_t *d = &w->dec[i];
#define DEC(field, suffix, elems) do { snprintf(n, sizeof(n), "layers.%d." suffix, i); NEED(d->field = bf16_ptr(c, n, (elems))); } while (0)
#define DECF(field, suffix, elems) do { snprintf(n, sizeof(n), "layers.%d." suffix, i); NEED(d->field = f32_copy(c, n, (elems))); } while (0)
        DEC(wq, "attention.wq.weight", 4096u * 3072); DEC(wk, "attention.wk.weight", 1024u * 3072);
        DEC(wv, "attention.wv.weight", 1024u * 3072); DEC(wo, "attention.wo.weight", 3072u * 4096);
        DEC(w1, "feed_forward.w1.weight", 9216u * 3072); DEC(w2, "feed_forward.w2.weight", 3072u * 9216);
        DEC(w3, "feed_forward.w3.weight", 9216u * 3072);
        DECF(attention_norm, "attention_norm.weight", 3072); DECF(ffn_norm, "ffn_norm.weight", 3072);
    }
    NEED(w->dec_norm = f32_copy(c, "norm.weight", 3072));
    if (compute_ada_scale(c)) goto fail;
    c->model = vox_cuda_model_create(w);
    if (c->model) c->state = vox_cuda_state_create(c->model);
    if (!c->model || !c->state) { fprintf(stderr, "vox_b200_load: %s\n", vox_cuda_last_error()); goto fail; }
    return c;
fail:
    vox_b200_free(c);
    return NULL;
}

void vox_b200_free(vox_b200_ctx *c) {
    if (!c) return;
    if (c->state) vox_cuda_state_free(c->state);
    if (c->model) vox_cuda_model_free(c->model);
    for (int i = 0; i < c->n_small; i++) free(c->small[i]);
    free(c->ada_scale);
    if (c->map) munmap(c->map, c->map_total);
    if (c->fd >= 0) close(c->fd);
    free(c);
}

int vox_b200_set_delay_tokens(vox_b200_ctx *c, int delay_tokens) {
    if (!c || delay_tokens < 0 || delay_tokens > 64) return -1;
    c->delay_tokens = delay_tokens;
    if (compute_ada_scale(c)) return -1;
    return vox_cuda_model_set_ada_scale(c->model, c->ada_scale);
}

/* ---- audio padding (what vox_mel_ctx sees for a whole utterance) ---------------------------
 * 200 + 32*1280 zeros | samples | zeros to a 1280 multiple + (delay + 1 + 10) tokens | the last
 * 200 samples mirrored (voxtral.c:1255,1291,1645-1658; voxtral_audio.c:545,602-629). */
static int right_pad_samples(int n_samples, int delay_tokens) {
    return (TOKEN_SAMPLES - n_samples % TOKEN_SAMPLES) % TOKEN_SAMPLES + (delay_tokens + 1 + 10) * TOKEN_SAMPLES;
}

static void reflect_tail(float *buf, size_t n) {           /* writes buf[n .. n+199] */
    for (int i = 0; i < 200; i++) {
        const long src = (long)n - 2 - i;
        buf[n + i] = src >= 0 ? buf[src] : 0.0f;
    }
}

static int truncate_at_eos(const int *ids, int n) {
    for (int i = 0; i < n; i++)
        if (ids[i] == VOX_TOKEN_EOS) return i;
    return n;
}

int vox_b200_transcribe_ids(vox_b200_ctx *c, const float *samples, int n_samples, int *ids_out, int max_ids,
                            float *device_ms) {
    if (!c || !samples || n_samples < 0 || !ids_out) return -1;
    const int tokens = LEFT_PAD_TOKENS + (n_samples + TOKEN_SAMPLES - 1) / TOKEN_SAMPLES + c->delay_tokens + 1 + 10;
    if (tokens > 1000) {                                    /* longer than the device adapter buffer: stream it */
        vox_b200_stream *s = vox_b200_stream_init(c);
        if (!s) return -1;
        int n = -1;
        if (vox_b200_stream_feed(s, samples, n_samples) == 0 && vox_b200_stream_finish(s) == 0)
            n = vox_b200_stream_get_ids(s, ids_out, max_ids);
        vox_b200_stream_free(s);
        if (device_ms) *device_ms = 0.f;
        return n < 0 ? -1 : truncate_at_eos(ids_out, n);
    }
    const size_t left = 200 + (size_t)LEFT_PAD_TOKENS * TOKEN_SAMPLES;
    const size_t body = left + (size_t)n_samples + (size_t)right_pad_samples(n_samples, c->delay_tokens);
    float *buf = (float *)vox_cuda_host_alloc((body + 200) * sizeof(float));      /* pinned: no staging copy */
    if (!buf) { fprintf(stderr, "vox_b200_transcribe_ids: %s\n", vox_cuda_last_error()); return -1; }
    memset(buf, 0, body * sizeof(float));
    memcpy(buf + left, samples, (size_t)n_samples * sizeof(float));
    reflect_tail(buf, body);
    const int n_frames = (int)((body + 200 - WIN) / HOP);   /* total frames - 1: the last frame is dropped */
    const int prompt_len = 1 + LEFT_PAD_TOKENS + c->delay_tokens;
    int n = vox_cuda_transcribe(c->state, buf, (int)(body + 200), n_frames, prompt_len, ids_out, max_ids, device_ms);
    vox_cuda_host_free(buf);
    if (n < 0) { fprintf(stderr, "vox_b200_transcribe_ids: %s\n", vox_cuda_last_error()); return -1; }
    return truncate_at_eos(ids_out, n);
}

/* ---- streaming ------------------------------------------------------------------------------ */
struct vox_b200_stream {
    vox_b200_ctx *ctx;
    float *buf;                 /* padded sample buffer: left pad, then everything fed so far */
    size_t n, cap;
    long real_samples;
    int frames_done;            /* mel frames already sent to the encoder */
    int interval_frames;        /* -I in mel frames (100 per second) */
    int first_chunk_done;
    int finished;
    float *mel;                 /* scratch [frames][128] */
    int mel_cap;
    int decoder_started, eos_seen, gen_pos, prev_token;
    int *ids;                   /* pending output queue */
    int n_ids, ids_cap, ids_read;
    double mel_ms, enc_ms, dec_ms, prefill_ms;
    int enc_calls, dec_steps, n_text;
    int stem_resid;             /* parity residual of the conv stem (voxtral.c:656-733), to count rows per call */
    int enc_rows[64];           /* encoder rows of the first 64 calls (diagnostics, -v) */
};

vox_b200_stream *vox_b200_stream_init(vox_b200_ctx *c) {
    if (!c) return NULL;
    vox_b200_stream *s = (vox_b200_stream *)calloc(1, sizeof(*s));
    if (!s) return NULL;
    s->ctx = c;
    s->cap = 200 + (size_t)LEFT_PAD_TOKENS * TOKEN_SAMPLES + 16000 * 4;
    s->buf = (float *)calloc(s->cap, sizeof(float));
    if (!s->buf) { free(s); return NULL; }
    s->n = 200 + (size_t)LEFT_PAD_TOKENS * TOKEN_SAMPLES;      /* starts as silence */
    s->interval_frames = 200;
    s->prev_token = VOX_TOKEN_STREAMING_PAD;
    vox_cuda_state_reset_encoder(c->state);
    vox_cuda_state_reset_decoder(c->state);
    return s;
}

void vox_b200_set_processing_interval(vox_b200_stream *s, float seconds) {
    if (!s) return;
    int f = (int)((seconds > 0 ? seconds : 0) * 100.0f);       /* voxtral.c:1669-1675 */
    s->interval_frames = f < 1 ? 1 : f;
}

static int append(vox_b200_stream *s, const float *x, size_t n, size_t zeros) {
    const size_t need = s->n + n + zeros + 200;
    if (need > s->cap) {
        size_t cap = s->cap;
        while (cap < need) cap *= 2;
        float *p = (float *)realloc(s->buf, cap * sizeof(float));
        if (!p) return -1;
        s->buf = p;
        s->cap = cap;
    }
    if (n) memcpy(s->buf + s->n, x, n * sizeof(float));
    if (zeros) memset(s->buf + s->n + n, 0, zeros * sizeof(float));
    s->n += n + zeros;
    return 0;
}

static int push_ids(vox_b200_stream *s, const int *ids, int n) {
    if (s->n_ids + n > s->ids_cap) {
        int cap = s->ids_cap ? s->ids_cap : 256;
        while (cap < s->n_ids + n) cap *= 2;
        int *p = (int *)realloc(s->ids, sizeof(int) * (size_t)cap);
        if (!p) return -1;
        s->ids = p;
        s->ids_cap = cap;
    }
    memcpy(s->ids + s->n_ids, ids, sizeof(int) * (size_t)n);
    s->n_ids += n;
    return 0;
}

/* prefill once the prompt is covered, then one step per available adapter token
 * (the control flow of stream_run_decoder, voxtral.c:1012-1145) */
static int run_decoder(vox_b200_stream *s) {
    vox_cuda_state_t *st = s->ctx->state;
    const int prompt_len = 1 + LEFT_PAD_TOKENS + s->ctx->delay_tokens;
    const int total = vox_cuda_adapter_count(st);
    const double t0 = now_ms();
    if (!s->decoder_started) {
        if (total < prompt_len) return 0;
        int prompt[1 + LEFT_PAD_TOKENS + 64];
        for (int i = 0; i < prompt_len; i++) prompt[i] = i == 0 ? VOX_TOKEN_BOS : VOX_TOKEN_STREAMING_PAD;
        if (vox_cuda_decoder_prefill_fused(st, prompt, 0, prompt_len - 1)) return -1;
        s->prefill_ms += now_ms() - t0;
        s->decoder_started = 1;
        s->gen_pos = prompt_len - 1;
        s->prev_token = prompt[prompt_len - 1];
    }
    if (!s->eos_seen && total > s->gen_pos) {
        const int k = total - s->gen_pos;
        int *out = (int *)malloc(sizeof(int) * (size_t)k);
        if (!out) return -1;
        if (vox_cuda_decoder_run(st, s->prev_token, s->gen_pos, k, out) != k) { free(out); return -1; }
        const int keep = truncate_at_eos(out, k);
        if (keep < k) s->eos_seen = 1;
        for (int i = 0; i < keep; i++) s->n_text += out[i] >= 1000;
        const int rc = push_ids(s, out, keep);
        s->prev_token = out[k - 1];
        s->gen_pos = total;
        s->dec_steps += k;
        free(out);
        if (rc) return -1;
        /* stream_adapter_compact (voxtral.c:762-775): consumed tokens are no longer needed */
        vox_cuda_adapter_discard_before(st, s->gen_pos);
    }
    /* after EOS nothing consumes adapter tokens any more: drop them as they arrive so that the device
     * buffer (1024 rows) cannot fill while audio keeps coming (the reference reallocs its host buffer
     * instead, voxtral.c:905-921) */
    if (s->eos_seen) vox_cuda_adapter_discard_before(st, total);
    s->dec_ms += now_ms() - t0;
    return 0;
}

/* mel + encoder on every frame that is complete, in chunks of at most 6656 frames (one internal pass of the backend) */
static int process_frames(vox_b200_stream *s, int upto_frames) {
    vox_cuda_state_t *st = s->ctx->state;
    while (s->frames_done < upto_frames) {
        int n = upto_frames - s->frames_done;
        if (n > 6656) n = 6656;
        if (n > s->mel_cap) {
            float *p = (float *)realloc(s->mel, sizeof(float) * 128 * (size_t)n);
            if (!p) return -1;
            s->mel = p;
            s->mel_cap = n;
        }
        double t0 = now_ms();
        if (vox_cuda_mel_frames(s->buf + (size_t)s->frames_done * HOP, n, s->mel)) return -1;
        s->mel_ms += now_ms() - t0;
        t0 = now_ms();
        if (vox_cuda_encode_chunk(st, s->mel, n) < 0) return -1;
        s->enc_ms += now_ms() - t0;
        if (s->enc_calls < 64) s->enc_rows[s->enc_calls] = (s->stem_resid + n) / 2;
        s->stem_resid = (s->stem_resid + n) & 1;
        s->enc_calls++;
        s->frames_done += n;
        s->first_chunk_done = 1;
        if (run_decoder(s)) return -1;
    }
    return 0;
}

int vox_b200_stream_feed(vox_b200_stream *s, const float *samples, int n_samples) {
    if (!s || s->finished || n_samples < 0 || (!samples && n_samples)) return -1;
    if (append(s, samples, (size_t)n_samples, 0)) return -1;
    s->real_samples += n_samples;
    const int avail = s->n >= WIN ? (int)((s->n - WIN) / HOP) + 1 : 0;
    int fresh = avail - s->frames_done;
    /* first chunk: 312 mel frames (the decoder prompt), later ones: the processing interval; every complete
     * frame goes in, an odd count leaves a conv0 row on the device (voxtral.c:837-840,405,1669-1675) */
    const int need = s->first_chunk_done ? s->interval_frames : 312;
    if (fresh < need) return 0;
    if (process_frames(s, s->frames_done + fresh)) { fprintf(stderr, "vox_b200_stream_feed: %s\n", vox_cuda_last_error()); return -1; }
    return 0;
}

int vox_b200_stream_finish(vox_b200_stream *s) {
    if (!s) return -1;
    if (s->finished) return 0;
    const int pad = right_pad_samples((int)(s->real_samples % TOKEN_SAMPLES), s->ctx->delay_tokens);
    if (append(s, NULL, 0, (size_t)pad)) return -1;
    /* vox_stream_flush: every frame the padding completed goes through the encoder first ... */
    const int before = s->n >= WIN ? (int)((s->n - WIN) / HOP) + 1 : 0;
    if (before > s->frames_done && process_frames(s, before)) {
        fprintf(stderr, "vox_b200_stream_finish: %s\n", vox_cuda_last_error());
        return -1;
    }
    /* ... then vox_mel_finish adds the reflected tail and the last frames follow (voxtral.c:1299-1317) */
    reflect_tail(s->buf, s->n);
    s->n += 200;
    const int total = (int)((s->n - WIN) / HOP) + 1;
    s->finished = 1;
    if (process_frames(s, total - 1)) { fprintf(stderr, "vox_b200_stream_finish: %s\n", vox_cuda_last_error()); return -1; }
    return 0;
}

int vox_b200_stream_get_ids(vox_b200_stream *s, int *ids_out, int max) {
    if (!s || !ids_out || max < 0) return -1;
    int n = s->n_ids - s->ids_read;
    if (n > max) n = max;
    memcpy(ids_out, s->ids + s->ids_read, sizeof(int) * (size_t)n);
    s->ids_read += n;
    return n;
}

void vox_b200_stream_stats(const vox_b200_stream *s, double *mel_ms, double *encoder_ms, double *decoder_ms,
                           int *encoder_calls, int *decoder_steps) {
    if (!s) return;
    if (mel_ms) *mel_ms = s->mel_ms;
    if (encoder_ms) *encoder_ms = s->enc_ms;
    if (decoder_ms) *decoder_ms = s->dec_ms;
    if (encoder_calls) *encoder_calls = s->enc_calls;
    if (decoder_steps) *decoder_steps = s->dec_steps;
}

/* The reference's own summary lines, byte-compatible with what vox_stream_free prints at verbosity 1
 * (voxtral.c:1359-1370) and /root/reference/benchmark.py:25-30 parses. */
void vox_b200_stream_report(const vox_b200_stream *s, FILE *f, int verbose) {
    if (!s || !f) return;
    fprintf(f, "Encoder: %d mel -> %d tokens (%.0f ms)\n", s->frames_done, s->gen_pos, s->enc_ms + s->mel_ms);
    if (s->n_text > 0) {
        const double gen_ms = s->dec_ms - s->prefill_ms;
        const int n_generated = s->dec_steps;
        fprintf(f, "Decoder: %d text tokens (%d steps) in %.0f ms (prefill %.0f ms + %.1f ms/step)\n",
                s->n_text, n_generated, s->dec_ms, s->prefill_ms, n_generated > 1 ? gen_ms / (n_generated - 1) : 0);
    }
    if (verbose) {
        fprintf(f, "encoder calls: %d rows:", s->enc_calls);
        for (int i = 0; i < s->enc_calls && i < 64; i++) fprintf(f, " %d", s->enc_rows[i]);
        fprintf(f, "\n");
    }
}

void vox_b200_stream_free(vox_b200_stream *s) {
    if (!s) return;
    free(s->buf);
    free(s->mel);
    free(s->ids);
    free(s);
}

/* ---- WAV: RIFF parsing on the host (vox_parse_wav_buffer, voxtral_audio.c:49-92), the sample arithmetic on the
 * device (vox_cuda_pcm_from_s16: channel mean, / 32768, linear-interpolation resampling to 16 kHz,
 * voxtral_audio.c:94-137).  16-bit PCM, any channel count and rate, like the reference. -------- */
float *vox_b200_load_wav(const char *path, int *n_samples) {
    FILE *f = fopen(path, "rb");
    if (!f) { fprintf(stderr, "vox_b200_load_wav: cannot open %s\n", path); return NULL; }
    fseek(f, 0, SEEK_END);
    const long size = ftell(f);
    fseek(f, 0, SEEK_SET);
    uint8_t *d = (uint8_t *)malloc((size_t)size);
    if (!d || fread(d, 1, (size_t)size, f) != (size_t)size) { fclose(f); free(d); return NULL; }
    fclose(f);
    float *out = NULL;
    int16_t *frames = NULL;
    if (size < 44 || memcmp(d, "RIFF", 4) || memcmp(d + 8, "WAVE", 4)) goto bad;
    {
        int fmt = 0, ch = 0, bits = 0;
        uint32_t rate = 0;
        long p = 12;
        while (p + 8 <= size) {
            uint32_t len;
            memcpy(&len, d + p + 4, 4);
            if (!memcmp(d + p, "fmt ", 4) && len >= 16 && p + 8 + (long)len <= size) {   /* voxtral_audio.c:66 */
                fmt = d[p + 8] | (d[p + 9] << 8);
                ch = d[p + 10] | (d[p + 11] << 8);
                memcpy(&rate, d + p + 12, 4);
                bits = d[p + 22] | (d[p + 23] << 8);
            } else if (!memcmp(d + p, "data", 4)) {
                if (fmt != 1 || ch < 1 || bits != 16 || rate == 0) goto bad;
                if ((long)len > size - p - 8 || len == 0) len = (uint32_t)(size - p - 8);   /* piped ffmpeg writes -1 */
                const int n_frames = (int)(len / (2u * (uint32_t)ch));
                const int n_out = vox_cuda_pcm_resampled_len(n_frames, (int)rate);
                if (n_frames <= 0 || n_out < 0) goto bad;
                frames = (int16_t *)malloc((size_t)n_frames * (size_t)ch * 2);
                out = (float *)malloc(sizeof(float) * (size_t)(n_out > 0 ? n_out : 1));
                if (!frames || !out) goto bad;
                memcpy(frames, d + p + 8, (size_t)n_frames * (size_t)ch * 2);                /* little-endian host */
                if (vox_cuda_pcm_from_s16(frames, n_frames, ch, (int)rate, out, n_out) != n_out) {
                    fprintf(stderr, "vox_b200_load_wav: %s\n", vox_cuda_last_error());
                    goto bad;
                }
                *n_samples = n_out;
                free(frames);
                free(d);
                return out;
            }
            p += 8 + (long)len + (len & 1);
        }
    }
bad:
    fprintf(stderr, "vox_b200_load_wav: %s is not a 16-bit PCM WAV file\n", path);
    free(d);
    free(out);
    free(frames);
    return NULL;
}
