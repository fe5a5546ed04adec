// api_dec.cu - per-stream state and the decoder entry points of the C ABI.
#include "vox_internal.h"
#include "rows.h"
#include "gemm_tc.h"
#include <cuda_fp16.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

void vox_mem_account(long delta);
const unsigned *dec_mega_error_record(vox_cuda_state *s);
int enc_state_create(vox_cuda_state *s);
void enc_state_free(vox_cuda_state *s);
void enc_state_reset(vox_cuda_state *s);
void batch_free(vox_cuda_state *s);

static int dmalloc(void **p, size_t bytes) {
    if (cudaMalloc(p, bytes) != cudaSuccess) {
        vox_set_error("cudaMalloc(%zu MB) failed", bytes >> 20);
        return -1;
    }
    vox_mem_account((long)bytes);
    return 0;
}
#define DM(ptr, bytes) do { if (dmalloc((void **)&(ptr), (bytes))) goto fail; } while (0)

static int push_params(vox_cuda_state *s) {
    CUDA_TRY(cudaMemcpyAsync(s->d_params, &s->h_params, sizeof(DecStepParams),
                             cudaMemcpyHostToDevice, s->stream));
    return 0;
}

extern "C" vox_cuda_state_t *vox_cuda_state_create(vox_cuda_model_t *m) { return vox_cuda_state_create_ex(m, 32); }

extern "C" int vox_cuda_state_kv_bits(const vox_cuda_state_t *s) { return s ? (s->kv16 ? 16 : 32) : -1; }

// kv_bits 32: fp32 K/V rings (the reference's CPU / BLAS layout: ids equal the reference); 16: fp16 rings, the
// reference's GPU default (voxtral.c:189-190) - half the K/V bytes per token (0.87 of 1.74 GB at the full window).
extern "C" vox_cuda_state_t *vox_cuda_state_create_ex(vox_cuda_model_t *m, int kv_bits) {
    if (!m) return nullptr;
    if (kv_bits != 32 && kv_bits != 16) { vox_set_error("state_create: kv_bits must be 32 or 16"); return nullptr; }
    vox_cuda_state *s = (vox_cuda_state *)calloc(1, sizeof(vox_cuda_state));
    s->m = m;
    s->kv16 = kv_bits == 16;
    const size_t kv = s->kv16 ? (size_t)m->n_dec * PF_KV_ROWS * DEC_KVDIM * 4 : (size_t)m->n_dec * DEC_WINDOW * DEC_KVDIM * 4;
    const size_t kvh = (size_t)m->n_dec * DEC_WINDOW * DEC_KVDIM * 2;
    if (cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking) != cudaSuccess) goto fail;
    DM(s->kv_k, kv); DM(s->kv_v, kv);
    if (s->kv16) {
        DM(s->kv_k16, kvh); DM(s->kv_v16, kvh);
        if (cudaMemsetAsync(s->kv_k16, 0, kvh, s->stream) != cudaSuccess) goto fail;
        if (cudaMemsetAsync(s->kv_v16, 0, kvh, s->stream) != cudaSuccess) goto fail;
    }
    DM(s->d_params, sizeof(DecStepParams));
    DM(s->d_x, DEC_DIM * 4); DM(s->d_q, DEC_QDIM * 4); DM(s->d_attn, DEC_QDIM * 4);
    DM(s->d_h, DEC_HIDDEN * 4); DM(s->d_logits, (size_t)VOCAB * 4); DM(s->d_embed_in, DEC_DIM * 4);
    DM(s->d_rope, DEC_HD * 4);
    {
        size_t part = (size_t)ATTN_NSPLIT * DEC_HEADS * (DEC_HD + 2);
        if ((size_t)dec_mega_attn_part_floats() > part) part = (size_t)dec_mega_attn_part_floats();
        DM(s->d_attn_part, part * 4);
    }
    DM(s->d_blockmax, (size_t)2 * (VOCAB / LOGITS_ROWS_PER_CTA) * 4);
    DM(s->d_tokens, (size_t)TOKENS_RING * 4);
    s->adapter_cap = 1024;
    DM(s->adapter_buf, (size_t)s->adapter_cap * DEC_DIM * 4);
    if (cudaMallocHost(&s->h_tokens, (size_t)TOKENS_RING * 4) != cudaSuccess) goto fail;
    if (cudaMallocHost(&s->h_pinned, (size_t)VOCAB * 4) != cudaSuccess) goto fail;
    if (cudaEventCreate(&s->ev0) != cudaSuccess || cudaEventCreate(&s->ev1) != cudaSuccess) goto fail;
    if (cudaMemsetAsync(s->kv_k, 0, kv, s->stream) != cudaSuccess) goto fail;
    if (cudaMemsetAsync(s->kv_v, 0, kv, s->stream) != cudaSuccess) goto fail;
    if (enc_state_create(s)) goto fail;
    {
        const char *env = getenv("VOX_CUDA_DEC_GRAPH");
        s->use_graph_steps = (env && atoi(env)) ? 1 : 0;
        if (s->use_graph_steps && s->kv16) { vox_set_error("state_create: fp16 K/V needs the default (persistent kernel) engine"); goto fail; }
        if (s->use_graph_steps && m->q8) { vox_set_error("state_create: a Q8 model needs the default (persistent kernel) engine"); goto fail; }
    }
    s->alt_n = 1;
    if (dec_mega_init(s)) goto fail;
    vox_cuda_state_reset_decoder(s);
    vox_cuda_state_reset_encoder(s);
    if (cudaStreamSynchronize(s->stream) != cudaSuccess) goto fail;
    return s;
fail:
    if (!*vox_cuda_last_error()) vox_set_error("vox_cuda_state_create failed: %s", cudaGetErrorString(cudaGetLastError()));
    vox_cuda_state_free(s);
    return nullptr;
}

extern "C" void vox_cuda_state_free(vox_cuda_state_t *s) {
    if (!s) return;
    if (s->stream) cudaStreamSynchronize(s->stream);
    if (s->step_graph) cudaGraphExecDestroy(s->step_graph);
    batch_free(s);
    enc_state_free(s);
    dec_mega_free(s);
    dec_prefill_wt_free(s);
    void *ptrs[] = {s->kv_k, s->kv_v, s->d_params, s->d_x, s->d_q, s->d_attn, s->d_h, s->d_logits,
                    s->d_embed_in, s->d_rope, s->d_attn_part, s->d_blockmax, s->d_tokens,
                    s->adapter_buf, s->pf_buf, s->d_alt, s->d_alt_cand, s->kv_k16, s->kv_v16};
    for (void *p : ptrs) if (p) cudaFree(p);
    if (s->m) {
        const size_t kv = s->kv16 ? (size_t)s->m->n_dec * (PF_KV_ROWS * 4 + DEC_WINDOW * 2) * DEC_KVDIM
                                  : (size_t)s->m->n_dec * DEC_WINDOW * DEC_KVDIM * 4;
        vox_mem_account(-(long)(2 * kv));
    }
    if (s->h_tokens) cudaFreeHost(s->h_tokens);
    if (s->h_pinned) cudaFreeHost(s->h_pinned);
    if (s->h_alt) cudaFreeHost(s->h_alt);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    if (s->stream) cudaStreamDestroy(s->stream);
    free(s);
}

extern "C" void vox_cuda_state_reset_decoder(vox_cuda_state_t *s) {
    if (!s) return;
    memset(&s->h_params, 0, sizeof(s->h_params));
    s->h_params.prev_token = 1;  // BOS, voxtral.c:788
    push_params(s);
}

extern "C" int vox_cuda_decoder_position(const vox_cuda_state_t *s) { return s ? s->h_params.pos : -1; }

extern "C" int vox_cuda_decoder_set_position(vox_cuda_state_t *s, int logical_pos, int n_valid) {
    if (!s || logical_pos < 0 || n_valid < 0 || n_valid > logical_pos) return -1;
    s->h_params.pos = logical_pos;
    s->h_params.n_valid = n_valid > DEC_WINDOW ? DEC_WINDOW : n_valid;
    return push_params(s);
}

// ---- KV ring <-> host (logical addressing) ------------------------------------------
static int kv_copy(vox_cuda_state *s, int layer, int logical_pos, int n, float *k, float *v, bool to_device) {
    if (!s || layer < 0 || layer >= s->m->n_dec || logical_pos < 0 || n < 0 || n > DEC_WINDOW || (!k && !v)) return -1;
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    const size_t layer_off = (size_t)layer * DEC_WINDOW * DEC_KVDIM;
    int done = 0;
    while (done < n) {
        const int slot = (logical_pos + done) & (DEC_WINDOW - 1);
        const int run = (n - done < DEC_WINDOW - slot) ? n - done : DEC_WINDOW - slot;   // up to the ring end
        const size_t bytes = (size_t)run * DEC_KVDIM * sizeof(float);
        const size_t doff = layer_off + (size_t)slot * DEC_KVDIM, hoff = (size_t)done * DEC_KVDIM;
        if (s->kv16) {                                   // host rows stay fp32 (the reference's layout): convert on the host
            const size_t cnt = (size_t)run * DEC_KVDIM;
            __half *tmp = (__half *)malloc(cnt * sizeof(__half));
            if (!tmp) return -1;
            for (int which = 0; which < 2; which++) {
                float *h = which ? v : k;
                __half *d = (__half *)(which ? s->kv_v16 : s->kv_k16) + doff;
                if (!h) continue;
                cudaError_t e;
                if (to_device) {
                    for (size_t i = 0; i < cnt; i++) tmp[i] = __float2half_rn(h[hoff + i]);
                    e = cudaMemcpy(d, tmp, cnt * sizeof(__half), cudaMemcpyHostToDevice);
                } else {
                    e = cudaMemcpy(tmp, d, cnt * sizeof(__half), cudaMemcpyDeviceToHost);
                    for (size_t i = 0; i < cnt; i++) h[hoff + i] = __half2float(tmp[i]);
                }
                if (e != cudaSuccess) { free(tmp); vox_set_error("decoder_kv copy: %s", cudaGetErrorString(e)); return -1; }
            }
            free(tmp);
            done += run;
            continue;
        }
        if (k) CUDA_TRY(to_device ? cudaMemcpy(s->kv_k + doff, k + hoff, bytes, cudaMemcpyHostToDevice)
                                  : cudaMemcpy(k + hoff, s->kv_k + doff, bytes, cudaMemcpyDeviceToHost));
        if (v) CUDA_TRY(to_device ? cudaMemcpy(s->kv_v + doff, v + hoff, bytes, cudaMemcpyHostToDevice)
                                  : cudaMemcpy(v + hoff, s->kv_v + doff, bytes, cudaMemcpyDeviceToHost));
        done += run;
    }
    return 0;
}
extern "C" int vox_cuda_decoder_kv_write(vox_cuda_state_t *s, int layer, int logical_pos, int n,
                                         const float *k, const float *v) {
    return kv_copy(s, layer, logical_pos, n, const_cast<float *>(k), const_cast<float *>(v), true);
}
extern "C" int vox_cuda_decoder_kv_read(vox_cuda_state_t *s, int layer, int logical_pos, int n, float *k, float *v) {
    return kv_copy(s, layer, logical_pos, n, k, v, false);
}

// ---- one step ----------------------------------------------------------------
static int ensure_graph(vox_cuda_state *s) {
    if (s->graph_ready) return 0;
    cudaGraph_t g;
    CUDA_TRY(cudaStreamBeginCapture(s->stream, cudaStreamCaptureModeThreadLocal));
    if (dec_launch_step(s, s->stream)) {
        cudaStreamEndCapture(s->stream, &g);
        return -1;
    }
    CUDA_TRY(cudaStreamEndCapture(s->stream, &g));
    CUDA_TRY(cudaGraphInstantiate(&s->step_graph, g, 0));
    cudaGraphDestroy(g);
    s->graph_ready = 1;
    return 0;
}

static inline int step_kernels(const vox_cuda_state *s) { return 3 + 6 * s->m->n_dec; }

// Enqueue n steps; tokens land in d_tokens[(step_count0 + i) & ring].
static int enqueue_steps(vox_cuda_state *s, int n) {
    if (!s->use_graph_steps) {
        // one persistent cooperative kernel runs up to MEGA_CHUNK steps back to back
        const int MEGA_CHUNK = 512;
        for (int i = 0; i < n; i += MEGA_CHUNK)
            if (dec_mega_launch(s, n - i < MEGA_CHUNK ? n - i : MEGA_CHUNK)) return -1;
        return 0;
    }
    if (ensure_graph(s)) return -1;
    for (int i = 0; i < n; i++) CUDA_TRY(cudaGraphLaunch(s->step_graph, s->stream));
    g_launch_count += (unsigned long long)n * step_kernels(s);
    return 0;
}

static void advance_host(vox_cuda_state *s, int n, int last_token) {
    s->h_params.pos += n;
    s->h_params.n_valid = s->h_params.n_valid + n > DEC_WINDOW ? DEC_WINDOW : s->h_params.n_valid + n;
    s->h_params.adapter_index += n;
    s->h_params.step_count += n;
    s->h_params.prev_token = last_token;
    s->h_params.out_token = last_token;
}

static int run_one(vox_cuda_state *s, float *logits) {
    const int slot = s->h_params.step_count & (TOKENS_RING - 1);
    if (push_params(s)) return -1;
    if (enqueue_steps(s, 1)) return -1;
    CUDA_TRY(cudaMemcpyAsync(s->h_tokens + slot, s->d_tokens + slot, 4, cudaMemcpyDeviceToHost, s->stream));
    if (logits)
        CUDA_TRY(cudaMemcpyAsync(s->h_pinned, s->d_logits, (size_t)VOCAB * 4, cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    if (logits) memcpy(logits, s->h_pinned, (size_t)VOCAB * 4);
    const int tok = s->h_tokens[slot];
    advance_host(s, 1, tok);
    return tok;
}

extern "C" int vox_cuda_decoder_step(vox_cuda_state_t *s, const float *embed, float *logits) {
    if (!s || !embed) return -1;
    memcpy(s->h_pinned, embed, DEC_DIM * 4);
    CUDA_TRY(cudaMemcpyAsync(s->d_embed_in, s->h_pinned, DEC_DIM * 4, cudaMemcpyHostToDevice, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));  // h_pinned is reused for the logits below
    s->h_params.use_host_embed = 1;
    return run_one(s, logits);
}

extern "C" int vox_cuda_decoder_step_fused(vox_cuda_state_t *s, int prev_token, int adapter_index,
                                           float *logits) {
    if (!s || prev_token < 0 || prev_token >= VOCAB) return -1;
    const int phys = adapter_index - s->adapter_offset;
    if (phys < 0 || adapter_index >= s->adapter_count) {
        vox_set_error("decoder_step_fused: adapter index %d outside [%d,%d)", adapter_index,
                      s->adapter_offset, s->adapter_count);
        return -1;
    }
    s->h_params.use_host_embed = 0;
    s->h_params.prev_token = prev_token;
    s->h_params.adapter_index = phys;
    return run_one(s, logits);
}

// Run n greedy steps back to back on the device (token i feeds step i+1) without
// any host round trip; used by the stream orchestrator and by bench.py.
extern "C" int vox_cuda_decoder_run(vox_cuda_state_t *s, int prev_token, int adapter_index, int n,
                                    int *tokens_out) {
    if (!s || n <= 0 || n > TOKENS_RING) return -1;
    const int phys = adapter_index - s->adapter_offset;
    if (phys < 0 || adapter_index + n > s->adapter_count) {
        vox_set_error("decoder_run: adapter range [%d,%d) outside [%d,%d)", adapter_index,
                      adapter_index + n, s->adapter_offset, s->adapter_count);
        return -1;
    }
    s->h_params.use_host_embed = 0;
    s->h_params.prev_token = prev_token;
    s->h_params.adapter_index = phys;
    const int c0 = s->h_params.step_count;
    if (push_params(s)) return -1;
    if (enqueue_steps(s, n)) return -1;
    for (int i = 0; i < n;) {  // ring may wrap: copy in contiguous pieces
        const int slot = (c0 + i) & (TOKENS_RING - 1);
        const int len = (n - i < TOKENS_RING - slot) ? n - i : TOKENS_RING - slot;
        CUDA_TRY(cudaMemcpyAsync(s->h_tokens + slot, s->d_tokens + slot, (size_t)len * 4,
                                 cudaMemcpyDeviceToHost, s->stream));
        i += len;
    }
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    for (int i = 0; i < n; i++) tokens_out[i] = s->h_tokens[(c0 + i) & (TOKENS_RING - 1)];
    advance_host(s, n, tokens_out[n - 1]);
    return n;
}

// ---- alternatives fused into the step (stream_fill_alts, voxtral.c:955-1010) --------------------
// The vocabulary phase of the decoder kernel tracks the four best text logits next to the argmax
// (dec_mega.cu); here the cutoff rule is applied to the per-step records: candidate i is kept while
// 1 - p_i / p_best <= cutoff, p_i / p_best = exp(l_i - l_best) (the softmax normaliser cancels).
extern "C" int vox_cuda_decoder_set_alt(vox_cuda_state_t *s, int n_alt, float cutoff) {
    if (!s) return -1;
    if (n_alt < 1) n_alt = 1;
    if (n_alt > VOX_CUDA_MAX_ALT) n_alt = VOX_CUDA_MAX_ALT;
    if (cutoff < 0.f) cutoff = 0.f;
    if (cutoff > 1.f) cutoff = 1.f;                  // vox_stream_set_alt, voxtral.c:1330-1339
    if (n_alt > 1 && s->use_graph_steps) { vox_set_error("decoder_set_alt: needs the default (persistent kernel) engine"); return -1; }
    if (n_alt > 1 && !s->d_alt) {
        CUDA_TRY(cudaMalloc((void **)&s->d_alt, sizeof(AltRecord) * TOKENS_RING));
        CUDA_TRY(cudaMalloc((void **)&s->d_alt_cand, sizeof(float2) * 4 * 1024));
        CUDA_TRY(cudaMallocHost((void **)&s->h_alt, sizeof(AltRecord) * TOKENS_RING));
    }
    s->alt_n = n_alt; s->alt_cutoff = cutoff;
    return 0;
}

// alts_out [n][VOX_CUDA_MAX_ALT] for the LAST n decoded steps: [0] = the greedy id, then alternatives, -1 = none.
// Only text tokens (id >= 1000) get alternatives, as in stream_run_decoder (voxtral.c:1064-1069,1122-1131).
extern "C" int vox_cuda_decoder_get_alts(vox_cuda_state_t *s, int n, int *alts_out) {
    if (!s || !alts_out || n < 0 || n > TOKENS_RING || n > s->h_params.step_count) return -1;
    if (s->alt_n <= 1 || !s->d_alt) { vox_set_error("decoder_get_alts: call vox_cuda_decoder_set_alt(n_alt > 1) before decoding"); return -1; }
    const int c0 = s->h_params.step_count - n;
    for (int i = 0; i < n;) {
        const int slot = (c0 + i) & (TOKENS_RING - 1);
        const int len = (n - i < TOKENS_RING - slot) ? n - i : TOKENS_RING - slot;
        CUDA_TRY(cudaMemcpyAsync(s->h_alt + slot, s->d_alt + slot, sizeof(AltRecord) * len, cudaMemcpyDeviceToHost, s->stream));
        CUDA_TRY(cudaMemcpyAsync(s->h_tokens + slot, s->d_tokens + slot, (size_t)len * 4, cudaMemcpyDeviceToHost, s->stream));
        i += len;
    }
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    for (int i = 0; i < n; i++) {
        const int slot = (c0 + i) & (TOKENS_RING - 1);
        const AltRecord &r = s->h_alt[slot];
        int *o = alts_out + (size_t)i * VOX_CUDA_MAX_ALT;
        o[0] = s->h_tokens[slot];
        for (int k = 1; k < VOX_CUDA_MAX_ALT; k++) o[k] = -1;
        if (o[0] < 1000) continue;
        for (int k = 0; k + 1 < s->alt_n && k < 3; k++) {
            if (r.id[k] < 0) break;
            const float ratio = expf(r.v[k] - r.best);
            if (1.0f - ratio > s->alt_cutoff) break;
            o[k + 1] = r.id[k];
        }
    }
    return n;
}

extern "C" int vox_cuda_decoder_read_logits(vox_cuda_state_t *s, float *logits) {
    if (!s || !logits) return -1;
    CUDA_TRY(cudaMemcpyAsync(s->h_pinned, s->d_logits, (size_t)VOCAB * 4, cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    memcpy(logits, s->h_pinned, (size_t)VOCAB * 4);
    return 0;
}

// Same as vox_cuda_decoder_run but returns the average device time per step in
// milliseconds, measured with CUDA events on the launch stream.
extern "C" float vox_cuda_bench_decoder_steps(vox_cuda_state_t *s, int prev_token, int adapter_index,
                                              int n, int *tokens_out) {
    if (!s || n <= 0 || n > TOKENS_RING) return -1.f;
    const int phys = adapter_index - s->adapter_offset;
    if (phys < 0 || adapter_index + n > s->adapter_count) {
        vox_set_error("bench_decoder_steps: adapter range outside buffer");
        return -1.f;
    }
    s->h_params.use_host_embed = 0;
    s->h_params.prev_token = prev_token;
    s->h_params.adapter_index = phys;
    const int c0 = s->h_params.step_count;
    if (push_params(s) || (s->use_graph_steps && ensure_graph(s))) return -1.f;
    if (cudaEventRecord(s->ev0, s->stream) != cudaSuccess) return -1.f;
    if (enqueue_steps(s, n)) return -1.f;
    if (cudaEventRecord(s->ev1, s->stream) != cudaSuccess) return -1.f;
    if (cudaMemcpyAsync(s->h_tokens, s->d_tokens, (size_t)TOKENS_RING * 4, cudaMemcpyDeviceToHost,
                        s->stream) != cudaSuccess) return -1.f;
    if (cudaStreamSynchronize(s->stream) != cudaSuccess) {
        const unsigned *er = dec_mega_error_record(s);
        vox_set_error("bench_decoder_steps: %s (mega watchdog code %u cta %u tid %u info 0x%x)",
                      cudaGetErrorString(cudaGetLastError()), er ? er[0] : 0, er ? er[1] : 0, er ? er[2] : 0, er ? er[3] : 0);
        return -1.f;
    }
    float ms = 0.f;
    cudaEventElapsedTime(&ms, s->ev0, s->ev1);
    int last = prev_token;
    for (int i = 0; i < n; i++) {
        last = s->h_tokens[(c0 + i) & (TOKENS_RING - 1)];
        if (tokens_out) tokens_out[i] = last;
    }
    advance_host(s, n, last);
    return ms / (float)n;
}

// ---- prefill (vox_decoder_prefill, voxtral_decoder.c:447-612) -------------------
// fp32 activations on the CUDA-core GEMM; rows are processed in blocks of
// PF_ROWS (causal attention makes blocked prefill exact).
constexpr int PF_ROWS = 64;

// Two arithmetic modes, like the encoder: default = tcgen05 GEMMs on bf16x2 split
// activations (fp32-class accuracy, weights streamed once at HBM speed);
// VOX_CUDA_PREFILL_F32=1 = fp32 CUDA-core GEMM.
static int prefill_f32_mode(void) {
    static int m = -1;
    if (m < 0) { const char *env = getenv("VOX_CUDA_PREFILL_F32"); m = (env && atoi(env)) ? 1 : 0; }
    return m;
}

// VOX_CUDA_PREFILL_TC=1 keeps the prompt rows on the large-tile GEMM (gemm_tc.cu; A/B runs: 5.05 ms per 38-row prompt)
static int prefill_wt_mode(void) {
    static int m = -1;
    if (m < 0) { const char *env = getenv("VOX_CUDA_PREFILL_TC"); m = (env && atoi(env)) ? 0 : 1; }
    return m;
}

static int ensure_pf(vox_cuda_state *s) {
    const size_t per_row = (size_t)(2 * DEC_DIM + DEC_QKV + DEC_QDIM + 2 * DEC_HIDDEN + DEC_HIDDEN) * 4;
    const size_t need = per_row * PF_ROWS;
    if (s->pf_bytes >= need) return 0;
    if (s->pf_buf) cudaFree(s->pf_buf);
    if (cudaMalloc((void **)&s->pf_buf, need) != cudaSuccess) { vox_set_error("prefill scratch alloc failed"); return -1; }
    s->pf_bytes = need;
    return 0;
}

static TcEpilogue pf_epi(float *of, long ldf, uint16_t *ob, long ldb, int flags, long split_off = 0,
                         const float *col_scale = nullptr) {
    TcEpilogue t;
    t.out_f32 = of; t.ldo_f32 = ldf; t.out_bf16 = ob; t.ldo_bf16 = ldb; t.bias = nullptr; t.flags = flags;
    t.split_off = split_off;
    t.col_scale = col_scale;             // Q8 model: row scales of the quantised matrix, else null
    return t;
}

// x_dev: [seq_len, 3072] device, consumed (used as the residual stream).  K/V rows go to the ring
// kbase/vbase ([n_dec] layers kv_layer floats apart, slot = logical position & mask); *pos / *n_valid
// are the logical position of the first row and the valid entries before it, advanced on return.
// Enqueues only (no synchronisation); scratch is the state's prefill buffer.
int dec_prefill_into(vox_cuda_state *s, float *x_dev, int seq_len, float *kbase, float *vbase, size_t kv_layer,
                     int mask, int *pos, int *n_valid) {
    if (ensure_pf(s)) return -1;
    vox_cuda_model *m = s->m;
    cudaStream_t st = s->stream;
    float *xn = s->pf_buf;
    float *qkv = xn + (size_t)PF_ROWS * DEC_DIM;
    float *att = qkv + (size_t)PF_ROWS * DEC_QKV;
    float *gu = att + (size_t)PF_ROWS * DEC_QDIM;
    float *hh = gu + (size_t)PF_ROWS * 2 * DEC_HIDDEN;
    // bf16x2 views of the same scratch (tensor-core mode): [hi | lo] rows
    uint16_t *xn_b = (uint16_t *)xn;      // [PF_ROWS][2*3072]   (same bytes as fp32 [PF_ROWS][3072])
    uint16_t *att_b = (uint16_t *)gu;     // [PF_ROWS][2*4096]   (gu is unused in this mode)
    uint16_t *hh_b = (uint16_t *)hh;      // [PF_ROWS][2*9216]
    const bool tcm = !prefill_f32_mode();
    if (!tcm && m->q8) { vox_set_error("prefill: VOX_CUDA_PREFILL_F32 is not available for a Q8 model"); return -1; }
    for (int r0 = 0; r0 < seq_len; r0 += PF_ROWS) {
        const int M = seq_len - r0 < PF_ROWS ? seq_len - r0 : PF_ROWS;
        float *x = x_dev + (size_t)r0 * DEC_DIM;
        const int pos0 = *pos;
        const int first_valid = pos0 - *n_valid;
        if (tcm && prefill_wt_mode()) {                     // default: skinny swap-AB GEMM with K-range splits (dec_prefill.cu)
            if (dec_prefill_block_wt(s, x, M, kbase, vbase, kv_layer, mask, pos0, first_valid)) return -1;
            *pos += M;
            *n_valid = *n_valid + M > DEC_WINDOW ? DEC_WINDOW : *n_valid + M;
            continue;
        }
        for (int l = 0; l < m->n_dec; l++) {
            const DecLayerDev &L = m->dec[l];
            float *kc = kbase + l * kv_layer, *vc = vbase + l * kv_layer;
            const float *ada = m->ada_scale + (size_t)l * DEC_DIM;
            if (tcm) {
                if (rows_rmsnorm_bf16(st, x, DEC_DIM, L.attn_norm, nullptr, xn_b, 2 * DEC_DIM, M, DEC_DIM, 1)) return -1;
                if (gemm_tc(st, M, DEC_QKV, DEC_DIM, 2 * DEC_DIM, xn_b, 2 * DEC_DIM, L.wqkv,
                            pf_epi(qkv, DEC_QKV, nullptr, 0, TC_OUT_F32, 0, L.s_wqkv))) return -1;
            } else {
                if (rows_rmsnorm(st, x, DEC_DIM, L.attn_norm, nullptr, xn, DEC_DIM, M, DEC_DIM)) return -1;
                if (sgemm_bf16w(st, M, DEC_QKV, DEC_DIM, xn, DEC_DIM, L.wqkv, qkv, DEC_QKV, nullptr, EPI_NONE)) return -1;
            }
            if (rows_rope_kv(st, qkv, M, DEC_QDIM, DEC_KVDIM, DEC_HD, m->inv_freq_dec, pos0,
                             mask, kc, vc)) return -1;
            if (rows_attention(st, DEC_HD, qkv, DEC_QKV, kc, vc, DEC_KVDIM, mask, DEC_HEADS,
                               DEC_HEADS / DEC_KVH, M, pos0, first_valid, DEC_WINDOW, att, DEC_QDIM)) return -1;
            if (tcm) {
                if (f32_to_bf16_rows(st, att, DEC_QDIM, att_b, 2 * DEC_QDIM, M, DEC_QDIM, 1)) return -1;
                if (gemm_tc(st, M, DEC_DIM, DEC_QDIM, 2 * DEC_QDIM, att_b, 2 * DEC_QDIM, L.wo,
                            pf_epi(x, DEC_DIM, nullptr, 0, TC_RESIDUAL, 0, L.s_wo))) return -1;
                if (rows_rmsnorm_bf16(st, x, DEC_DIM, L.ffn_norm, ada, xn_b, 2 * DEC_DIM, M, DEC_DIM, 1)) return -1;
                if (gemm_tc(st, M, 2 * DEC_HIDDEN, DEC_DIM, 2 * DEC_DIM, xn_b, 2 * DEC_DIM, L.w13,
                            pf_epi(nullptr, 0, hh_b, 2 * DEC_HIDDEN, TC_SWIGLU | TC_OUT_BF16 | TC_OUT_SPLIT, DEC_HIDDEN, L.s_w13))) return -1;
                if (gemm_tc(st, M, DEC_DIM, DEC_HIDDEN, 2 * DEC_HIDDEN, hh_b, 2 * DEC_HIDDEN, L.w2,
                            pf_epi(x, DEC_DIM, nullptr, 0, TC_RESIDUAL, 0, L.s_w2))) return -1;
            } else {
                if (sgemm_bf16w(st, M, DEC_DIM, DEC_QDIM, att, DEC_QDIM, L.wo, x, DEC_DIM, nullptr, EPI_RESIDUAL)) return -1;
                if (rows_rmsnorm(st, x, DEC_DIM, L.ffn_norm, ada, xn, DEC_DIM, M, DEC_DIM)) return -1;
                if (sgemm_bf16w(st, M, 2 * DEC_HIDDEN, DEC_DIM, xn, DEC_DIM, L.w13, gu, 2 * DEC_HIDDEN, nullptr, EPI_NONE)) return -1;
                if (rows_swiglu(st, gu, hh, M, DEC_HIDDEN)) return -1;
                if (sgemm_bf16w(st, M, DEC_DIM, DEC_HIDDEN, hh, DEC_HIDDEN, L.w2, x, DEC_DIM, nullptr, EPI_RESIDUAL)) return -1;
            }
        }
        *pos += M;
        *n_valid = *n_valid + M > DEC_WINDOW ? DEC_WINDOW : *n_valid + M;
    }
    return 0;
}

// prompt rows [n_dec][PF_KV_ROWS][1024] fp32 (rows 0 .. gridDim.x-1) -> slots 0 .. of the fp16 rings
__global__ void __launch_bounds__(256) kv_rows_to_half_kernel(const float *__restrict__ src_k, const float *__restrict__ src_v,
                                                              __half *__restrict__ dst_k, __half *__restrict__ dst_v) {
    const int row = blockIdx.x, l = blockIdx.y;
    const size_t so = ((size_t)l * PF_KV_ROWS + row) * DEC_KVDIM, d = ((size_t)l * DEC_WINDOW + row) * DEC_KVDIM;
    for (int i = threadIdx.x; i < DEC_KVDIM; i += 256) {
        dst_k[d + i] = __float2half_rn(src_k[so + i]);
        dst_v[d + i] = __float2half_rn(src_v[so + i]);
    }
}

static int prefill_device(vox_cuda_state *s, float *x_dev, int seq_len) {
    // a block writes all its K/V rows into the ring before its attention runs: once a row reaches logical
    // position 8191 the ring slot of the oldest key it needs has been overwritten by a later row of the block
    if (s->h_params.pos + seq_len > DEC_WINDOW) {
        vox_set_error("decoder_prefill: rows [%d,%d) reach the %d-position window; prefill only below it (step beyond)",
                      s->h_params.pos, s->h_params.pos + seq_len, DEC_WINDOW);
        return -1;
    }
    if (s->kv16) {
        // fp32 prefill into the PF_KV_ROWS-row buffer, then the rows are rounded into the fp16 ring (api_group.cu does the same)
        if (s->h_params.pos != 0 || seq_len > PF_KV_ROWS) {
            vox_set_error("decoder_prefill: an fp16-K/V state prefills at most %d rows, at position 0", PF_KV_ROWS);
            return -1;
        }
        if (dec_prefill_into(s, x_dev, seq_len, s->kv_k, s->kv_v, (size_t)PF_KV_ROWS * DEC_KVDIM, PF_KV_ROWS - 1,
                             &s->h_params.pos, &s->h_params.n_valid)) return -1;
        kv_rows_to_half_kernel<<<dim3(seq_len, s->m->n_dec), 256, 0, s->stream>>>(s->kv_k, s->kv_v, (__half *)s->kv_k16, (__half *)s->kv_v16);
        LAUNCHED();
    } else if (dec_prefill_into(s, x_dev, seq_len, s->kv_k, s->kv_v, (size_t)DEC_WINDOW * DEC_KVDIM, DEC_WINDOW - 1,
                         &s->h_params.pos, &s->h_params.n_valid)) return -1;
    if (push_params(s)) return -1;
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    return 0;
}

extern "C" int vox_cuda_decoder_prefill(vox_cuda_state_t *s, const float *x, int seq_len) {
    if (!s || !x || seq_len <= 0) return -1;
    float *xd = nullptr;
    CUDA_TRY(cudaMalloc((void **)&xd, (size_t)seq_len * DEC_DIM * 4));
    int rc = -1;
    if (cudaMemcpyAsync(xd, x, (size_t)seq_len * DEC_DIM * 4, cudaMemcpyHostToDevice, s->stream) == cudaSuccess)
        rc = prefill_device(s, xd, seq_len);
    cudaFree(xd);
    return rc;
}

// embed rows on device: x[i] = adapter[start+i] + tok_emb[tokens[i]]  (voxtral.c:1039-1048)
__global__ void __launch_bounds__(256) embed_rows_kernel(const int *__restrict__ tokens,
                                                         const float *__restrict__ adapter,
                                                         const uint16_t *__restrict__ tok_emb,
                                                         const int8_t *__restrict__ q_tok_emb,
                                                         const float *__restrict__ s_tok_emb,
                                                         float *__restrict__ x) {
    const int r = blockIdx.x;
    const float *a = adapter + (size_t)r * DEC_DIM;
    if (q_tok_emb) {                      // Q8: embedding row = q * scale (tok_embed_q8_to_f32, voxtral.c:443-452)
        const int8_t *e = q_tok_emb + (size_t)tokens[r] * DEC_DIM;
        const float sc = s_tok_emb[tokens[r]];
        const int swz = (tokens[r] & 1) << 6;             // odd rows: 64-byte halves swapped (q8_swizzle_kernel, api_core.cu)
        for (int i = threadIdx.x; i < DEC_DIM; i += 256)
            x[(size_t)r * DEC_DIM + i] = a[i] + (float)e[i ^ swz] * sc;
        return;
    }
    const uint16_t *e = tok_emb + (size_t)tokens[r] * DEC_DIM;
    for (int i = threadIdx.x; i < DEC_DIM; i += 256)
        x[(size_t)r * DEC_DIM + i] = a[i] + __uint_as_float(((uint32_t)e[i]) << 16);
}

extern "C" int vox_cuda_decoder_prefill_fused(vox_cuda_state_t *s, const int *tokens,
                                              int adapter_start, int seq_len) {
    if (!s || !tokens || seq_len <= 0) return -1;
    const int phys = adapter_start - s->adapter_offset;
    if (phys < 0 || adapter_start + seq_len > s->adapter_count) {
        vox_set_error("prefill_fused: adapter range outside buffer");
        return -1;
    }
    for (int i = 0; i < seq_len; i++) if (tokens[i] < 0 || tokens[i] >= VOCAB) return -1;
    float *xd = nullptr; int *td = nullptr;
    CUDA_TRY(cudaMalloc((void **)&xd, (size_t)seq_len * DEC_DIM * 4));
    if (cudaMalloc((void **)&td, (size_t)seq_len * 4) != cudaSuccess) { cudaFree(xd); return -1; }
    int rc = -1;
    if (cudaMemcpyAsync(td, tokens, (size_t)seq_len * 4, cudaMemcpyHostToDevice, s->stream) == cudaSuccess) {
        embed_rows_kernel<<<seq_len, 256, 0, s->stream>>>(td, s->adapter_buf + (size_t)phys * DEC_DIM,
                                                          s->m->tok_emb, s->m->q_tok_emb, s->m->s_tok_emb, xd);
        LAUNCHED();
        rc = prefill_device(s, xd, seq_len);
    }
    cudaFree(xd); cudaFree(td);
    return rc;
}

// ---- alternatives (stream_fill_alts, voxtral.c:955-1010) -------------------------
__global__ void __launch_bounds__(1024) alt_topk_kernel(const float *__restrict__ logits, int best,
                                                        int k, int *__restrict__ ids,
                                                        float *__restrict__ vals) {
    __shared__ float sv[32];
    __shared__ int si[32];
    __shared__ int picked[VOX_CUDA_MAX_ALT];
    const int t = threadIdx.x;
    for (int round = 0; round < k; round++) {
        float bv = -INFINITY; int bi = 0x7fffffff;
        for (int i = 1000 + t; i < VOCAB; i += 1024) {
            if (i == best) continue;
            bool skip = false;
            for (int j = 0; j < round; j++) skip |= (picked[j] == i);
            if (skip) continue;
            const float v = logits[i];
            if (v > bv || (v == bv && i < bi)) { bv = v; bi = i; }
        }
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if ((t & 31) == 0) { sv[t >> 5] = bv; si[t >> 5] = bi; }
        __syncthreads();
        if (t == 0) {
            for (int w = 1; w < 32; w++)
                if (sv[w] > bv || (sv[w] == bv && si[w] < bi)) { bv = sv[w]; bi = si[w]; }
            picked[round] = bi; ids[round] = bi; vals[round] = bv;
            if (round == 0) vals[k] = logits[best];
        }
        __syncthreads();
    }
}

extern "C" int vox_cuda_decoder_alternatives(vox_cuda_state_t *s, int max_alt, float cutoff, int *ids) {
    if (!s || !ids) return -1;
    if (max_alt > VOX_CUDA_MAX_ALT) max_alt = VOX_CUDA_MAX_ALT;
    const int k = max_alt - 1;
    if (k <= 0) return 0;
    int *d_ids = (int *)s->d_attn_part;            // scratch, idle between steps
    float *d_vals = (float *)(d_ids + 8);
    alt_topk_kernel<<<1, 1024, 0, s->stream>>>(s->d_logits, s->h_params.out_token, k, d_ids, d_vals);
    LAUNCHED();
    int h_ids[8]; float h_vals[8];
    CUDA_TRY(cudaMemcpyAsync(h_ids, d_ids, k * 4, cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaMemcpyAsync(h_vals, d_vals, (k + 1) * 4, cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    int n = 0;
    for (int i = 0; i < k; i++) {
        if (h_ids[i] == 0x7fffffff) break;
        // p_i / p_best = exp(l_i - l_best)
        const float r = 1.0f - expf(h_vals[i] - h_vals[k]);
        if (r > cutoff) break;
        ids[n++] = h_ids[i];
    }
    return n;
}
