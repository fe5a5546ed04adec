// api_group.cu - lock-step greedy decoding of B independent streams on one GPU
// ("within a GPU, streams are batched per decode step - weights read once for B streams").
//
// The reference decodes one stream per vox_ctx_t (vox_decoder_forward, /root/reference/
// voxtral_decoder.c:640-780); B streams cost B passes over the 6.86 GB of decoder weights.  A
// group keeps B streams' adapter rows, K/V rings and residual rows on the device and advances all
// of them by one token per step; every weight matrix crosses HBM once per step (gemm_wt.cu).  Each
// stream's arithmetic is that of vox_decoder_forward on its own rows - nothing is shared between
// streams but the weights - so the greedy ids of stream i equal those of a single-stream run.
//
// Per step and layer (7 launches):
//   group_norm   x += sum of the previous product's K-range partials (fixed order); RMSNorm (x (1+ada)
//                for the FFN norm, voxtral_decoder.c:742-745) -> split bf16 rows          [B CTAs]
//   gemm_wt      wqkv -> partials
//   group_attn   per (stream, kv head): sum partials, RoPE(pos_b) on q,k (voxtral_kernels.c:617-655),
//                append k,v to the stream's ring, online-softmax GQA attention over the ring
//                (voxtral_kernels.c:541-611) -> split bf16 rows                           [B x 8 CTAs]
//   gemm_wt      wo -> partials;  group_norm (ffn);  gemm_wt w1|w3 + SwiGLU;  gemm_wt w2 -> partials
// then group_norm (final norm), gemm_wt over the tied embedding (131072 logits per stream) and
// group_argmax (first maximum, voxtral_decoder.c:771-779), which also advances the stream's
// position / previous token on the device, so a whole clip decodes without a host round trip.
#include "vox_internal.h"
#include "gemm_wt.h"
#include "tc_ptx.cuh"
#include "group_dev.cuh"
#include "group.h"
#include <string.h>
#include <stdlib.h>
#include <math.h>

using namespace grp;

namespace {

__global__ void __launch_bounds__(256) group_norm_kernel(
    float *__restrict__ x, int mode, const float *__restrict__ part, int n_split, long split_stride,
    const float *__restrict__ adapter, long adapter_stride, const uint16_t *__restrict__ tok_emb,
    const float *__restrict__ tok_scale, const Slot *__restrict__ slots, const float *__restrict__ w, const float *__restrict__ ada,
    uint16_t *__restrict__ out) {
    __shared__ float red[8];
    tcx::pdl_launch_dependents();
    tcx::pdl_wait();
    norm_row(blockIdx.x, threadIdx.x, x, mode, part, n_split, split_stride, adapter, adapter_stride, tok_emb, tok_scale, slots, w, ada,
             out, red, SyncThreads());
}

template <class KV>
__global__ void __launch_bounds__(256, 2) group_attn_kernel(
    const float *__restrict__ part, int n_split, long split_stride, const Slot *__restrict__ slots,
    const float *__restrict__ inv_freq, KV *__restrict__ kv_k, KV *__restrict__ kv_v, size_t stream_stride,
    int ring_mask, uint16_t *__restrict__ out) {
    __shared__ __align__(16) AttnSmem sm;
    tcx::pdl_launch_dependents();
    tcx::pdl_wait();
    attn_task(blockIdx.x, blockIdx.y, threadIdx.x, part, n_split, split_stride, slots, inv_freq, kv_k, kv_v, stream_stride,
              ring_mask, out, sm, SyncThreads());
}

__global__ void __launch_bounds__(1024) group_argmax_kernel(const float *__restrict__ logits, Slot *__restrict__ slots,
                                                            int *__restrict__ tokens, int step, int n_streams) {
    __shared__ float sv[32];
    __shared__ int si[32];
    tcx::pdl_launch_dependents();
    tcx::pdl_wait();
    argmax_row<1024>(blockIdx.x, threadIdx.x, logits, slots, tokens, step, n_streams, sv, si, SyncThreads());
}

// kv16: prompt rows written by the prefill as fp32 [n_dec][128][1024] -> the stream's fp16 ring
__global__ void __launch_bounds__(256) group_kv_round_kernel(const float *__restrict__ src_k, const float *__restrict__ src_v,
                                                             __half *__restrict__ dst_k, __half *__restrict__ dst_v,
                                                             size_t kv_layer, int ring_mask) {
    const int l = blockIdx.y, row = blockIdx.x;
    const size_t s = ((size_t)l * 128 + row) * DEC_KVDIM, d = (size_t)l * kv_layer + (size_t)(row & ring_mask) * DEC_KVDIM;
    for (int i = threadIdx.x; i < DEC_KVDIM; i += 256) {
        dst_k[d + i] = __float2half_rn(src_k[s + i]);
        dst_v[d + i] = __float2half_rn(src_v[s + i]);
    }
}

__global__ void __launch_bounds__(256) group_embed_rows_kernel(const int *__restrict__ tokens, const float *__restrict__ adapter,
                                                               const uint16_t *__restrict__ tok_emb, const float *__restrict__ tok_scale,
                                                               float *__restrict__ x) {
    const int r = blockIdx.x;
    const float *a = adapter + (size_t)r * DEC_DIM;
    const uint16_t *e = tok_emb + (size_t)tokens[r] * DEC_DIM;
    const float es = tok_scale ? tok_scale[tokens[r]] : 1.0f;        // Q8 model: int-valued row * row scale
    for (int i = threadIdx.x; i < DEC_DIM; i += 256) {
        const float ev = __uint_as_float(((uint32_t)e[i]) << 16);
        x[(size_t)r * DEC_DIM + i] = a[i] + (tok_scale ? ev * es : ev);
    }
}

}  // namespace

extern "C" void vox_cuda_group_free(vox_cuda_group_t *g) {
    if (!g) return;
    cudaFree(g->kv_k); cudaFree(g->kv_v); cudaFree(g->kv_k16); cudaFree(g->kv_v16); cudaFree(g->pf_k); cudaFree(g->pf_v);
    cudaFree(g->adapter); cudaFree(g->d_slots); cudaFree(g->x);
    cudaFree(g->xn_b); cudaFree(g->att_b); cudaFree(g->hh_b); cudaFree(g->part); cudaFree(g->logits);
    cudaFree(g->d_tokens); cudaFree(g->pf_x); cudaFree(g->pf_tok);
    if (g->ev0) cudaEventDestroy(g->ev0);
    if (g->ev1) cudaEventDestroy(g->ev1);
    group_mega_free(g);
    if (g->eng) vox_cuda_state_free(g->eng);
    free(g);
}

static int galloc(void **p, size_t bytes, bool zero) {
    if (cudaMalloc(p, bytes) != cudaSuccess) { vox_set_error("group: cudaMalloc(%zu) failed", bytes); return -1; }
    if (zero && cudaMemset(*p, 0, bytes) != cudaSuccess) { vox_set_error("group: memset failed"); return -1; }
    return 0;
}

extern "C" vox_cuda_group_t *vox_cuda_group_create(vox_cuda_model_t *m, int n_streams, int max_positions) {
    return vox_cuda_group_create_ex(m, n_streams, max_positions, 32);
}

// kv_bits = 32: K/V rings in fp32 (the reference's CPU / BLAS layout: ids equal the reference);
// kv_bits = 16: fp16 rings, the reference's GPU default (voxtral.c:189-190, kv_cache_switch_to_fp32 is the
// opt-out there): half the K/V bytes per step; see DESIGN.md 4.8 for the tolerance statement.
extern "C" vox_cuda_group_t *vox_cuda_group_create_ex(vox_cuda_model_t *m, int n_streams, int max_positions, int kv_bits) {
    if (!m || n_streams < 1 || n_streams > MAX_STREAMS || max_positions < 2 || (kv_bits != 32 && kv_bits != 16)) {
        vox_set_error("group_create: need 1..%d streams, max_positions >= 2, kv_bits 32 or 16", MAX_STREAMS);
        return nullptr;
    }
    vox_cuda_group *g = (vox_cuda_group *)calloc(1, sizeof(vox_cuda_group));
    if (!g) return nullptr;
    g->m = m; g->B = n_streams; g->max_pos = max_positions; g->kv16 = kv_bits == 16;
    g->NB = n_streams <= 32 ? 32 : (n_streams <= 64 ? 64 : 128);
    int cap = 64;
    while (cap < max_positions && cap < DEC_WINDOW) cap <<= 1;      // ring of min(window, pow2 >= positions) rows
    g->cap = cap; g->mask = cap - 1;
    g->kv_layer = (size_t)cap * DEC_KVDIM;
    g->kv_stream = g->kv_layer * m->n_dec;
    g->eng = vox_cuda_state_create(m);
    bool ok = g->eng != nullptr;
    const size_t B = n_streams, NB = g->NB;
    if (g->kv16) {
        ok = ok && !galloc((void **)&g->kv_k16, g->kv_stream * B * 2, false) && !galloc((void **)&g->kv_v16, g->kv_stream * B * 2, false);
        ok = ok && !galloc((void **)&g->pf_k, (size_t)m->n_dec * 128 * DEC_KVDIM * 4, false) &&
             !galloc((void **)&g->pf_v, (size_t)m->n_dec * 128 * DEC_KVDIM * 4, false);
    } else {
        ok = ok && !galloc((void **)&g->kv_k, g->kv_stream * B * 4, false) && !galloc((void **)&g->kv_v, g->kv_stream * B * 4, false);
    }
    ok = ok && !galloc((void **)&g->adapter, B * max_positions * DEC_DIM * 4, false);
    ok = ok && !galloc((void **)&g->d_slots, B * sizeof(Slot), true);
    ok = ok && !galloc((void **)&g->x, B * DEC_DIM * 4, true);
    ok = ok && !galloc((void **)&g->xn_b, NB * 2 * DEC_DIM * 2, true) && !galloc((void **)&g->att_b, NB * 2 * DEC_QDIM * 2, true) &&
         !galloc((void **)&g->hh_b, NB * 2 * DEC_HIDDEN * 2, true);
    ok = ok && !galloc((void **)&g->part, (size_t)KS_MAX * B * DEC_QKV * 4, true);
    ok = ok && !galloc((void **)&g->logits, B * (size_t)VOCAB * 4, false);
    ok = ok && !galloc((void **)&g->pf_x, (size_t)128 * DEC_DIM * 4, false) && !galloc((void **)&g->pf_tok, 128 * 4, false);
    ok = ok && cudaEventCreate(&g->ev0) == cudaSuccess && cudaEventCreate(&g->ev1) == cudaSuccess;
    for (int l = 0; ok && l < m->n_dec; l++) {
        const DecLayerDev &L = m->dec[l];
        ok = ok && !gemm_wt_prepare(&g->op_qkv[l], g->NB, g->B, DEC_QKV, DEC_DIM, 1, g->xn_b, 2 * DEC_DIM, L.wqkv, KS_QKV);
        ok = ok && !gemm_wt_prepare(&g->op_wo[l], g->NB, g->B, DEC_DIM, DEC_QDIM, 1, g->att_b, 2 * DEC_QDIM, L.wo, KS_WO);
        ok = ok && !gemm_wt_prepare(&g->op_w13[l], g->NB, g->B, 2 * DEC_HIDDEN, DEC_DIM, 1, g->xn_b, 2 * DEC_DIM, L.w13, 1);
        ok = ok && !gemm_wt_prepare(&g->op_w2[l], g->NB, g->B, DEC_DIM, DEC_HIDDEN, 1, g->hh_b, 2 * DEC_HIDDEN, L.w2, KS_W2);
        if (!ok) break;
        g->op_qkv[l].out_f32 = g->part; g->op_qkv[l].ldo = DEC_QKV; g->op_qkv[l].split_stride = (long)B * DEC_QKV;
        g->op_wo[l].out_f32 = g->part; g->op_wo[l].ldo = DEC_DIM; g->op_wo[l].split_stride = (long)B * DEC_DIM;
        g->op_w13[l].mode = WT_OUT_SWIGLU_BF16X2; g->op_w13[l].out_bf16 = g->hh_b; g->op_w13[l].ldb = 2 * DEC_HIDDEN; g->op_w13[l].lo_off = DEC_HIDDEN;
        g->op_w2[l].out_f32 = g->part; g->op_w2[l].ldo = DEC_DIM; g->op_w2[l].split_stride = (long)B * DEC_DIM;
        // Q8 model: the weights are the int8 values as bf16, the row scale goes into the GEMM epilogue (null otherwise)
        g->op_qkv[l].row_scale = L.s_wqkv; g->op_wo[l].row_scale = L.s_wo; g->op_w13[l].row_scale = L.s_w13; g->op_w2[l].row_scale = L.s_w2;
    }
    if (ok) {
        ok = !gemm_wt_prepare(&g->op_logits, g->NB, g->B, VOCAB, DEC_DIM, 1, g->xn_b, 2 * DEC_DIM, m->tok_emb, 1);
        g->op_logits.out_f32 = g->logits; g->op_logits.ldo = VOCAB; g->op_logits.split_stride = 0;
        g->op_logits.row_scale = m->s_tok_emb;
    }
    if (!ok) { vox_cuda_group_free(g); return nullptr; }
    if (!g->kv16) group_mega_init(g);   // persistent step kernel, opt-in (VOX_CUDA_GROUP_MEGA=1, fp32 rings only); default: kernel per op + PDL
    return g;
}

extern "C" int vox_cuda_group_streams(const vox_cuda_group_t *g) { return g ? g->B : -1; }

// whole-clip front end of stream i: PCM -> mel -> encoder -> adapter rows, kept on the device
extern "C" int vox_cuda_group_encode(vox_cuda_group_t *g, int stream, const float *pcm, int n_samples, int n_frames) {
    if (!g || stream < 0 || stream >= g->B || !pcm || n_frames <= 0) return -1;
    if (batch_encode_clip(g->eng, pcm, n_samples, n_frames)) return -1;
    const int T = g->eng->adapter_count;
    if (T > g->max_pos) { vox_set_error("group_encode: %d adapter tokens > max_positions %d", T, g->max_pos); return -1; }
    CUDA_TRY(cudaMemcpyAsync(g->adapter + (size_t)stream * g->max_pos * DEC_DIM, g->eng->adapter_buf,
                             (size_t)T * DEC_DIM * 4, cudaMemcpyDeviceToDevice, g->eng->stream));
    g->count[stream] = T;
    return T;
}

// adapter rows computed elsewhere (e.g. by a streaming state): host fp32 [n][3072]
extern "C" int vox_cuda_group_set_adapter(vox_cuda_group_t *g, int stream, const float *rows, int n) {
    if (!g || stream < 0 || stream >= g->B || !rows || n < 1 || n > g->max_pos) return -1;
    CUDA_TRY(cudaMemcpyAsync(g->adapter + (size_t)stream * g->max_pos * DEC_DIM, rows, (size_t)n * DEC_DIM * 4,
                             cudaMemcpyHostToDevice, g->eng->stream));
    CUDA_TRY(cudaStreamSynchronize(g->eng->stream));
    g->count[stream] = n;
    return n;
}

static int group_step(vox_cuda_group *g, int step) {
    vox_cuda_model *m = g->m;
    cudaStream_t st = g->eng->stream;
    const int B = g->B;
    const long astride = (long)g->max_pos * DEC_DIM;
    for (int l = 0; l < m->n_dec; l++) {
        const DecLayerDev &L = m->dec[l];
        if (l == 0)
            CUDA_TRY(launch_pdl(group_norm_kernel, B, 256, 0, st, g->x, 0, nullptr, 0, 0L, g->adapter, astride, m->tok_emb, (const float *)(m->q8 ? m->s_tok_emb : nullptr), g->d_slots,
                                L.attn_norm, nullptr, g->xn_b));
        else
            CUDA_TRY(launch_pdl(group_norm_kernel, B, 256, 0, st, g->x, 1, g->part, KS_W2, (long)B * DEC_DIM, nullptr, 0L, nullptr, (const float *)nullptr, g->d_slots,
                                L.attn_norm, nullptr, g->xn_b));
        LAUNCHED();
        if (gemm_wt_launch(st, &g->op_qkv[l])) return -1;
        if (g->kv16)
            CUDA_TRY(launch_pdl(group_attn_kernel<__half>, dim3(B, DEC_KVH), 256, 0, st, (const float *)g->part, KS_QKV, (long)B * DEC_QKV,
                                (const Slot *)g->d_slots, (const float *)m->inv_freq_dec, g->kv_k16 + l * g->kv_layer,
                                g->kv_v16 + l * g->kv_layer, g->kv_stream, g->mask, g->att_b));
        else
            CUDA_TRY(launch_pdl(group_attn_kernel<float>, dim3(B, DEC_KVH), 256, 0, st, (const float *)g->part, KS_QKV, (long)B * DEC_QKV,
                                (const Slot *)g->d_slots, (const float *)m->inv_freq_dec, g->kv_k + l * g->kv_layer,
                                g->kv_v + l * g->kv_layer, g->kv_stream, g->mask, g->att_b));
        LAUNCHED();
        if (gemm_wt_launch(st, &g->op_wo[l])) return -1;
        CUDA_TRY(launch_pdl(group_norm_kernel, B, 256, 0, st, g->x, 1, g->part, KS_WO, (long)B * DEC_DIM, nullptr, 0L, nullptr, (const float *)nullptr, g->d_slots,
                            L.ffn_norm, m->ada_scale + (size_t)l * DEC_DIM, g->xn_b));
        LAUNCHED();
        if (gemm_wt_launch(st, &g->op_w13[l])) return -1;
        if (gemm_wt_launch(st, &g->op_w2[l])) return -1;
    }
    CUDA_TRY(launch_pdl(group_norm_kernel, B, 256, 0, st, g->x, 1, g->part, KS_W2, (long)B * DEC_DIM, nullptr, 0L, nullptr, (const float *)nullptr, g->d_slots,
                        m->dec_norm, nullptr, g->xn_b));
    LAUNCHED();
    if (gemm_wt_launch(st, &g->op_logits)) return -1;
    CUDA_TRY(launch_pdl(group_argmax_kernel, B, 1024, 0, st, g->logits, g->d_slots, g->d_tokens, step, B));
    LAUNCHED();
    return 0;
}

// Prompt prefill of every encoded stream (BOS + streaming pads, voxtral.c:1039-1051) and lock-step
// greedy decoding of all remaining adapter positions.  tokens_out: [n_streams][max_tokens] ids,
// counts_out: [n_streams] ids written per stream (T_i - (prompt_len - 1), capped at max_tokens).
// Returns the number of lock-step decode steps, or -1.  device_ms (optional): CUDA-event time.
extern "C" int vox_cuda_group_decode(vox_cuda_group_t *g, int prompt_len, int *tokens_out, int max_tokens,
                                     int *counts_out, float *device_ms) {
    if (!g || !tokens_out || !counts_out || prompt_len < 2 || prompt_len > 128 || max_tokens < 1) return -1;
    vox_cuda_model *m = g->m;
    cudaStream_t st = g->eng->stream;
    const int B = g->B, n_pre = prompt_len - 1;
    int prompt[128];
    prompt[0] = 1;
    for (int i = 1; i < prompt_len; i++) prompt[i] = 32;
    Slot h_slots[MAX_STREAMS];
    int n_steps = 0;
    for (int b = 0; b < B; b++) {
        const int T = g->count[b];
        if (T && T < prompt_len) { vox_set_error("group_decode: stream %d has %d adapter tokens (< prompt %d)", b, T, prompt_len); return -1; }
        int n = T ? T - n_pre : 0;
        if (n > max_tokens) n = max_tokens;
        counts_out[b] = n;
        if (n > n_steps) n_steps = n;
        h_slots[b].pos = n_pre; h_slots[b].prev = prompt[n_pre]; h_slots[b].aidx = T ? n_pre : 0; h_slots[b].remaining = n;
    }
    if (n_steps == 0) return 0;
    if ((size_t)n_steps * B > g->tok_cap) {
        cudaFree(g->d_tokens); g->d_tokens = nullptr; g->tok_cap = 0;
        if (galloc((void **)&g->d_tokens, (size_t)n_steps * B * 4, false)) return -1;
        g->tok_cap = (size_t)n_steps * B;
    }
    CUDA_TRY(cudaEventRecord(g->ev0, st));
    CUDA_TRY(cudaMemcpyAsync(g->pf_tok, prompt, (size_t)n_pre * 4, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(g->d_slots, h_slots, (size_t)B * sizeof(Slot), cudaMemcpyHostToDevice, st));
    for (int b = 0; b < B; b++) {
        float *kb = g->kv16 ? g->pf_k : g->kv_k + (size_t)b * g->kv_stream;
        float *vb = g->kv16 ? g->pf_v : g->kv_v + (size_t)b * g->kv_stream;
        if (!g->count[b]) {
            // idle slot: one zero K/V row per layer so that its (ignored) attention reads initialised memory
            for (int l = 0; l < m->n_dec; l++) {
                if (g->kv16) {
                    CUDA_TRY(cudaMemsetAsync(g->kv_k16 + (size_t)b * g->kv_stream + l * g->kv_layer, 0, (size_t)(n_pre + 1) * DEC_KVDIM * 2, st));
                    CUDA_TRY(cudaMemsetAsync(g->kv_v16 + (size_t)b * g->kv_stream + l * g->kv_layer, 0, (size_t)(n_pre + 1) * DEC_KVDIM * 2, st));
                } else {
                    CUDA_TRY(cudaMemsetAsync(kb + l * g->kv_layer, 0, (size_t)(n_pre + 1) * DEC_KVDIM * 4, st));
                    CUDA_TRY(cudaMemsetAsync(vb + l * g->kv_layer, 0, (size_t)(n_pre + 1) * DEC_KVDIM * 4, st));
                }
            }
            continue;
        }
        group_embed_rows_kernel<<<n_pre, 256, 0, st>>>(g->pf_tok, g->adapter + (size_t)b * g->max_pos * DEC_DIM, m->tok_emb, m->q8 ? m->s_tok_emb : nullptr, g->pf_x);
        LAUNCHED();
        int pos = 0, n_valid = 0;
        if (g->kv16) {
            // the prompt rows attend to each other in fp32 (38 rows; nothing is re-read from the ring), then are rounded once
            if (dec_prefill_into(g->eng, g->pf_x, n_pre, kb, vb, (size_t)128 * DEC_KVDIM, 127, &pos, &n_valid)) return -1;
            group_kv_round_kernel<<<dim3(n_pre, m->n_dec), 256, 0, st>>>(kb, vb, g->kv_k16 + (size_t)b * g->kv_stream,
                                                                         g->kv_v16 + (size_t)b * g->kv_stream, g->kv_layer, g->mask);
            LAUNCHED();
        } else if (dec_prefill_into(g->eng, g->pf_x, n_pre, kb, vb, g->kv_layer, g->mask, &pos, &n_valid)) return -1;
    }
    if (g->mega) {
        if (group_mega_run(g, n_steps)) return -1;
    } else {
        for (int step = 0; step < n_steps; step++)
            if (group_step(g, step)) return -1;
    }
    CUDA_TRY(cudaEventRecord(g->ev1, st));
    int *h_tok = (int *)malloc((size_t)n_steps * B * 4);
    if (!h_tok) return -1;
    if (cudaMemcpyAsync(h_tok, g->d_tokens, (size_t)n_steps * B * 4, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
        cudaStreamSynchronize(st) != cudaSuccess) {
        const unsigned *rec = g->mega ? group_mega_error_record(g) : nullptr;
        vox_set_error("group_decode: %s (watchdog code %u arg %u cta %u thread %u)", cudaGetErrorString(cudaGetLastError()),
                      rec ? rec[0] : 0u, rec ? rec[1] : 0u, rec ? rec[2] : 0u, rec ? rec[3] : 0u);
        free(h_tok);
        return -1;
    }
    if (g->mega) group_mega_dump_prof(g, n_steps);
    for (int b = 0; b < B; b++)
        for (int i = 0; i < counts_out[b]; i++) tokens_out[(size_t)b * max_tokens + i] = h_tok[(size_t)i * B + b];
    free(h_tok);
    if (device_ms) cudaEventElapsedTime(device_ms, g->ev0, g->ev1);
    return n_steps;
}
