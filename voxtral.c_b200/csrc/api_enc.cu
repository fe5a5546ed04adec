// api_enc.cu - encoder side of the C ABI: incremental conv stem, 32-layer
// causal encoder against a device KV ring, 4-alignment residual, adapter, and
// the device adapter-token buffer the decoder reads from.
//
// This file holds the PRECISE mode (fp32 activations, CUDA-core GEMM), whose
// arithmetic is the reference's up to summation order; the tcgen05 throughput
// mode lives in enc_fast.cu and shares all of the state below.
#include "vox_internal.h"
#include "rows.h"
#include "enc_state.h"
#include "gemm_tc.h"
#include <stdlib.h>
#include <string.h>

void vox_mem_account(long delta);
const uint16_t *vox_model_lookup(vox_cuda_model *m, const void *host, size_t bytes);

static int dm(void **p, size_t bytes) {
    if (cudaMalloc(p, bytes) != cudaSuccess) { vox_set_error("cudaMalloc(%zu MB) failed", bytes >> 20); return -1; }
    vox_mem_account((long)bytes);
    return 0;
}

int enc_state_create(vox_cuda_state *s) {
    EncState *e = (EncState *)calloc(1, sizeof(EncState));
    s->enc_priv = e;
    const int n_enc = s->m->n_enc;
    const size_t kv = (size_t)(n_enc > 0 ? n_enc : 1) * ENC_RING * ENC_QKV * 4;
    const size_t R = ENC_CHUNK + 4;
    if (dm((void **)&e->kv_k, kv) || dm((void **)&e->kv_v, kv)) return -1;
    e->kv_bytes = kv;
    if (dm((void **)&e->mel_in, (size_t)(2 + ENC_MAXMEL) * VOX_CUDA_MEL_BINS * 4)) return -1;
    if (dm((void **)&e->c0, (size_t)(2 + ENC_MAXMEL) * ENC_DIM * 4)) return -1;
    if (dm((void **)&e->x, R * ENC_DIM * 4) || dm((void **)&e->xn, R * ENC_DIM * 4)) return -1;
    if (dm((void **)&e->qkv, R * 3 * ENC_QKV * 4) || dm((void **)&e->att, R * ENC_QKV * 4)) return -1;
    if (dm((void **)&e->gu, R * 2 * ENC_HIDDEN * 4) || dm((void **)&e->hh, R * ENC_HIDDEN * 4)) return -1;
    if (dm((void **)&e->encout, (R + 4) * ENC_DIM * 4)) return -1;
    if (dm((void **)&e->mid, (R / 4 + 2) * DEC_DIM * 4)) return -1;
    if (cudaMallocHost((void **)&e->h_stage, (size_t)ENC_MAXMEL * ENC_DIM * 4 / 2 + 4096) != cudaSuccess) return -1;
    return 0;
}

void enc_fast_free(vox_cuda_state *s);
void enc_small_free(void *p);
int enc_layers_small(vox_cuda_state *s, float *x, int M, float *out);
// VOX_CUDA_ENC_SMALL=0 keeps streaming chunks (M <= 128) on the large-tile GEMM (A/B runs)
static int enc_small_mode(void) {
    static int v = -1;
    if (v < 0) { const char *e = getenv("VOX_CUDA_ENC_SMALL"); v = (e && !atoi(e)) ? 0 : 1; }
    return v;
}
int adapter_fast(vox_cuda_state *s, const float *enc_rows, int k_tokens, float *out);
static int enc_precise_mode(void);

void enc_state_free(vox_cuda_state *s) {
    EncState *e = (EncState *)s->enc_priv;
    if (!e) return;
    enc_fast_free(s);
    enc_small_free(e->small);
    void *ptrs[] = {e->kv_k, e->kv_v, e->mel_in, e->c0, e->x, e->xn, e->qkv, e->att, e->gu, e->hh, e->encout, e->mid,
                    e->mel_hi, e->mel_lo, e->c0_hi, e->c0_lo, e->attn_part, e->attn_cnt};
    for (void *p : ptrs) if (p) cudaFree(p);
    vox_mem_account(-(long)(2 * e->kv_bytes));
    if (e->h_stage) cudaFreeHost(e->h_stage);
    free(e);
    s->enc_priv = nullptr;
}

extern "C" void vox_cuda_state_reset_encoder(vox_cuda_state_t *s) {
    if (!s) return;
    EncState *e = (EncState *)s->enc_priv;
    e->pos = 0; e->pos_min = 0;
    e->stem_init = 0; e->c0_resid = 0; e->enc_resid = 0;
    // zero histories: rows 0..1 of mel_in, rows 0..1 of c0 (voxtral.c:807-817)
    cudaMemsetAsync(e->mel_in, 0, 2 * VOX_CUDA_MEL_BINS * 4, s->stream);
    cudaMemsetAsync(e->c0, 0, 2 * ENC_DIM * 4, s->stream);
    s->adapter_count = 0; s->adapter_offset = 0;
}

// ---- conv stem (stream_conv_stem, voxtral.c:581-759) -----------------------------
// mel rows are already at e->mel_in + 2 rows.  Produces *out_len rows in e->x.
__global__ void stem_shift_rows_kernel(float *buf, int width, int src0, int src1, int copy1) {
    // row0 <- row src0 ; row1 <- row src1 (if copy1).  Sources are rows >= 2.
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < width; i += gridDim.x * blockDim.x) {
        buf[i] = buf[(size_t)src0 * width + i];
        if (copy1) buf[width + i] = buf[(size_t)src1 * width + i];
    }
}
__global__ void stem_mel_hist_one_kernel(float *mel_in) {
    // n_mel == 1: history becomes [0, new]  (voxtral.c:613-618 with tc = 1)
    const int i = threadIdx.x;
    mel_in[i] = 0.f;
    mel_in[VOX_CUDA_MEL_BINS + i] = mel_in[2 * VOX_CUDA_MEL_BINS + i];
}

// bf16 planes for the tensor-core stem (allocated on first use)
static int stem_planes(EncState *e) {
    if (e->mel_hi) return 0;
    const size_t mel_b = (size_t)(2 + ENC_MAXMEL) * VOX_CUDA_MEL_BINS * 2, c0_b = (size_t)(2 + ENC_MAXMEL) * ENC_DIM * 2;
    if (cudaMalloc((void **)&e->mel_hi, mel_b) != cudaSuccess || cudaMalloc((void **)&e->mel_lo, mel_b) != cudaSuccess ||
        cudaMalloc((void **)&e->c0_hi, c0_b) != cudaSuccess || cudaMalloc((void **)&e->c0_lo, c0_b) != cudaSuccess) {
        vox_set_error("conv stem: plane buffers alloc failed");
        return -1;
    }
    vox_mem_account((long)(2 * mel_b + 2 * c0_b));
    return 0;
}

static TcEpilogue stem_epi(float *out, const float *bias) {
    TcEpilogue t;
    t.out_f32 = out; t.ldo_f32 = ENC_DIM; t.out_bf16 = nullptr; t.ldo_bf16 = 0; t.bias = bias;
    t.flags = TC_BIAS | TC_GELU | TC_OUT_F32; t.split_off = 0;
    return t;
}

static int stem_device(vox_cuda_state *s, int n_mel, int *out_len) {
    EncState *e = (EncState *)s->enc_priv;
    vox_cuda_model *m = s->m;
    cudaStream_t st = s->stream;
    *out_len = 0;
    if (n_mel <= 0) return 0;
    const int prev_res = e->c0_resid;
    // Both convolutions are GEMMs over overlapping [3 x channels] row windows (no im2col copy).  Default:
    // tcgen05 GEMM on split-bf16 planes of the rows (fp32-class); precise mode: fp32 CUDA-core GEMM.
    const bool tcm = !enc_precise_mode();
    if (tcm && stem_planes(e)) return -1;
    // conv0 (k=3, s=1, causal): out row t = gelu(W0 . mel[t-2..t] + b0); A rows overlap (lda = 128)
    if (tcm) {
        if (f32_to_bf16_planes(st, e->mel_in, e->mel_hi, e->mel_lo, (size_t)(n_mel + 2) * VOX_CUDA_MEL_BINS)) return -1;
        if (gemm_tc_planes(st, n_mel, ENC_DIM, 3 * VOX_CUDA_MEL_BINS, e->mel_hi, e->mel_lo, VOX_CUDA_MEL_BINS, m->conv0_w,
                           stem_epi(e->c0 + (size_t)(1 + prev_res) * ENC_DIM, m->conv0_b))) return -1;
    } else if (sgemm_bf16w(st, n_mel, ENC_DIM, 3 * VOX_CUDA_MEL_BINS, e->mel_in, VOX_CUDA_MEL_BINS, m->conv0_w,
                           e->c0 + (size_t)(1 + prev_res) * ENC_DIM, ENC_DIM, m->conv0_b, EPI_BIAS | EPI_GELU)) return -1;
    if (n_mel >= 2) {
        stem_shift_rows_kernel<<<1, 128, 0, st>>>(e->mel_in, VOX_CUDA_MEL_BINS, n_mel, n_mel + 1, 1);
    } else {
        stem_mel_hist_one_kernel<<<1, VOX_CUDA_MEL_BINS, 0, st>>>(e->mel_in);
    }
    LAUNCHED();
    e->stem_init = 1;
    // parity alignment: conv1 (stride 2) only ever sees an even number of rows
    const int total = prev_res + n_mel;
    const int new_res = total & 1;
    const int feed_total = total - new_res;
    if (feed_total <= 0) { e->c0_resid = new_res; return 0; }  // the lone row already sits at row 1
    // conv1 (k=3, s=2): out row j = gelu(W1 . c0[2j-1 .. 2j+1] + b1), row 0 of c0 is the history row
    const int L = feed_total / 2;
    if (tcm) {
        if (f32_to_bf16_planes(st, e->c0, e->c0_hi, e->c0_lo, (size_t)(feed_total + 1) * ENC_DIM)) return -1;
        if (gemm_tc_planes(st, L, ENC_DIM, 3 * ENC_DIM, e->c0_hi, e->c0_lo, 2 * ENC_DIM, m->conv1_w,
                           stem_epi(e->x, m->conv1_b))) return -1;
    } else if (sgemm_bf16w(st, L, ENC_DIM, 3 * ENC_DIM, e->c0, 2 * ENC_DIM, m->conv1_w, e->x, ENC_DIM,
                           m->conv1_b, EPI_BIAS | EPI_GELU)) return -1;
    stem_shift_rows_kernel<<<8, 256, 0, st>>>(e->c0, ENC_DIM, feed_total, feed_total + 1, new_res);
    LAUNCHED();
    e->c0_resid = new_res;
    *out_len = L;
    CUDA_TRY(cudaGetLastError());
    return 0;
}

// ---- transformer layers, precise mode (voxtral_encoder.c:562-686) -----------------
// x: [M, 1280] device, updated in place; final-normed rows are written to out.
int enc_layers_precise(vox_cuda_state *s, float *x, int M, float *out) {
    EncState *e = (EncState *)s->enc_priv;
    vox_cuda_model *m = s->m;
    cudaStream_t st = s->stream;
    const size_t kv_layer = (size_t)ENC_RING * ENC_QKV;
    const int pos0 = e->pos;
    for (int l = 0; l < m->n_enc; l++) {
        const EncLayerDev &L = m->enc[l];
        float *kc = e->kv_k + l * kv_layer, *vc = e->kv_v + l * kv_layer;
        if (rows_rmsnorm(st, x, ENC_DIM, L.attn_norm, nullptr, e->xn, ENC_DIM, M, ENC_DIM)) return -1;
        if (sgemm_bf16w(st, M, 3 * ENC_QKV, ENC_DIM, e->xn, ENC_DIM, L.wqkv, e->qkv, 3 * ENC_QKV, L.bqkv, EPI_BIAS)) return -1;
        if (rows_rope_kv(st, e->qkv, M, ENC_QKV, ENC_QKV, ENC_HD, m->inv_freq_enc, pos0, ENC_RING - 1, kc, vc)) return -1;
        if (rows_attention(st, ENC_HD, e->qkv, 3 * ENC_QKV, kc, vc, ENC_QKV, ENC_RING - 1, ENC_HEADS, 1, M,
                           pos0, e->pos_min, ENC_WINDOW, e->att, ENC_QKV)) return -1;
        if (sgemm_bf16w(st, M, ENC_DIM, ENC_QKV, e->att, ENC_QKV, L.wo, x, ENC_DIM, L.bo, EPI_BIAS | EPI_RESIDUAL)) return -1;
        if (rows_rmsnorm(st, x, ENC_DIM, L.ffn_norm, nullptr, e->xn, ENC_DIM, M, ENC_DIM)) return -1;
        if (sgemm_bf16w(st, M, 2 * ENC_HIDDEN, ENC_DIM, e->xn, ENC_DIM, L.w13, e->gu, 2 * ENC_HIDDEN, nullptr, EPI_NONE)) return -1;
        if (rows_swiglu(st, e->gu, e->hh, M, ENC_HIDDEN)) return -1;
        if (sgemm_bf16w(st, M, ENC_DIM, ENC_HIDDEN, e->hh, ENC_HIDDEN, L.w2, x, ENC_DIM, L.b2, EPI_BIAS | EPI_RESIDUAL)) return -1;
    }
    if (rows_rmsnorm(st, x, ENC_DIM, m->enc_norm, nullptr, out, ENC_DIM, M, ENC_DIM)) return -1;
    e->pos += M;
    return 0;
}

int enc_layers_fast(vox_cuda_state *s, float *x, int M, float *out);  // enc_fast.cu

static int enc_precise_mode(void) {
    static int mode = -1;
    if (mode < 0) {
        const char *env = getenv("VOX_CUDA_ENC_PRECISE");
        mode = (env && atoi(env)) ? 1 : 0;
    }
    return mode;
}

static int enc_layers(vox_cuda_state *s, float *x, int M, float *out) {
    if (enc_precise_mode() && s->m->q8) { vox_set_error("encoder: VOX_CUDA_ENC_PRECISE is not available for a Q8 model"); return -1; }
    if (enc_precise_mode()) return enc_layers_precise(s, x, M, out);
    return (M <= 128 && enc_small_mode()) ? enc_layers_small(s, x, M, out) : enc_layers_fast(s, x, M, out);
}

// ---- adapter (vox_adapter_forward, voxtral_encoder.c:699-737) -----------------------
static int adapter_device(vox_cuda_state *s, const float *enc_rows, int k_tokens, float *mid, float *out) {
    vox_cuda_model *m = s->m;
    if (!enc_precise_mode() || m->q8) return adapter_fast(s, enc_rows, k_tokens, out);
    if (sgemm_bf16w(s->stream, k_tokens, DEC_DIM, 4 * ENC_DIM, enc_rows, 4 * ENC_DIM, m->adapter_w0, mid,
                    DEC_DIM, nullptr, EPI_GELU)) return -1;
    return sgemm_bf16w(s->stream, k_tokens, DEC_DIM, DEC_DIM, mid, DEC_DIM, m->adapter_w1, out, DEC_DIM,
                       nullptr, EPI_NONE);
}

__global__ void move_rows_kernel(float *dst, const float *src, size_t n) {
    // forward copy of NON-overlapping ranges (the <= 3 leftover rows of encode_piece)
    for (size_t i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
}

// One internal piece: n_mel <= ENC_MAXMEL mel rows already on the device.
static int encode_piece(vox_cuda_state *s, int n_mel) {
    EncState *e = (EncState *)s->enc_priv;
    int L = 0;
    if (stem_device(s, n_mel, &L)) return -1;
    if (L <= 0) return 0;
    // final-normed rows go behind the 0..3 residual rows (voxtral.c:868-891)
    float *out = e->encout + (size_t)e->enc_resid * ENC_DIM;
    if (enc_layers(s, e->x, L, out)) return -1;
    const int total = e->enc_resid + L;
    const int usable = total / 4 * 4, leftover = total - usable;
    if (usable > 0) {
        const int k = usable / 4;
        if (s->adapter_count - s->adapter_offset + k > s->adapter_cap) {
            vox_set_error("adapter buffer full (%d tokens); call vox_cuda_adapter_discard_before", s->adapter_cap);
            return -1;
        }
        float *dst = s->adapter_buf + (size_t)(s->adapter_count - s->adapter_offset) * DEC_DIM;
        if (adapter_device(s, e->encout, k, e->mid, dst)) return -1;
        s->adapter_count += k;
    }
    if (leftover > 0 && usable > 0) {
        move_rows_kernel<<<1, 1024, 0, s->stream>>>(e->encout, e->encout + (size_t)usable * ENC_DIM,
                                                    (size_t)leftover * ENC_DIM);
        LAUNCHED();
    }
    e->enc_resid = leftover;
    return 0;
}

extern "C" int vox_cuda_encode_chunk(vox_cuda_state_t *s, const float *mel, int n_mel) {
    if (!s || (!mel && n_mel > 0) || n_mel < 0) return -1;
    EncState *e = (EncState *)s->enc_priv;
    for (int off = 0; off < n_mel; off += ENC_MAXMEL) {
        const int n = n_mel - off < ENC_MAXMEL ? n_mel - off : ENC_MAXMEL;
        const size_t bytes = (size_t)n * VOX_CUDA_MEL_BINS * 4;
        CUDA_TRY(cudaStreamSynchronize(s->stream));  // h_stage reuse
        memcpy(e->h_stage, mel + (size_t)off * VOX_CUDA_MEL_BINS, bytes);
        CUDA_TRY(cudaMemcpyAsync(e->mel_in + 2 * VOX_CUDA_MEL_BINS, e->h_stage, bytes, cudaMemcpyHostToDevice, s->stream));
        if (encode_piece(s, n)) return -1;
    }
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    return s->adapter_count;
}

// mel already on the device (written by the mel kernel into e->mel_in + 2 rows)
int encode_chunk_device_mel(vox_cuda_state *s, int n_mel) {
    if (n_mel > ENC_MAXMEL) return -1;
    return encode_piece(s, n_mel);
}

extern "C" int vox_cuda_adapter_count(const vox_cuda_state_t *s) { return s ? s->adapter_count : -1; }

extern "C" int vox_cuda_adapter_read(vox_cuda_state_t *s, int start, int n, float *out) {
    if (!s || !out || n < 0 || start < s->adapter_offset || start + n > s->adapter_count) return -1;
    if (!n) return 0;
    CUDA_TRY(cudaMemcpyAsync(out, s->adapter_buf + (size_t)(start - s->adapter_offset) * DEC_DIM,
                             (size_t)n * DEC_DIM * 4, cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    return n;
}

extern "C" int vox_cuda_adapter_append(vox_cuda_state_t *s, const float *tok, int n) {
    if (!s || !tok || n < 0) return -1;
    if (s->adapter_count - s->adapter_offset + n > s->adapter_cap) { vox_set_error("adapter buffer full"); return -1; }
    CUDA_TRY(cudaMemcpyAsync(s->adapter_buf + (size_t)(s->adapter_count - s->adapter_offset) * DEC_DIM, tok,
                             (size_t)n * DEC_DIM * 4, cudaMemcpyHostToDevice, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    s->adapter_count += n;
    return s->adapter_count;
}

extern "C" int vox_cuda_adapter_discard_before(vox_cuda_state_t *s, int keep_from) {
    if (!s) return -1;
    if (keep_from > s->adapter_count) keep_from = s->adapter_count;
    const int consumed = keep_from - s->adapter_offset;
    if (consumed <= 0) return 0;
    const int remaining = s->adapter_count - keep_from;
    // Slide the kept rows to the front.  Pieces of at most `consumed` rows never overlap their own
    // destination, and stream order keeps piece i+1 (which reads what piece i did not write) behind
    // piece i; when that would be many tiny copies the rows go through a scratch buffer instead.
    if (remaining > 0) {
        const size_t row = (size_t)DEC_DIM * 4;
        if ((remaining + consumed - 1) / consumed <= 8) {
            for (int r = 0; r < remaining; r += consumed) {
                const int n = remaining - r < consumed ? remaining - r : consumed;
                CUDA_TRY(cudaMemcpyAsync(s->adapter_buf + (size_t)r * DEC_DIM, s->adapter_buf + (size_t)(r + consumed) * DEC_DIM,
                                         n * row, cudaMemcpyDeviceToDevice, s->stream));
            }
        } else {
            float *tmp = nullptr;
            CUDA_TRY(cudaMalloc((void **)&tmp, remaining * row));
            cudaError_t e1 = cudaMemcpyAsync(tmp, s->adapter_buf + (size_t)consumed * DEC_DIM, remaining * row,
                                             cudaMemcpyDeviceToDevice, s->stream);
            cudaError_t e2 = cudaMemcpyAsync(s->adapter_buf, tmp, remaining * row, cudaMemcpyDeviceToDevice, s->stream);
            cudaStreamSynchronize(s->stream);
            cudaFree(tmp);
            if (e1 != cudaSuccess || e2 != cudaSuccess) { vox_set_error("adapter_discard_before: copy failed"); return -1; }
        }
    }
    s->adapter_offset += consumed;
    return 0;
}

// Host-owned encoder position (voxtral.h:218-224: enc_kv_pos_offset + enc_kv_cache_len).  The ring is
// addressed by logical position, so honouring a host reset or compaction is this one integer.
extern "C" int vox_cuda_encoder_position(const vox_cuda_state_t *s) {
    return s ? ((EncState *)s->enc_priv)->pos : -1;
}
extern "C" int vox_cuda_encoder_set_position(vox_cuda_state_t *s, int logical_pos) {
    if (!s || logical_pos < 0) return -1;
    EncState *e = (EncState *)s->enc_priv;
    if (logical_pos != e->pos) { e->pos = logical_pos; e->pos_min = logical_pos; }   // a jump = an empty cache at that position
    return 0;
}
// conv-stem tails only (conv_stem_initialized = 0, voxtral.c:807-817); the K/V ring position and the
// adapter buffer are left alone
extern "C" void vox_cuda_state_reset_stem(vox_cuda_state_t *s) {
    if (!s) return;
    EncState *e = (EncState *)s->enc_priv;
    e->stem_init = 0; e->c0_resid = 0;
    cudaMemsetAsync(e->mel_in, 0, 2 * VOX_CUDA_MEL_BINS * 4, s->stream);
    cudaMemsetAsync(e->c0, 0, 2 * ENC_DIM * 4, s->stream);
}

// ---- host-pointer variants of the individual stages (reference call sites) -------
extern "C" int vox_cuda_conv_stem(vox_cuda_state_t *s, const float *mel, int n_mel, float *out, int *out_len) {
    if (!s || !mel || !out || !out_len) return -1;
    EncState *e = (EncState *)s->enc_priv;
    *out_len = 0;
    int written = 0;
    for (int off = 0; off < n_mel; off += ENC_MAXMEL) {
        const int n = n_mel - off < ENC_MAXMEL ? n_mel - off : ENC_MAXMEL;
        CUDA_TRY(cudaMemcpyAsync(e->mel_in + 2 * VOX_CUDA_MEL_BINS, mel + (size_t)off * VOX_CUDA_MEL_BINS,
                                 (size_t)n * VOX_CUDA_MEL_BINS * 4, cudaMemcpyHostToDevice, s->stream));
        int L = 0;
        if (stem_device(s, n, &L)) return -1;
        if (L > 0)
            CUDA_TRY(cudaMemcpyAsync(out + (size_t)written * ENC_DIM, e->x, (size_t)L * ENC_DIM * 4,
                                     cudaMemcpyDeviceToHost, s->stream));
        CUDA_TRY(cudaStreamSynchronize(s->stream));
        written += L;
    }
    *out_len = written;
    return 0;
}

extern "C" int vox_cuda_encoder_step(vox_cuda_state_t *s, float *x, int new_len) {
    if (!s || !x || new_len <= 0) return -1;
    EncState *e = (EncState *)s->enc_priv;
    for (int off = 0; off < new_len; off += ENC_CHUNK) {
        const int M = new_len - off < ENC_CHUNK ? new_len - off : ENC_CHUNK;
        float *hx = x + (size_t)off * ENC_DIM;
        CUDA_TRY(cudaMemcpyAsync(e->x, hx, (size_t)M * ENC_DIM * 4, cudaMemcpyHostToDevice, s->stream));
        if (enc_layers(s, e->x, M, e->xn)) return -1;
        CUDA_TRY(cudaMemcpyAsync(hx, e->xn, (size_t)M * ENC_DIM * 4, cudaMemcpyDeviceToHost, s->stream));
        CUDA_TRY(cudaStreamSynchronize(s->stream));
    }
    return 0;
}

extern "C" int vox_cuda_adapter(vox_cuda_state_t *s, const float *enc_out, int enc_len, float *out) {
    if (!s || !enc_out || !out) return -1;
    EncState *e = (EncState *)s->enc_priv;
    const int k_all = enc_len / 4;
    for (int t0 = 0; t0 < k_all; t0 += ENC_CHUNK / 4) {
        const int k = k_all - t0 < ENC_CHUNK / 4 ? k_all - t0 : ENC_CHUNK / 4;
        CUDA_TRY(cudaMemcpyAsync(e->x, enc_out + (size_t)t0 * 4 * ENC_DIM, (size_t)k * 4 * ENC_DIM * 4,
                                 cudaMemcpyHostToDevice, s->stream));
        if (adapter_device(s, e->x, k, e->mid, e->xn)) return -1;
        CUDA_TRY(cudaMemcpyAsync(out + (size_t)t0 * DEC_DIM, e->xn, (size_t)k * DEC_DIM * 4,
                                 cudaMemcpyDeviceToHost, s->stream));
        CUDA_TRY(cudaStreamSynchronize(s->stream));
    }
    return 0;
}

// ---- op-level hook (vox_metal_sgemm_bf16, voxtral_metal.h:38) ------------------------
extern "C" int vox_cuda_sgemm_bf16(vox_cuda_model_t *m, int M, int N, int K, const float *A,
                                   const uint16_t *B_bf16, float *C) {
    if (!A || !B_bf16 || !C || M <= 0) return -1;
    const size_t wbytes = (size_t)N * K * 2;
    const uint16_t *dW = m ? vox_model_lookup(m, B_bf16, wbytes) : nullptr;
    uint16_t *tmpW = nullptr;
    float *dA = nullptr, *dC = nullptr;
    int rc = -1;
    if (cudaMalloc((void **)&dA, (size_t)M * K * 4) != cudaSuccess) goto done;
    if (cudaMalloc((void **)&dC, (size_t)M * N * 4) != cudaSuccess) goto done;
    if (!dW) {
        if (cudaMalloc((void **)&tmpW, wbytes) != cudaSuccess) goto done;
        if (cudaMemcpy(tmpW, B_bf16, wbytes, cudaMemcpyHostToDevice) != cudaSuccess) goto done;
        dW = tmpW;
    }
    if (cudaMemcpy(dA, A, (size_t)M * K * 4, cudaMemcpyHostToDevice) != cudaSuccess) goto done;
    if (sgemm_bf16w(0, M, N, K, dA, K, dW, dC, N, nullptr, EPI_NONE)) goto done;
    if (cudaMemcpy(C, dC, (size_t)M * N * 4, cudaMemcpyDeviceToHost) != cudaSuccess) goto done;
    rc = 0;
done:
    if (rc) vox_set_error("vox_cuda_sgemm_bf16(%d,%d,%d) failed: %s", M, N, K, cudaGetErrorString(cudaGetLastError()));
    cudaFree(dA); cudaFree(dC); cudaFree(tmpW);
    return rc;
}
