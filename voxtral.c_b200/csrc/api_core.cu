// api_core.cu - backend lifetime, error reporting, weight upload.
#include "vox_internal.h"
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <vector>

static char g_err[1024] = "";
static int g_ready = 0;
static size_t g_mem_used = 0;
unsigned long long g_launch_count = 0;

void vox_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    fprintf(stderr, "vox_cuda: %s\n", g_err);
}

void vox_mem_account(long delta) { g_mem_used = (size_t)((long)g_mem_used + delta); }

extern "C" const char *vox_cuda_last_error(void) { return g_err; }
extern "C" size_t vox_cuda_memory_used(void) { return g_mem_used; }
extern "C" unsigned long long vox_cuda_launch_count(void) { return g_launch_count; }
extern "C" int vox_cuda_available(void) { return g_ready; }

extern "C" int vox_cuda_init(void) {
    if (g_ready) return 1;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
        vox_set_error("no CUDA device visible (this backend has no CPU fallback)");
        return 0;
    }
    int dev = 0;
    const char *env = getenv("VOX_CUDA_DEVICE");
    if (env) dev = atoi(env);
    else if ((env = getenv("LOCAL_RANK"))) dev = atoi(env) % n;
    if (dev < 0 || dev >= n) { vox_set_error("VOX_CUDA_DEVICE=%d out of range (%d devices)", dev, n); return 0; }
    cudaDeviceProp prop;
    if (cudaSetDevice(dev) != cudaSuccess || cudaGetDeviceProperties(&prop, dev) != cudaSuccess) {
        vox_set_error("cannot select device %d", dev);
        return 0;
    }
    if (prop.major != 10) {
        vox_set_error("device %d is sm_%d%d; this library carries sm_100a code only", dev, prop.major, prop.minor);
        return 0;
    }
    cudaFree(0);
    g_ready = 1;
    return 1;
}

extern "C" void vox_cuda_shutdown(void) {
    if (!g_ready) return;
    cudaDeviceSynchronize();
    g_ready = 0;
}

// ---------------------------------------------------------------------------
// Weight upload.  One arena in execution order; merged QKV and row-interleaved
// w1|w3 replace the Metal backend's merged caches (voxtral_metal.m:342-400).
// bf16 stays bf16 (no f16 conversion as on Metal).
// ---------------------------------------------------------------------------
namespace {
struct Arena {
    uint8_t *base = nullptr;
    size_t off = 0, cap = 0;
    void *take(size_t bytes) {
        off = (off + 255) & ~(size_t)255;
        void *p = base ? base + off : nullptr;
        off += bytes;
        return p;
    }
};

inline uint16_t f2bf(float f, int *inexact) {
    uint32_t u; memcpy(&u, &f, 4);
    if (u & 0xFFFFu) *inexact = 1;
    uint32_t r = ((u >> 16) & 1u) + 0x7FFFu;
    return (uint16_t)((u + r) >> 16);
}

struct Uploader {
    vox_cuda_model *m;
    Arena a;
    bool dry;
    int fail = 0;
    void reg(const void *host, const void *dev, size_t bytes) {
        if (dry || !host || m->n_reg >= 512) return;
        m->reg[m->n_reg++] = {host, dev, bytes};
    }
    // plain copy of a host tensor
    const void *put(const void *host, size_t bytes, bool do_reg = false) {
        void *d = a.take(bytes);
        if (!dry) {
            if (!host) { fail = 1; return d; }
            if (cudaMemcpy(d, host, bytes, cudaMemcpyHostToDevice) != cudaSuccess) fail = 1;
            if (do_reg) reg(host, d, bytes);
        }
        return d;
    }
    // rows of several host matrices stacked
    const uint16_t *stack3(const uint16_t *h0, size_t r0, const uint16_t *h1, size_t r1,
                           const uint16_t *h2, size_t r2, size_t K) {
        uint8_t *d = (uint8_t *)a.take((r0 + r1 + r2) * K * 2);
        if (!dry) {
            if (!h0 || !h1 || !h2) { fail = 1; return (uint16_t *)d; }
            if (cudaMemcpy(d, h0, r0 * K * 2, cudaMemcpyHostToDevice) != cudaSuccess) fail = 1;
            if (cudaMemcpy(d + r0 * K * 2, h1, r1 * K * 2, cudaMemcpyHostToDevice) != cudaSuccess) fail = 1;
            if (cudaMemcpy(d + (r0 + r1) * K * 2, h2, r2 * K * 2, cudaMemcpyHostToDevice) != cudaSuccess) fail = 1;
            reg(h0, d, r0 * K * 2); reg(h1, d + r0 * K * 2, r1 * K * 2); reg(h2, d + (r0 + r1) * K * 2, r2 * K * 2);
        }
        return (uint16_t *)d;
    }
    // row-interleave two [R,K] matrices: out row 2i = a[i], 2i+1 = b[i]
    const uint16_t *interleave2(const uint16_t *ha, const uint16_t *hb, size_t R, size_t K) {
        uint8_t *d = (uint8_t *)a.take(2 * R * K * 2);
        if (!dry) {
            if (!ha || !hb) { fail = 1; return (uint16_t *)d; }
            if (cudaMemcpy2D(d, 2 * K * 2, ha, K * 2, K * 2, R, cudaMemcpyHostToDevice) != cudaSuccess) fail = 1;
            if (cudaMemcpy2D(d + K * 2, 2 * K * 2, hb, K * 2, K * 2, R, cudaMemcpyHostToDevice) != cudaSuccess) fail = 1;
        }
        return (uint16_t *)d;
    }
};

void layout_tail(vox_cuda_model *m, const vox_cuda_weights_t *w, Uploader &u);

// ---- Q8 (vox_cuda_model_create_q8) ----------------------------------------------------------------------------
// int8 rows already on the device -> the same values as bf16 (exact: |q| <= 128 has 8 significant bits) for the
// tensor-core products, and the row sums sum_k q[r][k] the single-token kernel needs (dec_mega.cu).  One CTA per row.
__global__ void __launch_bounds__(256) q8_expand_kernel(const int8_t *__restrict__ q, size_t K, uint16_t *__restrict__ out,
                                                        int *__restrict__ rowsum) {
    const size_t r = blockIdx.x;
    const uint4 *src = reinterpret_cast<const uint4 *>(q + r * K);
    int acc = 0;
    for (size_t i = threadIdx.x; i < K / 16; i += blockDim.x) {
        const uint4 v = src[i];
        const uint32_t wds[4] = {v.x, v.y, v.z, v.w};
        uint32_t o[8];
        #pragma unroll
        for (int j = 0; j < 4; j++) {
            const int a = (int)(int8_t)(wds[j] & 0xff), b = (int)(int8_t)((wds[j] >> 8) & 0xff);
            const int c = (int)(int8_t)((wds[j] >> 16) & 0xff), d = (int)(int8_t)(wds[j] >> 24);
            acc += a + b + c + d;
            o[2 * j] = (__float_as_uint((float)a) >> 16) | (__float_as_uint((float)b) & 0xffff0000u);
            o[2 * j + 1] = (__float_as_uint((float)c) >> 16) | (__float_as_uint((float)d) & 0xffff0000u);
        }
        if (out) {
            uint4 *dst = reinterpret_cast<uint4 *>(out + r * K) + 2 * i;
            dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
            dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
        }
    }
    if (rowsum) {
        __shared__ int red[8];
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            int t = 0;
            for (int k = 0; k < 8; k++) t += red[k];
            rowsum[r] = t;
        }
    }
}

// In place: swap the two 64-byte halves of every 128 bytes on odd rows (the layout dec_mega.cu's Q8 consumers and the
// embedding gathers expect: q8_swz).  One CTA per odd row.
__global__ void __launch_bounds__(256) q8_swizzle_kernel(int8_t *__restrict__ q, size_t K) {
    const size_t r = 2 * (size_t)blockIdx.x + 1;
    uint4 *row = reinterpret_cast<uint4 *>(q + r * K);
    for (size_t i = threadIdx.x; i < K / 32; i += blockDim.x) {          // i -> 16-byte unit (i % 4) of the even half of pair i / 4
        const size_t a = (i / 4) * 8 + (i % 4), b = a + 4;
        const uint4 t = row[a];
        row[a] = row[b];
        row[b] = t;
    }
}

// One merged Q8 matrix being assembled: parts are copied to rows row0, row0 + step, ... of q / scale, then finish()
// expands.  `q` lives in the arena when the single-token kernel streams it (decoder, vocabulary), else in `stage`.
struct Q8Mat {
    int8_t *q; float *s; int *qs; uint16_t *bf; size_t R, K;
};
struct Q8Up {
    Uploader &u;
    int8_t *stage;            // device staging for matrices whose int8 rows are not kept (encoder, adapter)
    Q8Mat begin(size_t R, size_t K, bool keep_q, bool want_bf) {
        Q8Mat t;
        t.R = R; t.K = K;
        t.q = keep_q ? (int8_t *)u.a.take(R * K) : stage;
        t.s = (float *)u.a.take(R * 4);
        t.qs = keep_q ? (int *)u.a.take(R * 4) : nullptr;
        t.bf = want_bf ? (uint16_t *)u.a.take(R * K * 2) : nullptr;
        return t;
    }
    void part(const Q8Mat &t, const vox_cuda_q8_t &src, size_t rows, size_t row0, size_t step) {
        if (u.dry) return;
        if (!src.q || !src.scale) { u.fail = 1; return; }
        if (cudaMemcpy2D(t.q + row0 * t.K, step * t.K, src.q, t.K, t.K, rows, cudaMemcpyHostToDevice) != cudaSuccess) u.fail = 1;
        if (cudaMemcpy2D(t.s + row0, step * 4, src.scale, 4, 4, rows, cudaMemcpyHostToDevice) != cudaSuccess) u.fail = 1;
    }
    void finish(const Q8Mat &t) {
        if (u.dry || u.fail) return;
        q8_expand_kernel<<<(unsigned)t.R, 256>>>(t.q, t.K, t.bf, t.qs);
        if (t.qs) q8_swizzle_kernel<<<(unsigned)(t.R / 2), 256>>>(t.q, t.K);   // kept int8 rows: the single-token kernel's layout
        if (cudaDeviceSynchronize() != cudaSuccess) u.fail = 1;      // the staging buffer is reused by the next matrix
    }
};

void layout_q8(vox_cuda_model *m, const vox_cuda_weights_t *w, const vox_cuda_weights_q8_t *qw, Uploader &u, int8_t *stage) {
    Q8Up q{u, stage};
    for (int l = 0; l < m->n_dec; l++) {
        const vox_cuda_dec_layer_w_t &s = w->dec[l];
        const vox_cuda_layer_q8_t &g = qw->dec[l];
        DecLayerDev &d = m->dec[l];
        d.attn_norm = (const float *)u.put(s.attention_norm, DEC_DIM * 4);
        Q8Mat t = q.begin(DEC_QKV, DEC_DIM, true, true);
        q.part(t, g.wq, DEC_QDIM, 0, 1); q.part(t, g.wk, DEC_KVDIM, DEC_QDIM, 1); q.part(t, g.wv, DEC_KVDIM, DEC_QDIM + DEC_KVDIM, 1);
        q.finish(t);
        d.wqkv = t.bf; d.q_wqkv = t.q; d.s_wqkv = t.s; d.qs_wqkv = t.qs;
        t = q.begin(DEC_DIM, DEC_QDIM, true, true);
        q.part(t, g.wo, DEC_DIM, 0, 1); q.finish(t);
        d.wo = t.bf; d.q_wo = t.q; d.s_wo = t.s; d.qs_wo = t.qs;
        d.ffn_norm = (const float *)u.put(s.ffn_norm, DEC_DIM * 4);
        t = q.begin(2 * DEC_HIDDEN, DEC_DIM, true, true);
        q.part(t, g.w1, DEC_HIDDEN, 0, 2); q.part(t, g.w3, DEC_HIDDEN, 1, 2); q.finish(t);
        d.w13 = t.bf; d.q_w13 = t.q; d.s_w13 = t.s; d.qs_w13 = t.qs;
        t = q.begin(DEC_DIM, DEC_HIDDEN, true, true);
        q.part(t, g.w2, DEC_DIM, 0, 1); q.finish(t);
        d.w2 = t.bf; d.q_w2 = t.q; d.s_w2 = t.s; d.qs_w2 = t.qs;
    }
    m->dec_norm = (const float *)u.put(w->dec_norm, DEC_DIM * 4);
    {
        // the bf16 image (805 MB) serves the stream groups' logits product and embedding rows (api_group.cu)
        Q8Mat t = q.begin(VOCAB, DEC_DIM, true, true);
        q.part(t, qw->tok_embeddings, VOCAB, 0, 1); q.finish(t);
        m->tok_emb = t.bf; m->q_tok_emb = t.q; m->s_tok_emb = t.s; m->qs_tok_emb = t.qs;
    }
    for (int l = 0; l < m->n_enc; l++) {
        const vox_cuda_enc_layer_w_t &s = w->enc[l];
        const vox_cuda_layer_q8_t &g = qw->enc[l];
        EncLayerDev &d = m->enc[l];
        d.attn_norm = (const float *)u.put(s.attention_norm, ENC_DIM * 4);
        Q8Mat t = q.begin(3 * ENC_QKV, ENC_DIM, false, true);
        q.part(t, g.wq, ENC_QKV, 0, 1); q.part(t, g.wk, ENC_QKV, ENC_QKV, 1); q.part(t, g.wv, ENC_QKV, 2 * ENC_QKV, 1);
        q.finish(t);
        d.wqkv = t.bf; d.s_wqkv = t.s;
        {   // merged bias, wk has none (voxtral_encoder.c:591-593)
            float *db = (float *)u.a.take(3 * ENC_QKV * 4);
            if (!u.dry) {
                std::vector<float> hb(3 * ENC_QKV, 0.f);
                if (s.wq_bias && s.wv_bias) {
                    memcpy(hb.data(), s.wq_bias, ENC_QKV * 4);
                    memcpy(hb.data() + 2 * ENC_QKV, s.wv_bias, ENC_QKV * 4);
                } else u.fail = 1;
                if (cudaMemcpy(db, hb.data(), hb.size() * 4, cudaMemcpyHostToDevice) != cudaSuccess) u.fail = 1;
            }
            d.bqkv = db;
        }
        t = q.begin(ENC_DIM, ENC_QKV, false, true);
        q.part(t, g.wo, ENC_DIM, 0, 1); q.finish(t);
        d.wo = t.bf; d.s_wo = t.s;
        d.bo = (const float *)u.put(s.wo_bias, ENC_DIM * 4);
        d.ffn_norm = (const float *)u.put(s.ffn_norm, ENC_DIM * 4);
        t = q.begin(2 * ENC_HIDDEN, ENC_DIM, false, true);
        q.part(t, g.w1, ENC_HIDDEN, 0, 2); q.part(t, g.w3, ENC_HIDDEN, 1, 2); q.finish(t);
        d.w13 = t.bf; d.s_w13 = t.s;
        t = q.begin(ENC_DIM, ENC_HIDDEN, false, true);
        q.part(t, g.w2, ENC_DIM, 0, 1); q.finish(t);
        d.w2 = t.bf; d.s_w2 = t.s;
        d.b2 = (const float *)u.put(s.w2_bias, ENC_DIM * 4);
    }
    m->enc_norm = (const float *)u.put(w->enc_norm, ENC_DIM * 4);
    {
        Q8Mat t = q.begin(DEC_DIM, 4 * ENC_DIM, false, true);
        q.part(t, qw->adapter_w0, DEC_DIM, 0, 1); q.finish(t);
        m->adapter_w0 = t.bf; m->s_adapter0 = t.s;
        t = q.begin(DEC_DIM, DEC_DIM, false, true);
        q.part(t, qw->adapter_w1, DEC_DIM, 0, 1); q.finish(t);
        m->adapter_w1 = t.bf; m->s_adapter1 = t.s;
    }
    layout_tail(m, w, u);
}

void layout(vox_cuda_model *m, const vox_cuda_weights_t *w, Uploader &u) {
    for (int l = 0; l < m->n_dec; l++) {
        const vox_cuda_dec_layer_w_t &s = w->dec[l];
        DecLayerDev &d = m->dec[l];
        d.attn_norm = (const float *)u.put(s.attention_norm, DEC_DIM * 4);
        d.wqkv = u.stack3(s.wq, DEC_QDIM, s.wk, DEC_KVDIM, s.wv, DEC_KVDIM, DEC_DIM);
        d.wo = (const uint16_t *)u.put(s.wo, (size_t)DEC_DIM * DEC_QDIM * 2, true);
        d.ffn_norm = (const float *)u.put(s.ffn_norm, DEC_DIM * 4);
        d.w13 = u.interleave2(s.w1, s.w3, DEC_HIDDEN, DEC_DIM);
        d.w2 = (const uint16_t *)u.put(s.w2, (size_t)DEC_DIM * DEC_HIDDEN * 2, true);
    }
    m->dec_norm = (const float *)u.put(w->dec_norm, DEC_DIM * 4);
    m->tok_emb = (const uint16_t *)u.put(w->tok_embeddings, (size_t)VOCAB * DEC_DIM * 2, true);
    for (int l = 0; l < m->n_enc; l++) {
        const vox_cuda_enc_layer_w_t &s = w->enc[l];
        EncLayerDev &d = m->enc[l];
        d.attn_norm = (const float *)u.put(s.attention_norm, ENC_DIM * 4);
        d.wqkv = u.stack3(s.wq, ENC_QKV, s.wk, ENC_QKV, s.wv, ENC_QKV, ENC_DIM);
        {   // merged bias, wk has none (voxtral_encoder.c:591-593)
            float *db = (float *)u.a.take(3 * ENC_QKV * 4);
            if (!u.dry) {
                std::vector<float> hb(3 * ENC_QKV, 0.f);
                if (s.wq_bias && s.wv_bias) {
                    memcpy(hb.data(), s.wq_bias, ENC_QKV * 4);
                    memcpy(hb.data() + 2 * ENC_QKV, s.wv_bias, ENC_QKV * 4);
                } else u.fail = 1;
                if (cudaMemcpy(db, hb.data(), hb.size() * 4, cudaMemcpyHostToDevice) != cudaSuccess) u.fail = 1;
            }
            d.bqkv = db;
        }
        d.wo = (const uint16_t *)u.put(s.wo, (size_t)ENC_DIM * ENC_QKV * 2, true);
        d.bo = (const float *)u.put(s.wo_bias, ENC_DIM * 4);
        d.ffn_norm = (const float *)u.put(s.ffn_norm, ENC_DIM * 4);
        d.w13 = u.interleave2(s.w1, s.w3, ENC_HIDDEN, ENC_DIM);
        d.w2 = (const uint16_t *)u.put(s.w2, (size_t)ENC_DIM * ENC_HIDDEN * 2, true);
        d.b2 = (const float *)u.put(s.w2_bias, ENC_DIM * 4);
    }
    m->enc_norm = (const float *)u.put(w->enc_norm, ENC_DIM * 4);
    m->adapter_w0 = (const uint16_t *)u.put(w->adapter_w0, (size_t)DEC_DIM * 4 * ENC_DIM * 2, true);
    m->adapter_w1 = (const uint16_t *)u.put(w->adapter_w1, (size_t)DEC_DIM * DEC_DIM * 2, true);
    layout_tail(m, w, u);
}

// conv stem, ada_scale, RoPE tables: fp32 sources, common to the bf16 and the Q8 model
void layout_tail(vox_cuda_model *m, const vox_cuda_weights_t *w, Uploader &u) {
    // conv stem: [oc][ic][3] f32 -> [oc][tap][ic] bf16 (exact: the file stores bf16)
    for (int cv = 0; cv < 2; cv++) {
        const int cin = cv ? ENC_DIM : VOX_CUDA_MEL_BINS;
        const float *hw = cv ? w->conv1_weight : w->conv0_weight;
        uint16_t *dw = (uint16_t *)u.a.take((size_t)ENC_DIM * 3 * cin * 2);
        if (!u.dry) {
            if (!hw) u.fail = 1;
            else {
                std::vector<uint16_t> t((size_t)ENC_DIM * 3 * cin);
                int inexact = 0;
                for (int oc = 0; oc < ENC_DIM; oc++)
                    for (int ic = 0; ic < cin; ic++)
                        for (int k = 0; k < 3; k++)
                            t[((size_t)oc * 3 + k) * cin + ic] = f2bf(hw[((size_t)oc * cin + ic) * 3 + k], &inexact);
                if (inexact)
                    fprintf(stderr, "vox_cuda: warning: conv%d weights are not bf16-exact; rounded\n", cv);
                if (cudaMemcpy(dw, t.data(), t.size() * 2, cudaMemcpyHostToDevice) != cudaSuccess) u.fail = 1;
            }
        }
        if (cv) m->conv1_w = dw; else m->conv0_w = dw;
    }
    m->conv0_b = (const float *)u.put(w->conv0_bias, ENC_DIM * 4);
    m->conv1_b = (const float *)u.put(w->conv1_bias, ENC_DIM * 4);
    m->ada_scale = (float *)u.put(w->ada_scale, (size_t)m->n_dec * DEC_DIM * 4);
    m->inv_freq_dec = (float *)u.a.take(DEC_HD / 2 * 4);
    m->inv_freq_enc = (float *)u.a.take(ENC_HD / 2 * 4);
    if (!u.dry) {
        // same float expressions as vox_compute_rope_freqs (voxtral_kernels.c:623)
        float fd[DEC_HD / 2], fe[ENC_HD / 2];
        for (int d = 0; d < DEC_HD / 2; d++) fd[d] = 1.0f / powf(ROPE_THETA, (float)(2 * d) / (float)DEC_HD);
        for (int d = 0; d < ENC_HD / 2; d++) fe[d] = 1.0f / powf(ROPE_THETA, (float)(2 * d) / (float)ENC_HD);
        if (cudaMemcpy(m->inv_freq_dec, fd, sizeof(fd), cudaMemcpyHostToDevice) != cudaSuccess) u.fail = 1;
        if (cudaMemcpy(m->inv_freq_enc, fe, sizeof(fe), cudaMemcpyHostToDevice) != cudaSuccess) u.fail = 1;
    }
}
}  // namespace

extern "C" vox_cuda_model_t *vox_cuda_model_create(const vox_cuda_weights_t *w) {
    if (!g_ready && !vox_cuda_init()) return nullptr;
    if (!w || w->n_enc_layers < 0 || w->n_enc_layers > VOX_CUDA_ENC_LAYERS ||
        w->n_dec_layers < 1 || w->n_dec_layers > VOX_CUDA_DEC_LAYERS) {
        vox_set_error("vox_cuda_model_create: bad weight table");
        return nullptr;
    }
    vox_cuda_model *m = (vox_cuda_model *)calloc(1, sizeof(vox_cuda_model));
    m->n_enc = w->n_enc_layers; m->n_dec = w->n_dec_layers;
    Uploader dry{m, Arena(), true};
    layout(m, w, dry);
    m->arena_bytes = dry.a.off + 256;
    if (cudaMalloc(&m->arena, m->arena_bytes) != cudaSuccess) {
        vox_set_error("vox_cuda_model_create: cudaMalloc(%zu MB) failed", m->arena_bytes >> 20);
        free(m);
        return nullptr;
    }
    vox_mem_account((long)m->arena_bytes);
    Uploader real{m, Arena(), false};
    real.a.base = m->arena;
    layout(m, w, real);
    if (real.fail || cudaDeviceSynchronize() != cudaSuccess) {
        vox_set_error("vox_cuda_model_create: weight upload failed (missing tensor or copy error: %s)",
                      cudaGetErrorString(cudaGetLastError()));
        cudaFree(m->arena);
        vox_mem_account(-(long)m->arena_bytes);
        free(m);
        return nullptr;
    }
    return m;
}

extern "C" vox_cuda_model_t *vox_cuda_model_create_q8(const vox_cuda_weights_t *w, const vox_cuda_weights_q8_t *q) {
    if (!g_ready && !vox_cuda_init()) return nullptr;
    if (!w || !q || w->n_enc_layers < 0 || w->n_enc_layers > VOX_CUDA_ENC_LAYERS ||
        w->n_dec_layers < 1 || w->n_dec_layers > VOX_CUDA_DEC_LAYERS) {
        vox_set_error("vox_cuda_model_create_q8: bad weight table");
        return nullptr;
    }
    vox_cuda_model *m = (vox_cuda_model *)calloc(1, sizeof(vox_cuda_model));
    m->n_enc = w->n_enc_layers; m->n_dec = w->n_dec_layers; m->q8 = 1;
    Uploader dry{m, Arena(), true};
    layout_q8(m, w, q, dry, nullptr);
    m->arena_bytes = dry.a.off + 256;
    int8_t *stage = nullptr;
    const size_t stage_bytes = (size_t)DEC_DIM * 4 * ENC_DIM;      // largest matrix that is not kept as int8 (adapter_w0)
    static_assert((size_t)DEC_DIM * 4 * ENC_DIM >= (size_t)2 * ENC_HIDDEN * ENC_DIM, "staging covers the encoder's w1|w3");
    if (cudaMalloc(&m->arena, m->arena_bytes) != cudaSuccess || cudaMalloc(&stage, stage_bytes) != cudaSuccess) {
        vox_set_error("vox_cuda_model_create_q8: cudaMalloc(%zu MB) failed", m->arena_bytes >> 20);
        cudaFree(m->arena);
        free(m);
        return nullptr;
    }
    vox_mem_account((long)m->arena_bytes);
    Uploader real{m, Arena(), false};
    real.a.base = m->arena;
    layout_q8(m, w, q, real, stage);
    const cudaError_t e = cudaDeviceSynchronize();
    cudaFree(stage);
    if (real.fail || e != cudaSuccess) {
        vox_set_error("vox_cuda_model_create_q8: weight upload failed (missing tensor or copy error: %s)",
                      cudaGetErrorString(cudaGetLastError()));
        cudaFree(m->arena);
        vox_mem_account(-(long)m->arena_bytes);
        free(m);
        return nullptr;
    }
    return m;
}

extern "C" int vox_cuda_model_is_q8(const vox_cuda_model_t *m) { return m ? m->q8 : -1; }

extern "C" int vox_cuda_model_set_ada_scale(vox_cuda_model_t *m, const float *ada_scale) {
    if (!m || !ada_scale) return -1;
    CUDA_TRY(cudaMemcpy(m->ada_scale, ada_scale, (size_t)m->n_dec * DEC_DIM * 4, cudaMemcpyHostToDevice));
    return 0;
}

extern "C" void vox_cuda_model_free(vox_cuda_model_t *m) {
    if (!m) return;
    cudaDeviceSynchronize();
    cudaFree(m->arena);
    vox_mem_account(-(long)m->arena_bytes);
    free(m);
}

// registry lookup used by vox_cuda_sgemm_bf16
const uint16_t *vox_model_lookup(vox_cuda_model *m, const void *host, size_t bytes) {
    for (int i = 0; i < m->n_reg; i++)
        if (m->reg[i].host == host && m->reg[i].bytes >= bytes) return (const uint16_t *)m->reg[i].dev;
    return nullptr;
}
