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
#include <string.h>
#include <stdlib.h>
#include <math.h>

namespace {

constexpr int MAX_STREAMS = 128;
constexpr int KS_QKV = 3, KS_WO = 6, KS_W2 = 6;     // K-range splits: 48x3, 24x6, 24x6 work items for 148 SMs
constexpr int KS_MAX = 6;

struct Slot {
    int pos;        // logical position of the token being processed (= valid K/V rows before it)
    int prev;       // token whose embedding is added to the adapter row (voxtral.c:1106-1113)
    int aidx;       // adapter row of this position
    int remaining;  // steps this stream still has to take; 0 = finished (rows are computed but ignored)
};

// ---- x (+= partials | = embedding) -> RMSNorm -> split bf16 rows --------------------------------
// mode 0: x[b] = adapter[b][aidx] + tok_emb[prev];  mode 1: x[b] += sum_s part[s][b]
__global__ void __launch_bounds__(256) group_norm_kernel(
    float *__restrict__ x, int mode, const float *__restrict__ part, int n_split, long split_stride,
    const float *__restrict__ adapter, long adapter_stride, const uint16_t *__restrict__ tok_emb,
    const Slot *__restrict__ slots, const float *__restrict__ w, const float *__restrict__ ada,
    uint16_t *__restrict__ out) {
    const int b = blockIdx.x, tid = threadIdx.x;
    float *xr = x + (size_t)b * DEC_DIM;
    float v[12];
    tcx::pdl_launch_dependents();
    tcx::pdl_wait();
    if (mode == 0) {
        const Slot sl = slots[b];
        const float *a = adapter + (size_t)b * adapter_stride + (size_t)sl.aidx * DEC_DIM;
        const uint16_t *e = tok_emb + (size_t)sl.prev * DEC_DIM;
        #pragma unroll
        for (int j = 0; j < 3; j++) {
            const int col = j * 1024 + tid * 4;
            const float4 av = *reinterpret_cast<const float4 *>(a + col);
            const uint2 ev = *reinterpret_cast<const uint2 *>(e + col);
            v[4 * j] = av.x + bf16_lo(ev.x); v[4 * j + 1] = av.y + bf16_hi(ev.x);
            v[4 * j + 2] = av.z + bf16_lo(ev.y); v[4 * j + 3] = av.w + bf16_hi(ev.y);
        }
    } else {
        #pragma unroll
        for (int j = 0; j < 3; j++) {
            const int col = j * 1024 + tid * 4;
            float4 t = *reinterpret_cast<const float4 *>(xr + col);
            for (int s = 0; s < n_split; s++) {
                const float4 pv = *reinterpret_cast<const float4 *>(part + (size_t)s * split_stride + (size_t)b * DEC_DIM + col);
                t.x += pv.x; t.y += pv.y; t.z += pv.z; t.w += pv.w;
            }
            v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
        }
    }
    float ss = 0.f;
    #pragma unroll
    for (int j = 0; j < 3; j++) {
        *reinterpret_cast<float4 *>(xr + j * 1024 + tid * 4) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        ss += v[4 * j] * v[4 * j] + v[4 * j + 1] * v[4 * j + 1] + v[4 * j + 2] * v[4 * j + 2] + v[4 * j + 3] * v[4 * j + 3];
    }
    __shared__ float red[8];
    ss = warp_sum(ss);
    if ((tid & 31) == 0) red[tid >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
    #pragma unroll
    for (int i = 0; i < 8; i++) tot += red[i];
    const float rinv = 1.0f / sqrtf(tot / (float)DEC_DIM + NORM_EPS);
    uint16_t *orow = out + (size_t)b * 2 * DEC_DIM;
    #pragma unroll
    for (int j = 0; j < 3; j++) {
        const int col = j * 1024 + tid * 4;
        const float4 wv = *reinterpret_cast<const float4 *>(w + col);
        float y[4] = {v[4 * j] * rinv * wv.x, v[4 * j + 1] * rinv * wv.y, v[4 * j + 2] * rinv * wv.z, v[4 * j + 3] * rinv * wv.w};
        if (ada) {
            const float4 av = *reinterpret_cast<const float4 *>(ada + col);
            y[0] *= (1.0f + av.x); y[1] *= (1.0f + av.y); y[2] *= (1.0f + av.z); y[3] *= (1.0f + av.w);
        }
        const __nv_bfloat162 h0 = __floats2bfloat162_rn(y[0], y[1]), h1 = __floats2bfloat162_rn(y[2], y[3]);
        const __nv_bfloat162 l0 = __floats2bfloat162_rn(y[0] - __bfloat162float(h0.x), y[1] - __bfloat162float(h0.y));
        const __nv_bfloat162 l1 = __floats2bfloat162_rn(y[2] - __bfloat162float(h1.x), y[3] - __bfloat162float(h1.y));
        uint2 hv, lv;
        hv.x = *reinterpret_cast<const uint32_t *>(&h0); hv.y = *reinterpret_cast<const uint32_t *>(&h1);
        lv.x = *reinterpret_cast<const uint32_t *>(&l0); lv.y = *reinterpret_cast<const uint32_t *>(&l1);
        *reinterpret_cast<uint2 *>(orow + col) = hv;
        *reinterpret_cast<uint2 *>(orow + DEC_DIM + col) = lv;
    }
}

// ---- RoPE + ring append + GQA attention of one (stream, kv head) -------------------------------
// Key groups of 8 per warp with all 16 row loads in flight (the single-stream kernel's scheme,
// dec_mega.cu); the 8 warps' online-softmax states are merged in shared memory.
constexpr int AT_WARPS = 8;
__global__ void __launch_bounds__(256, 2) group_attn_kernel(
    const float *__restrict__ part, int n_split, long split_stride, const Slot *__restrict__ slots,
    const float *__restrict__ inv_freq, float *__restrict__ kv_k, float *__restrict__ kv_v, size_t stream_stride,
    int ring_mask, uint16_t *__restrict__ out) {
    const int b = blockIdx.x, kvh = blockIdx.y;
    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
    __shared__ __align__(16) float qs[4][DEC_HD];
    __shared__ __align__(16) float wsm[AT_WARPS][4][DEC_HD + 4];
    tcx::pdl_launch_dependents();
    tcx::pdl_wait();
    const int pos = slots[b].pos;
    float *kc = kv_k + (size_t)b * stream_stride, *vc = kv_v + (size_t)b * stream_stride;
    const size_t slot_new = (size_t)(pos & ring_mask) * DEC_KVDIM + (size_t)kvh * DEC_HD;
    {
        // 256 q pairs (4 heads), 64 k pairs, 128 v values of this kv head; partial sums in split order
        const float *pr = part + (size_t)b * DEC_QKV;
        for (int i = tid; i < 320 + 128; i += 256) {
            if (i < 320) {
                const int col = i < 256 ? kvh * 4 * DEC_HD + 2 * i : DEC_QDIM + kvh * DEC_HD + 2 * (i - 256);
                float x0 = 0.f, x1 = 0.f;
                for (int s = 0; s < n_split; s++) {
                    const float2 t = *reinterpret_cast<const float2 *>(pr + (size_t)s * split_stride + col);
                    x0 += t.x; x1 += t.y;
                }
                const int d = (col % DEC_HD) >> 1;
                const float angle = (float)pos * inv_freq[d];
                const float c = cosf(angle), sn = sinf(angle);
                const float y0 = x0 * c - x1 * sn, y1 = x0 * sn + x1 * c;
                if (i < 256) { qs[i >> 6][2 * (i & 63)] = y0; qs[i >> 6][2 * (i & 63) + 1] = y1; }
                else { kc[slot_new + 2 * (i - 256)] = y0; kc[slot_new + 2 * (i - 256) + 1] = y1; }
            } else {
                const int d = i - 320;
                float x0 = 0.f;
                for (int s = 0; s < n_split; s++) x0 += pr[(size_t)s * split_stride + DEC_QDIM + DEC_KVDIM + kvh * DEC_HD + d];
                vc[slot_new + d] = x0;
            }
        }
    }
    __syncthreads();                                     // q in smem, new K/V row visible to the CTA
    const int n_keys = min(pos + 1, DEC_WINDOW);
    const int first = pos - n_keys + 1;
    const int ng = (n_keys + 7) >> 3;
    const float scale = 0.08838834764831845f;            // 1/sqrt(128)
    float4 q[4];
    #pragma unroll
    for (int h = 0; h < 4; h++) q[h] = reinterpret_cast<const float4 *>(&qs[h][0])[lane];
    float m = -INFINITY, l = 0.f;                        // running max / sum of head (lane & 3)
    float4 acc[4];
    #pragma unroll
    for (int h = 0; h < 4; h++) acc[h] = make_float4(0.f, 0.f, 0.f, 0.f);
    const size_t hoff = (size_t)kvh * DEC_HD + 4 * lane;
    for (int g = w; g < ng; g += AT_WARPS) {
        const int j0 = g * 8;
        const int nv = min(8, n_keys - j0);
        float4 kf[8], vf[8];
        #pragma unroll
        for (int i = 0; i < 8; i++) {
            const int slot = (first + j0 + min(i, nv - 1)) & ring_mask;     // clamp: always a valid row
            kf[i] = *reinterpret_cast<const float4 *>(kc + (size_t)slot * DEC_KVDIM + hoff);
            vf[i] = *reinterpret_cast<const float4 *>(vc + (size_t)slot * DEC_KVDIM + hoff);
        }
        float v[32];
        #pragma unroll
        for (int i = 0; i < 8; i++)
            #pragma unroll
            for (int h = 0; h < 4; h++)
                v[i * 4 + h] = kf[i].x * q[h].x + kf[i].y * q[h].y + kf[i].z * q[h].z + kf[i].w * q[h].w;
        // transposed reduction: score (key i, head h) ends in lane 4i + h
        #pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            const bool up = (lane & o) != 0;
            #pragma unroll
            for (int k = 0; k < o; k++) {
                const float keep = up ? v[k + o] : v[k];
                const float send = up ? v[k] : v[k + o];
                v[k] = keep + __shfl_xor_sync(0xffffffffu, send, o);
            }
        }
        const bool valid = (lane >> 2) < nv;
        const float sc = valid ? v[0] * scale : -INFINITY;
        float mx = sc;
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
        const float mn = fmaxf(m, mx);                   // finite: every group has >= 1 valid key
        const float corr = (m == -INFINITY) ? 0.f : expf(m - mn);
        const float p = valid ? expf(sc - mn) : 0.f;
        float ps = p;
        ps += __shfl_xor_sync(0xffffffffu, ps, 4);
        ps += __shfl_xor_sync(0xffffffffu, ps, 8);
        ps += __shfl_xor_sync(0xffffffffu, ps, 16);
        l = l * corr + ps;
        m = mn;
        #pragma unroll
        for (int h = 0; h < 4; h++) {
            const float ch = __shfl_sync(0xffffffffu, corr, h);
            acc[h].x *= ch; acc[h].y *= ch; acc[h].z *= ch; acc[h].w *= ch;
        }
        #pragma unroll
        for (int i = 0; i < 8; i++) {
            #pragma unroll
            for (int h = 0; h < 4; h++) {
                const float ph = __shfl_sync(0xffffffffu, p, i * 4 + h);    // 0 for padded keys
                acc[h].x = fmaf(ph, vf[i].x, acc[h].x); acc[h].y = fmaf(ph, vf[i].y, acc[h].y);
                acc[h].z = fmaf(ph, vf[i].z, acc[h].z); acc[h].w = fmaf(ph, vf[i].w, acc[h].w);
            }
        }
    }
    #pragma unroll
    for (int h = 0; h < 4; h++) reinterpret_cast<float4 *>(&wsm[w][h][0])[lane] = acc[h];
    if (lane < 4) { wsm[w][lane][DEC_HD] = m; wsm[w][lane][DEC_HD + 1] = l; }
    __syncthreads();
    {   // merge the 8 warps: thread -> (head, two dims); normalise; split bf16 out
        const int h = tid >> 6, d = (tid & 63) * 2;
        float M = -INFINITY;
        #pragma unroll
        for (int k = 0; k < AT_WARPS; k++) M = fmaxf(M, wsm[k][h][DEC_HD]);
        float n0 = 0.f, n1 = 0.f, den = 0.f;
        #pragma unroll
        for (int k = 0; k < AT_WARPS; k++) {
            const float mk_ = wsm[k][h][DEC_HD];
            const float e = (mk_ == -INFINITY) ? 0.f : expf(mk_ - M);
            n0 = fmaf(e, wsm[k][h][d], n0);
            n1 = fmaf(e, wsm[k][h][d + 1], n1);
            den = fmaf(e, wsm[k][h][DEC_HD + 1], den);
        }
        const float inv = den > 0.f ? 1.0f / den : 0.f;
        const float y0 = n0 * inv, y1 = n1 * inv;
        const __nv_bfloat162 hi = __floats2bfloat162_rn(y0, y1);
        const __nv_bfloat162 lo = __floats2bfloat162_rn(y0 - __bfloat162float(hi.x), y1 - __bfloat162float(hi.y));
        uint16_t *o = out + (size_t)b * 2 * DEC_QDIM + (kvh * 4 + h) * DEC_HD + d;
        *reinterpret_cast<__nv_bfloat162 *>(o) = hi;
        *reinterpret_cast<__nv_bfloat162 *>(o + DEC_QDIM) = lo;
    }
}

// ---- first-maximum argmax per stream + advance the stream ---------------------------------------
__global__ void __launch_bounds__(1024) group_argmax_kernel(const float *__restrict__ logits, Slot *__restrict__ slots,
                                                            int *__restrict__ tokens, int step, int n_streams) {
    const int b = blockIdx.x, tid = threadIdx.x;
    tcx::pdl_launch_dependents();
    tcx::pdl_wait();
    const float4 *row = reinterpret_cast<const float4 *>(logits + (size_t)b * VOCAB);
    float best = -INFINITY; int bi = 0x7fffffff;
    for (int i = tid; i < VOCAB / 4; i += 1024) {        // ascending indices per thread: strict > keeps the first
        const float4 t = row[i];
        if (t.x > best) { best = t.x; bi = 4 * i; }
        if (t.y > best) { best = t.y; bi = 4 * i + 1; }
        if (t.z > best) { best = t.z; bi = 4 * i + 2; }
        if (t.w > best) { best = t.w; bi = 4 * i + 3; }
    }
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    __shared__ float sv[32]; __shared__ int si[32];
    if ((tid & 31) == 0) { sv[tid >> 5] = best; si[tid >> 5] = bi; }
    __syncthreads();
    if (tid < 32) {
        best = sv[tid]; bi = si[tid];
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
        }
        if (tid == 0) {
            Slot sl = slots[b];
            if (bi == 0x7fffffff) bi = 0;                // all -inf / NaN row: the reference's loop leaves index 0
            tokens[(size_t)step * n_streams + b] = sl.remaining > 0 ? bi : -1;
            if (sl.remaining > 0) {
                sl.remaining--;
                if (sl.remaining > 0) { sl.pos++; sl.aidx++; sl.prev = bi; }   // a finished stream idles on its last row
                slots[b] = sl;
            }
        }
    }
}

__global__ void __launch_bounds__(256) group_embed_rows_kernel(const int *__restrict__ tokens, const float *__restrict__ adapter,
                                                               const uint16_t *__restrict__ tok_emb, float *__restrict__ x) {
    const int r = blockIdx.x;
    const float *a = adapter + (size_t)r * DEC_DIM;
    const uint16_t *e = tok_emb + (size_t)tokens[r] * DEC_DIM;
    for (int i = threadIdx.x; i < DEC_DIM; i += 256)
        x[(size_t)r * DEC_DIM + i] = a[i] + __uint_as_float(((uint32_t)e[i]) << 16);
}

}  // namespace

struct vox_cuda_group {
    vox_cuda_model *m;
    vox_cuda_state *eng;         // encoder engine + prefill scratch; its stream orders all group work
    int B, NB, max_pos, cap, mask;
    float *kv_k, *kv_v;          // [B][n_dec][cap][1024]
    size_t kv_layer, kv_stream;  // floats
    float *adapter;              // [B][max_pos][3072]
    int count[MAX_STREAMS];      // adapter rows per stream (0 = not encoded)
    Slot *d_slots;
    float *x;                    // [B][3072] residual rows
    uint16_t *xn_b, *att_b, *hh_b;   // split bf16 rows [NB][2*3072], [NB][2*4096], [NB][2*9216]
    float *part;                 // [KS_MAX][B][6144] K-range partials of the current product
    float *logits;               // [B][131072]
    int *d_tokens; size_t tok_cap;   // [steps][B]
    float *pf_x; int *pf_tok;    // prefill rows / prompt ids
    WtOp op_qkv[VOX_CUDA_DEC_LAYERS], op_wo[VOX_CUDA_DEC_LAYERS], op_w13[VOX_CUDA_DEC_LAYERS], op_w2[VOX_CUDA_DEC_LAYERS], op_logits;
    cudaEvent_t ev0, ev1;
};

extern "C" void vox_cuda_group_free(vox_cuda_group_t *g) {
    if (!g) return;
    cudaFree(g->kv_k); cudaFree(g->kv_v); cudaFree(g->adapter); cudaFree(g->d_slots); cudaFree(g->x);
    cudaFree(g->xn_b); cudaFree(g->att_b); cudaFree(g->hh_b); cudaFree(g->part); cudaFree(g->logits);
    cudaFree(g->d_tokens); cudaFree(g->pf_x); cudaFree(g->pf_tok);
    if (g->ev0) cudaEventDestroy(g->ev0);
    if (g->ev1) cudaEventDestroy(g->ev1);
    if (g->eng) vox_cuda_state_free(g->eng);
    free(g);
}

static int galloc(void **p, size_t bytes, bool zero) {
    if (cudaMalloc(p, bytes) != cudaSuccess) { vox_set_error("group: cudaMalloc(%zu) failed", bytes); return -1; }
    if (zero && cudaMemset(*p, 0, bytes) != cudaSuccess) { vox_set_error("group: memset failed"); return -1; }
    return 0;
}

extern "C" vox_cuda_group_t *vox_cuda_group_create(vox_cuda_model_t *m, int n_streams, int max_positions) {
    if (!m || n_streams < 1 || n_streams > MAX_STREAMS || max_positions < 2) {
        vox_set_error("group_create: need 1..%d streams and max_positions >= 2", MAX_STREAMS);
        return nullptr;
    }
    vox_cuda_group *g = (vox_cuda_group *)calloc(1, sizeof(vox_cuda_group));
    if (!g) return nullptr;
    g->m = m; g->B = n_streams; g->max_pos = max_positions;
    g->NB = n_streams <= 32 ? 32 : (n_streams <= 64 ? 64 : 128);
    int cap = 64;
    while (cap < max_positions && cap < DEC_WINDOW) cap <<= 1;      // ring of min(window, pow2 >= positions) rows
    g->cap = cap; g->mask = cap - 1;
    g->kv_layer = (size_t)cap * DEC_KVDIM;
    g->kv_stream = g->kv_layer * m->n_dec;
    g->eng = vox_cuda_state_create(m);
    bool ok = g->eng != nullptr;
    const size_t B = n_streams, NB = g->NB;
    ok = ok && !galloc((void **)&g->kv_k, g->kv_stream * B * 4, false) && !galloc((void **)&g->kv_v, g->kv_stream * B * 4, false);
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
    }
    if (ok) {
        ok = !gemm_wt_prepare(&g->op_logits, g->NB, g->B, VOCAB, DEC_DIM, 1, g->xn_b, 2 * DEC_DIM, m->tok_emb, 1);
        g->op_logits.out_f32 = g->logits; g->op_logits.ldo = VOCAB; g->op_logits.split_stride = 0;
    }
    if (!ok) { vox_cuda_group_free(g); return nullptr; }
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
            CUDA_TRY(launch_pdl(group_norm_kernel, B, 256, 0, st, g->x, 0, nullptr, 0, 0L, g->adapter, astride, m->tok_emb, g->d_slots,
                                L.attn_norm, nullptr, g->xn_b));
        else
            CUDA_TRY(launch_pdl(group_norm_kernel, B, 256, 0, st, g->x, 1, g->part, KS_W2, (long)B * DEC_DIM, nullptr, 0L, nullptr, g->d_slots,
                                L.attn_norm, nullptr, g->xn_b));
        LAUNCHED();
        if (gemm_wt_launch(st, &g->op_qkv[l])) return -1;
        CUDA_TRY(launch_pdl(group_attn_kernel, dim3(B, DEC_KVH), 256, 0, st, g->part, KS_QKV, (long)B * DEC_QKV, g->d_slots,
                            m->inv_freq_dec, g->kv_k + l * g->kv_layer, g->kv_v + l * g->kv_layer, g->kv_stream, g->mask, g->att_b));
        LAUNCHED();
        if (gemm_wt_launch(st, &g->op_wo[l])) return -1;
        CUDA_TRY(launch_pdl(group_norm_kernel, B, 256, 0, st, g->x, 1, g->part, KS_WO, (long)B * DEC_DIM, nullptr, 0L, nullptr, g->d_slots,
                            L.ffn_norm, m->ada_scale + (size_t)l * DEC_DIM, g->xn_b));
        LAUNCHED();
        if (gemm_wt_launch(st, &g->op_w13[l])) return -1;
        if (gemm_wt_launch(st, &g->op_w2[l])) return -1;
    }
    CUDA_TRY(launch_pdl(group_norm_kernel, B, 256, 0, st, g->x, 1, g->part, KS_W2, (long)B * DEC_DIM, nullptr, 0L, nullptr, g->d_slots,
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
        float *kb = g->kv_k + (size_t)b * g->kv_stream, *vb = g->kv_v + (size_t)b * g->kv_stream;
        if (!g->count[b]) {
            // idle slot: one zero K/V row per layer so that its (ignored) attention reads initialised memory
            for (int l = 0; l < m->n_dec; l++) {
                CUDA_TRY(cudaMemsetAsync(kb + l * g->kv_layer, 0, (size_t)(n_pre + 1) * DEC_KVDIM * 4, st));
                CUDA_TRY(cudaMemsetAsync(vb + l * g->kv_layer, 0, (size_t)(n_pre + 1) * DEC_KVDIM * 4, st));
            }
            continue;
        }
        group_embed_rows_kernel<<<n_pre, 256, 0, st>>>(g->pf_tok, g->adapter + (size_t)b * g->max_pos * DEC_DIM, m->tok_emb, g->pf_x);
        LAUNCHED();
        int pos = 0, n_valid = 0;
        if (dec_prefill_into(g->eng, g->pf_x, n_pre, kb, vb, g->kv_layer, g->mask, &pos, &n_valid)) return -1;
    }
    for (int step = 0; step < n_steps; step++)
        if (group_step(g, step)) return -1;
    CUDA_TRY(cudaEventRecord(g->ev1, st));
    int *h_tok = (int *)malloc((size_t)n_steps * B * 4);
    if (!h_tok) return -1;
    if (cudaMemcpyAsync(h_tok, g->d_tokens, (size_t)n_steps * B * 4, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
        cudaStreamSynchronize(st) != cudaSuccess) {
        vox_set_error("group_decode: %s", cudaGetErrorString(cudaGetLastError()));
        free(h_tok);
        return -1;
    }
    for (int b = 0; b < B; b++)
        for (int i = 0; i < counts_out[b]; i++) tokens_out[(size_t)b * max_tokens + i] = h_tok[(size_t)i * B + b];
    free(h_tok);
    if (device_ms) cudaEventElapsedTime(device_ms, g->ev0, g->ev1);
    return n_steps;
}
