// dec_kernels.cu - single-token decoder step, one kernel per fused stage.
//
// Mirrors vox_decoder_forward (/root/reference/voxtral_decoder.c:640-780) with
// the fusions SURVEY.md 2.3 lists: RMSNorm folded into the GEMV prologue, RoPE
// and the KV-ring write folded into the QKV epilogue, residual adds folded into
// wo / w2, SwiGLU folded into the w1|w3 GEMV, argmax folded into the logits GEMV.
// Activations and accumulation are fp32 (bf16 -> fp32 widening of the weights is
// exact), so the only difference to the oracle is summation order.
//
// All kernels read the step scalars from DecStepParams in HBM; the captured CUDA
// graph is therefore position independent and is replayed once per token.
#include "vox_internal.h"

// x vector in shared memory, split so that lane l reads conflict-free float4s:
// A[c] = x[8c..8c+3], B[c] = x[8c+4..8c+7]
template <int K>
struct XSmem {
    float4 A[K / 8];
    float4 B[K / 8];
};

template <int K, int NW>
__device__ __forceinline__ void warp_gemv(const uint16_t *__restrict__ W, size_t row0,
                                          const XSmem<K> &xs, float (&acc)[NW]) {
    constexpr int NCH = K / 256;
    const int lane = threadIdx.x & 31;
    #pragma unroll
    for (int r = 0; r < NW; r++) acc[r] = 0.f;
    const uint16_t *base = W + row0 * (size_t)K + lane * 8;
    #pragma unroll 4
    for (int j = 0; j < NCH; j++) {
        const int c = j * 32 + lane;
        uint4 w[NW];
        #pragma unroll
        for (int r = 0; r < NW; r++) w[r] = ldg_stream(base + (size_t)r * K + j * 256);
        const float4 a = xs.A[c], b = xs.B[c];
        #pragma unroll
        for (int r = 0; r < NW; r++) acc[r] = dot8(w[r], a, b, acc[r]);
    }
    #pragma unroll
    for (int r = 0; r < NW; r++) acc[r] = warp_sum(acc[r]);
}

// Load x[K] into XSmem, optionally applying RMSNorm (weight nw) and the
// (1 + ada) time-conditioning scale.  Block-wide; ends with __syncthreads().
// RMSNorm per /root/reference/voxtral_kernels.c:475-492.
template <int K, int NT>
__device__ __forceinline__ void load_x_norm(XSmem<K> &xs, const float *__restrict__ x,
                                            const float *__restrict__ nw,
                                            const float *__restrict__ ada, float *red) {
    const int t = threadIdx.x;
    float rinv = 1.0f;
    if (nw) {
        float ss = 0.f;
        for (int i = t; i < K; i += NT) { float v = x[i]; ss += v * v; }
        ss = warp_sum(ss);
        if ((t & 31) == 0) red[t >> 5] = ss;
        __syncthreads();
        float tot = 0.f;
        #pragma unroll
        for (int w = 0; w < NT / 32; w++) tot += red[w];
        rinv = 1.0f / sqrtf(tot / (float)K + NORM_EPS);
    }
    for (int c = t; c < K / 8; c += NT) {
        float4 a = *reinterpret_cast<const float4 *>(x + c * 8);
        float4 b = *reinterpret_cast<const float4 *>(x + c * 8 + 4);
        if (nw) {
            const float4 wa = *reinterpret_cast<const float4 *>(nw + c * 8);
            const float4 wb = *reinterpret_cast<const float4 *>(nw + c * 8 + 4);
            a.x = a.x * rinv * wa.x; a.y = a.y * rinv * wa.y; a.z = a.z * rinv * wa.z; a.w = a.w * rinv * wa.w;
            b.x = b.x * rinv * wb.x; b.y = b.y * rinv * wb.y; b.z = b.z * rinv * wb.z; b.w = b.w * rinv * wb.w;
            if (ada) {
                const float4 da = *reinterpret_cast<const float4 *>(ada + c * 8);
                const float4 db = *reinterpret_cast<const float4 *>(ada + c * 8 + 4);
                a.x *= (1.0f + da.x); a.y *= (1.0f + da.y); a.z *= (1.0f + da.z); a.w *= (1.0f + da.w);
                b.x *= (1.0f + db.x); b.y *= (1.0f + db.y); b.z *= (1.0f + db.z); b.w *= (1.0f + db.w);
            }
        }
        xs.A[c] = a; xs.B[c] = b;
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------
// K0: build the step input and the RoPE table.
//   x = adapter[idx] + bf16->f32(tok_emb[prev])   (voxtral.c:1106-1113)
//   rope[d] = (cos, sin)(pos * inv_freq[d])        (voxtral_kernels.c:617-629)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) dec_embed_kernel(
    const DecStepParams *__restrict__ p, const float *__restrict__ embed_in,
    const float *__restrict__ adapter, const uint16_t *__restrict__ tok_emb, const float *__restrict__ tok_scale,
    const float *__restrict__ inv_freq, float *__restrict__ x, float *__restrict__ rope) {
    const int t = threadIdx.x;
    if (p->use_host_embed) {
        for (int i = t; i < DEC_DIM; i += 256) x[i] = embed_in[i];
    } else {
        const float *a = adapter + (size_t)p->adapter_index * DEC_DIM;
        const uint16_t *e = tok_emb + (size_t)p->prev_token * DEC_DIM;
        if (tok_scale) {                  // Q8 model: int-valued row * row scale (tok_embed_q8_to_f32, voxtral.c:443-452)
            const float es = tok_scale[p->prev_token];
            for (int i = t; i < DEC_DIM; i += 256)
                x[i] = a[i] + __uint_as_float(((uint32_t)e[i]) << 16) * es;
        } else {
            for (int i = t; i < DEC_DIM; i += 256)
                x[i] = a[i] + __uint_as_float(((uint32_t)e[i]) << 16);
        }
    }
    if (t < DEC_HD / 2) {
        const float angle = (float)p->pos * inv_freq[t];
        rope[2 * t] = cosf(angle);
        rope[2 * t + 1] = sinf(angle);
    }
}

// ---------------------------------------------------------------------------
// K1: x_norm = rmsnorm(x) ; [q|k|v] = Wqkv x_norm ; RoPE(q,k) ; K,V -> ring slot
// (voxtral_decoder.c:710-725).  Each warp produces adjacent row pairs so that
// the interleaved-pair rotation (voxtral_kernels.c:631-655) is warp-local.
// ---------------------------------------------------------------------------
constexpr int QKV_WARPS = 8;
constexpr int QKV_ROWS_PER_WARP = 4;   // two rotation pairs
constexpr int QKV_ROWS_PER_CTA = QKV_WARPS * QKV_ROWS_PER_WARP;  // 32 -> 192 CTAs

__global__ void __launch_bounds__(QKV_WARPS * 32) dec_qkv_kernel(
    const DecStepParams *__restrict__ p, const float *__restrict__ x,
    const float *__restrict__ attn_norm, const uint16_t *__restrict__ wqkv,
    const float *__restrict__ rope, float *__restrict__ q_out,
    float *__restrict__ kc, float *__restrict__ vc) {
    __shared__ XSmem<DEC_DIM> xs;
    __shared__ float red[QKV_WARPS];
    load_x_norm<DEC_DIM, QKV_WARPS * 32>(xs, x, attn_norm, nullptr, red);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row0 = blockIdx.x * QKV_ROWS_PER_CTA + warp * QKV_ROWS_PER_WARP;
    float acc[QKV_ROWS_PER_WARP];
    warp_gemv<DEC_DIM, QKV_ROWS_PER_WARP>(wqkv, row0, xs, acc);
    if (lane == 0) {
        const int slot = p->pos & (DEC_WINDOW - 1);
        #pragma unroll
        for (int r = 0; r < QKV_ROWS_PER_WARP; r += 2) {
            const int row = row0 + r;
            float v0 = acc[r], v1 = acc[r + 1];
            if (row < DEC_QDIM + DEC_KVDIM) {
                const int d = (row & (DEC_HD - 1)) >> 1;
                const float c = rope[2 * d], s = rope[2 * d + 1];
                const float r0 = v0 * c - v1 * s, r1 = v0 * s + v1 * c;
                v0 = r0; v1 = r1;
            }
            float *dst;
            if (row < DEC_QDIM) dst = q_out + row;
            else if (row < DEC_QDIM + DEC_KVDIM) dst = kc + (size_t)slot * DEC_KVDIM + (row - DEC_QDIM);
            else dst = vc + (size_t)slot * DEC_KVDIM + (row - DEC_QDIM - DEC_KVDIM);
            dst[0] = v0; dst[1] = v1;
        }
    }
}

// ---------------------------------------------------------------------------
// K2: GQA attention for one query over the KV ring, split over keys
// (flash-decoding).  Semantics: vox_causal_attention with seq_q=1
// (voxtral_kernels.c:541-611, call voxtral_decoder.c:731-733): keys are the
// last min(pos+1, 8192) logical positions.  grid = (8 kv heads, ATTN_NSPLIT).
// Phase 1: thread-per-key scores for the 4 query heads sharing the kv head.
// Phase 2: each warp accumulates p*V for its 32 keys, lanes own 4 dims.
// ---------------------------------------------------------------------------
constexpr int ATT_T = 128;

__global__ void __launch_bounds__(ATT_T) dec_attn_kernel(
    const DecStepParams *__restrict__ p, const float *__restrict__ q,
    const float *__restrict__ kc, const float *__restrict__ vc,
    float *__restrict__ part, int kvh0) {
    const int kvh = kvh0 + blockIdx.x, split = blockIdx.y;
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    __shared__ float4 qs[4][DEC_HD / 4];
    __shared__ float ps[4][ATT_T];
    __shared__ float m_run[4], l_run[4], corr[4];
    __shared__ float4 accs[4][4][DEC_HD / 4];  // [warp][head][dim/4]

    const int pos = p->pos;
    const int n_keys = min(p->n_valid, DEC_WINDOW - 1) + 1;
    const int first = pos - n_keys + 1;           // logical position of key 0
    int per = (n_keys + ATTN_NSPLIT - 1) / ATTN_NSPLIT;
    per = (per + ATT_T - 1) / ATT_T * ATT_T;
    const int k0 = split * per, k1 = min(n_keys, k0 + per);
    float *out = part + ((size_t)split * DEC_HEADS + kvh * 4) * (DEC_HD + 2);
    if (k0 >= k1) {
        if (t < 4) { out[t * (DEC_HD + 2) + DEC_HD] = -INFINITY; out[t * (DEC_HD + 2) + DEC_HD + 1] = 0.f; }
        return;
    }
    {
        const float4 *qg = reinterpret_cast<const float4 *>(q + kvh * 4 * DEC_HD);
        for (int i = t; i < 4 * DEC_HD / 4; i += ATT_T) qs[i / (DEC_HD / 4)][i % (DEC_HD / 4)] = qg[i];
        if (t < 4) { m_run[t] = -INFINITY; l_run[t] = 0.f; }
    }
    float4 acc[4];
    #pragma unroll
    for (int h = 0; h < 4; h++) acc[h] = make_float4(0, 0, 0, 0);
    const float scale = 0.08838834764831845f;  // 1/sqrt(128)
    __syncthreads();

    for (int base = k0; base < k1; base += ATT_T) {
        const int j = base + t;
        float s[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        if (j < k1) {
            const int slot = (first + j) & (DEC_WINDOW - 1);
            const float4 *kr = reinterpret_cast<const float4 *>(kc + (size_t)slot * DEC_KVDIM + kvh * DEC_HD);
            float d0 = 0, d1 = 0, d2 = 0, d3 = 0;
            #pragma unroll 8
            for (int i = 0; i < DEC_HD / 4; i++) {
                const float4 kv = kr[i];
                const float4 q0 = qs[0][i], q1 = qs[1][i], q2 = qs[2][i], q3 = qs[3][i];
                d0 += kv.x * q0.x + kv.y * q0.y + kv.z * q0.z + kv.w * q0.w;
                d1 += kv.x * q1.x + kv.y * q1.y + kv.z * q1.z + kv.w * q1.w;
                d2 += kv.x * q2.x + kv.y * q2.y + kv.z * q2.z + kv.w * q2.w;
                d3 += kv.x * q3.x + kv.y * q3.y + kv.z * q3.z + kv.w * q3.w;
            }
            s[0] = d0 * scale; s[1] = d1 * scale; s[2] = d2 * scale; s[3] = d3 * scale;
        }
        #pragma unroll
        for (int h = 0; h < 4; h++) ps[h][t] = s[h];
        __syncthreads();
        {   // warp h: tile max for head h, new running max and correction
            const int h = warp;
            float mx = fmaxf(fmaxf(ps[h][lane], ps[h][lane + 32]), fmaxf(ps[h][lane + 64], ps[h][lane + 96]));
            mx = warp_max(mx);
            if (lane == 0) {
                const float mo = m_run[h], mn = fmaxf(mo, mx);
                corr[h] = (mo == -INFINITY) ? 0.f : expf(mo - mn);
                m_run[h] = mn;
            }
        }
        __syncthreads();
        #pragma unroll
        for (int h = 0; h < 4; h++) ps[h][t] = (j < k1) ? expf(s[h] - m_run[h]) : 0.f;
        __syncthreads();
        {   // warp h: running sum for head h
            const int h = warp;
            float sm = ps[h][lane] + ps[h][lane + 32] + ps[h][lane + 64] + ps[h][lane + 96];
            sm = warp_sum(sm);
            if (lane == 0) l_run[h] = l_run[h] * corr[h] + sm;
        }
        // phase 2: this warp's 32 keys, all 4 heads, lane owns dims 4*lane..4*lane+3
        #pragma unroll
        for (int h = 0; h < 4; h++) {
            const float c = corr[h];
            acc[h].x *= c; acc[h].y *= c; acc[h].z *= c; acc[h].w *= c;
        }
        const int kb = base + warp * 32;
        const int kn = min(32, k1 - kb);
        for (int i = 0; i < kn; i++) {
            const int slot = (first + kb + i) & (DEC_WINDOW - 1);
            const float4 v = reinterpret_cast<const float4 *>(vc + (size_t)slot * DEC_KVDIM + kvh * DEC_HD)[lane];
            #pragma unroll
            for (int h = 0; h < 4; h++) {
                const float w = ps[h][warp * 32 + i];
                acc[h].x = fmaf(w, v.x, acc[h].x); acc[h].y = fmaf(w, v.y, acc[h].y);
                acc[h].z = fmaf(w, v.z, acc[h].z); acc[h].w = fmaf(w, v.w, acc[h].w);
            }
        }
        __syncthreads();
    }
    #pragma unroll
    for (int h = 0; h < 4; h++) accs[warp][h][lane] = acc[h];
    __syncthreads();
    {   // warp h reduces head h over the 4 warps
        const int h = warp;
        float4 a = accs[0][h][lane];
        #pragma unroll
        for (int w = 1; w < 4; w++) {
            const float4 b = accs[w][h][lane];
            a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
        }
        float *o = out + h * (DEC_HD + 2);
        reinterpret_cast<float2 *>(o)[lane * 2] = make_float2(a.x, a.y);
        reinterpret_cast<float2 *>(o)[lane * 2 + 1] = make_float2(a.z, a.w);
        if (lane == 0) { o[DEC_HD] = m_run[h]; o[DEC_HD + 1] = l_run[h]; }
    }
}

// K3: merge the split partials: out = sum_s e^{m_s-M} acc_s / sum_s e^{m_s-M} l_s
__global__ void __launch_bounds__(DEC_HD) dec_attn_combine_kernel(
    const float *__restrict__ part, float *__restrict__ attn, int h0) {
    const int h = h0 + blockIdx.x, d = threadIdx.x;
    float M = -INFINITY;
    #pragma unroll
    for (int s = 0; s < ATTN_NSPLIT; s++)
        M = fmaxf(M, part[((size_t)s * DEC_HEADS + h) * (DEC_HD + 2) + DEC_HD]);
    float num = 0.f, den = 0.f;
    #pragma unroll
    for (int s = 0; s < ATTN_NSPLIT; s++) {
        const float *o = part + ((size_t)s * DEC_HEADS + h) * (DEC_HD + 2);
        const float m = o[DEC_HD];
        if (m == -INFINITY) continue;
        const float w = expf(m - M);
        num += w * o[d];
        den += w * o[DEC_HD + 1];
    }
    attn[h * DEC_HD + d] = den > 0.f ? num / den : 0.f;
}

// ---------------------------------------------------------------------------
// K4 / K6: x += W v   (wo: K=4096, voxtral_decoder.c:738-739;
//                      w2: K=9216, voxtral_decoder.c:758-760)
// ---------------------------------------------------------------------------
constexpr int RES_WARPS = 8;
constexpr int RES_ROWS_PER_WARP = 2;
constexpr int RES_ROWS_PER_CTA = RES_WARPS * RES_ROWS_PER_WARP;  // 16 -> 192 CTAs

template <int K>
__global__ void __launch_bounds__(RES_WARPS * 32) dec_gemv_residual_kernel(
    const float *__restrict__ v, const uint16_t *__restrict__ W, float *__restrict__ x) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    XSmem<K> &xs = *reinterpret_cast<XSmem<K> *>(smem_raw);
    load_x_norm<K, RES_WARPS * 32>(xs, v, nullptr, nullptr, nullptr);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row0 = blockIdx.x * RES_ROWS_PER_CTA + warp * RES_ROWS_PER_WARP;
    float acc[RES_ROWS_PER_WARP];
    warp_gemv<K, RES_ROWS_PER_WARP>(W, row0, xs, acc);
    if (lane == 0) {
        #pragma unroll
        for (int r = 0; r < RES_ROWS_PER_WARP; r++) x[row0 + r] += acc[r];
    }
}

// ---------------------------------------------------------------------------
// K5: x_norm = rmsnorm(x, ffn_norm) * (1 + ada) ; h = silu(w1 x_norm) * (w3 x_norm)
// (voxtral_decoder.c:741-757).  w13 rows are interleaved (2i = w1[i], 2i+1 = w3[i]).
// ---------------------------------------------------------------------------
constexpr int FFN_WARPS = 8;
constexpr int FFN_ROWS_PER_WARP = 4;  // two (gate, up) pairs
constexpr int FFN_ROWS_PER_CTA = FFN_WARPS * FFN_ROWS_PER_WARP;  // 32 -> 576 CTAs

__global__ void __launch_bounds__(FFN_WARPS * 32) dec_ffn_gate_kernel(
    const float *__restrict__ x, const float *__restrict__ ffn_norm,
    const float *__restrict__ ada, const uint16_t *__restrict__ w13,
    float *__restrict__ h) {
    __shared__ XSmem<DEC_DIM> xs;
    __shared__ float red[FFN_WARPS];
    load_x_norm<DEC_DIM, FFN_WARPS * 32>(xs, x, ffn_norm, ada, red);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row0 = blockIdx.x * FFN_ROWS_PER_CTA + warp * FFN_ROWS_PER_WARP;
    float acc[FFN_ROWS_PER_WARP];
    warp_gemv<DEC_DIM, FFN_ROWS_PER_WARP>(w13, row0, xs, acc);
    if (lane == 0) {
        #pragma unroll
        for (int r = 0; r < FFN_ROWS_PER_WARP; r += 2)
            h[(row0 + r) >> 1] = silu(acc[r]) * acc[r + 1];
    }
}

// ---------------------------------------------------------------------------
// K7: logits = tok_emb . rmsnorm(x, norm), per-CTA argmax (first max wins,
// voxtral_decoder.c:765-779).  K8: final reduce + advance the step scalars.
// ---------------------------------------------------------------------------
constexpr int LOG_WARPS = 8;
constexpr int LOG_ROWS_PER_WARP = 4;
static_assert(LOGITS_ROWS_PER_CTA % (LOG_WARPS * LOG_ROWS_PER_WARP) == 0, "");

__device__ __forceinline__ void argmax_merge(float &bv, int &bi, float v, int i) {
    // strict '>' with lowest index on ties; NaN never wins (comparison false)
    if (v > bv || (v == bv && i < bi)) { bv = v; bi = i; }
}

__global__ void __launch_bounds__(LOG_WARPS * 32) dec_logits_kernel(
    const float *__restrict__ x, const float *__restrict__ norm,
    const uint16_t *__restrict__ tok_emb, float *__restrict__ logits,
    float *__restrict__ blockmax) {
    __shared__ XSmem<DEC_DIM> xs;
    __shared__ float red[LOG_WARPS];
    __shared__ float bvs[LOG_WARPS];
    __shared__ int bis[LOG_WARPS];
    load_x_norm<DEC_DIM, LOG_WARPS * 32>(xs, x, norm, nullptr, red);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float bv = -INFINITY; int bi = 0x7fffffff;
    for (int it = 0; it < LOGITS_ROWS_PER_CTA / (LOG_WARPS * LOG_ROWS_PER_WARP); it++) {
        const int row0 = blockIdx.x * LOGITS_ROWS_PER_CTA +
                         (it * LOG_WARPS + warp) * LOG_ROWS_PER_WARP;
        float acc[LOG_ROWS_PER_WARP];
        warp_gemv<DEC_DIM, LOG_ROWS_PER_WARP>(tok_emb, row0, xs, acc);
        if (lane == 0) {
            #pragma unroll
            for (int r = 0; r < LOG_ROWS_PER_WARP; r++) {
                logits[row0 + r] = acc[r];
                argmax_merge(bv, bi, acc[r], row0 + r);
            }
        }
    }
    if (lane == 0) { bvs[warp] = bv; bis[warp] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        #pragma unroll
        for (int w = 1; w < LOG_WARPS; w++) argmax_merge(bv, bi, bvs[w], bis[w]);
        blockmax[2 * blockIdx.x] = bv;
        blockmax[2 * blockIdx.x + 1] = __int_as_float(bi);
    }
}

__global__ void __launch_bounds__(1024) dec_argmax_final_kernel(
    DecStepParams *__restrict__ p, const float *__restrict__ blockmax, int nblk,
    int *__restrict__ tokens) {
    __shared__ float bvs[32];
    __shared__ int bis[32];
    const int t = threadIdx.x;
    float bv = -INFINITY; int bi = 0x7fffffff;
    for (int i = t; i < nblk; i += 1024)
        argmax_merge(bv, bi, blockmax[2 * i], __float_as_int(blockmax[2 * i + 1]));
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        argmax_merge(bv, bi, ov, oi);
    }
    if ((t & 31) == 0) { bvs[t >> 5] = bv; bis[t >> 5] = bi; }
    __syncthreads();
    if (t == 0) {
        for (int w = 1; w < 32; w++) argmax_merge(bv, bi, bvs[w], bis[w]);
        if (bi == 0x7fffffff) bi = 0;  // all-NaN logits: reference returns index 0
        p->out_token = bi;
        p->prev_token = bi;
        tokens[p->step_count & (TOKENS_RING - 1)] = bi;
        p->step_count += 1;
        p->pos += 1;
        p->n_valid = min(p->n_valid + 1, DEC_WINDOW);
        p->adapter_index += 1;
    }
}

// ---------------------------------------------------------------------------
// Launchers shared with the tensor-parallel step (dec_tp.cu): the step input and
// the attention of a contiguous range of kv heads (a rank's shard).
// ---------------------------------------------------------------------------
void dec_launch_embed(cudaStream_t st, const DecStepParams *p, const float *embed_in, const float *adapter,
                      const uint16_t *tok_emb, const float *tok_scale, const float *inv_freq, float *x, float *rope) {
    dec_embed_kernel<<<1, 256, 0, st>>>(p, embed_in, adapter, tok_emb, tok_scale, inv_freq, x, rope);
    LAUNCHED();
}

void dec_launch_attn(cudaStream_t st, const DecStepParams *p, const float *q, const float *kc, const float *vc,
                     float *part, float *attn, int kvh0, int n_kvh) {
    dec_attn_kernel<<<dim3(n_kvh, ATTN_NSPLIT), ATT_T, 0, st>>>(p, q, kc, vc, part, kvh0);
    dec_attn_combine_kernel<<<n_kvh * (DEC_HEADS / DEC_KVH), DEC_HD, 0, st>>>(part, attn, kvh0 * (DEC_HEADS / DEC_KVH));
    g_launch_count += 2;
}

// ---------------------------------------------------------------------------
// Host: enqueue one decoder step (captured once into a CUDA graph by the caller)
// ---------------------------------------------------------------------------
int dec_launch_step(vox_cuda_state *s, cudaStream_t st) {
    vox_cuda_model *m = s->m;
    const size_t kv_layer = (size_t)DEC_WINDOW * DEC_KVDIM;
    dec_embed_kernel<<<1, 256, 0, st>>>(s->d_params, s->d_embed_in, s->adapter_buf, m->tok_emb, (const float *)nullptr,
                                        m->inv_freq_dec, s->d_x, s->d_rope);
    LAUNCHED();
    static bool attr_set = false;
    if (!attr_set) {
        CUDA_TRY(cudaFuncSetAttribute(dec_gemv_residual_kernel<DEC_HIDDEN>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)sizeof(XSmem<DEC_HIDDEN>)));
        attr_set = true;
    }
    for (int l = 0; l < m->n_dec; l++) {
        const DecLayerDev &L = m->dec[l];
        float *kc = s->kv_k + l * kv_layer, *vc = s->kv_v + l * kv_layer;
        dec_qkv_kernel<<<DEC_QKV / QKV_ROWS_PER_CTA, QKV_WARPS * 32, 0, st>>>(
            s->d_params, s->d_x, L.attn_norm, L.wqkv, s->d_rope, s->d_q, kc, vc);
        dec_attn_kernel<<<dim3(DEC_KVH, ATTN_NSPLIT), ATT_T, 0, st>>>(
            s->d_params, s->d_q, kc, vc, s->d_attn_part, 0);
        dec_attn_combine_kernel<<<DEC_HEADS, DEC_HD, 0, st>>>(s->d_attn_part, s->d_attn, 0);
        dec_gemv_residual_kernel<DEC_QDIM><<<DEC_DIM / RES_ROWS_PER_CTA, RES_WARPS * 32,
                                             sizeof(XSmem<DEC_QDIM>), st>>>(s->d_attn, L.wo, s->d_x);
        dec_ffn_gate_kernel<<<2 * DEC_HIDDEN / FFN_ROWS_PER_CTA, FFN_WARPS * 32, 0, st>>>(
            s->d_x, L.ffn_norm, m->ada_scale + (size_t)l * DEC_DIM, L.w13, s->d_h);
        dec_gemv_residual_kernel<DEC_HIDDEN><<<DEC_DIM / RES_ROWS_PER_CTA, RES_WARPS * 32,
                                               sizeof(XSmem<DEC_HIDDEN>), st>>>(s->d_h, L.w2, s->d_x);
        g_launch_count += 6;
    }
    const int nblk = VOCAB / LOGITS_ROWS_PER_CTA;
    dec_logits_kernel<<<nblk, LOG_WARPS * 32, 0, st>>>(s->d_x, m->dec_norm, m->tok_emb,
                                                       s->d_logits, s->d_blockmax);
    dec_argmax_final_kernel<<<1, 1024, 0, st>>>(s->d_params, s->d_blockmax, nblk, s->d_tokens);
    g_launch_count += 2;
    CUDA_TRY(cudaGetLastError());
    return 0;
}
