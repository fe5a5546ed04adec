// dec_mega.cu - the single-token decoder step as ONE persistent cooperative kernel.
//
// vox_decoder_forward (/root/reference/voxtral_decoder.c:640-780) streams 6.86 GB
// of bf16 weights per token through 26 x {qkv, attention, wo, w1|w3, w2} + logits.
// Each matrix is only 25-113 MB, i.e. 4-17 us of HBM time, so a kernel per op
// spends a large part of the step on launch gaps, ramp-up and tails.  Here one
// CTA per SM stays resident for the whole step (or for n steps):
//
//   * warp 8 (one elected lane) is the PRODUCER: it walks the static list of
//     weight tiles this CTA owns - contiguous row ranges of every matrix, in
//     consumption order, across phase, layer and token boundaries - and streams
//     them into an 8 x 24 KB shared-memory ring with cp.async.bulk (TMA 1-D
//     bulk copies, mbarrier complete_tx, L2 evict_first).  It never waits for
//     activations, only for ring space, so the HBM pipe stays busy while the
//     consumers sit in a grid barrier (192 KB/SM = 4.4 us of HBM time).
//   * warps 0-7 are CONSUMERS: each warp owns a fixed K-slice of every row
//     (the activation slice lives in registers), multiplies tiles as they land,
//     and the per-tile epilogue applies RoPE + KV-ring write, residual add,
//     SwiGLU, or logits + running argmax.
//   * phases are separated by a grid-wide barrier (release/acquire counter in
//     global memory); attention is split over (kv head, key range) CTAs and its
//     partials are merged in the wo prologue.
//
// Arithmetic is fp32 on exactly-widened bf16 weights, like the reference; only
// the summation order differs.
#include "vox_internal.h"
#include <stdlib.h>

namespace mk {

constexpr int CW = 8;                       // consumer warps
constexpr int CT = CW * 32;                 // consumer threads
constexpr int NT = CT + 32;                 // + producer warp
constexpr int STAGE_BYTES = 24576;
constexpr int NSTAGE = 8;
constexpr int MAXTR = 4;                    // rows per tile at K = 3072
constexpr int PART_LD = DEC_HD + 4;         // acc[128], m, l, pad (16-byte rows)
constexpr int MAX_SPLITS = 32;

template <int K> struct TileRows { static constexpr int v = STAGE_BYTES / (K * 2); };
static_assert(TileRows<3072>::v == 4 && TileRows<4096>::v == 3 && TileRows<9216>::v == 1, "");

struct Smem {
    alignas(128) unsigned char ring[NSTAGE][STAGE_BYTES];
    alignas(8) unsigned long long full[NSTAGE];
    alignas(8) unsigned long long empty[NSTAGE];
    float part[2][MAXTR][CW];
    float red[CW];
    float rope[DEC_HD];                     // (cos, sin) x 64 for this step's position
    alignas(16) float qs[4][DEC_HD];
    alignas(16) float wsm[CW][4][PART_LD];
    float bv[CW];
    int bi[CW];
    int token;
};

struct Params {
    const DecLayerDev *layers;              // device array [n_dec]
    int n_dec;
    const uint16_t *tok_emb;
    const float *dec_norm, *ada_scale, *inv_freq;
    float *kv_k, *kv_v;
    float *x, *q, *attn_part, *h, *logits, *blockmax;
    const float *embed_in, *adapter;
    DecStepParams *params;
    int *tokens;
    unsigned *bar;
    unsigned bar_base;
    int n_steps;
    unsigned *err;
};

// ---- PTX helpers ---------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long *b, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *b, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(unsigned long long *b, unsigned parity) {
    unsigned ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(smem_u32(b)), "r"(parity) : "memory");
    return ok != 0;
}
// bounded wait: a protocol bug must end in a trap, never in a hung GPU
__device__ __forceinline__ void mbar_wait(unsigned long long *b, unsigned parity, unsigned *err) {
    if (mbar_try_wait(b, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try_wait(b, parity)) {
        if (clock64() - t0 > 8000000000LL) { atomicExch(err, 1u); __trap(); }
    }
}
__device__ __forceinline__ unsigned long long policy_evict_first() {
    unsigned long long p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *bar,
                                         unsigned long long policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
        ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy) : "memory");
}
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(CT) : "memory"); }

__device__ __forceinline__ unsigned ld_acquire(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_add(unsigned *p, unsigned v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Grid-wide barrier for the consumer threads of all CTAs.  `target` is the
// counter value at which this barrier opens (wrap-safe signed compare).
__device__ __forceinline__ void grid_sync(unsigned *bar, unsigned target, unsigned *err) {
    consumer_sync();
    if (threadIdx.x == 0) {
        red_release_add(bar, 1u);
        const long long t0 = clock64();
        while ((int)(ld_acquire(bar) - target) < 0) {
            if (clock64() - t0 > 8000000000LL) { atomicExch(err, 2u); __trap(); }
        }
    }
    consumer_sync();
}

struct Pipe {
    int stage;
    unsigned phase;
    __device__ __forceinline__ void advance() {
        if (++stage == NSTAGE) { stage = 0; phase ^= 1u; }
    }
};

// rows [r0, r1) of an R-row matrix owned by CTA c; boundaries are multiples of `unit`
__device__ __forceinline__ void cta_rows(int R, int unit, int c, int G, int &r0, int &r1) {
    const int U = R / unit;
    r0 = (int)(((long long)c * U) / G) * unit;
    r1 = (int)(((long long)(c + 1) * U) / G) * unit;
}

// ---- producer ----------------------------------------------------------------------
template <int K>
__device__ __forceinline__ void produce(Smem &sm, Pipe &pp, const uint16_t *W, int R, int unit, int c, int G,
                                        unsigned long long policy, unsigned *err) {
    constexpr int TR = TileRows<K>::v;
    int r0, r1;
    cta_rows(R, unit, c, G, r0, r1);
    for (int r = r0; r < r1; r += TR) {
        const int n = min(TR, r1 - r);
        const unsigned bytes = (unsigned)n * K * 2u;
        mbar_wait(&sm.empty[pp.stage], pp.phase ^ 1u, err);
        mbar_expect_tx(&sm.full[pp.stage], bytes);
        bulk_g2s(sm.ring[pp.stage], W + (size_t)r * K, bytes, &sm.full[pp.stage], policy);
        pp.advance();
    }
}

// ---- consumer: tile dot products -------------------------------------------------------
// Lane l of warp w owns columns (j*256 + w*32 + l)*4 .. +3, j < K/1024, of every row.
template <int K, int TR>
__device__ __forceinline__ void tile_dot(const unsigned char *tile, int nrows, const float (&xr)[K / 256],
                                         float (&acc)[TR], int w, int lane) {
    constexpr int J = K / 1024;
    #pragma unroll
    for (int r = 0; r < TR; r++) {
        float a = 0.f;
        if (r < nrows) {
            const uint2 *rowp = reinterpret_cast<const uint2 *>(tile + (size_t)r * K * 2) + w * 32 + lane;
            uint2 v[J];
            #pragma unroll
            for (int j = 0; j < J; j++) v[j] = rowp[j * 256];
            #pragma unroll
            for (int j = 0; j < J; j++) {
                a = fmaf(bf16_lo(v[j].x), xr[4 * j + 0], a);
                a = fmaf(bf16_hi(v[j].x), xr[4 * j + 1], a);
                a = fmaf(bf16_lo(v[j].y), xr[4 * j + 2], a);
                a = fmaf(bf16_hi(v[j].y), xr[4 * j + 3], a);
            }
        }
        acc[r] = a;
    }
}

// One GEMV phase for the consumers.  EPI(row, v0, v1, npair_or_row) is called by
// one thread per output row (PAIR = false) or per adjacent row pair (PAIR = true).
template <int K, bool PAIR, typename Epi>
__device__ __forceinline__ void consume(Smem &sm, Pipe &cp, int &pbuf, int R, int unit, int c, int G,
                                        const float (&xr)[K / 256], unsigned *err, Epi epi) {
    constexpr int TR = TileRows<K>::v;
    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
    int r0, r1;
    cta_rows(R, unit, c, G, r0, r1);
    for (int r = r0; r < r1; r += TR) {
        const int n = min(TR, r1 - r);
        mbar_wait(&sm.full[cp.stage], cp.phase, err);
        float acc[TR];
        tile_dot<K, TR>(sm.ring[cp.stage], n, xr, acc, w, lane);
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.empty[cp.stage]);   // this warp is done with the stage
        cp.advance();
        #pragma unroll
        for (int i = 0; i < TR; i++) acc[i] = warp_sum(acc[i]);
        if (lane == 0) {
            #pragma unroll
            for (int i = 0; i < TR; i++) sm.part[pbuf][i][w] = acc[i];
        }
        consumer_sync();
        if (PAIR) {
            if (tid < (n >> 1)) {
                float v0 = 0.f, v1 = 0.f;
                #pragma unroll
                for (int k = 0; k < CW; k++) { v0 += sm.part[pbuf][2 * tid][k]; v1 += sm.part[pbuf][2 * tid + 1][k]; }
                epi(r + 2 * tid, v0, v1);
            }
        } else {
            if (tid < n) {
                float v0 = 0.f;
                #pragma unroll
                for (int k = 0; k < CW; k++) v0 += sm.part[pbuf][tid][k];
                epi(r + tid, v0, 0.f);
            }
        }
        pbuf ^= 1;
    }
}

__device__ __forceinline__ float block_sum(Smem &sm, float v) {
    v = warp_sum(v);
    consumer_sync();                       // previous readers of red[] are done
    if ((threadIdx.x & 31) == 0) sm.red[threadIdx.x >> 5] = v;
    consumer_sync();
    float t = 0.f;
    #pragma unroll
    for (int k = 0; k < CW; k++) t += sm.red[k];
    return t;
}

__device__ __forceinline__ void argmax_merge(float &bv, int &bi, float v, int i) {
    // strict '>' with lowest index on ties; NaN never wins (voxtral_decoder.c:771-778)
    if (v > bv || (v == bv && i < bi)) { bv = v; bi = i; }
}

// ---- attention for one (kv head, key range) ------------------------------------------------
// vox_causal_attention with seq_q = 1 (voxtral_kernels.c:541-611): the query at
// logical position pos sees the last n_keys positions of the ring.
__device__ __forceinline__ void attention_phase(Smem &sm, const Params &P, const float *kc, const float *vc,
                                                int pos, int n_keys, int c, int G) {
    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
    const int kvh = c & 7, split = c >> 3;
    const int s_max = min(G >> 3, MAX_SPLITS);
    const int s_eff = min(s_max, (n_keys + 255) >> 8);
    if (split >= s_eff) return;
    const int first = pos - n_keys + 1;
    const int nb = (n_keys + 31) >> 5;
    const int b0 = (int)(((long long)split * nb) / s_eff), b1 = (int)(((long long)(split + 1) * nb) / s_eff);
    {
        const float4 *qg = reinterpret_cast<const float4 *>(P.q + kvh * 4 * DEC_HD);
        if (tid < 4 * DEC_HD / 4) reinterpret_cast<float4 *>(&sm.qs[0][0])[tid] = __ldcg(qg + tid);
    }
    consumer_sync();
    const float scale = 0.08838834764831845f;  // 1/sqrt(128)
    float m[4], l[4];
    float4 acc[4];
    #pragma unroll
    for (int h = 0; h < 4; h++) { m[h] = -INFINITY; l[h] = 0.f; acc[h] = make_float4(0.f, 0.f, 0.f, 0.f); }
    for (int b = b0 + w; b < b1; b += CW) {
        const int j = b * 32 + lane;           // key index within the window
        const bool valid = j < n_keys;
        float s[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        if (valid) {
            const int slot = (first + j) & (DEC_WINDOW - 1);
            const float4 *kr = reinterpret_cast<const float4 *>(kc + (size_t)slot * DEC_KVDIM + kvh * DEC_HD);
            float d0 = 0.f, d1 = 0.f, d2 = 0.f, d3 = 0.f;
            #pragma unroll 8
            for (int i = 0; i < DEC_HD / 4; i++) {
                const float4 kv = kr[i];
                const float4 q0 = reinterpret_cast<const float4 *>(sm.qs[0])[i];
                const float4 q1 = reinterpret_cast<const float4 *>(sm.qs[1])[i];
                const float4 q2 = reinterpret_cast<const float4 *>(sm.qs[2])[i];
                const float4 q3 = reinterpret_cast<const float4 *>(sm.qs[3])[i];
                d0 += kv.x * q0.x + kv.y * q0.y + kv.z * q0.z + kv.w * q0.w;
                d1 += kv.x * q1.x + kv.y * q1.y + kv.z * q1.z + kv.w * q1.w;
                d2 += kv.x * q2.x + kv.y * q2.y + kv.z * q2.z + kv.w * q2.w;
                d3 += kv.x * q3.x + kv.y * q3.y + kv.z * q3.z + kv.w * q3.w;
            }
            s[0] = d0 * scale; s[1] = d1 * scale; s[2] = d2 * scale; s[3] = d3 * scale;
        }
        float p[4];
        #pragma unroll
        for (int h = 0; h < 4; h++) {
            const float mn = fmaxf(m[h], warp_max(s[h]));      // block has >= 1 valid key
            const float corr = (m[h] == -INFINITY) ? 0.f : expf(m[h] - mn);
            p[h] = valid ? expf(s[h] - mn) : 0.f;
            l[h] = l[h] * corr + warp_sum(p[h]);
            acc[h].x *= corr; acc[h].y *= corr; acc[h].z *= corr; acc[h].w *= corr;
            m[h] = mn;
        }
        const int kn = min(32, n_keys - b * 32);
        for (int i = 0; i < kn; i++) {
            const int slot = (first + b * 32 + i) & (DEC_WINDOW - 1);
            const float4 v = reinterpret_cast<const float4 *>(vc + (size_t)slot * DEC_KVDIM + kvh * DEC_HD)[lane];
            #pragma unroll
            for (int h = 0; h < 4; h++) {
                const float ph = __shfl_sync(0xffffffffu, p[h], i);
                acc[h].x = fmaf(ph, v.x, acc[h].x); acc[h].y = fmaf(ph, v.y, acc[h].y);
                acc[h].z = fmaf(ph, v.z, acc[h].z); acc[h].w = fmaf(ph, v.w, acc[h].w);
            }
        }
    }
    #pragma unroll
    for (int h = 0; h < 4; h++) {
        reinterpret_cast<float4 *>(&sm.wsm[w][h][0])[lane] = acc[h];
        if (lane == 0) { sm.wsm[w][h][DEC_HD] = m[h]; sm.wsm[w][h][DEC_HD + 1] = l[h]; }
    }
    consumer_sync();
    {   // merge the 8 warps: thread -> (head, two dims)
        const int h = tid >> 6, d = (tid & 63) * 2;
        float M = -INFINITY;
        #pragma unroll
        for (int k = 0; k < CW; k++) M = fmaxf(M, sm.wsm[k][h][DEC_HD]);
        float n0 = 0.f, n1 = 0.f, den = 0.f;
        #pragma unroll
        for (int k = 0; k < CW; k++) {
            const float mk_ = sm.wsm[k][h][DEC_HD];
            const float e = (mk_ == -INFINITY) ? 0.f : expf(mk_ - M);
            n0 = fmaf(e, sm.wsm[k][h][d], n0);
            n1 = fmaf(e, sm.wsm[k][h][d + 1], n1);
            den = fmaf(e, sm.wsm[k][h][DEC_HD + 1], den);
        }
        float *o = P.attn_part + ((size_t)split * DEC_HEADS + kvh * 4 + h) * PART_LD;
        *reinterpret_cast<float2 *>(o + d) = make_float2(n0, n1);
        if (d == 0) { o[DEC_HD] = M; o[DEC_HD + 1] = den; }
    }
}

// ---- the kernel ----------------------------------------------------------------------------
__global__ void __launch_bounds__(NT, 1) dec_mega_kernel(const Params P) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem &sm = *reinterpret_cast<Smem *>(smem_raw);
    const int tid = threadIdx.x;
    const int c = blockIdx.x, G = gridDim.x;

    if (tid == 0) {
        for (int i = 0; i < NSTAGE; i++) { mbar_init(&sm.full[i], 1); mbar_init(&sm.empty[i], CW); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const int n_dec = P.n_dec;

    if (tid >= CT) {
        // =========================== PRODUCER ===========================
        if (tid != CT) return;
        const unsigned long long policy = policy_evict_first();
        Pipe pp{0, 0u};
        for (int step = 0; step < P.n_steps; step++) {
            for (int l = 0; l < n_dec; l++) {
                const DecLayerDev L = P.layers[l];
                produce<DEC_DIM>(sm, pp, L.wqkv, DEC_QKV, 2, c, G, policy, P.err);
                produce<DEC_QDIM>(sm, pp, L.wo, DEC_DIM, 1, c, G, policy, P.err);
                produce<DEC_DIM>(sm, pp, L.w13, 2 * DEC_HIDDEN, 2, c, G, policy, P.err);
                produce<DEC_HIDDEN>(sm, pp, L.w2, DEC_DIM, 1, c, G, policy, P.err);
            }
            produce<DEC_DIM>(sm, pp, P.tok_emb, VOCAB, 1, c, G, policy, P.err);
        }
        return;
    }

    // =============================== CONSUMERS ===============================
    const int w = tid >> 5, lane = tid & 31;
    Pipe cp{0, 0u};
    int pbuf = 0;
    unsigned bar_target = P.bar_base;
    // step scalars: every CTA tracks them in registers
    int pos = P.params->pos, n_valid = P.params->n_valid, prev_token = P.params->prev_token;
    int adapter_index = P.params->adapter_index, step_count = P.params->step_count;
    const int use_host_embed = P.params->use_host_embed;
    const size_t kv_layer = (size_t)DEC_WINDOW * DEC_KVDIM;
    // column base of this lane for unit j: (j*256 + w*32 + lane) * 4
    const int col0 = (w * 32 + lane) * 4;

    for (int step = 0; step < P.n_steps; step++) {
        const int n_keys = min(n_valid, DEC_WINDOW - 1) + 1;
        const int slot = pos & (DEC_WINDOW - 1);
        const int s_eff = min(min(G >> 3, MAX_SPLITS), (n_keys + 255) >> 8);
        // RoPE table of this position (voxtral_kernels.c:617-629)
        if (tid < DEC_HD / 2) {
            const float angle = (float)pos * P.inv_freq[tid];
            sm.rope[2 * tid] = cosf(angle);
            sm.rope[2 * tid + 1] = sinf(angle);
        }
        for (int l = 0; l < n_dec; l++) {
            const DecLayerDev L = P.layers[l];
            float *kc = P.kv_k + l * kv_layer, *vc = P.kv_v + l * kv_layer;
            // ---------------- phase 1: rmsnorm -> qkv -> rope -> q / KV ring ----------------
            {
                float xr[DEC_DIM / 256];
                float ss = 0.f;
                #pragma unroll
                for (int j = 0; j < DEC_DIM / 1024; j++) {
                    const int col = j * 1024 + col0;
                    float4 v;
                    if (l == 0) {
                        // x = adapter[idx] + bf16->f32(tok_emb[prev])  (voxtral.c:1106-1113)
                        if (use_host_embed) v = __ldcg(reinterpret_cast<const float4 *>(P.embed_in + col));
                        else {
                            v = __ldcg(reinterpret_cast<const float4 *>(P.adapter + (size_t)adapter_index * DEC_DIM + col));
                            const uint2 e = *reinterpret_cast<const uint2 *>(P.tok_emb + (size_t)prev_token * DEC_DIM + col);
                            v.x += bf16_lo(e.x); v.y += bf16_hi(e.x); v.z += bf16_lo(e.y); v.w += bf16_hi(e.y);
                        }
                        if (c == 0) *reinterpret_cast<float4 *>(P.x + col) = v;
                    } else {
                        v = __ldcg(reinterpret_cast<const float4 *>(P.x + col));
                    }
                    xr[4 * j] = v.x; xr[4 * j + 1] = v.y; xr[4 * j + 2] = v.z; xr[4 * j + 3] = v.w;
                    ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
                }
                const float tot = block_sum(sm, ss);
                const float rinv = 1.0f / sqrtf(tot / (float)DEC_DIM + NORM_EPS);
                #pragma unroll
                for (int j = 0; j < DEC_DIM / 1024; j++) {
                    const float4 nw = *reinterpret_cast<const float4 *>(L.attn_norm + j * 1024 + col0);
                    xr[4 * j] = xr[4 * j] * rinv * nw.x; xr[4 * j + 1] = xr[4 * j + 1] * rinv * nw.y;
                    xr[4 * j + 2] = xr[4 * j + 2] * rinv * nw.z; xr[4 * j + 3] = xr[4 * j + 3] * rinv * nw.w;
                }
                consume<DEC_DIM, true>(sm, cp, pbuf, DEC_QKV, 2, c, G, xr, P.err,
                    [&](int row, float v0, float v1) {
                        if (row < DEC_QDIM + DEC_KVDIM) {
                            const int d = (row & (DEC_HD - 1)) >> 1;
                            const float cs = sm.rope[2 * d], sn = sm.rope[2 * d + 1];
                            const float r0 = v0 * cs - v1 * sn, r1 = v0 * sn + v1 * cs;
                            v0 = r0; v1 = r1;
                        }
                        float *dst;
                        if (row < DEC_QDIM) dst = P.q + row;
                        else if (row < DEC_QDIM + DEC_KVDIM) dst = kc + (size_t)slot * DEC_KVDIM + (row - DEC_QDIM);
                        else dst = vc + (size_t)slot * DEC_KVDIM + (row - DEC_QDIM - DEC_KVDIM);
                        *reinterpret_cast<float2 *>(dst) = make_float2(v0, v1);
                    });
            }
            grid_sync(P.bar, bar_target += G, P.err);
            // ---------------- phase 2: attention partials ----------------
            attention_phase(sm, P, kc, vc, pos, n_keys, c, G);
            grid_sync(P.bar, bar_target += G, P.err);
            // ---------------- phase 3: merge partials -> wo -> residual ----------------
            {
                float xr[DEC_QDIM / 256];
                #pragma unroll
                for (int j = 0; j < DEC_QDIM / 1024; j++) {
                    const int head = j * 8 + w;          // column j*1024 + w*128 + 4*lane
                    float M = -INFINITY;
                    for (int s = 0; s < s_eff; s++)
                        M = fmaxf(M, __ldcg(P.attn_part + ((size_t)s * DEC_HEADS + head) * PART_LD + DEC_HD));
                    float4 num = make_float4(0.f, 0.f, 0.f, 0.f);
                    float den = 0.f;
                    for (int s = 0; s < s_eff; s++) {
                        const float *o = P.attn_part + ((size_t)s * DEC_HEADS + head) * PART_LD;
                        const float ms = __ldcg(o + DEC_HD);
                        const float e = (ms == -INFINITY) ? 0.f : expf(ms - M);
                        const float4 a = __ldcg(reinterpret_cast<const float4 *>(o) + lane);
                        num.x = fmaf(e, a.x, num.x); num.y = fmaf(e, a.y, num.y);
                        num.z = fmaf(e, a.z, num.z); num.w = fmaf(e, a.w, num.w);
                        den = fmaf(e, __ldcg(o + DEC_HD + 1), den);
                    }
                    const float inv = den > 0.f ? 1.0f / den : 0.f;
                    xr[4 * j] = num.x * inv; xr[4 * j + 1] = num.y * inv;
                    xr[4 * j + 2] = num.z * inv; xr[4 * j + 3] = num.w * inv;
                }
                consume<DEC_QDIM, false>(sm, cp, pbuf, DEC_DIM, 1, c, G, xr, P.err,
                    [&](int row, float v0, float) { P.x[row] = __ldcg(P.x + row) + v0; });
            }
            grid_sync(P.bar, bar_target += G, P.err);
            // ---------------- phase 4: rmsnorm * (1 + ada) -> w1|w3 -> SwiGLU ----------------
            {
                float xr[DEC_DIM / 256];
                float ss = 0.f;
                #pragma unroll
                for (int j = 0; j < DEC_DIM / 1024; j++) {
                    const float4 v = __ldcg(reinterpret_cast<const float4 *>(P.x + j * 1024 + col0));
                    xr[4 * j] = v.x; xr[4 * j + 1] = v.y; xr[4 * j + 2] = v.z; xr[4 * j + 3] = v.w;
                    ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
                }
                const float tot = block_sum(sm, ss);
                const float rinv = 1.0f / sqrtf(tot / (float)DEC_DIM + NORM_EPS);
                const float *ada = P.ada_scale + (size_t)l * DEC_DIM;
                #pragma unroll
                for (int j = 0; j < DEC_DIM / 1024; j++) {
                    const float4 nw = *reinterpret_cast<const float4 *>(L.ffn_norm + j * 1024 + col0);
                    const float4 ad = *reinterpret_cast<const float4 *>(ada + j * 1024 + col0);
                    xr[4 * j] = xr[4 * j] * rinv * nw.x * (1.0f + ad.x);
                    xr[4 * j + 1] = xr[4 * j + 1] * rinv * nw.y * (1.0f + ad.y);
                    xr[4 * j + 2] = xr[4 * j + 2] * rinv * nw.z * (1.0f + ad.z);
                    xr[4 * j + 3] = xr[4 * j + 3] * rinv * nw.w * (1.0f + ad.w);
                }
                consume<DEC_DIM, true>(sm, cp, pbuf, 2 * DEC_HIDDEN, 2, c, G, xr, P.err,
                    [&](int row, float v0, float v1) { P.h[row >> 1] = silu(v0) * v1; });
            }
            grid_sync(P.bar, bar_target += G, P.err);
            // ---------------- phase 5: w2 -> residual ----------------
            {
                float xr[DEC_HIDDEN / 256];
                #pragma unroll
                for (int j = 0; j < DEC_HIDDEN / 1024; j++) {
                    const float4 v = __ldcg(reinterpret_cast<const float4 *>(P.h + j * 1024 + col0));
                    xr[4 * j] = v.x; xr[4 * j + 1] = v.y; xr[4 * j + 2] = v.z; xr[4 * j + 3] = v.w;
                }
                consume<DEC_HIDDEN, false>(sm, cp, pbuf, DEC_DIM, 1, c, G, xr, P.err,
                    [&](int row, float v0, float) { P.x[row] = __ldcg(P.x + row) + v0; });
            }
            grid_sync(P.bar, bar_target += G, P.err);
        }
        // ---------------- final norm -> logits -> argmax (voxtral_decoder.c:765-779) ----------------
        {
            float xr[DEC_DIM / 256];
            float ss = 0.f;
            #pragma unroll
            for (int j = 0; j < DEC_DIM / 1024; j++) {
                const float4 v = __ldcg(reinterpret_cast<const float4 *>(P.x + j * 1024 + col0));
                xr[4 * j] = v.x; xr[4 * j + 1] = v.y; xr[4 * j + 2] = v.z; xr[4 * j + 3] = v.w;
                ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
            }
            const float tot = block_sum(sm, ss);
            const float rinv = 1.0f / sqrtf(tot / (float)DEC_DIM + NORM_EPS);
            #pragma unroll
            for (int j = 0; j < DEC_DIM / 1024; j++) {
                const float4 nw = *reinterpret_cast<const float4 *>(P.dec_norm + j * 1024 + col0);
                xr[4 * j] = xr[4 * j] * rinv * nw.x; xr[4 * j + 1] = xr[4 * j + 1] * rinv * nw.y;
                xr[4 * j + 2] = xr[4 * j + 2] * rinv * nw.z; xr[4 * j + 3] = xr[4 * j + 3] * rinv * nw.w;
            }
            float bv = -INFINITY;
            int bi = 0x7fffffff;
            consume<DEC_DIM, false>(sm, cp, pbuf, VOCAB, 1, c, G, xr, P.err,
                [&](int row, float v0, float) { P.logits[row] = v0; argmax_merge(bv, bi, v0, row); });
            if (tid < MAXTR) { sm.bv[tid] = bv; sm.bi[tid] = bi; }
            consumer_sync();
            if (tid == 0) {
                #pragma unroll
                for (int k = 1; k < MAXTR; k++) argmax_merge(bv, bi, sm.bv[k], sm.bi[k]);
                P.blockmax[2 * c] = bv;
                P.blockmax[2 * c + 1] = __int_as_float(bi);
            }
        }
        grid_sync(P.bar, bar_target += G, P.err);
        // every CTA derives the token; CTA 0 records it
        {
            float bv = -INFINITY;
            int bi = 0x7fffffff;
            for (int i = tid; i < G; i += CT)
                argmax_merge(bv, bi, __ldcg(P.blockmax + 2 * i), __float_as_int(__ldcg(P.blockmax + 2 * i + 1)));
            #pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                argmax_merge(bv, bi, ov, oi);
            }
            consumer_sync();               // sm.bv / sm.bi readers above are done
            if (lane == 0) { sm.bv[w] = bv; sm.bi[w] = bi; }
            consumer_sync();
            if (tid == 0) {
                for (int k = 1; k < CW; k++) argmax_merge(bv, bi, sm.bv[k], sm.bi[k]);
                if (bi == 0x7fffffff) bi = 0;   // all-NaN logits: the reference returns index 0
                sm.token = bi;
            }
            consumer_sync();
            prev_token = sm.token;
            if (c == 0 && tid == 0) P.tokens[step_count & (TOKENS_RING - 1)] = prev_token;
        }
        step_count += 1;
        pos += 1;
        n_valid = min(n_valid + 1, DEC_WINDOW);
        adapter_index += 1;
    }
    if (c == 0 && tid == 0) {
        DecStepParams *p = P.params;
        p->out_token = prev_token; p->prev_token = prev_token; p->step_count = step_count;
        p->pos = pos; p->n_valid = n_valid; p->adapter_index = adapter_index;
    }
}

}  // namespace mk

// ---- host ---------------------------------------------------------------------------------------
struct MegaHost {
    DecLayerDev *d_layers;
    unsigned *d_bar, *d_err;
    unsigned bar_count;      // value of *d_bar after all launches enqueued so far
    int grid;
    int barriers_per_step;
};

int dec_mega_init(vox_cuda_state *s) {
    MegaHost *h = (MegaHost *)calloc(1, sizeof(MegaHost));
    s->mega = h;
    int dev = 0, sms = 0, occ = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    CUDA_TRY(cudaFuncSetAttribute(mk::dec_mega_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)sizeof(mk::Smem)));
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, mk::dec_mega_kernel, mk::NT, sizeof(mk::Smem)));
    if (occ < 1 || sms < 8) { vox_set_error("dec_mega: kernel does not fit (occupancy %d, %d SMs)", occ, sms); return -1; }
    h->grid = sms;
    h->barriers_per_step = 5 * s->m->n_dec + 1;
    CUDA_TRY(cudaMalloc((void **)&h->d_layers, sizeof(DecLayerDev) * VOX_CUDA_DEC_LAYERS));
    CUDA_TRY(cudaMemcpy(h->d_layers, s->m->dec, sizeof(DecLayerDev) * VOX_CUDA_DEC_LAYERS, cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMalloc((void **)&h->d_bar, 8));
    CUDA_TRY(cudaMemset(h->d_bar, 0, 8));
    h->d_err = h->d_bar + 1;
    h->bar_count = 0;
    return 0;
}

void dec_mega_free(vox_cuda_state *s) {
    MegaHost *h = (MegaHost *)s->mega;
    if (!h) return;
    cudaFree(h->d_layers);
    cudaFree(h->d_bar);
    free(h);
    s->mega = nullptr;
}

// Enqueue n decoder steps as one cooperative launch.  The step scalars must
// already be in s->d_params (push_params).
int dec_mega_launch(vox_cuda_state *s, int n_steps) {
    MegaHost *h = (MegaHost *)s->mega;
    vox_cuda_model *m = s->m;
    mk::Params P;
    P.layers = h->d_layers; P.n_dec = m->n_dec;
    P.tok_emb = m->tok_emb; P.dec_norm = m->dec_norm; P.ada_scale = m->ada_scale; P.inv_freq = m->inv_freq_dec;
    P.kv_k = s->kv_k; P.kv_v = s->kv_v;
    P.x = s->d_x; P.q = s->d_q; P.attn_part = s->d_attn_part; P.h = s->d_h; P.logits = s->d_logits;
    P.blockmax = s->d_blockmax;
    P.embed_in = s->d_embed_in; P.adapter = s->adapter_buf;
    P.params = s->d_params; P.tokens = s->d_tokens;
    P.bar = h->d_bar; P.bar_base = h->bar_count; P.n_steps = n_steps; P.err = h->d_err;
    void *args[] = {&P};
    CUDA_TRY(cudaLaunchCooperativeKernel((const void *)mk::dec_mega_kernel, dim3(h->grid), dim3(mk::NT), args,
                                         sizeof(mk::Smem), s->stream));
    h->bar_count += (unsigned)h->grid * (unsigned)h->barriers_per_step * (unsigned)n_steps;
    LAUNCHED();
    return 0;
}

int dec_mega_attn_part_floats(void) { return mk::MAX_SPLITS * DEC_HEADS * mk::PART_LD; }
