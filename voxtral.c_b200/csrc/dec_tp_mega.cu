// dec_tp_mega.cu - the tensor-parallel decoder step as ONE persistent cooperative kernel per rank.
//
// Same arithmetic, shards and exchange windows as the kernel-per-op step in dec_tp.cu (vox_decoder_forward,
// /root/reference/voxtral_decoder.c:640-780, cut as SURVEY.md 8(e) lays out); what changes is the control
// flow.  148 CTAs (one per SM) run n_steps whole tokens.  Every CTA carries the residual row x[3072] in its
// own shared memory (all CTAs of all ranks compute it redundantly and bit-identically), so x never makes a
// round trip through HBM.  Per layer:
//
//   P1  RMSNorm(x) -> this rank's rows of wq|wk|wv, RoPE, K/V ring write            --- grid barrier B1
//   P2  attention tasks (local kv head, key split); last arriver per head merges    --- grid barrier B2
//   P3  wo over the rank's column range -> {value, seq} words into every rank's window (NVLink)
//   P4  poll the G x 3072 words of that exchange, x += sum (rank order); RMSNorm (1 + ada);
//       rows of w1|w3, SiLU * up -> h                                                --- grid barrier B3
//   P5  w2 over the rank's column range -> exchange words;  next layer's P1 polls them
//
// so a layer costs three LOCAL grid barriers; the two exchanges need no barrier at all, on this GPU or
// across GPUs - the self-flagged words are the synchronisation (a CTA proceeds when the words it needs
// have arrived).  The logits phase ends with a last-arriver reduction of the per-CTA maxima, one
// (max, index) word pair per rank, and every CTA derives the token from the G pairs: no host round trip
// and no barrier between tokens.  A warp issues the weight loads of the NEXT phase's first row batch
// before it waits at a barrier or polls an exchange, so HBM keeps streaming through every wait.
//
// Why slot reuse is safe here: a rank writes exchange s+2 into a peer's slot (same parity as s) only after
// it consumed that peer's s+1, which the peer produced after a local barrier that all of its CTAs reach
// only after consuming s.
#include "dec_tp_dev.cuh"
#include <stdlib.h>

namespace tpm {

template <int G> struct Cfg;
template <> struct Cfg<1> { static constexpr int WO_NW = 2, WO_NCH = 16, W2_NW = 1, W2_NCH = 36, WMAX = 36; };
template <> struct Cfg<2> { static constexpr int WO_NW = 4, WO_NCH = 8,  W2_NW = 2, W2_NCH = 18, WMAX = 36; };
template <> struct Cfg<4> { static constexpr int WO_NW = 8, WO_NCH = 4,  W2_NW = 4, W2_NCH = 9,  WMAX = 36; };
template <> struct Cfg<8> { static constexpr int WO_NW = 8, WO_NCH = 2,  W2_NW = 8, W2_NCH = 5,  WMAX = 40; };
constexpr int F_NW = 2, F_NCH = DEC_DIM / 256;            // K = 3072 products: 2 rows x 12 chunks per warp batch

struct Args {
    TpDev td;
    DecStepParams *p;
    int n_steps, n_layers, att_splits;
    const DecLayerDev *layers;        // device copy of the model's layer table
    const float *ada_scale, *dec_norm, *adapter, *inv_freq;
    const uint16_t *tok_emb;
    float *kv_k, *kv_v;               // [layers][8192][1024] ring
    float *q, *attn, *h, *logits, *attn_part, *blockmax;
    int *tokens;
    unsigned long long *bar;          // monotonic arrival counter of the local grid barrier
    unsigned long long bar_base;      // its value when this launch starts (host-tracked)
};

struct Smem {
    float4 xrA[DEC_DIM / 8], xrB[DEC_DIM / 8];     // residual row x (fp32), A/B halves of 8-float chunks
    float4 vA[DEC_HIDDEN / 8], vB[DEC_HIDDEN / 8]; // the GEMV's input vector (normalised x, attention slice or h slice)
    float rope[DEC_HD];
    float red[TP_WARPS];
    float bvs[TP_WARPS]; int bis[TP_WARPS];
    float vals[TP_MAX]; int idxs[TP_MAX];
    int token, last;
    TpAttnSmem at;
};

// flat register file of a row batch: w[r * NCH + j].  Always writes every entry (zeros when the warp has no batch:
// `active` false): a conditional call would leave the old contents "maybe live" for the register allocator across
// the whole layer loop, and 36-40 x 16 B per thread do not fit beside anything else.
template <int NW, int NCH, int WMAX>
__device__ __forceinline__ void wload(uint4 (&w)[WMAX], const uint16_t *__restrict__ W, size_t stride, int ks, bool active) {
    const int lane = threadIdx.x & 31;
    if (!active) ks = 0;
    #pragma unroll
    for (int j = 0; j < NCH; j++) {
        const int col = j * 256 + lane * 8;
        #pragma unroll
        for (int r = 0; r < NW; r++)
            w[r * NCH + j] = col < ks ? ldg_stream(W + (size_t)r * stride + col) : make_uint4(0u, 0u, 0u, 0u);
    }
}
template <int NW, int NCH, int WMAX>
__device__ __forceinline__ void wdot(const uint4 (&w)[WMAX], const float4 *xA, const float4 *xB, int ks, float (&acc)[NW]) {
    const int lane = threadIdx.x & 31;
    #pragma unroll
    for (int r = 0; r < NW; r++) acc[r] = 0.f;
    #pragma unroll
    for (int j = 0; j < NCH; j++) {
        const int col = j * 256 + lane * 8;
        if (col < ks) {
            const float4 a = xA[col >> 3], b = xB[col >> 3];
            #pragma unroll
            for (int r = 0; r < NW; r++) acc[r] = dot8(w[r * NCH + j], a, b, acc[r]);
        }
    }
    #pragma unroll
    for (int r = 0; r < NW; r++) acc[r] = warp_sum(acc[r]);
}

__device__ __forceinline__ unsigned long long ld_acquire_gpu_u64(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// local grid barrier: everything the CTA wrote before is visible to every CTA after
__device__ __forceinline__ void grid_barrier(const Args &a, unsigned long long &target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        target += gridDim.x;
        asm volatile("red.release.gpu.global.add.u64 [%0], %1;" :: "l"(a.bar), "l"(1ull) : "memory");
        if (ld_acquire_gpu_u64(a.bar) < target) {
            const unsigned long long t0 = globaltimer_ns();
            while (ld_acquire_gpu_u64(a.bar) < target)
                if (globaltimer_ns() - t0 > 8000000000ull) __trap();      // a CTA never arrived (bounded: no hung GPU)
        }
    }
    __syncthreads();
}

// 4 consecutive values of one source rank: v holds a first read; spins until all 4 words carry seq
__device__ __forceinline__ void ll_settle4(const TpDev &td, const uint2 *src, int p, unsigned seq, uint4 (&v)[2]) {
    if (v[0].y == seq && v[0].w == seq && v[1].y == seq && v[1].w == seq) return;
    const unsigned long long t0 = globaltimer_ns();
    for (;;) {
        v[0] = ld_ll2(src); v[1] = ld_ll2(src + 2);
        if (v[0].y == seq && v[0].w == seq && v[1].y == seq && v[1].w == seq) return;
        if (globaltimer_ns() - t0 > 8000000000ull) tp_timeout(td.rank, p, seq);
    }
}

// x (shared, kept) += the G partials of exchange a_idx (rank order); RMSNorm * w (* (1 + ada)) -> vA / vB.
// A thread owns three 4-float pieces of x (768 pieces over 256 threads) and reads the exchange words of all
// its pieces from up to 4 source ranks in ONE round of loads (24 x 16 B in flight per thread, which is what
// fits beside a prefetched row batch of 96 registers); 8 ranks take two rounds, still in rank order.
template <bool REDUCE, int G>
__device__ __forceinline__ void x_reduce_norm(Smem &sm, const TpDev &td, int a_idx, unsigned seq,
                                              const float *__restrict__ nw, const float *__restrict__ ada) {
    constexpr int NP = DEC_DIM / 4 / TP_THREADS;       // 3 pieces per thread
    constexpr int PG = G < 4 ? G : 4;                  // source ranks per round
    const int t = threadIdx.x;
    float4 x[NP];
    #pragma unroll
    for (int i = 0; i < NP; i++) {
        const int hc = t + i * TP_THREADS;
        x[i] = (hc & 1) ? sm.xrB[hc >> 1] : sm.xrA[hc >> 1];
    }
    if (REDUCE) {
        #pragma unroll
        for (int p0 = 0; p0 < G; p0 += PG) {
            uint4 v[NP][PG][2];
            #pragma unroll
            for (int i = 0; i < NP; i++)
                #pragma unroll
                for (int pp = 0; pp < PG; pp++) {
                    const uint2 *src = td.self->part[a_idx & 1][p0 + pp] + (t + i * TP_THREADS) * 4;
                    v[i][pp][0] = ld_ll2(src); v[i][pp][1] = ld_ll2(src + 2);
                }
            #pragma unroll
            for (int i = 0; i < NP; i++)
                #pragma unroll
                for (int pp = 0; pp < PG; pp++) {
                    ll_settle4(td, td.self->part[a_idx & 1][p0 + pp] + (t + i * TP_THREADS) * 4, p0 + pp, seq, v[i][pp]);
                    x[i].x += __uint_as_float(v[i][pp][0].x); x[i].y += __uint_as_float(v[i][pp][0].z);
                    x[i].z += __uint_as_float(v[i][pp][1].x); x[i].w += __uint_as_float(v[i][pp][1].z);
                }
        }
        #pragma unroll
        for (int i = 0; i < NP; i++) {
            const int hc = t + i * TP_THREADS;
            if (hc & 1) sm.xrB[hc >> 1] = x[i]; else sm.xrA[hc >> 1] = x[i];
        }
    }
    // norm weights after the poll (registers); their L2 latency hides behind the two block syncs below
    float4 wn[NP], wd[NP];
    #pragma unroll
    for (int i = 0; i < NP; i++) {
        wn[i] = *reinterpret_cast<const float4 *>(nw + (t + i * TP_THREADS) * 4);
        if (ada) wd[i] = *reinterpret_cast<const float4 *>(ada + (t + i * TP_THREADS) * 4);
    }
    float ss = 0.f;
    #pragma unroll
    for (int i = 0; i < NP; i++) ss += f4sq(x[i]);
    ss = warp_sum(ss);
    __syncthreads();                                   // sm.red and vA / vB of the previous phase are no longer read
    if ((t & 31) == 0) sm.red[t >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
    #pragma unroll
    for (int w = 0; w < TP_WARPS; w++) tot += sm.red[w];
    const float rinv = 1.0f / sqrtf(tot / (float)DEC_DIM + NORM_EPS);
    #pragma unroll
    for (int i = 0; i < NP; i++) {
        const int hc = t + i * TP_THREADS;
        float4 u = x[i];
        u.x = u.x * rinv * wn[i].x; u.y = u.y * rinv * wn[i].y; u.z = u.z * rinv * wn[i].z; u.w = u.w * rinv * wn[i].w;
        if (ada) { u.x *= (1.0f + wd[i].x); u.y *= (1.0f + wd[i].y); u.z *= (1.0f + wd[i].z); u.w *= (1.0f + wd[i].w); }
        if (hc & 1) sm.vB[hc >> 1] = u; else sm.vA[hc >> 1] = u;
    }
    __syncthreads();
}

// a slice of a vector another CTA wrote (attention output, h) -> vA / vB, through L2
__device__ __forceinline__ void load_slice(Smem &sm, const float *__restrict__ v, int ks) {
    __syncthreads();                                   // the previous phase's readers of vA / vB are done
    for (int c = threadIdx.x; c < ks / 8; c += TP_THREADS) {
        sm.vA[c] = __ldcg(reinterpret_cast<const float4 *>(v + c * 8));
        sm.vB[c] = __ldcg(reinterpret_cast<const float4 *>(v + c * 8 + 4));
    }
    __syncthreads();
}

template <int G>
__global__ void __launch_bounds__(TP_THREADS, 1) tp_mega_kernel(const Args a) {
    using C = Cfg<G>;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem &sm = *reinterpret_cast<Smem *>(smem_raw);
    const TpDev &td = a.td;
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int gw = blockIdx.x * TP_WARPS + warp, nwarps = gridDim.x * TP_WARPS;
    const int L = a.n_layers;
    const size_t kv_layer = (size_t)DEC_WINDOW * DEC_KVDIM;
    constexpr int NQ = DEC_QDIM / G, NK = DEC_KVDIM / G;          // this rank's q rows / k (= v) rows
    constexpr int QKV_BATCHES = (NQ + 2 * NK) / F_NW;
    constexpr int FFN_ROWS = 2 * DEC_HIDDEN / G, FFN_BATCHES = FFN_ROWS / F_NW;
    constexpr int VOC_ROWS = VOCAB / G, VOC_BATCHES = VOC_ROWS / F_NW;
    constexpr int WO_KS = DEC_QDIM / G, W2_KS = DEC_HIDDEN / G;
    constexpr int WO_BATCHES = DEC_DIM / C::WO_NW, W2_BATCHES = DEC_DIM / C::W2_NW;
    auto qkv_row = [&](int b) {
        const int lrow = b * F_NW;
        if (lrow < NQ) return td.rank * NQ + lrow;
        if (lrow < NQ + NK) return DEC_QDIM + td.rank * NK + (lrow - NQ);
        return DEC_QDIM + DEC_KVDIM + td.rank * NK + (lrow - NQ - NK);
    };
    // step scalars: every CTA tracks them; CTA 0 writes them back when the launch ends
    int pos = a.p->pos, n_valid = a.p->n_valid, prev = a.p->prev_token, aidx = a.p->adapter_index, step_count = a.p->step_count;
    unsigned epoch = *td.epoch;
    unsigned long long bar_target = a.bar_base;
    uint4 w[C::WMAX];
    uint2 *ll_dst[2];                                  // lane p < G: rank p's two slots for this rank's sums
    ll_dst[0] = lane < G ? td.peer[lane]->part[0][td.rank] : nullptr;
    ll_dst[1] = lane < G ? td.peer[lane]->part[1][td.rank] : nullptr;

    wload<F_NW, F_NCH, C::WMAX>(w, a.layers[0].wqkv + (size_t)qkv_row(gw) * DEC_DIM, DEC_DIM, DEC_DIM, gw < QKV_BATCHES);
    for (int step = 0; step < a.n_steps; step++) {
        const unsigned seq0 = epoch * (unsigned)td.ars;           // exchange i of this step carries seq0 + i + 1
        // ---- x = adapter[aidx] + tok_emb[prev] (voxtral.c:1106-1113); RoPE table (voxtral_kernels.c:617-629)
        for (int c = t; c < DEC_DIM / 8; c += TP_THREADS) {
            const float *ad = a.adapter + (size_t)aidx * DEC_DIM + c * 8;
            const uint4 e = *reinterpret_cast<const uint4 *>(a.tok_emb + (size_t)prev * DEC_DIM + c * 8);
            const float4 a0 = *reinterpret_cast<const float4 *>(ad), a1 = *reinterpret_cast<const float4 *>(ad + 4);
            sm.xrA[c] = make_float4(a0.x + bf16_lo(e.x), a0.y + bf16_hi(e.x), a0.z + bf16_lo(e.y), a0.w + bf16_hi(e.y));
            sm.xrB[c] = make_float4(a1.x + bf16_lo(e.z), a1.y + bf16_hi(e.z), a1.z + bf16_lo(e.w), a1.w + bf16_hi(e.w));
        }
        if (t < DEC_HD / 2) {
            const float angle = (float)pos * a.inv_freq[t];
            sm.rope[2 * t] = cosf(angle);
            sm.rope[2 * t + 1] = sinf(angle);
        }
        __syncthreads();
        const int slot = pos & (DEC_WINDOW - 1);
        for (int l = 0; l < L; l++) {
            const DecLayerDev &Wl = a.layers[l];
            float *kc = a.kv_k + l * kv_layer, *vc = a.kv_v + l * kv_layer;
            // ---- P1: qkv (voxtral_decoder.c:710-725) ----
            if (l == 0) x_reduce_norm<false, G>(sm, td, 0, 0u, Wl.attn_norm, nullptr);
            else        x_reduce_norm<true, G>(sm, td, 2 * l - 1, seq0 + 2 * l, Wl.attn_norm, nullptr);
            for (int b = gw; b < QKV_BATCHES;) {
                float acc[F_NW];
                wdot<F_NW, F_NCH, C::WMAX>(w, sm.vA, sm.vB, DEC_DIM, acc);
                const int row = qkv_row(b);
                b += nwarps;
                wload<F_NW, F_NCH, C::WMAX>(w, Wl.wqkv + (size_t)qkv_row(b) * DEC_DIM, DEC_DIM, DEC_DIM, b < QKV_BATCHES);
                if (lane == 0) {
                    float v0 = acc[0], v1 = acc[1];
                    if (row < DEC_QDIM + DEC_KVDIM) {     // interleaved-pair rotation, voxtral_kernels.c:631-655
                        const int d = (row & (DEC_HD - 1)) >> 1;
                        const float c = sm.rope[2 * d], s = sm.rope[2 * d + 1];
                        const float r0 = v0 * c - v1 * s, r1 = v0 * s + v1 * c;
                        v0 = r0; v1 = r1;
                    }
                    float *dst;
                    if (row < DEC_QDIM) dst = a.q + row;
                    else if (row < DEC_QDIM + DEC_KVDIM) dst = kc + (size_t)slot * DEC_KVDIM + (row - DEC_QDIM);
                    else dst = vc + (size_t)slot * DEC_KVDIM + (row - DEC_QDIM - DEC_KVDIM);
                    *reinterpret_cast<float2 *>(dst) = make_float2(v0, v1);
                }
            }
            grid_barrier(a, bar_target);                                                  // B1: q, K, V rows are out
            // ---- P2: attention (voxtral_decoder.c:731-733) ----
            for (int task = blockIdx.x; task < (DEC_KVH / G) * a.att_splits; task += gridDim.x)
                tp_attn_task(td, task / a.att_splits, task % a.att_splits, a.att_splits, pos, n_valid, a.q, kc, vc,
                             a.attn_part, a.attn, sm.at);
            wload<C::WO_NW, C::WO_NCH, C::WMAX>(w, Wl.wo + (size_t)gw * C::WO_NW * DEC_QDIM + (size_t)td.rank * WO_KS, DEC_QDIM, WO_KS, gw < WO_BATCHES);
            grid_barrier(a, bar_target);                                                  // B2: attention output is out
            // ---- P3: wo partial sums -> exchange 2l (voxtral_decoder.c:738-739) ----
            load_slice(sm, a.attn + (size_t)td.rank * WO_KS, WO_KS);
            for (int b = gw; b < WO_BATCHES;) {
                float acc[C::WO_NW];
                wdot<C::WO_NW, C::WO_NCH, C::WMAX>(w, sm.vA, sm.vB, WO_KS, acc);
                const int row0 = b * C::WO_NW;
                b += nwarps;
                wload<C::WO_NW, C::WO_NCH, C::WMAX>(w, Wl.wo + (size_t)b * C::WO_NW * DEC_QDIM + (size_t)td.rank * WO_KS, DEC_QDIM, WO_KS, b < WO_BATCHES);
                if (lane < G) {
                    #pragma unroll
                    for (int r = 0; r < C::WO_NW; r++) st_ll(ll_dst[0] + row0 + r, __float_as_uint(acc[r]), seq0 + 2 * l + 1);
                }
            }
            // ---- P4: FFN gate / up (voxtral_decoder.c:741-757) ----
            wload<F_NW, F_NCH, C::WMAX>(w, Wl.w13 + (size_t)(td.rank * FFN_ROWS + gw * F_NW) * DEC_DIM, DEC_DIM, DEC_DIM, gw < FFN_BATCHES);
            x_reduce_norm<true, G>(sm, td, 2 * l, seq0 + 2 * l + 1, Wl.ffn_norm, a.ada_scale + (size_t)l * DEC_DIM);
            for (int b = gw; b < FFN_BATCHES;) {
                float acc[F_NW];
                wdot<F_NW, F_NCH, C::WMAX>(w, sm.vA, sm.vB, DEC_DIM, acc);
                const int row = td.rank * FFN_ROWS + b * F_NW;
                b += nwarps;
                wload<F_NW, F_NCH, C::WMAX>(w, Wl.w13 + (size_t)(td.rank * FFN_ROWS + b * F_NW) * DEC_DIM, DEC_DIM, DEC_DIM, b < FFN_BATCHES);
                if (lane == 0) a.h[row >> 1] = silu(acc[0]) * acc[1];
            }
            wload<C::W2_NW, C::W2_NCH, C::WMAX>(w, Wl.w2 + (size_t)gw * C::W2_NW * DEC_HIDDEN + (size_t)td.rank * W2_KS, DEC_HIDDEN, W2_KS, gw < W2_BATCHES);
            grid_barrier(a, bar_target);                                                  // B3: h is out
            // ---- P5: w2 partial sums -> exchange 2l + 1 (voxtral_decoder.c:758-760) ----
            load_slice(sm, a.h + (size_t)td.rank * W2_KS, W2_KS);
            for (int b = gw; b < W2_BATCHES;) {
                float acc[C::W2_NW];
                wdot<C::W2_NW, C::W2_NCH, C::WMAX>(w, sm.vA, sm.vB, W2_KS, acc);
                const int row0 = b * C::W2_NW;
                b += nwarps;
                wload<C::W2_NW, C::W2_NCH, C::WMAX>(w, Wl.w2 + (size_t)b * C::W2_NW * DEC_HIDDEN + (size_t)td.rank * W2_KS, DEC_HIDDEN, W2_KS, b < W2_BATCHES);
                if (lane < G) {
                    #pragma unroll
                    for (int r = 0; r < C::W2_NW; r++) st_ll(ll_dst[1] + row0 + r, __float_as_uint(acc[r]), seq0 + 2 * l + 2);
                }
            }
            // prefetch for the next consumer of x: next layer's qkv rows, or the first vocabulary rows
            {
                const bool more = l + 1 < L;
                const uint16_t *nxt = more ? a.layers[more ? l + 1 : l].wqkv + (size_t)qkv_row(gw < QKV_BATCHES ? gw : 0) * DEC_DIM
                                           : a.tok_emb + (size_t)(td.rank * VOC_ROWS + (gw < VOC_BATCHES ? gw : 0) * F_NW) * DEC_DIM;
                wload<F_NW, F_NCH, C::WMAX>(w, nxt, DEC_DIM, DEC_DIM, more ? gw < QKV_BATCHES : gw < VOC_BATCHES);
            }
        }
        // ---- logits of this rank's vocabulary rows + first maximum (voxtral_decoder.c:765-779) ----
        x_reduce_norm<true, G>(sm, td, 2 * L - 1, seq0 + 2 * L, a.dec_norm, nullptr);
        float bv = -INFINITY; int bi = 0x7fffffff;
        for (int b = gw; b < VOC_BATCHES;) {
            float acc[F_NW];
            wdot<F_NW, F_NCH, C::WMAX>(w, sm.vA, sm.vB, DEC_DIM, acc);
            const int row = td.rank * VOC_ROWS + b * F_NW;
            b += nwarps;
            wload<F_NW, F_NCH, C::WMAX>(w, a.tok_emb + (size_t)(td.rank * VOC_ROWS + b * F_NW) * DEC_DIM, DEC_DIM, DEC_DIM, b < VOC_BATCHES);
            if (lane == 0) {
                *reinterpret_cast<float2 *>(a.logits + row) = make_float2(acc[0], acc[1]);
                tp_argmax_merge(bv, bi, acc[0], row);
                tp_argmax_merge(bv, bi, acc[1], row + 1);
            }
        }
        // next token's first qkv batch streams in while the maxima are merged and exchanged
        wload<F_NW, F_NCH, C::WMAX>(w, a.layers[0].wqkv + (size_t)qkv_row(gw) * DEC_DIM, DEC_DIM, DEC_DIM, step + 1 < a.n_steps && gw < QKV_BATCHES);
        if (lane == 0) { sm.bvs[warp] = bv; sm.bis[warp] = bi; }
        __syncthreads();
        const unsigned seq_amax = seq0 + (unsigned)td.ars;
        if (t == 0) {
            #pragma unroll
            for (int k = 1; k < TP_WARPS; k++) tp_argmax_merge(bv, bi, sm.bvs[k], sm.bis[k]);
            a.blockmax[2 * blockIdx.x] = bv;
            a.blockmax[2 * blockIdx.x + 1] = __int_as_float(bi);
            __threadfence();
            sm.last = (atomicAdd(td.done, 1u) == gridDim.x - 1);
        }
        __syncthreads();
        if (sm.last && warp == 0) {                     // last CTA to arrive: this rank's maximum -> every rank
            if (lane == 0) { *td.done = 0; __threadfence(); }
            float mv = -INFINITY; int mi = 0x7fffffff;
            for (int i = lane; i < (int)gridDim.x; i += 32)
                tp_argmax_merge(mv, mi, __ldcg(a.blockmax + 2 * i), __float_as_int(__ldcg(a.blockmax + 2 * i + 1)));
            #pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, mv, o);
                const int oi = __shfl_xor_sync(0xffffffffu, mi, o);
                tp_argmax_merge(mv, mi, ov, oi);
            }
            if (lane < G) {
                uint2 *dst = td.peer[lane]->amax[seq_amax & 1][td.rank];
                st_ll(dst, __float_as_uint(mv), seq_amax);
                st_ll(dst + 1, (unsigned)mi, seq_amax);
            }
        }
        if (t < G) {                                    // every CTA: the G (max, index) pairs -> the token
            const uint2 *src = td.self->amax[seq_amax & 1][t];
            uint4 v = ld_ll2(src);
            if (v.y != seq_amax || v.w != seq_amax) {
                const unsigned long long t0 = globaltimer_ns();
                do {
                    v = ld_ll2(src);
                    if (globaltimer_ns() - t0 > 8000000000ull) tp_timeout(td.rank, t, seq_amax);
                } while (v.y != seq_amax || v.w != seq_amax);
            }
            sm.vals[t] = __uint_as_float(v.x);
            sm.idxs[t] = (int)v.z;
        }
        __syncthreads();
        if (t == 0) {
            float fv = -INFINITY; int fi = 0x7fffffff;
            for (int r = 0; r < G; r++) tp_argmax_merge(fv, fi, sm.vals[r], sm.idxs[r]);
            if (fi == 0x7fffffff) fi = 0;               // all-NaN logits: the reference returns index 0
            sm.token = fi;
            if (blockIdx.x == 0) a.tokens[step_count & (TOKENS_RING - 1)] = fi;
        }
        __syncthreads();
        prev = sm.token;
        pos += 1; n_valid = min(n_valid + 1, DEC_WINDOW); aidx += 1; step_count += 1; epoch += 1;
    }
    if (blockIdx.x == 0 && t == 0) {
        a.p->pos = pos; a.p->n_valid = n_valid; a.p->prev_token = prev; a.p->out_token = prev;
        a.p->adapter_index = aidx; a.p->step_count = step_count;
        *td.epoch = epoch;
    }
}

}  // namespace tpm

// ---- host ----------------------------------------------------------------------------------------------------
struct TpMegaHost {
    DecLayerDev *d_layers;
    unsigned long long *d_bar;
    unsigned long long bar_count;
    int grid;
    size_t smem;
};

template <int G>
static int tp_mega_prepare(TpMegaHost *h) {
    h->smem = sizeof(tpm::Smem);
    CUDA_TRY(cudaFuncSetAttribute(tpm::tp_mega_kernel<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem));
    int dev = 0, sms = 0, per_sm = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, tpm::tp_mega_kernel<G>, TP_THREADS, h->smem));
    if (per_sm < 1) { vox_set_error("tp mega: kernel does not fit an SM"); return -1; }
    h->grid = sms;
    return 0;
}

void *tp_mega_create(const vox_cuda_model *m, int world) {
    TpMegaHost *h = (TpMegaHost *)calloc(1, sizeof(TpMegaHost));
    if (!h) return nullptr;
    bool ok = cudaMalloc((void **)&h->d_layers, sizeof(DecLayerDev) * VOX_CUDA_DEC_LAYERS) == cudaSuccess &&
              cudaMalloc((void **)&h->d_bar, 64) == cudaSuccess &&
              cudaMemcpy(h->d_layers, m->dec, sizeof(DecLayerDev) * VOX_CUDA_DEC_LAYERS, cudaMemcpyHostToDevice) == cudaSuccess &&
              cudaMemset(h->d_bar, 0, 64) == cudaSuccess;
    if (ok) {
        int rc = world == 1 ? tp_mega_prepare<1>(h) : world == 2 ? tp_mega_prepare<2>(h) : world == 4 ? tp_mega_prepare<4>(h) : tp_mega_prepare<8>(h);
        ok = rc == 0;
    }
    if (!ok) { cudaFree(h->d_layers); cudaFree(h->d_bar); free(h); cudaGetLastError(); return nullptr; }
    return h;
}

void tp_mega_destroy(void *hp) {
    TpMegaHost *h = (TpMegaHost *)hp;
    if (!h) return;
    cudaFree(h->d_layers); cudaFree(h->d_bar);
    free(h);
}

// enqueue n steps on st; td / buffers as the kernel-per-op step uses them
int tp_mega_run(void *hp, cudaStream_t st, const TpDev &td, vox_cuda_model *m, vox_cuda_state *s, DecStepParams *d_params,
                float *attn_part, int att_splits, float *blockmax, int *d_tokens, int n) {
    TpMegaHost *h = (TpMegaHost *)hp;
    tpm::Args a;
    a.td = td; a.p = d_params; a.n_steps = n; a.n_layers = m->n_dec; a.att_splits = att_splits;
    a.layers = h->d_layers; a.ada_scale = m->ada_scale; a.dec_norm = m->dec_norm; a.adapter = s->adapter_buf;
    a.inv_freq = m->inv_freq_dec; a.tok_emb = m->tok_emb; a.kv_k = s->kv_k; a.kv_v = s->kv_v;
    a.q = s->d_q; a.attn = s->d_attn; a.h = s->d_h; a.logits = s->d_logits; a.attn_part = attn_part; a.blockmax = blockmax;
    a.tokens = d_tokens; a.bar = h->d_bar; a.bar_base = h->bar_count;
    void *args[] = {&a};
    const void *fn = td.world == 1 ? (const void *)tpm::tp_mega_kernel<1> : td.world == 2 ? (const void *)tpm::tp_mega_kernel<2>
                   : td.world == 4 ? (const void *)tpm::tp_mega_kernel<4> : (const void *)tpm::tp_mega_kernel<8>;
    CUDA_TRY(cudaLaunchCooperativeKernel(fn, dim3(h->grid), dim3(TP_THREADS), args, h->smem, st));
    h->bar_count += (unsigned long long)n * 3ull * (unsigned long long)m->n_dec * (unsigned long long)h->grid;
    LAUNCHED();
    return 0;
}
