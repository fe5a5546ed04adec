// dec_tp.cu - tensor-parallel single-stream decoder step (BASELINE.json north_star:
// "an optional tensor-parallel single-stream decoder ... allreduce over NVLink after wo and w2").
//
// Arithmetic is vox_decoder_forward (/root/reference/voxtral_decoder.c:640-780) with the matrices cut the
// way SURVEY 8(e) lays out: rank r of G owns kv heads [r*8/G, (r+1)*8/G) and their 4 query heads each
// (rows of wq|wk|wv), rows [r*18432/G, ...) of the interleaved w1|w3, the matching COLUMN range of wo
// and w2 (whose products are therefore partial sums of all 3072 outputs), and vocabulary rows
// [r*131072/G, ...) of the tied embedding.  Weights are replicated (8.86 GB of 180 GB); a shard is a view.
//
// The reduction after wo / w2 is not a separate collective: the GEMV epilogue stores its 16 partial rows
// straight into every peer's exchange window over NVLink (peer memory opened through CUDA IPC), the last
// CTA publishes a sequence number in each peer's flag word, and the NEXT kernel's prologue (the RMSNorm of
// the FFN / of the next layer / of the logits) waits for the G flags and adds the G partials in rank order
// - so every rank holds bit-identical residual rows and no kernel is launched for the exchange.
// The greedy token is exchanged the same way as (max, index) pairs of the vocabulary slices.
#include "vox_internal.h"
#include <cuda.h>
#include <stdlib.h>
#include <string.h>

constexpr int TP_MAX = VOX_CUDA_TP_MAX;   // 8
constexpr int TP_WARPS = 8;
constexpr int TP_THREADS = TP_WARPS * 32;
constexpr int TP_RES_RPW = 2;                              // rows per warp of the wo / w2 partial GEMV
constexpr int TP_RES_ROWS = TP_WARPS * TP_RES_RPW;         // 16 rows per CTA -> 192 CTAs

// One rank's exchange window.  Same layout on every rank; peers write into it, the owner only reads.
struct TpComm {
    float part[2][TP_MAX][DEC_DIM];   // [sequence parity][source rank][row]  partial sums of wo / w2
    float amax[2][TP_MAX][2];         // [step parity][source rank] (max logit, index bits) of a vocabulary slice
    unsigned flag[TP_MAX];            // flag[src] = sequence number of the last exchange src published here
    unsigned pad[24];
};

struct TpDev {                        // by-value kernel argument
    int rank, world;
    int ars;                          // exchanges per step = 2 * layers + 1
    TpComm *self;
    TpComm *peer[TP_MAX];             // peer[p] = rank p's window as mapped here (peer[rank] == self)
    unsigned *done;                   // local arrival counter of the producing kernel's CTAs
    unsigned *epoch;                  // steps completed since creation (device resident; graphs are replayed)
};

__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned *p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// Wait until every rank has published exchange `seq` in this rank's window.  Bounded: a peer that died
// must not leave this GPU spinning (the pool counts a hung box as a strike).
__device__ __forceinline__ void tp_wait_all(const TpDev &td, unsigned seq) {
    if ((int)threadIdx.x < td.world) {
        const unsigned *f = &td.self->flag[threadIdx.x];
        if ((int)(ld_acquire_sys(f) - seq) < 0) {
            const unsigned long long t0 = globaltimer_ns();
            while ((int)(ld_acquire_sys(f) - seq) < 0) {
                __nanosleep(64);
                if (globaltimer_ns() - t0 > 8000000000ull) {
                    printf("vox_cuda tp: rank %d waited 8 s for rank %d, exchange %u (flag %u)\n", td.rank,
                           (int)threadIdx.x, seq, *(volatile const unsigned *)f);
                    __trap();
                }
            }
        }
    }
    __syncthreads();
}

// Publish: every thread that stored into peer windows has fenced; the CTA that arrives last raises the flags.
__device__ __forceinline__ void tp_publish(const TpDev &td, unsigned seq, unsigned n_cta) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        const unsigned old = atomicAdd(td.done, 1u);
        if (old == n_cta - 1) {
            *td.done = 0;                       // every CTA of this kernel has arrived; the next producer is a later kernel
            __threadfence_system();
            for (int p = 0; p < td.world; p++) st_release_sys(&td.peer[p]->flag[td.rank], seq);
        }
    }
}

// y[r] = sum_k W[row r][k] * x[k] over a slice of ks columns (ks % 8 == 0); x is in shared memory as
// A[c] = x[8c..8c+3], B[c] = x[8c+4..8c+7].  W points at the slice start of the first row.
template <int NW>
__device__ __forceinline__ void tp_warp_gemv(const uint16_t *__restrict__ W, size_t stride, int ks,
                                             const float4 *xA, const float4 *xB, float (&acc)[NW]) {
    const int lane = threadIdx.x & 31;
    #pragma unroll
    for (int r = 0; r < NW; r++) acc[r] = 0.f;
    const uint16_t *base = W + lane * 8;
    #pragma unroll 4
    for (int col = lane * 8; col < ks; col += 256) {
        uint4 w[NW];
        #pragma unroll
        for (int r = 0; r < NW; r++) w[r] = ldg_stream(base + (size_t)r * stride);
        base += 256;
        const float4 a = xA[col >> 3], b = xB[col >> 3];
        #pragma unroll
        for (int r = 0; r < NW; r++) acc[r] = dot8(w[r], a, b, acc[r]);
    }
    #pragma unroll
    for (int r = 0; r < NW; r++) acc[r] = warp_sum(acc[r]);
}

__device__ __forceinline__ float4 f4add(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ float f4sq(float4 a) { return a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w; }

// Prologue of the three consuming kernels: x = x_in (+ the G partials of exchange a_idx, in rank order),
// CTA 0 keeps the new residual row in x_out, then RMSNorm (voxtral_kernels.c:475-492) and the (1 + ada)
// scale (voxtral_decoder.c:742-745) in shared memory.
template <bool REDUCE>
__device__ __forceinline__ void tp_load_x_norm(float4 *xA, float4 *xB, float *red, const TpDev &td, int a_idx,
                                               const float *__restrict__ x_in, float *__restrict__ x_out,
                                               const float *__restrict__ nw, const float *__restrict__ ada) {
    const int t = threadIdx.x;
    unsigned seq = 0;
    if (REDUCE) {
        seq = *td.epoch * (unsigned)td.ars + (unsigned)a_idx + 1u;
        tp_wait_all(td, seq);
    }
    float ss = 0.f;
    for (int c = t; c < DEC_DIM / 8; c += TP_THREADS) {
        float4 a = *reinterpret_cast<const float4 *>(x_in + c * 8);
        float4 b = *reinterpret_cast<const float4 *>(x_in + c * 8 + 4);
        if (REDUCE) {
            for (int p = 0; p < td.world; p++) {
                const float *src = td.self->part[seq & 1][p] + c * 8;
                a = f4add(a, __ldcg(reinterpret_cast<const float4 *>(src)));
                b = f4add(b, __ldcg(reinterpret_cast<const float4 *>(src + 4)));
            }
            if (blockIdx.x == 0 && x_out) {
                *reinterpret_cast<float4 *>(x_out + c * 8) = a;
                *reinterpret_cast<float4 *>(x_out + c * 8 + 4) = b;
            }
        }
        xA[c] = a; xB[c] = b;
        ss += f4sq(a) + f4sq(b);
    }
    ss = warp_sum(ss);
    if ((t & 31) == 0) red[t >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
    #pragma unroll
    for (int w = 0; w < TP_WARPS; w++) tot += red[w];
    const float rinv = 1.0f / sqrtf(tot / (float)DEC_DIM + NORM_EPS);
    for (int c = t; c < DEC_DIM / 8; c += TP_THREADS) {
        float4 a = xA[c], b = xB[c];
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
        xA[c] = a; xB[c] = b;
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------
// T1: this rank's rows of wq|wk|wv on rmsnorm(x); RoPE; q -> q_out, K/V -> ring slot
// (voxtral_decoder.c:710-725).  Buffers keep GLOBAL row / head indices; a rank touches its slice only.
// ---------------------------------------------------------------------------
template <int RPW, bool REDUCE>
__global__ void __launch_bounds__(TP_THREADS) tp_qkv_kernel(
    TpDev td, int a_idx, const DecStepParams *__restrict__ p, const float *__restrict__ x_in,
    float *__restrict__ x_out, const float *__restrict__ attn_norm, const uint16_t *__restrict__ wqkv,
    const float *__restrict__ rope, float *__restrict__ q_out, float *__restrict__ kc, float *__restrict__ vc) {
    __shared__ float4 xA[DEC_DIM / 8], xB[DEC_DIM / 8];
    __shared__ float red[TP_WARPS];
    tp_load_x_norm<REDUCE>(xA, xB, red, td, a_idx, x_in, x_out, attn_norm, nullptr);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nq = DEC_QDIM / td.world, nk = DEC_KVDIM / td.world;
    const int lrow0 = blockIdx.x * (TP_WARPS * RPW) + warp * RPW;
    int grow0;
    if (lrow0 < nq) grow0 = td.rank * nq + lrow0;
    else if (lrow0 < nq + nk) grow0 = DEC_QDIM + td.rank * nk + (lrow0 - nq);
    else grow0 = DEC_QDIM + DEC_KVDIM + td.rank * nk + (lrow0 - nq - nk);
    float acc[RPW];
    tp_warp_gemv<RPW>(wqkv + (size_t)grow0 * DEC_DIM, DEC_DIM, DEC_DIM, xA, xB, acc);
    if (lane == 0) {
        const int slot = p->pos & (DEC_WINDOW - 1);
        #pragma unroll
        for (int r = 0; r < RPW; r += 2) {
            const int row = grow0 + r;
            float v0 = acc[r], v1 = acc[r + 1];
            if (row < DEC_QDIM + DEC_KVDIM) {     // interleaved-pair rotation, voxtral_kernels.c:631-655
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
// T2 / T4: partial[row] = W[row][r*ks .. (r+1)*ks) . v[r*ks ..)  for all 3072 rows (wo: K = 4096,
// voxtral_decoder.c:738-739; w2: K = 9216, :758-760), stored into every rank's window, then published.
// The residual add happens in the consumer's prologue, so x is not touched here.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(TP_THREADS) tp_gemv_partial_kernel(
    TpDev td, int a_idx, const float *__restrict__ v, const uint16_t *__restrict__ W, int K) {
    extern __shared__ __align__(16) unsigned char tp_smem[];
    __shared__ __align__(16) float rows[TP_RES_ROWS];
    const int ks = K / td.world;
    float4 *xA = reinterpret_cast<float4 *>(tp_smem);
    float4 *xB = xA + ks / 8;
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const float *vs = v + (size_t)td.rank * ks;
    for (int c = t; c < ks / 8; c += TP_THREADS) {
        xA[c] = *reinterpret_cast<const float4 *>(vs + c * 8);
        xB[c] = *reinterpret_cast<const float4 *>(vs + c * 8 + 4);
    }
    __syncthreads();
    const int row0 = blockIdx.x * TP_RES_ROWS + warp * TP_RES_RPW;
    float acc[TP_RES_RPW];
    tp_warp_gemv<TP_RES_RPW>(W + (size_t)row0 * K + (size_t)td.rank * ks, (size_t)K, ks, xA, xB, acc);
    if (lane == 0) {
        #pragma unroll
        for (int r = 0; r < TP_RES_RPW; r++) rows[warp * TP_RES_RPW + r] = acc[r];
    }
    __syncthreads();
    const unsigned seq = *td.epoch * (unsigned)td.ars + (unsigned)a_idx + 1u;
    if (t < td.world * (TP_RES_ROWS / 4)) {          // 64 B of rows per peer, one float4 per thread
        const int peer = t / (TP_RES_ROWS / 4), q4 = t % (TP_RES_ROWS / 4);
        float *dst = td.peer[peer]->part[seq & 1][td.rank] + blockIdx.x * TP_RES_ROWS + q4 * 4;
        *reinterpret_cast<float4 *>(dst) = reinterpret_cast<const float4 *>(rows)[q4];
        __threadfence_system();
    }
    tp_publish(td, seq, gridDim.x);
}

// ---------------------------------------------------------------------------
// T3: h[i] = silu(w1[i] . xn) * (w3[i] . xn) for this rank's rows of the interleaved w1|w3, xn =
// rmsnorm(x + sum of the wo partials) * (1 + ada)   (voxtral_decoder.c:741-757)
// ---------------------------------------------------------------------------
template <int RPW>
__global__ void __launch_bounds__(TP_THREADS) tp_ffn_kernel(
    TpDev td, int a_idx, const float *__restrict__ x_in, float *__restrict__ x_out,
    const float *__restrict__ ffn_norm, const float *__restrict__ ada, const uint16_t *__restrict__ w13,
    float *__restrict__ h) {
    __shared__ float4 xA[DEC_DIM / 8], xB[DEC_DIM / 8];
    __shared__ float red[TP_WARPS];
    tp_load_x_norm<true>(xA, xB, red, td, a_idx, x_in, x_out, ffn_norm, ada);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row0 = td.rank * (2 * DEC_HIDDEN / td.world) + blockIdx.x * (TP_WARPS * RPW) + warp * RPW;
    float acc[RPW];
    tp_warp_gemv<RPW>(w13 + (size_t)row0 * DEC_DIM, DEC_DIM, DEC_DIM, xA, xB, acc);
    if (lane == 0) {
        #pragma unroll
        for (int r = 0; r < RPW; r += 2) h[(row0 + r) >> 1] = silu(acc[r]) * acc[r + 1];
    }
}

// ---------------------------------------------------------------------------
// T5: logits of this rank's vocabulary rows + per-CTA first maximum (voxtral_decoder.c:765-779)
// ---------------------------------------------------------------------------
__device__ __forceinline__ void tp_argmax_merge(float &bv, int &bi, float v, int i) {
    if (v > bv || (v == bv && i < bi)) { bv = v; bi = i; }   // first maximum; NaN never wins
}

__global__ void __launch_bounds__(TP_THREADS) tp_logits_kernel(
    TpDev td, int a_idx, const float *__restrict__ x_in, const float *__restrict__ norm,
    const uint16_t *__restrict__ tok_emb, float *__restrict__ logits, float *__restrict__ blockmax) {
    __shared__ float4 xA[DEC_DIM / 8], xB[DEC_DIM / 8];
    __shared__ float red[TP_WARPS];
    __shared__ float bvs[TP_WARPS];
    __shared__ int bis[TP_WARPS];
    tp_load_x_norm<true>(xA, xB, red, td, a_idx, x_in, nullptr, norm, nullptr);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float bv = -INFINITY; int bi = 0x7fffffff;
    const int base = td.rank * (VOCAB / td.world) + blockIdx.x * LOGITS_ROWS_PER_CTA;
    for (int it = 0; it < LOGITS_ROWS_PER_CTA / (TP_WARPS * 4); it++) {
        const int row0 = base + (it * TP_WARPS + warp) * 4;
        float acc[4];
        tp_warp_gemv<4>(tok_emb + (size_t)row0 * DEC_DIM, DEC_DIM, DEC_DIM, xA, xB, acc);
        if (lane == 0) {
            #pragma unroll
            for (int r = 0; r < 4; r++) { logits[row0 + r] = acc[r]; tp_argmax_merge(bv, bi, acc[r], row0 + r); }
        }
    }
    if (lane == 0) { bvs[warp] = bv; bis[warp] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        #pragma unroll
        for (int w = 1; w < TP_WARPS; w++) tp_argmax_merge(bv, bi, bvs[w], bis[w]);
        blockmax[2 * blockIdx.x] = bv;
        blockmax[2 * blockIdx.x + 1] = __int_as_float(bi);
    }
}

// T6: first maximum of this rank's slice -> every rank's window (exchange index 2 * layers)
__global__ void __launch_bounds__(1024) tp_argmax_pub_kernel(TpDev td, const float *__restrict__ blockmax, int nblk) {
    __shared__ float bvs[32];
    __shared__ int bis[32];
    const int t = threadIdx.x;
    float bv = -INFINITY; int bi = 0x7fffffff;
    for (int i = t; i < nblk; i += 1024) tp_argmax_merge(bv, bi, blockmax[2 * i], __float_as_int(blockmax[2 * i + 1]));
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        tp_argmax_merge(bv, bi, ov, oi);
    }
    if ((t & 31) == 0) { bvs[t >> 5] = bv; bis[t >> 5] = bi; }
    __syncthreads();
    if (t == 0) {
        for (int w = 1; w < 32; w++) tp_argmax_merge(bv, bi, bvs[w], bis[w]);
        const unsigned e = *td.epoch;
        const unsigned seq = e * (unsigned)td.ars + (unsigned)td.ars;     // a_idx = ars - 1
        for (int p = 0; p < td.world; p++) {
            float *dst = td.peer[p]->amax[e & 1][td.rank];
            *reinterpret_cast<float2 *>(dst) = make_float2(bv, __int_as_float(bi));
        }
        __threadfence_system();
        for (int p = 0; p < td.world; p++) st_release_sys(&td.peer[p]->flag[td.rank], seq);
    }
}

// T7: merge the G slice maxima (same order on every rank), write the token, advance the step scalars
__global__ void __launch_bounds__(32) tp_argmax_fin_kernel(TpDev td, DecStepParams *__restrict__ p,
                                                           int *__restrict__ tokens) {
    const unsigned e = *td.epoch;
    tp_wait_all(td, e * (unsigned)td.ars + (unsigned)td.ars);
    if (threadIdx.x == 0) {
        float bv = -INFINITY; int bi = 0x7fffffff;
        for (int r = 0; r < td.world; r++) {
            const float2 m = __ldcg(reinterpret_cast<const float2 *>(td.self->amax[e & 1][r]));
            tp_argmax_merge(bv, bi, m.x, __float_as_int(m.y));
        }
        if (bi == 0x7fffffff) bi = 0;    // all-NaN logits: the reference returns index 0
        p->out_token = bi;
        p->prev_token = bi;
        tokens[p->step_count & (TOKENS_RING - 1)] = bi;
        p->step_count += 1;
        p->pos += 1;
        p->n_valid = min(p->n_valid + 1, DEC_WINDOW);
        p->adapter_index += 1;
        *td.epoch = e + 1;
    }
}

// ---------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------
struct TpRank {
    TpDev dev;
    TpComm *comm;
    unsigned *ctr;            // [0] = done, [1] = epoch
    float *xbuf;              // [2][3072] residual row, ping-pong across exchanges
    float *blockmax;          // (value, index) per logits CTA
    DecStepParams *d_params;
    int *d_tokens;            // [TOKENS_RING]
    void *opened[TP_MAX];     // peer windows opened through CUDA IPC (closed in free)
};

struct vox_cuda_tp {
    vox_cuda_model *m;
    vox_cuda_state *s;        // the stream's replicated state: encoder, adapter rows, prefill, K/V ring
    int world, n_local, connected;
    TpRank r[TP_MAX];
    cudaGraphExec_t graph;
    int *h_check;             // pinned, second token ring for the lock-step self check
};

struct TpHandle {             // VOX_CUDA_TP_HANDLE_BYTES
    cudaIpcMemHandle_t ipc;   // 64 B
    unsigned long long offset;// of the window inside the exported allocation
    unsigned long long magic;
};
static_assert(sizeof(TpHandle) == VOX_CUDA_TP_HANDLE_BYTES, "handle size is part of the ABI");
constexpr unsigned long long TP_MAGIC = 0x564f585450303031ull;

static int tp_valid_world(int world) { return world == 1 || world == 2 || world == 4 || world == 8; }

static void tp_rank_free(TpRank &R) {
    for (int p = 0; p < TP_MAX; p++) if (R.opened[p]) cudaIpcCloseMemHandle(R.opened[p]);
    cudaFree(R.comm); cudaFree(R.ctr); cudaFree(R.xbuf); cudaFree(R.blockmax); cudaFree(R.d_params); cudaFree(R.d_tokens);
    memset(&R, 0, sizeof(R));
}

static int tp_rank_init(vox_cuda_tp *tp, TpRank &R, int rank) {
    memset(&R, 0, sizeof(R));
    CUDA_TRY(cudaMalloc((void **)&R.comm, sizeof(TpComm)));
    CUDA_TRY(cudaMalloc((void **)&R.ctr, 64));
    CUDA_TRY(cudaMalloc((void **)&R.xbuf, 2 * DEC_DIM * 4));
    CUDA_TRY(cudaMalloc((void **)&R.blockmax, (size_t)2 * (VOCAB / LOGITS_ROWS_PER_CTA) * 4));
    CUDA_TRY(cudaMalloc((void **)&R.d_params, sizeof(DecStepParams)));
    CUDA_TRY(cudaMalloc((void **)&R.d_tokens, (size_t)TOKENS_RING * 4));
    CUDA_TRY(cudaMemset(R.comm, 0, sizeof(TpComm)));
    CUDA_TRY(cudaMemset(R.ctr, 0, 64));
    CUDA_TRY(cudaMemset(R.d_tokens, 0, (size_t)TOKENS_RING * 4));
    R.dev.rank = rank; R.dev.world = tp->world; R.dev.ars = 2 * tp->m->n_dec + 1;
    R.dev.self = R.comm; R.dev.done = R.ctr; R.dev.epoch = R.ctr + 1;
    R.dev.peer[rank] = R.comm;
    return 0;
}

static vox_cuda_tp *tp_alloc(vox_cuda_model *m, int world) {
    if (!m || !tp_valid_world(world)) { vox_set_error("tp: world size must be 1, 2, 4 or 8"); return nullptr; }
    vox_cuda_tp *tp = (vox_cuda_tp *)calloc(1, sizeof(vox_cuda_tp));
    tp->m = m; tp->world = world;
    tp->s = vox_cuda_state_create(m);
    if (!tp->s) { free(tp); return nullptr; }
    return tp;
}

extern "C" void vox_cuda_tp_free(vox_cuda_tp_t *tp) {
    if (!tp) return;
    if (tp->s) cudaStreamSynchronize(tp->s->stream);
    if (tp->graph) cudaGraphExecDestroy(tp->graph);
    for (int i = 0; i < tp->n_local; i++) tp_rank_free(tp->r[i]);
    if (tp->h_check) cudaFreeHost(tp->h_check);
    vox_cuda_state_free(tp->s);
    free(tp);
}

// One rank of a world of `world` processes (one per GPU).  Peers are attached with export / connect.
extern "C" vox_cuda_tp_t *vox_cuda_tp_create(vox_cuda_model_t *m, int rank, int world) {
    vox_cuda_tp *tp = tp_alloc(m, world);
    if (!tp) return nullptr;
    if (rank < 0 || rank >= world) { vox_set_error("tp: rank %d outside world %d", rank, world); vox_cuda_tp_free(tp); return nullptr; }
    tp->n_local = 1;
    if (tp_rank_init(tp, tp->r[0], rank)) { vox_cuda_tp_free(tp); return nullptr; }
    tp->connected = (world == 1);
    return tp;
}

// All `world` ranks on this device, advanced kernel by kernel in lock step on one stream: the shard
// arithmetic, the exchange windows and the sequence numbers are those of the multi-GPU run, so one GPU is
// enough to check them (tests/test_gpu_tp.py).  Not a fast path: every kernel is launched `world` times.
extern "C" vox_cuda_tp_t *vox_cuda_tp_create_local(vox_cuda_model_t *m, int world) {
    vox_cuda_tp *tp = tp_alloc(m, world);
    if (!tp) return nullptr;
    for (int i = 0; i < world; i++) {
        if (tp_rank_init(tp, tp->r[i], i)) { tp->n_local = i + 1; vox_cuda_tp_free(tp); return nullptr; }
        tp->n_local = i + 1;
    }
    for (int i = 0; i < world; i++)
        for (int p = 0; p < world; p++) tp->r[i].dev.peer[p] = tp->r[p].comm;
    if (cudaMallocHost((void **)&tp->h_check, (size_t)TOKENS_RING * 4) != cudaSuccess) { vox_cuda_tp_free(tp); return nullptr; }
    tp->connected = 1;
    return tp;
}

extern "C" vox_cuda_state_t *vox_cuda_tp_state(vox_cuda_tp_t *tp) { return tp ? tp->s : nullptr; }

// handle_out: VOX_CUDA_TP_HANDLE_BYTES describing this rank's exchange window for the other processes
extern "C" int vox_cuda_tp_export(vox_cuda_tp_t *tp, void *handle_out) {
    if (!tp || !handle_out || tp->n_local != 1) return -1;
    TpHandle h; memset(&h, 0, sizeof(h));
    CUdeviceptr base = 0; size_t size = 0;
    {   // the IPC handle names the whole allocation; the window may sit at an offset inside it
        typedef CUresult (*range_fn)(CUdeviceptr *, size_t *, CUdeviceptr);
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &fn, cudaEnableDefault, &qres) != cudaSuccess || !fn ||
            ((range_fn)fn)(&base, &size, (CUdeviceptr)tp->r[0].comm) != CUDA_SUCCESS) {
            vox_set_error("tp_export: cuMemGetAddressRange failed"); return -1;
        }
    }
    CUDA_TRY(cudaIpcGetMemHandle(&h.ipc, (void *)base));
    h.offset = (unsigned long long)((CUdeviceptr)tp->r[0].comm - base);
    h.magic = TP_MAGIC;
    memcpy(handle_out, &h, sizeof(h));
    return 0;
}

// handles: world x VOX_CUDA_TP_HANDLE_BYTES in rank order (entry [rank] is ignored)
extern "C" int vox_cuda_tp_connect(vox_cuda_tp_t *tp, const void *handles) {
    if (!tp || !handles || tp->n_local != 1) return -1;
    TpRank &R = tp->r[0];
    for (int p = 0; p < tp->world; p++) {
        if (p == R.dev.rank) continue;
        TpHandle h; memcpy(&h, (const char *)handles + (size_t)p * sizeof(TpHandle), sizeof(h));
        if (h.magic != TP_MAGIC) { vox_set_error("tp_connect: entry %d is not a vox_cuda_tp_export handle", p); return -1; }
        void *base = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&base, h.ipc, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            vox_set_error("tp_connect: cannot map rank %d's window (%s): peer access over NVLink is required", p, cudaGetErrorString(e));
            return -1;
        }
        R.opened[p] = base;
        R.dev.peer[p] = (TpComm *)((char *)base + h.offset);
    }
    tp->connected = 1;
    return 0;
}

static inline int tp_step_kernels(const vox_cuda_tp *tp) { return 1 + 6 * tp->m->n_dec + 3; }

static int tp_enqueue_step(vox_cuda_tp *tp, cudaStream_t st) {
    vox_cuda_model *m = tp->m;
    vox_cuda_state *s = tp->s;
    const int G = tp->world, NL = tp->n_local, L = m->n_dec;
    const size_t kv_layer = (size_t)DEC_WINDOW * DEC_KVDIM;
    const int qkv_rows = DEC_QKV / G, ffn_rows = 2 * DEC_HIDDEN / G;
    const int rpw = G <= 1 ? 4 : 2;               // keep >= 96 CTAs per GEMV as the shard shrinks
    for (int i = 0; i < NL; i++) {
        TpRank &R = tp->r[i];
        dec_launch_embed(st, R.d_params, s->d_embed_in, s->adapter_buf, m->tok_emb, m->inv_freq_dec, R.xbuf, s->d_rope);
    }
    for (int l = 0; l < L; l++) {
        const DecLayerDev &W = m->dec[l];
        float *kc = s->kv_k + l * kv_layer, *vc = s->kv_v + l * kv_layer;
        for (int i = 0; i < NL; i++) {
            TpRank &R = tp->r[i];
            const int a = 2 * l - 1;              // consumes the w2 exchange of the previous layer
            float *xin = l == 0 ? R.xbuf : R.xbuf + DEC_DIM, *xout = R.xbuf;
            const int grid = qkv_rows / (TP_WARPS * rpw);
            if (l == 0) {
                if (rpw == 4) tp_qkv_kernel<4, false><<<grid, TP_THREADS, 0, st>>>(R.dev, a, R.d_params, xin, nullptr, W.attn_norm, W.wqkv, s->d_rope, s->d_q, kc, vc);
                else          tp_qkv_kernel<2, false><<<grid, TP_THREADS, 0, st>>>(R.dev, a, R.d_params, xin, nullptr, W.attn_norm, W.wqkv, s->d_rope, s->d_q, kc, vc);
            } else {
                if (rpw == 4) tp_qkv_kernel<4, true><<<grid, TP_THREADS, 0, st>>>(R.dev, a, R.d_params, xin, xout, W.attn_norm, W.wqkv, s->d_rope, s->d_q, kc, vc);
                else          tp_qkv_kernel<2, true><<<grid, TP_THREADS, 0, st>>>(R.dev, a, R.d_params, xin, xout, W.attn_norm, W.wqkv, s->d_rope, s->d_q, kc, vc);
            }
        }
        for (int i = 0; i < NL; i++) {
            TpRank &R = tp->r[i];
            dec_launch_attn(st, R.d_params, s->d_q, kc, vc, s->d_attn_part, s->d_attn, R.dev.rank * (DEC_KVH / G), DEC_KVH / G);
        }
        for (int i = 0; i < NL; i++)
            tp_gemv_partial_kernel<<<DEC_DIM / TP_RES_ROWS, TP_THREADS, (size_t)(DEC_QDIM / G) * 4, st>>>(
                tp->r[i].dev, 2 * l, s->d_attn, W.wo, DEC_QDIM);
        for (int i = 0; i < NL; i++) {
            TpRank &R = tp->r[i];
            const int grid = ffn_rows / (TP_WARPS * rpw);
            const float *ada = m->ada_scale + (size_t)l * DEC_DIM;
            if (rpw == 4) tp_ffn_kernel<4><<<grid, TP_THREADS, 0, st>>>(R.dev, 2 * l, R.xbuf, R.xbuf + DEC_DIM, W.ffn_norm, ada, W.w13, s->d_h);
            else          tp_ffn_kernel<2><<<grid, TP_THREADS, 0, st>>>(R.dev, 2 * l, R.xbuf, R.xbuf + DEC_DIM, W.ffn_norm, ada, W.w13, s->d_h);
        }
        for (int i = 0; i < NL; i++)
            tp_gemv_partial_kernel<<<DEC_DIM / TP_RES_ROWS, TP_THREADS, (size_t)(DEC_HIDDEN / G) * 4, st>>>(
                tp->r[i].dev, 2 * l + 1, s->d_h, W.w2, DEC_HIDDEN);
    }
    const int nblk = VOCAB / G / LOGITS_ROWS_PER_CTA;
    for (int i = 0; i < NL; i++) {
        TpRank &R = tp->r[i];
        tp_logits_kernel<<<nblk, TP_THREADS, 0, st>>>(R.dev, 2 * L - 1, R.xbuf + DEC_DIM, m->dec_norm, m->tok_emb, s->d_logits, R.blockmax);
    }
    for (int i = 0; i < NL; i++) tp_argmax_pub_kernel<<<1, 1024, 0, st>>>(tp->r[i].dev, tp->r[i].blockmax, nblk);
    for (int i = 0; i < NL; i++) tp_argmax_fin_kernel<<<1, 32, 0, st>>>(tp->r[i].dev, tp->r[i].d_params, tp->r[i].d_tokens);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

static int tp_ensure_graph(vox_cuda_tp *tp) {
    if (tp->graph) return 0;
    cudaStream_t st = tp->s->stream;
    cudaGraph_t g;
    const unsigned long long before = g_launch_count;
    CUDA_TRY(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    const int rc = tp_enqueue_step(tp, st);
    cudaError_t e = cudaStreamEndCapture(st, &g);
    g_launch_count = before;                       // capture is not execution; replays are counted in tp_decode
    if (rc) return -1;
    CUDA_TRY(e);
    CUDA_TRY(cudaGraphInstantiate(&tp->graph, g, 0));
    cudaGraphDestroy(g);
    return 0;
}

// n greedy steps of the stream held by vox_cuda_tp_state(tp), which must have been encoded and prefilled
// through the usual calls ON EVERY RANK (replicated; the same arguments everywhere).  Same contract as
// vox_cuda_decoder_run; every rank returns the same ids.  ms_per_step (optional): CUDA-event time of the
// n steps on this rank's stream / n.  All ranks must call this together: a rank waits (bounded, 8 s) for
// its peers inside the kernels.
extern "C" int vox_cuda_tp_decode(vox_cuda_tp_t *tp, int prev_token, int adapter_index, int n, int *tokens_out,
                                  float *ms_per_step) {
    if (!tp || n <= 0 || n > TOKENS_RING || !tokens_out) return -1;
    if (!tp->connected) { vox_set_error("tp_decode: peers not attached (vox_cuda_tp_connect)"); return -1; }
    if (prev_token < 0 || prev_token >= VOCAB) return -1;
    vox_cuda_state *s = tp->s;
    const int phys = adapter_index - s->adapter_offset;
    if (phys < 0 || adapter_index + n > s->adapter_count) {
        vox_set_error("tp_decode: adapter range [%d,%d) outside [%d,%d)", adapter_index, adapter_index + n,
                      s->adapter_offset, s->adapter_count);
        return -1;
    }
    cudaStream_t st = s->stream;
    s->h_params.use_host_embed = 0;
    s->h_params.prev_token = prev_token;
    s->h_params.adapter_index = phys;
    const int c0 = s->h_params.step_count;
    if (tp_ensure_graph(tp)) return -1;
    for (int i = 0; i < tp->n_local; i++)
        CUDA_TRY(cudaMemcpyAsync(tp->r[i].d_params, &s->h_params, sizeof(DecStepParams), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaEventRecord(s->ev0, st));
    for (int i = 0; i < n; i++) CUDA_TRY(cudaGraphLaunch(tp->graph, st));
    CUDA_TRY(cudaEventRecord(s->ev1, st));
    g_launch_count += (unsigned long long)n * tp_step_kernels(tp) * tp->n_local;
    CUDA_TRY(cudaMemcpyAsync(s->h_tokens, tp->r[0].d_tokens, (size_t)TOKENS_RING * 4, cudaMemcpyDeviceToHost, st));
    if (tp->n_local > 1)
        CUDA_TRY(cudaMemcpyAsync(tp->h_check, tp->r[tp->n_local - 1].d_tokens, (size_t)TOKENS_RING * 4, cudaMemcpyDeviceToHost, st));
    if (cudaStreamSynchronize(st) != cudaSuccess) {
        vox_set_error("tp_decode: %s (a rank that waits 8 s for a peer's exchange traps)", cudaGetErrorString(cudaGetLastError()));
        return -1;
    }
    if (ms_per_step) { float ms = 0.f; cudaEventElapsedTime(&ms, s->ev0, s->ev1); *ms_per_step = ms / (float)n; }
    for (int i = 0; i < n; i++) {
        const int slot = (c0 + i) & (TOKENS_RING - 1);
        tokens_out[i] = s->h_tokens[slot];
        if (tp->n_local > 1 && tp->h_check[slot] != tokens_out[i]) {
            vox_set_error("tp_decode: ranks disagree at step %d (%d vs %d)", i, tokens_out[i], tp->h_check[slot]);
            return -1;
        }
    }
    s->h_params.pos += n;
    s->h_params.n_valid = s->h_params.n_valid + n > DEC_WINDOW ? DEC_WINDOW : s->h_params.n_valid + n;
    s->h_params.adapter_index += n;
    s->h_params.step_count += n;
    s->h_params.prev_token = s->h_params.out_token = tokens_out[n - 1];
    CUDA_TRY(cudaMemcpyAsync(s->d_params, &s->h_params, sizeof(DecStepParams), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return n;
}

// Whole clip like vox_cuda_transcribe (api_batch.cu; vox_transcribe_audio, voxtral.c:1390-1396): encoder and
// prompt prefill replicated on every rank, the greedy steps tensor-parallel.  Same arguments on all ranks.
extern "C" int vox_cuda_tp_transcribe(vox_cuda_tp_t *tp, const float *pcm, int n_samples, int n_frames,
                                      int prompt_len, int *tokens_out, int max_tokens, float *ms_per_step) {
    if (!tp || !tokens_out || n_frames <= 0 || prompt_len < 2 || prompt_len > 128) return -1;
    vox_cuda_state *s = tp->s;
    if (batch_encode_clip(s, pcm, n_samples, n_frames)) return -1;
    const int T = s->adapter_count, n_pre = prompt_len - 1;
    if (T < prompt_len) { vox_set_error("tp_transcribe: only %d adapter tokens (< prompt %d)", T, prompt_len); return -1; }
    int prompt[128];
    prompt[0] = 1;                               // BOS, then [STREAMING_PAD] (voxtral.c:396-399,1039-1048)
    for (int i = 1; i < prompt_len; i++) prompt[i] = 32;
    if (vox_cuda_decoder_prefill_fused(s, prompt, 0, n_pre)) return -1;
    int n = T - n_pre;
    if (n > max_tokens) n = max_tokens;
    return vox_cuda_tp_decode(tp, prompt[n_pre], n_pre, n, tokens_out, ms_per_step);
}
