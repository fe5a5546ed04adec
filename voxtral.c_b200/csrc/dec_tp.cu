// dec_tp.cu - tensor-parallel single-stream decoder step (BASELINE.json north_star:
// "an optional tensor-parallel single-stream decoder ... allreduce over NVLink after wo and w2").
//
// Arithmetic is vox_decoder_forward (/root/reference/voxtral_decoder.c:640-780) with the matrices cut the
// way SURVEY 8(e) lays out: rank r of G owns kv heads [r*8/G, (r+1)*8/G) and their 4 query heads each
// (rows of wq|wk|wv), rows [r*18432/G, ...) of the interleaved w1|w3, the matching COLUMN range of wo
// and w2 (whose products are therefore partial sums of all 3072 outputs), and vocabulary rows
// [r*131072/G, ...) of the tied embedding.  Weights are replicated (8.86 GB of 180 GB); a shard is a view.
//
// The reduction after wo / w2 is not a separate collective: as soon as a warp of the GEMV kernel has a row
// sum, lanes 0..G-1 store it into the G ranks' exchange windows over NVLink (peer memory opened through
// CUDA IPC) as ONE 8-byte word {value, sequence number} - the flag travels inside the data (the
// low-latency protocol NCCL calls LL), so there is no fence, no arrival counter and no flag round trip.
// The NEXT kernel's prologue (the RMSNorm of the FFN / of the next layer / of the logits) polls the words
// until they carry the expected sequence number and adds the G partials in rank order - every rank holds
// bit-identical residual rows and no kernel is launched for the exchange.  The greedy token is exchanged
// the same way as (max, index) words of the vocabulary slices.
//
// Every kernel is latency-bound at these sizes (a rank's share of a matrix is 2-28 MB), so: each warp
// issues ALL weight loads of its row batch before anything else (before the programmatic-dependent-launch
// wait and before the exchange wait, which then overlap the HBM fetch); kernels are chained with PDL;
// attention is one kernel (key splits across CTAs, the last CTA of a kv head merges the splits).
//
// PDL ordering.  Kernels that read what the previous kernel of the stream wrote (attention: q/K/V; the
// wo / w2 GEMVs: attention output / h; layer 0's qkv: the embedding) call griddepcontrol.wait and only
// THEN griddepcontrol.launch_dependents.  The three consumers of an exchange (ffn, qkv of layers >= 1,
// logits) do not wait at all: their only input from the previous kernel (the partial GEMV) are the
// exchange words, which carry their own flags, so they run beside it - weights fetched, x loaded, polling
// - and the NVLink hop overlaps the GEMV's tail.  They cannot start before every CTA of that GEMV has
// passed ITS wait, i.e. before the kernel two back (which wrote their x_in) has completed; every buffer
// they write (x_out, q/K/V, h, logits) was last read by a kernel at least that old.
#include "dec_tp_dev.cuh"
#include <cuda.h>
#include <stdlib.h>
#include <string.h>

// Prologue of the three consuming kernels: x = x_in (+ the G partials of exchange a_idx, in rank order),
// CTA 0 keeps the new residual row in x_out, then RMSNorm (voxtral_kernels.c:475-492) and the (1 + ada)
// scale (voxtral_decoder.c:742-745) into shared memory.  A thread owns three 4-float pieces of x (768 pieces
// over 256 threads) and reads the exchange words of all its pieces from up to 4 source ranks in ONE round of
// loads (24 x 16 B in flight per thread); 8 ranks take two rounds, still added in rank order.  Every
// independent load (x, norm weight, ada, the first read of the exchange words) is issued before anything is
// waited for.
__device__ __forceinline__ void tp_ll_settle4(const TpDev &td, const uint2 *src, int p, unsigned seq, uint4 (&v)[2]) {
    if (v[0].y == seq && v[0].w == seq && v[1].y == seq && v[1].w == seq) return;
    const unsigned long long t0 = globaltimer_ns();
    for (;;) {
        v[0] = ld_ll2(src); v[1] = ld_ll2(src + 2);
        if (v[0].y == seq && v[0].w == seq && v[1].y == seq && v[1].w == seq) return;
        if (globaltimer_ns() - t0 > 8000000000ull) tp_timeout(td.rank, p, seq);
    }
}

// Cluster gather (round 2, opt-in): with every CTA polling all G x 24 KB of exchange words the read volume grows with
// G while the GEMV shrinks (8 GPUs measured SLOWER than 4) - the hypothesis this tests; see tp_cluster_size for the result.  Launched as clusters of TP_CS CTAs, each CTA polls and sums
// only 1 / TP_CS of the row (G x 6 KB), stores its quarter into the shared memory of every cluster member (DSMEM), and
// one cluster barrier later every CTA has the whole reduced row.  Same summation order (rank order) as the plain path.
constexpr int TP_CS = 4;
__device__ __forceinline__ unsigned tp_cluster_rank() {
    unsigned r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void tp_st_cluster(float4 *local_slot, unsigned cta, float4 v) {
    unsigned la = (unsigned)__cvta_generic_to_shared(local_slot), ra;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(la), "r"(cta));
    asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(ra), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void tp_cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

template <bool REDUCE>
__device__ __forceinline__ void tp_load_x_norm(float4 *xA, float4 *xB, float *red, const TpDev &td, int a_idx,
                                               const float *__restrict__ x_in, float *__restrict__ x_out,
                                               const float *__restrict__ nw, const float *__restrict__ ada,
                                               float4 *xs = nullptr /* [768] shared: cluster gather */) {
    constexpr int NP = DEC_DIM / 4 / TP_THREADS;       // 3 pieces per thread
    const int t = threadIdx.x;
    float4 x[NP], wn[NP], wd[NP];
    const bool clustered = REDUCE && xs != nullptr;
    #pragma unroll
    for (int i = 0; i < NP; i++) {
        const int hc = t + i * TP_THREADS;
        if (!clustered) x[i] = *reinterpret_cast<const float4 *>(x_in + hc * 4);
        wn[i] = *reinterpret_cast<const float4 *>(nw + hc * 4);
        if (ada) wd[i] = *reinterpret_cast<const float4 *>(ada + hc * 4);
    }
    if (clustered) {
        constexpr int PPC = DEC_DIM / 4 / TP_CS;          // 192 pieces of 4 floats per cluster member
        const unsigned seq = *td.epoch * (unsigned)td.ars + (unsigned)a_idx + 1u;
        if (t < PPC) {
            const int hc = (int)tp_cluster_rank() * PPC + t;
            float4 xv = *reinterpret_cast<const float4 *>(x_in + hc * 4);
            uint4 v[TP_MAX][2];
            #pragma unroll
            for (int pp = 0; pp < TP_MAX; pp++)
                if (pp < td.world) {
                    const uint2 *src = td.self->part[a_idx & 1][pp] + hc * 4;
                    v[pp][0] = ld_ll2(src); v[pp][1] = ld_ll2(src + 2);
                }
            #pragma unroll
            for (int pp = 0; pp < TP_MAX; pp++)
                if (pp < td.world) {
                    tp_ll_settle4(td, td.self->part[a_idx & 1][pp] + hc * 4, pp, seq, v[pp]);
                    xv.x += __uint_as_float(v[pp][0].x); xv.y += __uint_as_float(v[pp][0].z);
                    xv.z += __uint_as_float(v[pp][1].x); xv.w += __uint_as_float(v[pp][1].z);
                }
            #pragma unroll
            for (unsigned c = 0; c < (unsigned)TP_CS; c++) tp_st_cluster(xs + hc, c, xv);
        }
        tp_cluster_sync();
        #pragma unroll
        for (int i = 0; i < NP; i++) x[i] = xs[t + i * TP_THREADS];
        if (blockIdx.x == 0 && x_out) {
            #pragma unroll
            for (int i = 0; i < NP; i++) *reinterpret_cast<float4 *>(x_out + (t + i * TP_THREADS) * 4) = x[i];
        }
    } else if (REDUCE) {
        const unsigned seq = *td.epoch * (unsigned)td.ars + (unsigned)a_idx + 1u;
        for (int p0 = 0; p0 < td.world; p0 += 4) {
            uint4 v[NP][4][2];
            #pragma unroll
            for (int i = 0; i < NP; i++)
                #pragma unroll
                for (int pp = 0; pp < 4; pp++)
                    if (p0 + pp < td.world) {
                        const uint2 *src = td.self->part[a_idx & 1][p0 + pp] + (t + i * TP_THREADS) * 4;
                        v[i][pp][0] = ld_ll2(src); v[i][pp][1] = ld_ll2(src + 2);
                    }
            #pragma unroll
            for (int i = 0; i < NP; i++)
                #pragma unroll
                for (int pp = 0; pp < 4; pp++)
                    if (p0 + pp < td.world) {
                        tp_ll_settle4(td, td.self->part[a_idx & 1][p0 + pp] + (t + i * TP_THREADS) * 4, p0 + pp, seq, v[i][pp]);
                        x[i].x += __uint_as_float(v[i][pp][0].x); x[i].y += __uint_as_float(v[i][pp][0].z);
                        x[i].z += __uint_as_float(v[i][pp][1].x); x[i].w += __uint_as_float(v[i][pp][1].z);
                    }
        }
        if (blockIdx.x == 0 && x_out) {
            #pragma unroll
            for (int i = 0; i < NP; i++) *reinterpret_cast<float4 *>(x_out + (t + i * TP_THREADS) * 4) = x[i];
        }
    }
    float ss = 0.f;
    #pragma unroll
    for (int i = 0; i < NP; i++) ss += f4sq(x[i]);
    ss = warp_sum(ss);
    if ((t & 31) == 0) red[t >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
    #pragma unroll
    for (int w = 0; w < TP_WARPS; w++) tot += red[w];
    const float rinv = 1.0f / sqrtf(tot / (float)DEC_DIM + NORM_EPS);
    #pragma unroll
    for (int i = 0; i < NP; i++) {
        const int hc = t + i * TP_THREADS;
        float4 u = x[i];
        u.x = u.x * rinv * wn[i].x; u.y = u.y * rinv * wn[i].y; u.z = u.z * rinv * wn[i].z; u.w = u.w * rinv * wn[i].w;
        if (ada) { u.x *= (1.0f + wd[i].x); u.y *= (1.0f + wd[i].y); u.z *= (1.0f + wd[i].z); u.w *= (1.0f + wd[i].w); }
        if (hc & 1) xB[hc >> 1] = u; else xA[hc >> 1] = u;
    }
    __syncthreads();
}

constexpr int FULL_NW = 2, FULL_NCH = DEC_DIM / 256;      // K = 3072 kernels: 2 rows x 12 chunks = 24 loads per lane

// ---------------------------------------------------------------------------
// T1: this rank's rows of wq|wk|wv on rmsnorm(x); RoPE; q -> q_out, K/V -> ring slot
// (voxtral_decoder.c:710-725).  Buffers keep GLOBAL row / head indices; a rank touches its slice only.
// ---------------------------------------------------------------------------
template <bool REDUCE, class KV>
__global__ void __launch_bounds__(TP_THREADS) tp_qkv_kernel(
    TpDev td, int a_idx, const DecStepParams *__restrict__ p, const float *__restrict__ x_in,
    float *__restrict__ x_out, const float *__restrict__ attn_norm, const uint16_t *__restrict__ wqkv,
    const float *__restrict__ rope, float *__restrict__ q_out, KV *__restrict__ kc, KV *__restrict__ vc, int cs,
    const float *__restrict__ row_scale /* Q8 model: scale of matrix row n (the weights are the int8 values as bf16), else null */) {
    __shared__ float4 xA[DEC_DIM / 8], xB[DEC_DIM / 8];
    __shared__ float4 xs[DEC_DIM / 4];               // cluster gather buffer (cs > 1)
    __shared__ float red[TP_WARPS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nq = DEC_QDIM / td.world, nk = DEC_KVDIM / td.world;
    const int nbatch = (nq + 2 * nk) / FULL_NW, nwarps = gridDim.x * TP_WARPS;
    auto grow_of = [&](int b) {
        const int lrow = b * FULL_NW;
        if (lrow < nq) return td.rank * nq + lrow;
        if (lrow < nq + nk) return DEC_QDIM + td.rank * nk + (lrow - nq);
        return DEC_QDIM + DEC_KVDIM + td.rank * nk + (lrow - nq - nk);
    };
    int b = blockIdx.x * TP_WARPS + warp;
    uint4 w[FULL_NW][FULL_NCH];
    if (b < nbatch) tp_wload<FULL_NW, FULL_NCH>(w, wqkv + (size_t)grow_of(b) * DEC_DIM, DEC_DIM, DEC_DIM);
    if (!REDUCE) tcx::pdl_wait();        // layer 0: x comes from the embed kernel; else see "PDL ordering" above
    tcx::pdl_launch_dependents();
    tp_load_x_norm<REDUCE>(xA, xB, red, td, a_idx, x_in, x_out, attn_norm, nullptr, cs > 1 ? xs : nullptr);
    const int slot = p->pos & (DEC_WINDOW - 1);
    while (b < nbatch) {
        float acc[FULL_NW];
        tp_wdot<FULL_NW, FULL_NCH>(w, xA, xB, DEC_DIM, acc);
        const int row = grow_of(b);
        b += nwarps;
        if (b < nbatch) tp_wload<FULL_NW, FULL_NCH>(w, wqkv + (size_t)grow_of(b) * DEC_DIM, DEC_DIM, DEC_DIM);
        if (lane == 0) {
            float v0 = acc[0], v1 = acc[1];
            if (row_scale) { v0 *= row_scale[row]; v1 *= row_scale[row + 1]; }
            if (row < DEC_QDIM + DEC_KVDIM) {     // interleaved-pair rotation, voxtral_kernels.c:631-655
                const int d = (row & (DEC_HD - 1)) >> 1;
                const float c = rope[2 * d], s = rope[2 * d + 1];
                const float r0 = v0 * c - v1 * s, r1 = v0 * s + v1 * c;
                v0 = r0; v1 = r1;
            }
            if (row < DEC_QDIM) *reinterpret_cast<float2 *>(q_out + row) = make_float2(v0, v1);
            else if (row < DEC_QDIM + DEC_KVDIM) tp_st_kv2(kc + (size_t)slot * DEC_KVDIM + (row - DEC_QDIM), v0, v1);
            else tp_st_kv2(vc + (size_t)slot * DEC_KVDIM + (row - DEC_QDIM - DEC_KVDIM), v0, v1);
        }
    }
}

// ---------------------------------------------------------------------------
// T2: GQA attention of this rank's kv heads over the ring (vox_causal_attention with seq_q = 1,
// voxtral_kernels.c:541-611; call voxtral_decoder.c:731-733).  grid = (local kv heads, key splits).
// A warp takes 8 keys at a time with all 16 K/V row loads in flight (lanes own 4 of the 128 dims; the
// 8 x 4 scores are reduced across lanes by a transposed butterfly) - the scheme of the single-stream
// kernel (dec_mega.cu) and of group_dev.cuh.  The CTA that arrives last for a kv head merges the splits.
// ---------------------------------------------------------------------------
template <class KV>
__global__ void __launch_bounds__(TP_THREADS) tp_attn_kernel(
    TpDev td, const DecStepParams *__restrict__ p, const float *__restrict__ q, const KV *__restrict__ kc,
    const KV *__restrict__ vc, float *__restrict__ part, float *__restrict__ attn_out) {
    __shared__ __align__(16) TpAttnSmem sm;
    tcx::pdl_wait();
    tcx::pdl_launch_dependents();
    tp_attn_task(td, blockIdx.x, blockIdx.y, gridDim.y, p->pos, p->n_valid, q, kc, vc, part, attn_out, sm);
}

// ---------------------------------------------------------------------------
// T3 / T5: partial[row] = W[row][r*ks .. (r+1)*ks) . v[r*ks ..)  for all 3072 rows (wo: K = 4096,
// voxtral_decoder.c:738-739; w2: K = 9216, :758-760).  Lane p of the warp that owns a row stores the sum
// into rank p's window (its own included) the moment it exists.  The residual add happens in the
// consumer's prologue, so x is not touched here.
// ---------------------------------------------------------------------------
template <int NW, int NCH>
__global__ void __launch_bounds__(TP_THREADS) tp_gemv_partial_kernel(
    TpDev td, int a_idx, const float *__restrict__ v, const uint16_t *__restrict__ W, int K, const float *__restrict__ row_scale) {
    extern __shared__ __align__(16) unsigned char tp_smem[];
    const int ks = K / td.world;
    float4 *xA = reinterpret_cast<float4 *>(tp_smem);
    float4 *xB = xA + ks / 8;
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int nbatch = DEC_DIM / NW, nwarps = gridDim.x * TP_WARPS;
    const uint16_t *Ws = W + (size_t)td.rank * ks;
    int b = blockIdx.x * TP_WARPS + warp;
    uint4 w[NW][NCH];
    if (b < nbatch) tp_wload<NW, NCH>(w, Ws + (size_t)b * NW * K, (size_t)K, ks);
    tcx::pdl_wait();
    tcx::pdl_launch_dependents();
    const float *vs = v + (size_t)td.rank * ks;
    for (int c = t; c < ks / 8; c += TP_THREADS) {
        xA[c] = *reinterpret_cast<const float4 *>(vs + c * 8);
        xB[c] = *reinterpret_cast<const float4 *>(vs + c * 8 + 4);
    }
    __syncthreads();
    const unsigned seq = *td.epoch * (unsigned)td.ars + (unsigned)a_idx + 1u;
    uint2 *dst = lane < td.world ? td.peer[lane]->part[a_idx & 1][td.rank] : nullptr;
    while (b < nbatch) {
        float acc[NW];
        tp_wdot<NW, NCH>(w, xA, xB, ks, acc);
        const int row0 = b * NW;
        b += nwarps;
        if (b < nbatch) tp_wload<NW, NCH>(w, Ws + (size_t)b * NW * K, (size_t)K, ks);
        if (dst) {                              // lane p sends the sums to rank p: value and flag in one 8-byte word
            #pragma unroll
            for (int r = 0; r < NW; r++) st_ll(dst + row0 + r, __float_as_uint(row_scale ? acc[r] * row_scale[row0 + r] : acc[r]), seq);
        }
    }
}

// ---------------------------------------------------------------------------
// T4: h[i] = silu(w1[i] . xn) * (w3[i] . xn) for this rank's rows of the interleaved w1|w3, xn =
// rmsnorm(x + sum of the wo partials) * (1 + ada)   (voxtral_decoder.c:741-757)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(TP_THREADS) tp_ffn_kernel(
    TpDev td, int a_idx, const float *__restrict__ x_in, float *__restrict__ x_out,
    const float *__restrict__ ffn_norm, const float *__restrict__ ada, const uint16_t *__restrict__ w13,
    float *__restrict__ h, int cs, const float *__restrict__ row_scale) {
    __shared__ float4 xA[DEC_DIM / 8], xB[DEC_DIM / 8];
    __shared__ float4 xs[DEC_DIM / 4];
    __shared__ float red[TP_WARPS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nrows = 2 * DEC_HIDDEN / td.world, row_base = td.rank * nrows;
    const int nbatch = nrows / FULL_NW, nwarps = gridDim.x * TP_WARPS;
    int b = blockIdx.x * TP_WARPS + warp;
    uint4 w[FULL_NW][FULL_NCH];
    if (b < nbatch) tp_wload<FULL_NW, FULL_NCH>(w, w13 + (size_t)(row_base + b * FULL_NW) * DEC_DIM, DEC_DIM, DEC_DIM);
    tcx::pdl_launch_dependents();        // no PDL wait: the exchange words are the dependency ("PDL ordering" above)
    tp_load_x_norm<true>(xA, xB, red, td, a_idx, x_in, x_out, ffn_norm, ada, cs > 1 ? xs : nullptr);
    while (b < nbatch) {
        float acc[FULL_NW];
        tp_wdot<FULL_NW, FULL_NCH>(w, xA, xB, DEC_DIM, acc);
        const int row = row_base + b * FULL_NW;
        b += nwarps;
        if (b < nbatch) tp_wload<FULL_NW, FULL_NCH>(w, w13 + (size_t)(row_base + b * FULL_NW) * DEC_DIM, DEC_DIM, DEC_DIM);
        if (lane == 0) {
            if (row_scale) { acc[0] *= row_scale[row]; acc[1] *= row_scale[row + 1]; }
            h[row >> 1] = silu(acc[0]) * acc[1];
        }
    }
}

// ---------------------------------------------------------------------------
// T6: logits of this rank's vocabulary rows + per-CTA first maximum (voxtral_decoder.c:765-779)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(TP_THREADS) tp_logits_kernel(
    TpDev td, int a_idx, const float *__restrict__ x_in, const float *__restrict__ norm,
    const uint16_t *__restrict__ tok_emb, float *__restrict__ logits, float *__restrict__ blockmax, int cs,
    const float *__restrict__ row_scale) {
    __shared__ float4 xA[DEC_DIM / 8], xB[DEC_DIM / 8];
    __shared__ float4 xs[DEC_DIM / 4];
    __shared__ float red[TP_WARPS];
    __shared__ float bvs[TP_WARPS];
    __shared__ int bis[TP_WARPS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nrows = VOCAB / td.world, row_base = td.rank * nrows;
    const int nbatch = nrows / FULL_NW, nwarps = gridDim.x * TP_WARPS;
    int b = blockIdx.x * TP_WARPS + warp;
    uint4 w[FULL_NW][FULL_NCH];
    if (b < nbatch) tp_wload<FULL_NW, FULL_NCH>(w, tok_emb + (size_t)(row_base + b * FULL_NW) * DEC_DIM, DEC_DIM, DEC_DIM);
    tcx::pdl_launch_dependents();        // no PDL wait: the exchange words are the dependency
    tp_load_x_norm<true>(xA, xB, red, td, a_idx, x_in, nullptr, norm, nullptr, cs > 1 ? xs : nullptr);
    float bv = -INFINITY; int bi = 0x7fffffff;
    while (b < nbatch) {
        float acc[FULL_NW];
        tp_wdot<FULL_NW, FULL_NCH>(w, xA, xB, DEC_DIM, acc);
        const int row = row_base + b * FULL_NW;
        b += nwarps;
        if (b < nbatch) tp_wload<FULL_NW, FULL_NCH>(w, tok_emb + (size_t)(row_base + b * FULL_NW) * DEC_DIM, DEC_DIM, DEC_DIM);
        if (lane == 0) {
            if (row_scale) { acc[0] *= row_scale[row]; acc[1] *= row_scale[row + 1]; }
            *reinterpret_cast<float2 *>(logits + row) = make_float2(acc[0], acc[1]);
            tp_argmax_merge(bv, bi, acc[0], row);
            tp_argmax_merge(bv, bi, acc[1], row + 1);
        }
    }
    if (lane == 0) { bvs[warp] = bv; bis[warp] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        #pragma unroll
        for (int k = 1; k < TP_WARPS; k++) tp_argmax_merge(bv, bi, bvs[k], bis[k]);
        blockmax[2 * blockIdx.x] = bv;
        blockmax[2 * blockIdx.x + 1] = __int_as_float(bi);
    }
}

// T7: first maximum of this rank's slice -> every rank's window (exchange index 2 * layers)
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
            uint2 *dst = td.peer[p]->amax[seq & 1][td.rank];
            st_ll(dst, __float_as_uint(bv), seq);
            st_ll(dst + 1, (unsigned)bi, seq);
        }
    }
}

// T8: merge the G slice maxima (same order on every rank), write the token, advance the step scalars
__global__ void __launch_bounds__(32) tp_argmax_fin_kernel(TpDev td, DecStepParams *__restrict__ p,
                                                           int *__restrict__ tokens) {
    __shared__ float vals[TP_MAX];
    __shared__ int idxs[TP_MAX];
    const unsigned e = *td.epoch;
    const unsigned seq = e * (unsigned)td.ars + (unsigned)td.ars;
    if ((int)threadIdx.x < td.world) {
        const uint2 *src = td.self->amax[seq & 1][threadIdx.x];
        uint4 v = ld_ll2(src);
        if (v.y != seq || v.w != seq) {
            const unsigned long long t0 = globaltimer_ns();
            do {
                v = ld_ll2(src);
                if (globaltimer_ns() - t0 > 8000000000ull) tp_timeout(td.rank, threadIdx.x, seq);
            } while (v.y != seq || v.w != seq);
        }
        vals[threadIdx.x] = __uint_as_float(v.x);
        idxs[threadIdx.x] = (int)v.z;
    }
    __syncwarp();
    if (threadIdx.x == 0) {
        float bv = -INFINITY; int bi = 0x7fffffff;
        for (int r = 0; r < td.world; r++) tp_argmax_merge(bv, bi, vals[r], idxs[r]);
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
    unsigned *ctr;            // [1..8] = attention arrivals per local kv head, [16] = epoch
    float *xbuf;              // [2][3072] residual row, ping-pong across exchanges
    float *blockmax;          // (value, index) per logits CTA
    DecStepParams *d_params;
    int *d_tokens;            // [TOKENS_RING]
    void *opened[TP_MAX];     // peer windows opened through CUDA IPC (closed in free)
};

struct vox_cuda_tp {
    vox_cuda_model *m;
    vox_cuda_state *s;        // the stream's replicated state: encoder, adapter rows, prefill, K/V ring
    float *attn_part;         // [8 kv heads][TP_ATT_SPLITS][4][TP_ATT_PART] split partials (global head index)
    int att_splits;
    int world, n_local, connected;
    TpRank r[TP_MAX];
    cudaGraphExec_t graph;
    void *mega;               // TpMegaHost*: persistent step kernel (dec_tp_mega.cu); null = kernel-per-op graph
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
    CUDA_TRY(cudaMalloc((void **)&R.ctr, 128));
    CUDA_TRY(cudaMalloc((void **)&R.xbuf, 2 * DEC_DIM * 4));
    CUDA_TRY(cudaMalloc((void **)&R.blockmax, (size_t)2 * (VOCAB / LOGITS_ROWS_PER_CTA) * 4));
    CUDA_TRY(cudaMalloc((void **)&R.d_params, sizeof(DecStepParams)));
    CUDA_TRY(cudaMalloc((void **)&R.d_tokens, (size_t)TOKENS_RING * 4));
    CUDA_TRY(cudaMemset(R.comm, 0, sizeof(TpComm)));
    CUDA_TRY(cudaMemset(R.ctr, 0, 128));
    CUDA_TRY(cudaMemset(R.d_tokens, 0, (size_t)TOKENS_RING * 4));
    R.dev.rank = rank; R.dev.world = tp->world; R.dev.ars = 2 * tp->m->n_dec + 1;
    R.dev.self = R.comm; R.dev.done = R.ctr; R.dev.epoch = R.ctr + 16;
    R.dev.peer[rank] = R.comm;
    return 0;
}

static vox_cuda_tp *tp_alloc(vox_cuda_model *m, int world, int kv_bits = 32) {
    if (!m || !tp_valid_world(world)) { vox_set_error("tp: world size must be 1, 2, 4 or 8"); return nullptr; }
    if (kv_bits != 32 && kv_bits != 16) { vox_set_error("tp: kv_bits must be 32 or 16"); return nullptr; }
    vox_cuda_tp *tp = (vox_cuda_tp *)calloc(1, sizeof(vox_cuda_tp));
    tp->m = m; tp->world = world;
    tp->s = vox_cuda_state_create_ex(m, kv_bits);
    if (!tp->s) { free(tp); return nullptr; }
    // key splits: enough CTAs to pull a long context at HBM speed, few enough that the merge stays short
    const int kh = DEC_KVH / world;
    tp->att_splits = kh >= 4 ? 16 : TP_ATT_SPLITS;
    if (cudaMalloc((void **)&tp->attn_part, (size_t)DEC_KVH * TP_ATT_SPLITS * 4 * TP_ATT_PART * 4) != cudaSuccess) {
        vox_set_error("tp: attention scratch alloc failed");
        vox_cuda_state_free(tp->s); free(tp); return nullptr;
    }
    return tp;
}

extern "C" void vox_cuda_tp_free(vox_cuda_tp_t *tp) {
    if (!tp) return;
    if (tp->s) cudaStreamSynchronize(tp->s->stream);
    if (tp->graph) cudaGraphExecDestroy(tp->graph);
    tp_mega_destroy(tp->mega);
    for (int i = 0; i < tp->n_local; i++) tp_rank_free(tp->r[i]);
    if (tp->h_check) cudaFreeHost(tp->h_check);
    cudaFree(tp->attn_part);
    vox_cuda_state_free(tp->s);
    free(tp);
}

// One rank of a world of `world` processes (one per GPU).  Peers are attached with export / connect.
extern "C" vox_cuda_tp_t *vox_cuda_tp_create(vox_cuda_model_t *m, int rank, int world) {
    return vox_cuda_tp_create_ex(m, rank, world, 32);
}

extern "C" vox_cuda_tp_t *vox_cuda_tp_create_ex(vox_cuda_model_t *m, int rank, int world, int kv_bits) {
    vox_cuda_tp *tp = tp_alloc(m, world, kv_bits);
    if (!tp) return nullptr;
    if (rank < 0 || rank >= world) { vox_set_error("tp: rank %d outside world %d", rank, world); vox_cuda_tp_free(tp); return nullptr; }
    tp->n_local = 1;
    if (tp_rank_init(tp, tp->r[0], rank)) { vox_cuda_tp_free(tp); return nullptr; }
    tp->connected = (world == 1);
    {   // Two step engines with the same arithmetic and exchange protocol: the kernel-per-op CUDA graph above (default)
        // and one persistent cooperative kernel per rank (dec_tp_mega.cu, VOX_CUDA_TP_MEGA=1).  Measured
        // (profiles/r01_j_tp_decode.md): 1.067 vs 1.077 ms per token at 2 GPUs, 0.810 vs 0.868 at 4 - a grid barrier
        // costs more than a PDL kernel boundary whose successor already runs beside its predecessor.
        const char *env = getenv("VOX_CUDA_TP_MEGA");
        if (env && atoi(env) != 0 && !tp->s->kv16 && !m->q8) tp->mega = tp_mega_create(m, world);   // the persistent engine: fp32 rings, bf16 models
    }
    return tp;
}

// All `world` ranks on this device, advanced kernel by kernel in lock step on one stream: the shard
// arithmetic, the exchange windows and the sequence numbers are those of the multi-GPU run, so one GPU is
// enough to check them (tests/test_gpu_tp.py).  Not a fast path: every kernel is launched `world` times.
extern "C" vox_cuda_tp_t *vox_cuda_tp_create_local(vox_cuda_model_t *m, int world) {
    return vox_cuda_tp_create_local_ex(m, world, 32);
}

extern "C" vox_cuda_tp_t *vox_cuda_tp_create_local_ex(vox_cuda_model_t *m, int world, int kv_bits) {
    vox_cuda_tp *tp = tp_alloc(m, world, kv_bits);
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

static inline int tp_step_kernels(const vox_cuda_tp *tp) { return 1 + 5 * tp->m->n_dec + 3; }

// wo / w2 slices: (rows per warp, 256-column chunks) so that a lane has 32-40 loads of 16 B in flight
template <int NW, int NCH>
static cudaError_t tp_launch_partial(cudaStream_t st, const TpDev &td, int a_idx, const float *v, const uint16_t *W, int K, const float *row_scale) {
    const int ks = K / td.world;
    const size_t smem = (size_t)ks * 4;
    static bool attr = false;
    if (!attr) {
        cudaFuncSetAttribute(tp_gemv_partial_kernel<NW, NCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, 40 * 1024);
        attr = true;
    }
    int grid = (DEC_DIM / NW + TP_WARPS - 1) / TP_WARPS;
    if (grid > 148) grid = 148;     // 170-240 registers per thread: one CTA per SM, warps loop over row batches
    return launch_pdl(tp_gemv_partial_kernel<NW, NCH>, grid, TP_THREADS, smem, st, td, a_idx, v, W, K, row_scale);
}

static cudaError_t tp_partial(cudaStream_t st, const TpDev &td, int a_idx, const float *v, const uint16_t *W, int K, const float *row_scale) {
    const int G = td.world;
    if (K == DEC_QDIM) {
        if (G == 1) return tp_launch_partial<2, 16>(st, td, a_idx, v, W, K, row_scale);
        if (G == 2) return tp_launch_partial<4, 8>(st, td, a_idx, v, W, K, row_scale);
        if (G == 4) return tp_launch_partial<8, 4>(st, td, a_idx, v, W, K, row_scale);
        return tp_launch_partial<8, 2>(st, td, a_idx, v, W, K, row_scale);
    }
    if (G == 1) return tp_launch_partial<1, 36>(st, td, a_idx, v, W, K, row_scale);
    if (G == 2) return tp_launch_partial<2, 18>(st, td, a_idx, v, W, K, row_scale);
    if (G == 4) return tp_launch_partial<4, 9>(st, td, a_idx, v, W, K, row_scale);
    return tp_launch_partial<8, 5>(st, td, a_idx, v, W, K, row_scale);
}

// PDL launch of an exchange consumer, as clusters of TP_CS CTAs when the cluster gather is on (cs > 1)
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl_cs(void (*kernel)(KArgs...), int grid, int cs, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(TP_THREADS); cfg.dynamicSmemBytes = 0; cfg.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    at[1].id = cudaLaunchAttributeClusterDimension;
    at[1].val.clusterDim.x = (unsigned)cs; at[1].val.clusterDim.y = 1; at[1].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = cs > 1 ? 2 : 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
// cluster gather: OFF by default - measured on 8 GPUs it is slower (1.002 vs 0.942 ms per token, profiles/r02_tp8.md): the
// exchange read volume is not what limits 8 GPUs, and the cluster barrier + DSMEM hop add latency to each of the 53
// exchanges.  VOX_CUDA_TP_CLUSTER=1 turns it on for every world > 1 (parity-tested in lock step); grids are rounded down
// to a multiple of TP_CS
static int tp_cluster_size(int world, int grid) {
    static int mode = -2;
    if (mode == -2) { const char *e = getenv("VOX_CUDA_TP_CLUSTER"); mode = e ? (atoi(e) ? 1 : 0) : -1; }
    const bool on = mode == 1 && world > 1;
    return (on && grid >= TP_CS) ? TP_CS : 1;
}

static int tp_grid_cap(void) {
    static int cap = 0;
    if (!cap) { const char *e = getenv("VOX_CUDA_TP_GRID"); cap = e && atoi(e) > 0 ? atoi(e) : 148; }
    return cap;
}
static inline int tp_grid(int rows) {
    int g = (rows / FULL_NW + TP_WARPS - 1) / TP_WARPS;
    return g > tp_grid_cap() ? tp_grid_cap() : g;
}

static int tp_enqueue_step(vox_cuda_tp *tp, cudaStream_t st) {
    vox_cuda_model *m = tp->m;
    vox_cuda_state *s = tp->s;
    const int G = tp->world, NL = tp->n_local, L = m->n_dec;
    const size_t kv_layer = (size_t)DEC_WINDOW * DEC_KVDIM;
    for (int i = 0; i < NL; i++) {
        TpRank &R = tp->r[i];
        dec_launch_embed(st, R.d_params, s->d_embed_in, s->adapter_buf, m->tok_emb, m->q8 ? m->s_tok_emb : nullptr, m->inv_freq_dec, R.xbuf, s->d_rope);
    }
    for (int l = 0; l < L; l++) {
        const DecLayerDev &W = m->dec[l];
        float *kc = s->kv_k + l * kv_layer, *vc = s->kv_v + l * kv_layer;
        __half *kc16 = (__half *)s->kv_k16 + l * kv_layer, *vc16 = (__half *)s->kv_v16 + l * kv_layer;   // fp16 rings (kv16 states)
        for (int i = 0; i < NL; i++) {
            TpRank &R = tp->r[i];
            const int a = 2 * l - 1;              // consumes the w2 exchange of the previous layer
            const DecStepParams *pp = R.d_params;
            if (l == 0) {
                if (s->kv16)
                    CUDA_TRY(launch_pdl(tp_qkv_kernel<false, __half>, tp_grid(DEC_QKV / G), TP_THREADS, 0, st, R.dev, a, pp, (const float *)R.xbuf,
                                        (float *)nullptr, W.attn_norm, W.wqkv, (const float *)s->d_rope, s->d_q, kc16, vc16, 1, W.s_wqkv));
                else
                    CUDA_TRY(launch_pdl(tp_qkv_kernel<false, float>, tp_grid(DEC_QKV / G), TP_THREADS, 0, st, R.dev, a, pp, (const float *)R.xbuf,
                                        (float *)nullptr, W.attn_norm, W.wqkv, (const float *)s->d_rope, s->d_q, kc, vc, 1, W.s_wqkv));
            } else {
                int g = tp_grid(DEC_QKV / G);
                const int cs = tp_cluster_size(G, g);
                g -= g % cs;
                if (s->kv16)
                    CUDA_TRY(launch_pdl_cs(tp_qkv_kernel<true, __half>, g, cs, st, R.dev, a, pp, (const float *)(R.xbuf + DEC_DIM),
                                           R.xbuf, W.attn_norm, W.wqkv, (const float *)s->d_rope, s->d_q, kc16, vc16, cs, W.s_wqkv));
                else
                    CUDA_TRY(launch_pdl_cs(tp_qkv_kernel<true, float>, g, cs, st, R.dev, a, pp, (const float *)(R.xbuf + DEC_DIM),
                                           R.xbuf, W.attn_norm, W.wqkv, (const float *)s->d_rope, s->d_q, kc, vc, cs, W.s_wqkv));
            }
        }
        for (int i = 0; i < NL; i++) {
            TpRank &R = tp->r[i];
            if (s->kv16)
                CUDA_TRY(launch_pdl(tp_attn_kernel<__half>, dim3(DEC_KVH / G, tp->att_splits), TP_THREADS, 0, st, R.dev, (const DecStepParams *)R.d_params,
                                    (const float *)s->d_q, (const __half *)kc16, (const __half *)vc16, tp->attn_part, s->d_attn));
            else
                CUDA_TRY(launch_pdl(tp_attn_kernel<float>, dim3(DEC_KVH / G, tp->att_splits), TP_THREADS, 0, st, R.dev, (const DecStepParams *)R.d_params,
                                    (const float *)s->d_q, (const float *)kc, (const float *)vc, tp->attn_part, s->d_attn));
        }
        for (int i = 0; i < NL; i++) CUDA_TRY(tp_partial(st, tp->r[i].dev, 2 * l, s->d_attn, W.wo, DEC_QDIM, W.s_wo));
        for (int i = 0; i < NL; i++) {
            TpRank &R = tp->r[i];
            int g = tp_grid(2 * DEC_HIDDEN / G);
            const int cs = tp_cluster_size(G, g);
            g -= g % cs;
            CUDA_TRY(launch_pdl_cs(tp_ffn_kernel, g, cs, st, R.dev, 2 * l, (const float *)R.xbuf, R.xbuf + DEC_DIM,
                                   W.ffn_norm, (const float *)(m->ada_scale + (size_t)l * DEC_DIM), W.w13, s->d_h, cs, W.s_w13));
        }
        for (int i = 0; i < NL; i++) CUDA_TRY(tp_partial(st, tp->r[i].dev, 2 * l + 1, s->d_h, W.w2, DEC_HIDDEN, W.s_w2));
    }
    int nblk = (VOCAB / G / FULL_NW + TP_WARPS - 1) / TP_WARPS;      // 100-400 MB: bandwidth-bound, one CTA per SM
    if (nblk > 148) nblk = 148;
    const int lcs = tp_cluster_size(G, nblk);
    nblk -= nblk % lcs;
    for (int i = 0; i < NL; i++) {
        TpRank &R = tp->r[i];
        CUDA_TRY(launch_pdl_cs(tp_logits_kernel, nblk, lcs, st, R.dev, 2 * L - 1, (const float *)(R.xbuf + DEC_DIM), m->dec_norm,
                               m->tok_emb, s->d_logits, R.blockmax, lcs, (const float *)(m->q8 ? m->s_tok_emb : nullptr)));
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
    if (!tp->mega && tp_ensure_graph(tp)) return -1;
    for (int i = 0; i < tp->n_local; i++)
        CUDA_TRY(cudaMemcpyAsync(tp->r[i].d_params, &s->h_params, sizeof(DecStepParams), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaEventRecord(s->ev0, st));
    if (tp->mega) {
        TpRank &R = tp->r[0];
        if (tp_mega_run(tp->mega, st, R.dev, tp->m, s, R.d_params, tp->attn_part, tp->att_splits, R.blockmax, R.d_tokens, n)) return -1;
    } else {
        for (int i = 0; i < n; i++) CUDA_TRY(cudaGraphLaunch(tp->graph, st));
        g_launch_count += (unsigned long long)n * tp_step_kernels(tp) * tp->n_local;
    }
    CUDA_TRY(cudaEventRecord(s->ev1, st));
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
