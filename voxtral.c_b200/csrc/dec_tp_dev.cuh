// dec_tp_dev.cuh - device side of the tensor-parallel decoder shared by the kernel-per-op step (dec_tp.cu) and
// the persistent step kernel (dec_tp_mega.cu): exchange window layout and words, row-batch GEMV, attention task.
#pragma once
#include <cuda_fp16.h>
#include "vox_internal.h"
#include "tc_ptx.cuh"

constexpr int TP_MAX = VOX_CUDA_TP_MAX;   // 8
constexpr int TP_WARPS = 8;
constexpr int TP_THREADS = TP_WARPS * 32;
constexpr int TP_ATT_SPLITS = 32;                          // max key splits per kv head of the attention kernel
constexpr int TP_ATT_PART = DEC_HD + 4;                    // floats per (split, head) partial: acc[128], m, l, pad

// One rank's exchange window.  Same layout on every rank; peers write into it, the owner only reads.
// Every element is an 8-byte word {payload bits, sequence number of the exchange that wrote it}: 8-byte
// aligned stores are single transactions, locally and over NVLink, so a reader that sees the sequence
// number sees the payload.  Two slots alternate (part: parity of the exchange's index in the step - 52 per step, so
// consecutive exchanges always differ; amax: parity of the sequence number): a writer of exchange s can only run
// ahead of a reader by one exchange (it needs every rank's s-1 to get there), so the slot s-2 used is free.
struct TpComm {
    uint2 part[2][TP_MAX][DEC_DIM];   // [exchange index & 1][source rank][row]  partial sums of wo / w2
    uint2 amax[2][TP_MAX][2];         // [sequence parity][source rank] {max logit bits, seq}, {index, seq}
};

struct TpDev {                        // by-value kernel argument
    int rank, world;
    int ars;                          // exchanges per step = 2 * layers + 1
    TpComm *self;
    TpComm *peer[TP_MAX];             // peer[p] = rank p's window as mapped here (peer[rank] == self)
    unsigned *done;                   // done[1 + k]: arrival counter of the attention CTAs of local kv head k
    unsigned *epoch;                  // steps completed since creation (device resident; graphs are replayed)
};

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// two exchange words per load / one per store; volatile = served by L2, where the peers' writes land
__device__ __forceinline__ uint4 ld_ll2(const uint2 *p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_ll(uint2 *p, unsigned payload, unsigned seq) {
    asm volatile("st.volatile.global.v2.u32 [%0], {%1,%2};" :: "l"(p), "r"(payload), "r"(seq) : "memory");
}
// Bounded wait: a peer that died must not leave this GPU spinning (the pool counts a hung box as a strike).
// A bare trap: a printf here is a call, and the registers it makes the compiler save around every polling loop
// cost more than the message is worth (the host reports which call failed).
__device__ __forceinline__ void tp_timeout(int rank, int peer, unsigned seq) {
    (void)rank; (void)peer; (void)seq;
    __trap();
}
// 8 consecutive values of source rank p, exchange seq: v holds a first read; spins until all 8 words carry seq
__device__ __forceinline__ void tp_ll_settle8(const TpDev &td, const uint2 *src, int p, unsigned seq, uint4 (&v)[4]) {
    bool ok = true;
    #pragma unroll
    for (int k = 0; k < 4; k++) ok = ok && v[k].y == seq && v[k].w == seq;
    if (ok) return;
    const unsigned long long t0 = globaltimer_ns();
    for (;;) {
        ok = true;
        #pragma unroll
        for (int k = 0; k < 4; k++) { v[k] = ld_ll2(src + 2 * k); ok = ok && v[k].y == seq && v[k].w == seq; }
        if (ok) return;
        if (globaltimer_ns() - t0 > 8000000000ull) tp_timeout(td.rank, p, seq);
    }
}

// ---- row-batch GEMV: a warp owns NW rows x NCH chunks of 256 columns, all loads issued at once ----------
// x is in shared memory as A[c] = x[8c..8c+3], B[c] = x[8c+4..8c+7]; ks = slice length (ks % 8 == 0,
// ks <= NCH * 256; lanes past the slice load nothing).  W points at the slice start of the batch's first row.
template <int NW, int NCH>
__device__ __forceinline__ void tp_wload(uint4 (&w)[NW][NCH], const uint16_t *__restrict__ W, size_t stride, int ks) {
    const int lane = threadIdx.x & 31;
    #pragma unroll
    for (int j = 0; j < NCH; j++) {
        const int col = j * 256 + lane * 8;
        #pragma unroll
        for (int r = 0; r < NW; r++)
            w[r][j] = col < ks ? ldg_stream(W + (size_t)r * stride + col) : make_uint4(0u, 0u, 0u, 0u);
    }
}
template <int NW, int NCH>
__device__ __forceinline__ void tp_wdot(const uint4 (&w)[NW][NCH], const float4 *xA, const float4 *xB, int ks, float (&acc)[NW]) {
    const int lane = threadIdx.x & 31;
    #pragma unroll
    for (int r = 0; r < NW; r++) acc[r] = 0.f;
    #pragma unroll
    for (int j = 0; j < NCH; j++) {
        const int col = j * 256 + lane * 8;
        if (col < ks) {
            const float4 a = xA[col >> 3], b = xB[col >> 3];
            #pragma unroll
            for (int r = 0; r < NW; r++) acc[r] = dot8(w[r][j], a, b, acc[r]);
        }
    }
    #pragma unroll
    for (int r = 0; r < NW; r++) acc[r] = warp_sum(acc[r]);
}

__device__ __forceinline__ float4 f4add(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ float f4sq(float4 a) { return a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w; }

__device__ __forceinline__ void tp_argmax_merge(float &bv, int &bi, float v, int i) {
    if (v > bv || (v == bv && i < bi)) { bv = v; bi = i; }   // first maximum; NaN never wins
}

struct __align__(16) TpAttnSmem {
    float qs[4][DEC_HD];
    float wsm[TP_WARPS][4][DEC_HD + 4];
    int last;
};

// One (local kv head lk, key split) task of the attention on 256 threads: see tp_attn_kernel in dec_tp.cu.
// q is read through L2 (another CTA of a persistent kernel may have written it).  Ends with the block in sync.
// K/V ring element: float, or __half (vox_cuda_tp_create_ex(.., 16): rows rounded once when appended, widened when read -
// the arrangement of vox_cuda_state_create_ex, dec_mega.cu)
__device__ __forceinline__ float4 tp_ld_kv4(const float *p) { return __ldcg(reinterpret_cast<const float4 *>(p)); }
__device__ __forceinline__ float4 tp_ld_kv4(const __half *p) {
    const uint2 u = __ldcg(reinterpret_cast<const uint2 *>(p));
    const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&u.x));
    const float2 b = __half22float2(*reinterpret_cast<const __half2 *>(&u.y));
    return make_float4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ void tp_st_kv2(float *p, float a, float b) { *reinterpret_cast<float2 *>(p) = make_float2(a, b); }
__device__ __forceinline__ void tp_st_kv2(__half *p, float a, float b) { *reinterpret_cast<__half2 *>(p) = __floats2half2_rn(a, b); }

template <class KV>
__device__ __forceinline__ void tp_attn_task(const TpDev &td, int lk, int split, int ns, int pos, int n_valid,
                                             const float *__restrict__ q, const KV *__restrict__ kc,
                                             const KV *__restrict__ vc, float *__restrict__ part,
                                             float *__restrict__ attn_out, TpAttnSmem &sm) {
    float (*qs)[DEC_HD] = sm.qs;
    float (*wsm)[4][DEC_HD + 4] = sm.wsm;
    int *s_last = &sm.last;
    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
    const int kvh = td.rank * (DEC_KVH / td.world) + lk;
    const int n_keys = min(n_valid, DEC_WINDOW - 1) + 1;
    const int first = pos - n_keys + 1;                  // logical position of key 0
    const int ng = (n_keys + 7) >> 3;
    int gps = (ng + ns - 1) / ns;
    gps = (gps + TP_WARPS - 1) / TP_WARPS * TP_WARPS;    // whole rounds of the 8 warps
    const int n_active = (ng + gps - 1) / gps;           // splits that own keys: 1 at a short context, ns at a full window
    if (split >= n_active) return;                       // whole CTA (uniform): idle splits neither write nor arrive
    const int g0 = split * gps, g1 = min(ng, g0 + gps);
    if (tid < 4 * DEC_HD / 4)
        reinterpret_cast<float4 *>(&qs[0][0])[tid] = __ldcg(reinterpret_cast<const float4 *>(q + (size_t)kvh * 4 * DEC_HD) + tid);
    __syncthreads();
    const float scale = 0.08838834764831845f;            // 1/sqrt(128)
    float4 qv[4];
    #pragma unroll
    for (int h = 0; h < 4; h++) qv[h] = reinterpret_cast<const float4 *>(&qs[h][0])[lane];
    float m = -INFINITY, l = 0.f;                        // lanes 0..3 matter: running max / sum of head (lane & 3)
    float4 acc[4];
    #pragma unroll
    for (int h = 0; h < 4; h++) acc[h] = make_float4(0.f, 0.f, 0.f, 0.f);
    const size_t hoff = (size_t)kvh * DEC_HD + 4 * lane;
    for (int g = g0 + w; g < g1; g += TP_WARPS) {
        const int j0 = g * 8;
        const int nv = min(8, n_keys - j0);
        float4 kf[8], vf[8];
        #pragma unroll
        for (int i = 0; i < 8; i++) {
            const int slot = (first + j0 + min(i, nv - 1)) & (DEC_WINDOW - 1);   // clamped: always a valid row
            kf[i] = tp_ld_kv4(kc + (size_t)slot * DEC_KVDIM + hoff);
            vf[i] = tp_ld_kv4(vc + (size_t)slot * DEC_KVDIM + hoff);
        }
        float v[32];
        #pragma unroll
        for (int i = 0; i < 8; i++)
            #pragma unroll
            for (int h = 0; h < 4; h++)
                v[i * 4 + h] = kf[i].x * qv[h].x + kf[i].y * qv[h].y + kf[i].z * qv[h].z + kf[i].w * qv[h].w;
        #pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {              // score (key i, head h) ends in lane 4i + h
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
        const float pr = valid ? expf(sc - mn) : 0.f;
        float ps = pr;
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
                const float ph = __shfl_sync(0xffffffffu, pr, i * 4 + h);       // 0 for padded keys
                acc[h].x = fmaf(ph, vf[i].x, acc[h].x); acc[h].y = fmaf(ph, vf[i].y, acc[h].y);
                acc[h].z = fmaf(ph, vf[i].z, acc[h].z); acc[h].w = fmaf(ph, vf[i].w, acc[h].w);
            }
        }
    }
    #pragma unroll
    for (int h = 0; h < 4; h++) reinterpret_cast<float4 *>(&wsm[w][h][0])[lane] = acc[h];
    if (lane < 4) { wsm[w][lane][DEC_HD] = m; wsm[w][lane][DEC_HD + 1] = l; }
    __syncthreads();
    const int h = tid >> 6, d = (tid & 63) * 2;          // thread -> (head, two dims)
    {   // merge the 8 warps into this split's partial (relative to its own maximum, not normalised)
        float M = -INFINITY;
        #pragma unroll
        for (int k = 0; k < TP_WARPS; k++) M = fmaxf(M, wsm[k][h][DEC_HD]);
        float n0 = 0.f, n1 = 0.f, den = 0.f;
        #pragma unroll
        for (int k = 0; k < TP_WARPS; k++) {
            const float mk = wsm[k][h][DEC_HD];
            const float e = (mk == -INFINITY) ? 0.f : expf(mk - M);
            n0 = fmaf(e, wsm[k][h][d], n0);
            n1 = fmaf(e, wsm[k][h][d + 1], n1);
            den = fmaf(e, wsm[k][h][DEC_HD + 1], den);
        }
        float *o = part + ((size_t)(kvh * TP_ATT_SPLITS + split) * 4 + h) * TP_ATT_PART;
        *reinterpret_cast<float2 *>(o + d) = make_float2(n0, n1);
        if ((tid & 63) == 0) { o[DEC_HD] = M; o[DEC_HD + 1] = den; }
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const unsigned old = atomicAdd(&td.done[1 + lk], 1u);
        *s_last = (old == (unsigned)n_active - 1u);
        if (*s_last) td.done[1 + lk] = 0;
    }
    __syncthreads();
    if (*s_last) {
        __threadfence();
        const float *base = part + ((size_t)kvh * TP_ATT_SPLITS * 4 + h) * TP_ATT_PART;
        float M = -INFINITY;
        #pragma unroll 4
        for (int s = 0; s < n_active; s++) M = fmaxf(M, __ldcg(base + (size_t)s * 4 * TP_ATT_PART + DEC_HD));
        float n0 = 0.f, n1 = 0.f, den = 0.f;
        #pragma unroll 4
        for (int s = 0; s < n_active; s++) {
            const float *o = base + (size_t)s * 4 * TP_ATT_PART;
            const float ms = __ldcg(o + DEC_HD);
            if (ms == -INFINITY) continue;
            const float e = expf(ms - M);
            const float2 a = __ldcg(reinterpret_cast<const float2 *>(o + d));
            n0 = fmaf(e, a.x, n0); n1 = fmaf(e, a.y, n1);
            den = fmaf(e, __ldcg(o + DEC_HD + 1), den);
        }
        const float inv = den > 0.f ? 1.0f / den : 0.f;
        *reinterpret_cast<float2 *>(attn_out + (size_t)(kvh * 4 + h) * DEC_HD + d) = make_float2(n0 * inv, n1 * inv);
    }
    __syncthreads();                                     // shared memory is free for the next task
}

// dec_tp_mega.cu: the step as one persistent cooperative kernel per rank (default when a rank has the GPU to itself)
void *tp_mega_create(const vox_cuda_model *m, int world);
void tp_mega_destroy(void *h);
int tp_mega_run(void *h, cudaStream_t st, const TpDev &td, vox_cuda_model *m, vox_cuda_state *s, DecStepParams *d_params,
                float *attn_part, int att_splits, float *blockmax, int *d_tokens, int n);
