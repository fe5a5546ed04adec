// attn_tc.cu - sliding-window causal attention of the encoder on tensor cores.
//
// vox_causal_attention as the encoder calls it (/root/reference/voxtral_kernels.c:541-611,
// voxtral_encoder.c:622-624): 32 heads, head_dim 64, query at logical position g attends
// to keys [max(first, g - 749), g] of the KV ring.  Flash-style: a CTA owns 64 queries of
// one head (4 warps x 16 rows), walks the <= 13 key tiles of 64 keys, keeps the running
// max / sum / output in registers.  S = Q K^T and O = P V run on mma.sync.m16n8k16 bf16
// with fp32 accumulation; fp32 operands are split into two bf16 terms (x = hi + lo) and
// the three significant products hi*hi + hi*lo + lo*hi are accumulated (TERMS = 3), which
// keeps fp32-class accuracy; TERMS = 1 is plain bf16.  K/V tiles are converted from the
// fp32 ring on their way into shared memory.
//
// (The projections around it use tcgen05/TMEM, gemm_tc.cu; moving this kernel to
// tcgen05 with P kept in TMEM is the follow-up - attention is < 10 % of encoder FLOPs.)
#include "vox_internal.h"
#include "enc_state.h"
#include "rows.h"
#include "gemm_tc.h"
#include "tc_ptx.cuh"
#include <stdlib.h>

namespace at {

constexpr int BQ = 64, BKEY = 64, HD = ENC_HD, LDS = 72;    // smem row stride in bf16 (144 B: conflict-free ldmatrix)
constexpr int NT = 128;
constexpr int TILE = BQ * LDS;                               // elements per smem tile
constexpr int SMEM_BYTES = 6 * TILE * 2;                     // Q, K, V x (hi, lo)

__device__ __forceinline__ void ldsm4(uint32_t (&r)[4], const void *p) {
    const uint32_t a = (uint32_t)__cvta_generic_to_shared(p);
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
}
__device__ __forceinline__ void ldsm4t(uint32_t (&r)[4], const void *p) {
    const uint32_t a = (uint32_t)__cvta_generic_to_shared(p);
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
}
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack2(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&v);
}
// split two floats into packed bf16 hi and lo
__device__ __forceinline__ void split2(float a, float b, uint32_t &hi, uint32_t &lo) {
    hi = pack2(a, b);
    lo = pack2(a - bf16_lo(hi), b - bf16_hi(hi));
}

// Load a 64 x 64 fp32 tile (row r at src(r), or zeros if src(r) == nullptr), scale, split, store to smem.
template <typename RowPtr>
__device__ __forceinline__ void load_tile(uint16_t *s_hi, uint16_t *s_lo, float scale, RowPtr rowptr) {
    const int c4 = threadIdx.x & 15, r0 = threadIdx.x >> 4;
    #pragma unroll
    for (int pass = 0; pass < 8; pass++) {
        const int r = pass * 8 + r0;
        const float *src = rowptr(r);
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (src) v = *reinterpret_cast<const float4 *>(src + c4 * 4);
        uint32_t h0, l0, h1, l1;
        split2(v.x * scale, v.y * scale, h0, l0);
        split2(v.z * scale, v.w * scale, h1, l1);
        *reinterpret_cast<uint2 *>(s_hi + r * LDS + c4 * 4) = make_uint2(h0, h1);
        *reinterpret_cast<uint2 *>(s_lo + r * LDS + c4 * 4) = make_uint2(l0, l1);
    }
}

template <int TERMS>
__global__ void __launch_bounds__(NT) enc_attn_kernel(
    const float *__restrict__ qkv, long ldq, const float *__restrict__ kc, const float *__restrict__ vc,
    int M, int pos0, int first_valid, int window, int ring_mask, uint16_t *__restrict__ out, long ldo, int split,
    const int *__restrict__ pos_dev) {
    extern __shared__ __align__(16) uint16_t sm[];
    tcx::pdl_launch_dependents();
    tcx::pdl_wait();
    if (pos_dev) { pos0 = pos_dev[0]; first_valid = pos_dev[1]; }   // replayed CUDA graphs keep {position, oldest valid key} in device memory
    uint16_t *Qh = sm, *Ql = sm + TILE, *Kh = sm + 2 * TILE, *Kl = sm + 3 * TILE, *Vh = sm + 4 * TILE, *Vl = sm + 5 * TILE;
    const int head = blockIdx.y, q0 = blockIdx.x * BQ;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;

    // Q tile, pre-scaled by 1/sqrt(64) (exact power of two)
    load_tile(Qh, Ql, 0.125f, [&](int r) -> const float * {
        return (q0 + r < M) ? qkv + (size_t)(q0 + r) * ldq + head * HD : nullptr;
    });
    __syncthreads();
    uint32_t qh[4][4], ql[4][4];
    {
        const int mi = lane >> 3, r = lane & 7;
        #pragma unroll
        for (int kk = 0; kk < 4; kk++) {
            const int off = (warp * 16 + 8 * (mi & 1) + r) * LDS + kk * 16 + 8 * (mi >> 1);
            ldsm4(qh[kk], Qh + off);
            if (TERMS > 1) ldsm4(ql[kk], Ql + off);
        }
    }
    const int gq0 = pos0 + q0 + warp * 16 + g, gq1 = gq0 + 8;      // logical positions of this thread's two rows
    const int gmin = pos0 + q0, gmax = pos0 + min(q0 + BQ - 1, M - 1);
    int k_lo = gmin - (window - 1);
    if (k_lo < first_valid) k_lo = first_valid;

    float o[8][4];
    #pragma unroll
    for (int n = 0; n < 8; n++) { o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f; }
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;

    for (int kt = k_lo; kt <= gmax; kt += BKEY) {
        __syncthreads();                                   // previous tile fully consumed
        load_tile(Kh, Kl, 1.0f, [&](int r) -> const float * {
            const int j = kt + r;
            return (j <= gmax) ? kc + (size_t)(j & ring_mask) * ENC_QKV + head * HD : nullptr;
        });
        load_tile(Vh, Vl, 1.0f, [&](int r) -> const float * {
            const int j = kt + r;
            return (j <= gmax) ? vc + (size_t)(j & ring_mask) * ENC_QKV + head * HD : nullptr;
        });
        __syncthreads();

        // ---- S = Q K^T (16 x 64 per warp) ----
        float s[8][4];
        #pragma unroll
        for (int n = 0; n < 8; n++) { s[n][0] = s[n][1] = s[n][2] = s[n][3] = 0.f; }
        {
            const int mi = lane >> 3, r = lane & 7;
            #pragma unroll
            for (int n = 0; n < 8; n++) {
                #pragma unroll
                for (int kp = 0; kp < 2; kp++) {
                    const int off = (n * 8 + r) * LDS + kp * 32 + 8 * mi;
                    uint32_t bh[4], bl[4];
                    ldsm4(bh, Kh + off);
                    if (TERMS > 1) {
                        ldsm4(bl, Kl + off);
                        mma16816(s[n], ql[2 * kp], bh[0], bh[1]);
                        mma16816(s[n], ql[2 * kp + 1], bh[2], bh[3]);
                        mma16816(s[n], qh[2 * kp], bl[0], bl[1]);
                        mma16816(s[n], qh[2 * kp + 1], bl[2], bl[3]);
                    }
                    mma16816(s[n], qh[2 * kp], bh[0], bh[1]);
                    mma16816(s[n], qh[2 * kp + 1], bh[2], bh[3]);
                }
            }
        }
        // ---- mask + online softmax ----
        float mx0 = -INFINITY, mx1 = -INFINITY;
        #pragma unroll
        for (int n = 0; n < 8; n++) {
            #pragma unroll
            for (int e = 0; e < 2; e++) {
                const int jk = kt + n * 8 + 2 * t + e;
                const bool v0 = jk <= gq0 && jk > gq0 - window && jk >= first_valid;
                const bool v1 = jk <= gq1 && jk > gq1 - window && jk >= first_valid;
                s[n][e] = v0 ? s[n][e] : -INFINITY;
                s[n][2 + e] = v1 ? s[n][2 + e] : -INFINITY;
                mx0 = fmaxf(mx0, s[n][e]);
                mx1 = fmaxf(mx1, s[n][2 + e]);
            }
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
        const float mu0 = (mn0 == -INFINITY) ? 0.f : mn0, mu1 = (mn1 == -INFINITY) ? 0.f : mn1;
        const float c0 = (m0 == -INFINITY) ? 0.f : expf(m0 - mu0), c1 = (m1 == -INFINITY) ? 0.f : expf(m1 - mu1);
        m0 = mn0; m1 = mn1;
        l0 *= c0; l1 *= c1;
        #pragma unroll
        for (int n = 0; n < 8; n++) { o[n][0] *= c0; o[n][1] *= c0; o[n][2] *= c1; o[n][3] *= c1; }
        uint32_t ph[4][4], pl[4][4];
        #pragma unroll
        for (int n = 0; n < 8; n++) {
            const float p0 = expf(s[n][0] - mu0), p1 = expf(s[n][1] - mu0);
            const float p2 = expf(s[n][2] - mu1), p3 = expf(s[n][3] - mu1);
            l0 += p0 + p1; l1 += p2 + p3;
            // C fragment of n-tiles (2k, 2k+1) = A fragment of key step k
            split2(p0, p1, ph[n >> 1][(n & 1) * 2], pl[n >> 1][(n & 1) * 2]);
            split2(p2, p3, ph[n >> 1][(n & 1) * 2 + 1], pl[n >> 1][(n & 1) * 2 + 1]);
        }
        // ---- O += P V ----
        {
            const int mi = lane >> 3, r = lane & 7;
            #pragma unroll
            for (int kk = 0; kk < 4; kk++) {
                #pragma unroll
                for (int dp = 0; dp < 4; dp++) {
                    const int off = (kk * 16 + 8 * (mi & 1) + r) * LDS + dp * 16 + 8 * (mi >> 1);
                    uint32_t bh[4], bl[4];
                    ldsm4t(bh, Vh + off);
                    if (TERMS > 1) {
                        ldsm4t(bl, Vl + off);
                        mma16816(o[2 * dp], pl[kk], bh[0], bh[1]);
                        mma16816(o[2 * dp + 1], pl[kk], bh[2], bh[3]);
                        mma16816(o[2 * dp], ph[kk], bl[0], bl[1]);
                        mma16816(o[2 * dp + 1], ph[kk], bl[2], bl[3]);
                    }
                    mma16816(o[2 * dp], ph[kk], bh[0], bh[1]);
                    mma16816(o[2 * dp + 1], ph[kk], bh[2], bh[3]);
                }
            }
        }
    }
    // ---- normalise and store [hi | lo] ----
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float i0 = l0 > 0.f ? 1.0f / l0 : 0.f, i1 = l1 > 0.f ? 1.0f / l1 : 0.f;
    const int row0 = q0 + warp * 16 + g, row1 = row0 + 8;
    #pragma unroll
    for (int n = 0; n < 8; n++) {
        const int col = head * HD + n * 8 + 2 * t;
        uint32_t h, l;
        if (row0 < M) {
            split2(o[n][0] * i0, o[n][1] * i0, h, l);
            *reinterpret_cast<uint32_t *>(out + (size_t)row0 * ldo + col) = h;
            if (split) *reinterpret_cast<uint32_t *>(out + (size_t)row0 * ldo + ENC_QKV + col) = l;
        }
        if (row1 < M) {
            split2(o[n][2] * i1, o[n][3] * i1, h, l);
            *reinterpret_cast<uint32_t *>(out + (size_t)row1 * ldo + col) = h;
            if (split) *reinterpret_cast<uint32_t *>(out + (size_t)row1 * ldo + ENC_QKV + col) = l;
        }
    }
}

}  // namespace at

// ======================================================================================================
// tcgen05 version (round 2): S = Q K^T and O_tile = P V on the 5th-generation tensor cores, accumulators in TMEM.
//
//   CTA = 128 queries of one head, 256 threads; threads t and t + 128 share query row t & 127 (= TMEM lane), each takes 32 of
//   the 64 key columns of a tile and 32 of the 64 output dims: the online softmax needs no shuffles, only one row-maximum
//   exchange through shared memory that rides on a barrier the pipeline has anyway.  Key tiles of 64 keys.  Shared memory holds K-major SWIZZLE_128B operand tiles of 64 bf16
//   columns (128-byte rows, 8-row groups 1024 B apart - the layout TMA would produce; here the tiles are written by the
//   threads, because K / V live in the fp32 ring and are split into hi + lo bf16 terms on the way in):
//       Q  [128 q   x 64 d]   hi, lo   (A operand of S)        K  [64 keys x 64 d]    hi, lo  (B operand of S)
//       P  [128 q   x 64 key] hi, lo   (A operand of P V)      Vt [64 d    x 64 keys] hi, lo  (B operand of P V: V transposed)
//   TMEM: columns 0..63 / 64..127 = two S tiles, 128..191 = O tile (fp32).  Software pipeline per key tile t: tcgen05.ld S(t);
//   K(t+1) -> smem and S(t+1) = Q K^T (hi*hi + lo*hi + hi*lo) issued into the other S buffer; mask + online softmax of tile t in
//   registers while it runs; wait for P V of tile t-1 (it had the whole softmax to finish), o = o * corr + O tile; P(t), Vt(t) ->
//   smem, P V(t) issued (fresh accumulator).  K / V rows of the next tiles travel in registers.  96 KB of smem, 256 TMEM
//   columns: two CTAs per SM.  Same arithmetic contract as enc_attn_kernel<3> (fp32-class products, fp32 softmax).
namespace at5 {

using namespace tcx;

constexpr int BQ = 128, BK = 64, HD = ENC_HD, NT = 256;         // two threads per query row: 32 key columns / 32 output dims each
constexpr int Q_BYTES = BQ * 128, K_BYTES = BK * 128;              // one bf16 plane of a tile
constexpr int OFF_QH = 0, OFF_QL = Q_BYTES, OFF_KH = 2 * Q_BYTES, OFF_KL = OFF_KH + K_BYTES, OFF_VH = OFF_KL + K_BYTES,
              OFF_VL = OFF_VH + K_BYTES, OFF_PH = OFF_VL + K_BYTES, OFF_PL = OFF_PH + Q_BYTES, OFF_END = OFF_PL + Q_BYTES;
constexpr int OFF_XM = OFF_END + 64;                                // row maxima exchanged between the two threads of a row: [2][128] floats
constexpr int SMEM_BYTES = OFF_XM + 2 * BQ * 4 + 1024;             // + alignment slack
constexpr uint32_t IDESC = umma_idesc(128, 64);
constexpr uint32_t TMEM_COLS = 256;                                // two S buffers (64 + 64) + O tile (64), power of two

// byte offset of the 16-byte chunk c (8 bf16 columns) of row r inside a SWIZZLE_128B K-major plane
__device__ __forceinline__ uint32_t sw_off(int r, int c) { return (uint32_t)(r * 128 + ((c ^ (r & 7)) << 4)); }

__device__ __forceinline__ uint32_t pk(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&v);
}
__device__ __forceinline__ void split_pk(float a, float b, uint32_t &hi, uint32_t &lo) {
    hi = pk(a, b);
    lo = pk(a - bf16_lo(hi), b - bf16_hi(hi));
}
// 8 consecutive floats -> one 16-byte chunk of the hi plane and one of the lo plane
__device__ __forceinline__ void store_chunk(unsigned char *ph, unsigned char *pl, int r, int c, const float4 &a, const float4 &b, float scale) {
    uint4 h, l;
    split_pk(a.x * scale, a.y * scale, h.x, l.x); split_pk(a.z * scale, a.w * scale, h.y, l.y);
    split_pk(b.x * scale, b.y * scale, h.z, l.z); split_pk(b.z * scale, b.w * scale, h.w, l.w);
    *reinterpret_cast<uint4 *>(ph + sw_off(r, c)) = h;
    *reinterpret_cast<uint4 *>(pl + sw_off(r, c)) = l;
}

// K tile (256 threads): thread t stages dims (t & 3) * 16 .. +15 of key (t >> 2): four threads read one 256-byte row.
// Load and store are separate so that the loads of the next tile are in flight while the current one is worked on.
__device__ __forceinline__ void load_k(float4 (&v)[4], const float *__restrict__ kc, int head, int kt, int gmax, int ring_mask) {
    const int key = threadIdx.x >> 2, q4 = threadIdx.x & 3, j = kt + key;
    if (j <= gmax) {
        const float4 *src = reinterpret_cast<const float4 *>(kc + (size_t)(j & ring_mask) * ENC_QKV + head * HD + q4 * 16);
        #pragma unroll
        for (int i = 0; i < 4; i++) v[i] = __ldg(src + i);
    } else {
        #pragma unroll
        for (int i = 0; i < 4; i++) v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
}
__device__ __forceinline__ void store_k(unsigned char *kh, unsigned char *kl, const float4 (&v)[4]) {
    const int key = threadIdx.x >> 2, q4 = threadIdx.x & 3;
    #pragma unroll
    for (int c = 0; c < 2; c++) store_chunk(kh, kl, key, q4 * 2 + c, v[2 * c], v[2 * c + 1], 1.0f);
}

// V tile, transposed (256 threads): thread t owns keys 2 * (t >> 3), +1 and dims i * 32 + (t & 7) * 4 .. +3 (i < 2); it writes, for
// each of its 8 dims, one 4-byte word = that dim of its two keys into row (dim) of Vt
__device__ __forceinline__ void load_v(float4 (&a)[2], float4 (&b)[2], const float *__restrict__ vc, int head, int kt, int gmax, int ring_mask) {
    const int kp = threadIdx.x >> 3, s8 = threadIdx.x & 7;
    const int j0 = kt + 2 * kp, j1 = j0 + 1;
    #pragma unroll
    for (int i = 0; i < 2; i++) {
        a[i] = (j0 <= gmax) ? __ldg(reinterpret_cast<const float4 *>(vc + (size_t)(j0 & ring_mask) * ENC_QKV + head * HD + i * 32 + s8 * 4))
                            : make_float4(0.f, 0.f, 0.f, 0.f);
        b[i] = (j1 <= gmax) ? __ldg(reinterpret_cast<const float4 *>(vc + (size_t)(j1 & ring_mask) * ENC_QKV + head * HD + i * 32 + s8 * 4))
                            : make_float4(0.f, 0.f, 0.f, 0.f);
    }
}
__device__ __forceinline__ void store_vt(unsigned char *vh, unsigned char *vl, const float4 (&a)[2], const float4 (&b)[2]) {
    const int kp = threadIdx.x >> 3, s8 = threadIdx.x & 7;
    const int col = 2 * kp;                                        // key column of the pair inside the 64-key row
    #pragma unroll
    for (int i = 0; i < 2; i++) {
        const float av[4] = {a[i].x, a[i].y, a[i].z, a[i].w}, bv[4] = {b[i].x, b[i].y, b[i].z, b[i].w};
        #pragma unroll
        for (int e = 0; e < 4; e++) {
            const int d = i * 32 + s8 * 4 + e;
            uint32_t h, l;
            split_pk(av[e], bv[e], h, l);
            const uint32_t off = sw_off(d, col >> 3) + (uint32_t)((col & 7) * 2);
            *reinterpret_cast<uint32_t *>(vh + off) = h;
            *reinterpret_cast<uint32_t *>(vl + off) = l;
        }
    }
}
__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

template <int TERMS>
__global__ void __launch_bounds__(NT, 2) enc_attn_tc5_kernel(
    const float *__restrict__ qkv, long ldq, const float *__restrict__ kc, const float *__restrict__ vc,
    int M, int pos0, int first_valid, int window, int ring_mask, uint16_t *__restrict__ out, long ldo, int split,
    const int *__restrict__ pos_dev, float *__restrict__ part, unsigned *__restrict__ cnt, int n_splits) {
    // n_splits == 0: blockIdx.x = tile of 128 queries, the CTA walks all its key tiles.
    // n_splits  > 0 (streaming chunks, M <= 128: ONE query tile against a full window would be 32 CTAs walking 13 key tiles
    // one after the other): blockIdx.x = key split; every CTA takes its share of the key tiles and leaves an unnormalised
    // partial (o, m, l) in `part`; the CTA that arrives last for a head (cnt[head]) merges the partials and writes the rows.
    extern __shared__ unsigned char smraw[];
    unsigned char *sm = reinterpret_cast<unsigned char *>(((uintptr_t)smraw + 1023) & ~(uintptr_t)1023);   // swizzle atoms are 1024-byte aligned
    uint64_t *bar_s = reinterpret_cast<uint64_t *>(sm + OFF_END), *bar_o = bar_s + 1;      // S products done / P V products done
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(sm + OFF_END + 16);
    int *s_last = reinterpret_cast<int *>(sm + OFF_END + 32);
    float *xm = reinterpret_cast<float *>(sm + OFF_XM);                                    // [2][128]
    pdl_launch_dependents();
    pdl_wait();
    if (pos_dev) { pos0 = pos_dev[0]; first_valid = pos_dev[1]; }
    // thread t: query row r = t & 127 (= TMEM lane; warp w reads the lane quarter w % 4), half hf = t >> 7: key columns
    // hf * 32 .. +31 of every S tile, output dims hf * 32 .. +31
    const int tid = threadIdx.x, warp = tid >> 5, r = tid & (BQ - 1), hf = tid >> 7;
    const int head = blockIdx.y, q0 = n_splits ? 0 : blockIdx.x * BQ;
    const int gq = pos0 + q0 + r;                                                    // logical position of this thread's query
    const int gmin = pos0 + q0, gmax = pos0 + min(q0 + BQ - 1, M - 1);
    int k_lo = gmin - (window - 1);
    if (k_lo < first_valid) k_lo = first_valid;
    int k_end = gmax, n_active = 1;                                                  // last key of this CTA's share; splits that have keys
    if (n_splits) {
        const int n_tiles = (gmax - k_lo) / BK + 1, tps = (n_tiles + n_splits - 1) / n_splits;
        const int t0 = blockIdx.x * tps;
        if (t0 >= n_tiles) return;                                                   // whole CTA: this split has no keys (it does not arrive either)
        n_active = (n_tiles + tps - 1) / tps;
        k_end = min(gmax, k_lo + (t0 + tps) * BK - 1);
        k_lo += t0 * BK;
    }

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(tmem_slot)), "n"(TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (tid == 0) {
        mb_init(bar_s, 1);
        mb_init(bar_o, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // Q tile: this thread's half row, pre-scaled by 1/sqrt(64) (an exact power of two)
    {
        const int row = q0 + r;
        float4 v[8];
        if (row < M) {
            const float4 *src = reinterpret_cast<const float4 *>(qkv + (size_t)row * ldq + head * HD + hf * 32);
            #pragma unroll
            for (int i = 0; i < 8; i++) v[i] = __ldg(src + i);
        } else {
            #pragma unroll
            for (int i = 0; i < 8; i++) v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        #pragma unroll
        for (int c = 0; c < 4; c++) store_chunk(sm + OFF_QH, sm + OFF_QL, r, hf * 4 + c, v[2 * c], v[2 * c + 1], 0.125f);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint32_t t_lane = tmem + ((uint32_t)((warp & 3) * 32) << 16);             // this warp's 32 lanes
    const uint32_t t_s = t_lane + (uint32_t)(hf * 32), t_o = t_lane + 128 + (uint32_t)(hf * 32);   // S buffers at columns 0 / 64, O at 128

    float o[32];
    #pragma unroll
    for (int i = 0; i < 32; i++) o[i] = 0.f;
    float m = -INFINITY, l = 0.f;                                                    // l: this thread's 32 key columns only
    const uint32_t a_q[2] = {s32(sm + OFF_QH), s32(sm + OFF_QL)}, a_k[2] = {s32(sm + OFF_KH), s32(sm + OFF_KL)};
    const uint32_t a_p[2] = {s32(sm + OFF_PH), s32(sm + OFF_PL)}, a_v[2] = {s32(sm + OFF_VH), s32(sm + OFF_VL)};
    constexpr float LOG2E = 1.4426950408889634f;

    // S = Q K^T of one key tile into S buffer `buf` (TMEM columns buf * 64 ..): hi*hi, lo*hi, hi*lo
    auto issue_s = [&](int buf) {
        tc_fence_after();
        #pragma unroll
        for (int term = 0; term < (TERMS > 1 ? 3 : 1); term++) {
            const uint64_t da = umma_desc(a_q[term == 1 ? 1 : 0]), db = umma_desc(a_k[term == 2 ? 1 : 0]);
            #pragma unroll
            for (int k = 0; k < HD / 16; k++)
                umma_f16(tmem + (uint32_t)(buf * 64), da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), IDESC, (term | k) ? 1u : 0u);
        }
        umma_commit(bar_s);
    };

    // Software pipeline over the key tiles: while the threads do the softmax of tile t, the tensor core computes S of tile
    // t+1 (second S buffer) and finishes P V of tile t-1; the K / V rows of the next tile travel in registers.
    float4 kreg[4], va[2], vb[2];
    load_k(kreg, kc, head, k_lo, gmax, ring_mask);
    load_v(va, vb, vc, head, k_lo, gmax, ring_mask);
    store_k(sm + OFF_KH, sm + OFF_KL, kreg);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy stores -> visible to the tensor core's reads
    tc_fence_before();
    __syncthreads();
    if (tid == 0) issue_s(0);
    if (k_lo + BK <= k_end) load_k(kreg, kc, head, k_lo + BK, gmax, ring_mask);

    unsigned ph_s = 0, ph_o = 0;
    float corr_prev = 0.f;
    int buf = 0;
    for (int kt = k_lo; kt <= k_end; kt += BK, buf ^= 1) {
        const bool has_next = kt + BK <= k_end, first = kt == k_lo;
        // ---- this thread's 32 columns of the S row -> registers; mask (boundary tiles only); local maximum ----
        mb_wait(bar_s, ph_s); ph_s ^= 1u;
        tc_fence_after();
        uint32_t sr[32];
        tmem_ld32(t_s + (uint32_t)(buf * 64), sr);
        const bool interior = kt + BK - 1 <= gmin && kt > gmax - window && kt >= first_valid;    // uniform over the CTA
        float mx = -INFINITY;
        if (interior) {
            #pragma unroll
            for (int i = 0; i < 32; i++) mx = fmaxf(mx, __uint_as_float(sr[i]));
        } else {
            #pragma unroll
            for (int i = 0; i < 32; i++) {
                const int jk = kt + hf * 32 + i;
                const bool ok = jk <= gq && jk > gq - window && jk >= first_valid;
                const float v = ok ? __uint_as_float(sr[i]) : -INFINITY;
                sr[i] = __float_as_uint(v);
                mx = fmaxf(mx, v);
            }
        }
        xm[hf * BQ + r] = mx;                                           // for the other thread of this row
        // ---- next K tile -> smem (S of this tile has retired, so the K planes are free), S(t+1) starts now ----
        if (has_next) {
            store_k(sm + OFF_KH, sm + OFF_KL, kreg);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        }
        tc_fence_before();
        __syncthreads();
        if (has_next && tid == 0) issue_s(buf ^ 1);
        // ---- online softmax in registers; both threads of a row use the same maximum ----
        mx = fmaxf(mx, xm[(hf ^ 1) * BQ + r]);
        const float mn = fmaxf(m, mx);
        const float mu = (mn == -INFINITY) ? 0.f : mn;
        const float corr = (m == -INFINITY) ? 0.f : ex2((m - mu) * LOG2E);
        const float mu2 = mu * LOG2E;
        m = mn;
        float ps = 0.f;
        #pragma unroll
        for (int i = 0; i < 32; i++) {
            const float pv = ex2(fmaf(__uint_as_float(sr[i]), LOG2E, -mu2));     // 2^-inf = 0 for masked keys
            sr[i] = __float_as_uint(pv);
            ps += pv;
        }
        l = l * corr + ps;
        // ---- P V of the previous tile has had the whole softmax to finish: o = o * corr + O tile ----
        if (!first) {
            mb_wait(bar_o, ph_o); ph_o ^= 1u;
            tc_fence_after();
            uint32_t rr[32];
            tmem_ld32(t_o, rr);
            #pragma unroll
            for (int i = 0; i < 32; i++) o[i] = fmaf(o[i], corr_prev, __uint_as_float(rr[i]));
        }
        corr_prev = corr;
        // ---- P and the transposed V tile -> smem (their planes are free: P V of the previous tile has retired) ----
        #pragma unroll
        for (int c = 0; c < 4; c++) {
            uint4 h, lo4;
            #define PV_(e) __uint_as_float(sr[c * 8 + (e)])
            split_pk(PV_(0), PV_(1), h.x, lo4.x); split_pk(PV_(2), PV_(3), h.y, lo4.y);
            split_pk(PV_(4), PV_(5), h.z, lo4.z); split_pk(PV_(6), PV_(7), h.w, lo4.w);
            #undef PV_
            *reinterpret_cast<uint4 *>(sm + OFF_PH + sw_off(r, hf * 4 + c)) = h;
            if (TERMS > 1) *reinterpret_cast<uint4 *>(sm + OFF_PL + sw_off(r, hf * 4 + c)) = lo4;
        }
        store_vt(sm + OFF_VH, sm + OFF_VL, va, vb);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        tc_fence_before();
        __syncthreads();
        if (tid == 0) {
            tc_fence_after();
            #pragma unroll
            for (int term = 0; term < (TERMS > 1 ? 3 : 1); term++) {
                const uint64_t da = umma_desc(a_p[term == 1 ? 1 : 0]), db = umma_desc(a_v[term == 2 ? 1 : 0]);
                #pragma unroll
                for (int k = 0; k < BK / 16; k++) umma_f16(tmem + 128, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), IDESC, (term | k) ? 1u : 0u);
            }
            umma_commit(bar_o);
        }
        // ---- rows of the tiles after next: in flight during the next iteration ----
        if (has_next) {
            load_v(va, vb, vc, head, kt + BK, gmax, ring_mask);
            if (kt + 2 * BK <= k_end) load_k(kreg, kc, head, kt + 2 * BK, gmax, ring_mask);
        }
    }
    {   // last tile's P V
        mb_wait(bar_o, ph_o);
        tc_fence_after();
        uint32_t rr[32];
        tmem_ld32(t_o, rr);
        #pragma unroll
        for (int i = 0; i < 32; i++) o[i] = fmaf(o[i], corr_prev, __uint_as_float(rr[i]));
    }
    // ---- the row sum is the sum of the two threads' halves (same reference maximum m) ----
    xm[hf * BQ + r] = l;
    __syncthreads();
    l += xm[(hf ^ 1) * BQ + r];

    const int row = q0 + r;
    if (n_splits && n_active > 1) {
        // ---- key-split mode: leave the partial, the last CTA of this head merges ----
        constexpr int PL = HD + 2;                                                   // o[64], m, l - stored [item][row]: coalesced over threads
        float *mine = part + ((size_t)head * n_splits + blockIdx.x) * PL * BQ;
        if (row < M) {
            #pragma unroll
            for (int i = 0; i < 32; i++) mine[(hf * 32 + i) * BQ + r] = o[i];
            if (hf == 0) { mine[HD * BQ + r] = m; mine[(HD + 1) * BQ + r] = l; }
        }
        __threadfence();
        __syncthreads();
        if (tid == 0) {
            const unsigned old = atomicAdd(cnt + head, 1u);
            *s_last = (old + 1u == (unsigned)n_active) ? 1 : 0;
            if (*s_last) cnt[head] = 0u;                                             // clean for the next launch
        }
        __syncthreads();
        if (*s_last && row < M) {
            __threadfence();
            const float *base = part + (size_t)head * n_splits * PL * BQ;
            float mm = -INFINITY;
            for (int sp = 0; sp < n_active; sp++) mm = fmaxf(mm, __ldcg(base + (size_t)sp * PL * BQ + HD * BQ + r));
            #pragma unroll
            for (int i = 0; i < 32; i++) o[i] = 0.f;
            l = 0.f;
            for (int sp = 0; sp < n_active; sp++) {
                const float *ps_ = base + (size_t)sp * PL * BQ;
                const float ms = __ldcg(ps_ + HD * BQ + r);
                const float e = (ms == -INFINITY) ? 0.f : ex2((ms - mm) * LOG2E);
                l = fmaf(e, __ldcg(ps_ + (HD + 1) * BQ + r), l);
                #pragma unroll
                for (int i = 0; i < 32; i++) o[i] = fmaf(e, __ldcg(ps_ + (hf * 32 + i) * BQ + r), o[i]);
            }
        }
        if (!*s_last) l = -1.f;                                                      // not the merger: nothing to store
    }
    // ---- normalise, split, store this thread's 32 dims of the row ----
    if (row < M && l >= 0.f) {
        const float inv = l > 0.f ? 1.0f / l : 0.f;
        uint16_t *oh = out + (size_t)row * ldo + head * HD + hf * 32;
        #pragma unroll
        for (int c = 0; c < 4; c++) {
            uint4 h, lo4;
            split_pk(o[8 * c] * inv, o[8 * c + 1] * inv, h.x, lo4.x); split_pk(o[8 * c + 2] * inv, o[8 * c + 3] * inv, h.y, lo4.y);
            split_pk(o[8 * c + 4] * inv, o[8 * c + 5] * inv, h.z, lo4.z); split_pk(o[8 * c + 6] * inv, o[8 * c + 7] * inv, h.w, lo4.w);
            *reinterpret_cast<uint4 *>(oh + 8 * c) = h;
            if (split) *reinterpret_cast<uint4 *>(oh + ENC_QKV + 8 * c) = lo4;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TMEM_COLS));
    }
}

}  // namespace at5
// 0: tcgen05 kernel (default), 1: the mma.sync kernel (VOX_CUDA_ENC_ATTN_MMA=1, kept as the cross-check)
static int attn_mma_mode(void) {
    static int v = -1;
    if (v < 0) {
        const char *e = getenv("VOX_CUDA_ENC_ATTN_MMA");
        v = (e && atoi(e)) ? 1 : 0;
        cudaFuncSetAttribute(at5::enc_attn_tc5_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, at5::SMEM_BYTES);
        cudaFuncSetAttribute(at5::enc_attn_tc5_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, at5::SMEM_BYTES);
        cudaFuncSetAttribute(at::enc_attn_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, at::SMEM_BYTES);
        cudaFuncSetAttribute(at::enc_attn_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, at::SMEM_BYTES);
    }
    return v;
}

// VOX_CUDA_ENC_ATTN_KSPLIT=0: streaming chunks walk their key tiles in one CTA per head (A/B runs)
static int attn_ksplit_mode(void) {
    static int v = -1;
    if (v < 0) { const char *e = getenv("VOX_CUDA_ENC_ATTN_KSPLIT"); v = (e && !atoi(e)) ? 0 : 1; }
    return v;
}

static int enc_attention_launch(vox_cuda_state *s, const float *qkv, const float *kc, const float *vc, int M, int pos0,
                                uint16_t *out_b, long ldo, int split, int terms, const int *pos_dev) {
    EncState *e = (EncState *)s->enc_priv;
    cudaStream_t st = s->stream;
    if (!attn_mma_mode()) {
        // one query tile (streaming chunks): split the keys over CTAs instead of walking them in one
        int n_splits = 0;
        if (M <= at5::BQ && attn_ksplit_mode()) {
            if (!e->attn_part) {                  // first use (never under graph capture: a chunk shape runs eagerly first)
                const size_t pb = (size_t)ENC_HEADS * ATTN_KSPLITS * (ENC_HD + 2) * at5::BQ * sizeof(float);
                if (cudaMalloc((void **)&e->attn_part, pb) != cudaSuccess || cudaMalloc((void **)&e->attn_cnt, ENC_HEADS * sizeof(unsigned)) != cudaSuccess ||
                    cudaMemset(e->attn_cnt, 0, ENC_HEADS * sizeof(unsigned)) != cudaSuccess) {
                    vox_set_error("enc_attention: key-split scratch alloc failed");
                    return -1;
                }
            }
            n_splits = ATTN_KSPLITS;
        }
        dim3 grid(n_splits ? n_splits : (M + at5::BQ - 1) / at5::BQ, ENC_HEADS);
        if (terms == 3)
            CUDA_TRY(launch_pdl(at5::enc_attn_tc5_kernel<3>, grid, at5::NT, at5::SMEM_BYTES, st, qkv, 3L * ENC_QKV, kc, vc, M, pos0, e->pos_min,
                                ENC_WINDOW, ENC_RING - 1, out_b, ldo, split, pos_dev, e->attn_part, e->attn_cnt, n_splits));
        else
            CUDA_TRY(launch_pdl(at5::enc_attn_tc5_kernel<1>, grid, at5::NT, at5::SMEM_BYTES, st, qkv, 3L * ENC_QKV, kc, vc, M, pos0, e->pos_min,
                                ENC_WINDOW, ENC_RING - 1, out_b, ldo, split, pos_dev, e->attn_part, e->attn_cnt, n_splits));
        LAUNCHED();
        return 0;
    }
    dim3 grid((M + at::BQ - 1) / at::BQ, ENC_HEADS);
    if (terms == 3)
        CUDA_TRY(launch_pdl(at::enc_attn_kernel<3>, grid, at::NT, at::SMEM_BYTES, st, qkv, 3L * ENC_QKV, kc, vc, M, pos0, e->pos_min, ENC_WINDOW,
                            ENC_RING - 1, out_b, ldo, split, pos_dev));
    else
        CUDA_TRY(launch_pdl(at::enc_attn_kernel<1>, grid, at::NT, at::SMEM_BYTES, st, qkv, 3L * ENC_QKV, kc, vc, M, pos0, e->pos_min, ENC_WINDOW,
                            ENC_RING - 1, out_b, ldo, split, pos_dev));
    LAUNCHED();
    return 0;
}

// qkv: fp32 [M][3*2048] (q rotated in place by rows_rope_kv); kc / vc: this layer's fp32 ring.
int enc_attention_dev(vox_cuda_state *s, const float *qkv, const float *kc, const float *vc, int M, int pos0,
                      uint16_t *out_b, long ldo, int split, const int *pos_dev);
int enc_attention(vox_cuda_state *s, const float *qkv, const float *kc, const float *vc, int M, int pos0,
                  uint16_t *out_b, long ldo, int split) {
    return enc_attention_dev(s, qkv, kc, vc, M, pos0, out_b, ldo, split, nullptr);
}
// pos_dev (optional): device int holding pos0 (tensor-core path only)
int enc_attention_dev(vox_cuda_state *s, const float *qkv, const float *kc, const float *vc, int M, int pos0,
                      uint16_t *out_b, long ldo, int split, const int *pos_dev) {
    EncState *e = (EncState *)s->enc_priv;
    cudaStream_t st = s->stream;
    static int mode = -1;     // 0: tensor cores, 1: fp32 CUDA-core kernel + conversion (A/B debugging)
    if (mode < 0) {
        const char *env = getenv("VOX_CUDA_ENC_ATTN_F32");
        mode = (env && atoi(env)) ? 1 : 0;
        cudaFuncSetAttribute(at::enc_attn_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, at::SMEM_BYTES);
        cudaFuncSetAttribute(at::enc_attn_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, at::SMEM_BYTES);
    }
    if (mode == 1) {
        if (pos_dev) { vox_set_error("enc_attention: device position needs the tensor-core path"); return -1; }
        if (rows_attention(st, ENC_HD, qkv, 3 * ENC_QKV, kc, vc, ENC_QKV, ENC_RING - 1, ENC_HEADS, 1, M, pos0, e->pos_min,
                           ENC_WINDOW, e->att, ENC_QKV)) return -1;
        return f32_to_bf16_rows(st, e->att, ENC_QKV, out_b, ldo, M, ENC_QKV, split);
    }
    return enc_attention_launch(s, qkv, kc, vc, M, pos0, out_b, ldo, split, split ? 3 : 1, pos_dev);
}

int enc_attention_terms(vox_cuda_state *s, const float *qkv, const float *kc, const float *vc, int M, int pos0,
                        uint16_t *out_b, long ldo, int split, int terms) {
    static int init = 0;
    if (!init) {
        cudaFuncSetAttribute(at::enc_attn_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, at::SMEM_BYTES);
        cudaFuncSetAttribute(at::enc_attn_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, at::SMEM_BYTES);
        init = 1;
    }
    return enc_attention_launch(s, qkv, kc, vc, M, pos0, out_b, ldo, split, terms, nullptr);
}
