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
//     them into a 6 x 24 KB shared-memory ring with cp.async.bulk (TMA 1-D
//     bulk copies, mbarrier complete_tx, L2 evict_first).  It never waits for
//     activations, only for ring space, so the HBM pipe stays busy while the
//     consumers sit in a grid barrier (144 KB/SM = 3.3 us of HBM time; deeper rings measured slower).
//   * warps 0-7 are CONSUMERS: each warp owns a fixed K-slice of every row (its
//     slice of the activation vector lives in registers), all warps multiply each
//     tile as it lands (packed FFMA2), row sums are combined once per group of
//     <= 16 rows, and the group's epilogue lanes apply RoPE + KV-ring write,
//     residual add, SwiGLU, or logits + running argmax.
//   * phases are separated by a grid-wide barrier (release/acquire counter in
//     global memory); attention is split over (kv head, key range) CTAs and its
//     partials are merged in the wo prologue.
//
// Arithmetic is fp32 on exactly-widened bf16 weights, like the reference; only
// the summation order differs.
#include "vox_internal.h"
#include <cuda_fp16.h>
#include <stdlib.h>
#include <string.h>

namespace mk {

#ifndef MK_PROF
#define MK_PROF 0                           // 1: per-tile wait timing of the consumers (VOX_CUDA_MEGA_PROF readout); `make prof`
#endif
#ifndef MK_Q8_TPW
#define MK_Q8_TPW 0                         // Q8, K = 3072 / 4096 phases: 1 = tile-per-warp consumer (consume_q8_w), 0 = K-split (consume_q8)
#endif
#ifndef MK_Q8_GROWS
#define MK_Q8_GROWS 16                      // Q8 K-split consumer: rows per CTA barrier + epilogue (16 / 32 rows: 1.16 / 1.26 ms per token)
#endif
#ifndef MK_NSTAGE
#define MK_NSTAGE 6                         // ring depth: 4 / 6 / 8 stages measured 1.434 / 1.427 / 1.470 ms per token
#endif

constexpr int CW = 8;                       // consumer warps
constexpr int CT = CW * 32;                 // consumer threads
constexpr int NT = CT + 128;                // + one warpgroup: producer warp, L2 prefetch warp, two idle warps (setmaxnreg needs whole warpgroups)
constexpr int STAGE_BYTES = 24576;
constexpr int NSTAGE = MK_NSTAGE;
constexpr int MAXTR = 4;                    // rows per tile at K = 3072
constexpr int PART_LD = DEC_HD + 4;         // acc[128], m, l, pad (16-byte rows)
constexpr int MAX_SPLITS = 32;
constexpr int KMAX = DEC_HIDDEN;            // longest activation vector (w2 input)

template <int K, int EB = 2> struct TileRows { static constexpr int v = STAGE_BYTES / (K * EB); };   // EB = bytes per weight
static_assert(TileRows<3072>::v == 4 && TileRows<4096>::v == 3 && TileRows<9216>::v == 1, "");
static_assert(TileRows<3072, 1>::v == 8 && TileRows<4096, 1>::v == 6 && TileRows<9216, 1>::v == 2, "");

// Activation vector of the current phase, split so that lane l reads conflict-free
// float4s: xa[c] = x[8c..8c+3], xb[c] = x[8c+4..8c+7].  The attention phase (which
// has no activation vector) reuses the space for its per-warp partials.
struct AttnScratch { float qs[4][DEC_HD]; float wsm[CW][4][PART_LD]; };
constexpr int GROWS = 16;                   // rows per consumer sync group
template <int K> struct GroupTiles { static constexpr int v = (K == 9216) ? 8 : 4; };

struct Smem {
    alignas(128) unsigned char ring[NSTAGE][STAGE_BYTES];
    alignas(16) struct { AttnScratch at; } u;
    float part[2][GROWS][CW];               // per-warp partial row sums of a sync group (double buffered)
    alignas(8) unsigned long long full[NSTAGE];
    alignas(8) unsigned long long empty[NSTAGE];
    float red[CW];
    float rope[DEC_HD];                     // (cos, sin) x 64 for this step's position
    float bv[CW];
    int bi[CW];
    int token;
    int last_split;                         // this CTA was the last split of its kv head to finish (it merges)
    int owners;                             // CTAs that own qkv rows of this CTA's kv head
    volatile unsigned issued;               // tiles the producer has requested so far (read by the L2 prefetcher)
    float resid[64];                        // residual values of the rows this CTA owns in wo / w2
    float altv[CW][4];                      // per-warp top-4 text logits of the vocabulary phase (--alt)
    int alti[CW][4];
};

// Q8 weights (vox_cuda_model_create_q8): the activation vector of a phase as four planes of unsigned 8-bit digits
// (see prep_digits) and the int32 partial sums of a row group.  Only the Q8 instantiations allocate this part.
constexpr int QDIG = 4;                     // digit planes: 32-bit offset-binary fixed point
constexpr int QGROWS = MK_Q8_GROWS;         // rows per sync group of the Q8 K-split consumer (<= 32: one epilogue lane per row)
constexpr int QROWS = 896;                  // rows of one phase per CTA: ceil(131072 / 148) = 886 with 148 CTAs (checked at launch)
constexpr int QPLANE = KMAX + 64;           // plane pitch in bytes (+64: the four planes start in different banks)
struct SmemQ : Smem {
    alignas(16) unsigned char dig[QDIG][QPLANE];
    int partq[2][QDIG][CW][QGROWS];          // [buffer][digit plane][warp][row of the group]
    float redm[CW];
    // scale and row sum of the rows this CTA owns in the current phase (<= 886: the vocabulary), staged with cp.async at the
    // start of the phase so that no epilogue waits for an L2 round trip
    float esc[QROWS];
    int ers[QROWS];
};
static_assert(NSTAGE > 6 || sizeof(SmemQ) <= 232448, "Q8 kernel: shared memory (deeper measurement rings: bf16 models only)");

struct Top4 {
    float v[4]; int i[4];
    __device__ __forceinline__ void clear() {
        #pragma unroll
        for (int k = 0; k < 4; k++) { v[k] = -INFINITY; i[k] = 0x7fffffff; }
    }
    // order: larger logit first, lower id on ties; NaN never enters (same rule as the argmax)
    __device__ __forceinline__ void insert(float nv, int ni) {
        if (!(nv > v[3] || (nv == v[3] && ni < i[3]))) return;
        v[3] = nv; i[3] = ni;
        #pragma unroll
        for (int k = 3; k > 0; k--) {
            const bool up = v[k] > v[k - 1] || (v[k] == v[k - 1] && i[k] < i[k - 1]);
            const float tv = up ? v[k - 1] : v[k]; const int ti = up ? i[k - 1] : i[k];
            v[k - 1] = up ? v[k] : v[k - 1]; i[k - 1] = up ? i[k] : i[k - 1];
            v[k] = tv; i[k] = ti;
        }
    }
    __device__ __forceinline__ void merge_warp() {
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            float ov[4]; int oi[4];
            #pragma unroll
            for (int k = 0; k < 4; k++) { ov[k] = __shfl_xor_sync(0xffffffffu, v[k], o); oi[k] = __shfl_xor_sync(0xffffffffu, i[k], o); }
            #pragma unroll
            for (int k = 0; k < 4; k++) insert(ov[k], oi[k]);
        }
    }
};

struct Params {
    const DecLayerDev *layers;              // device array [n_dec]
    int n_dec;
    const uint16_t *tok_emb;
    const int8_t *q_tok_emb;                // Q8 model: int8 rows, row scales, row sums of the tied embedding / LM head
    const float *s_tok_emb;
    const int *qs_tok_emb;
    const float *dec_norm, *ada_scale, *inv_freq;
    void *kv_k, *kv_v;                      // rings of float or __half rows (kernel template parameter KV)
    float *x, *q, *attn_part, *h, *logits, *blockmax;
    float *attn;                            // [4096] merged, normalised attention output of the current layer
    unsigned *attn_cnt;                     // hand-over counters, SYNC_* layout (zeroed by the host before each launch)
    const float *embed_in, *adapter;
    DecStepParams *params;
    int *tokens;
    unsigned *bar;
    unsigned bar_base;
    int n_steps;
    int inflight;                           // max tiles requested but not yet landed (1..NSTAGE)
    int l2_ahead;                           // tiles the L2 prefetcher runs ahead of the producer (0 = off)
    int spin_ns;                            // back-off of the grid-barrier poll loop (0 = tight loop)
    unsigned *err;
    unsigned long long *prof;               // optional [G][8] cycle counters of consumer thread 0 (debug)
    // --alt (stream_fill_alts, voxtral.c:955-1010): the vocabulary phase keeps, beside the running argmax, the
    // four best logits among ids >= 1000; alt_cand [G][4] (value, id) per CTA, alt_out [TOKENS_RING] records
    float2 *alt_cand;
    AltRecord *alt_out;
    int alt_on;
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
__device__ __noinline__ void die(volatile unsigned *err, unsigned code, unsigned info) {
    if (atomicCAS((unsigned *)err, 0u, code) == 0u) {
        err[1] = blockIdx.x; err[2] = threadIdx.x; err[3] = info;
        __threadfence_system();
    }
    __trap();
}
__device__ __forceinline__ void mbar_wait(unsigned long long *b, unsigned parity, unsigned *err, unsigned info = 0) {
    if (mbar_try_wait(b, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try_wait(b, parity)) {
        if (clock64() - t0 > 3000000000LL) die(err, 1u, info);
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
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(CT) : "memory"); }

__device__ __forceinline__ unsigned ld_relaxed(const unsigned *p) {
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
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
// bar.sync orders the CTA's writes before thread 0's release; thread 0's acquire
// fence followed by bar.sync orders the other CTAs' writes before our reads.
__device__ __forceinline__ unsigned atom_release_add(unsigned *p, unsigned v) {
    unsigned old;
    asm volatile("atom.release.gpu.global.add.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
    return old;
}
__device__ __forceinline__ void grid_sync(unsigned *bar, unsigned target, unsigned *err, long long *t_own = nullptr,
                                          long long *t_atom = nullptr, unsigned spin_ns = 0) {
    consumer_sync();
    if (threadIdx.x == 0) {
        if (t_atom) {                           // profiling variant: returning atomic = measurable round trip
            const long long ta_ = clock64();
            const unsigned old = atom_release_add(bar, 1u);
            *t_atom += (clock64() - ta_) + (long long)(old & 0u);
        } else {
            red_release_add(bar, 1u);
        }
        const long long t0 = clock64();
        // acquire loads: the one that sees the last arrival's release orders everything after it, no separate fence
        while ((int)(ld_acquire(bar) - target) < 0) {
            if (spin_ns) __nanosleep(spin_ns);
            if (clock64() - t0 > 3000000000LL) die(err, 2u, target);
        }
        if (t_own) *t_own += clock64() - t0;    // spin time
    }
    consumer_sync();
}

// Targeted waits that replace two of the five grid barriers of a layer (all counters are zeroed by the host before
// every launch and only grow): the attention CTAs of kv head g wait for the CTAs that own rows of q / k / v of g
// (SYNC_QKV + g), every CTA waits for the 8 merged kv heads before wo (SYNC_ATTN_DONE).
constexpr int SYNC_ATTN_ARRIVE = 0;          // [8] splits finished per kv head
constexpr int SYNC_QKV = 8;                  // [8] qkv row owners finished per kv head
constexpr int SYNC_ATTN_DONE = 16;           // kv heads merged
constexpr int SYNC_WORDS = 32;
__device__ __forceinline__ void wait_count(const unsigned *p, unsigned target, unsigned *err, unsigned code) {
    // thread 0 polls, everyone else parks in the CTA barrier behind it
    if (threadIdx.x == 0) {
        const long long t0 = clock64();
        while ((int)(ld_acquire(p) - target) < 0) {
            if (clock64() - t0 > 3000000000LL) die(err, code, target);
        }
    }
    consumer_sync();
}
// does CTA row range [a, b) of the 6144 qkv rows hold rows of kv head g (its 4 query heads, its k, its v)?
__device__ __forceinline__ bool qkv_rows_touch(int a, int b, int g) {
    const int q0 = g * 4 * DEC_HD, k0 = DEC_QDIM + g * DEC_HD, v0 = DEC_QDIM + DEC_KVDIM + g * DEC_HD;
    return (a < q0 + 4 * DEC_HD && b > q0) || (a < k0 + DEC_HD && b > k0) || (a < v0 + DEC_HD && b > v0);
}

// rows [r0, r1) of an R-row matrix owned by CTA c; boundaries are multiples of `unit`
__device__ __forceinline__ void cta_rows(int R, int unit, int c, int G, int &r0, int &r1) {
    const int U = R / unit;
    r0 = (int)(((long long)c * U) / G) * unit;
    r1 = (int)(((long long)(c + 1) * U) / G) * unit;
}

// Tiles are numbered globally per CTA (g = 0, 1, ...) in consumption order;
// tile g lives in ring stage g % NSTAGE during mbarrier phase (g / NSTAGE) & 1.

// ---- producer ----------------------------------------------------------------------
template <int K, int EB = 2>
__device__ __forceinline__ void produce(Smem &sm, unsigned &g, const void *W, int R, int unit, int c, int G,
                                        unsigned long long policy, unsigned inflight, unsigned *err) {
    constexpr int TR = TileRows<K, EB>::v;
    int r0, r1;
    cta_rows(R, unit, c, G, r0, r1);
    for (int r = r0; r < r1; r += TR, g++) {
        const int n = min(TR, r1 - r);
        const unsigned bytes = (unsigned)n * K * (unsigned)EB;
        const unsigned stage = g % NSTAGE, ph = (g / NSTAGE) & 1u;
        mbar_wait(&sm.empty[stage], ph ^ 1u, err, 0x80000000u | g);
        if (g >= inflight) {
            // throttle: a full ring of outstanding requests (25 MB chip-wide) adds microseconds of
            // queueing to every grid barrier and prologue load; landed tiles cost nothing
            const unsigned f = g - inflight;
            mbar_wait(&sm.full[f % NSTAGE], (f / NSTAGE) & 1u, err, 0x40000000u | g);
        }
        mbar_expect_tx(&sm.full[stage], bytes);
        bulk_g2s(sm.ring[stage], reinterpret_cast<const unsigned char *>(W) + (size_t)r * K * EB, bytes, &sm.full[stage], policy);
        sm.issued = g + 1;
    }
}

// ---- L2 prefetcher ---------------------------------------------------------------------
// The smem ring holds 3.8 us of HBM time; attention plus two grid barriers stall the
// consumers for longer.  A second elected thread walks the same tile list `ahead`
// tiles in front of the producer and pulls those tiles into L2 (cp.async.bulk.prefetch),
// so HBM keeps streaming through the stall and the ring later refills at L2 speed.
template <int K>
__device__ __forceinline__ void prefetch_l2(Smem &sm, unsigned &g, const uint16_t *W, int R, int unit, int c, int G,
                                            unsigned ahead) {
    constexpr int TR = TileRows<K>::v;
    int r0, r1;
    cta_rows(R, unit, c, G, r0, r1);
    for (int r = r0; r < r1; r += TR, g++) {
        unsigned is = sm.issued;
        while (g >= is + NSTAGE + ahead) { __nanosleep(200); is = sm.issued; }
        if (g < is + NSTAGE) continue;           // the producer is (nearly) there already
        const unsigned bytes = (unsigned)min(TR, r1 - r) * K * 2u;
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(W + (size_t)r * K), "r"(bytes) : "memory");
    }
}

// Packed fp32 FMA (Blackwell FFMA2): both halves of a bf16 pair in one instruction.
// Halves the FMA-pipe work of the GEMV inner loop (2 ALU unpacks + 1 FFMA2 per weight pair).
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    unsigned long long A, B, C, D;
    asm("mov.b64 %0, {%1, %2};" : "=l"(A) : "f"(a.x), "f"(a.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(B) : "f"(b.x), "f"(b.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(C) : "f"(c.x), "f"(c.y));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(D) : "l"(A), "l"(B), "l"(C));
    float2 d;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(D));
    return d;
}
__device__ __forceinline__ float2 dot8x2(const uint4 &w, const float4 &a, const float4 &b, float2 acc) {
    acc = ffma2(make_float2(bf16_lo(w.x), bf16_hi(w.x)), make_float2(a.x, a.y), acc);
    acc = ffma2(make_float2(bf16_lo(w.y), bf16_hi(w.y)), make_float2(a.z, a.w), acc);
    acc = ffma2(make_float2(bf16_lo(w.z), bf16_hi(w.z)), make_float2(b.x, b.y), acc);
    acc = ffma2(make_float2(bf16_lo(w.w), bf16_hi(w.w)), make_float2(b.z, b.w), acc);
    return acc;
}

// ---- consumers: every warp owns a K-slice of every row -----------------------------------------
// Lane l of warp w owns columns (j*256 + w*32 + l)*4 .. +3 (j < K/1024) of every row; its
// slice of the activation vector stays in registers for the whole phase (no re-reads from
// shared memory), and all 8 warps work on each tile, so a phase ends within one tile-time
// of its last byte landing.  Partial row sums are reduced across lanes with a transposed
// shuffle reduction and across warps through shared memory once per GROUP of tiles
// (<= 16 rows), not per tile.  epi(row, v0, v1) is called by one lane per output row
// (PAIR = false) or per adjacent row pair (PAIR = true).  g0 = global index of the phase's first tile.
template <int K, bool PAIR, typename Epi>
__device__ __forceinline__ void consume(Smem &sm, unsigned &g0, int &pbuf, int R, int unit, int c, int G,
                                        const float (&xr)[K / 256], unsigned *err, long long &t_wait, Epi epi) {
    constexpr int TR = TileRows<K>::v, GT = GroupTiles<K>::v, J = K / 1024;
    static_assert(GT * TR <= GROWS, "");
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int r0, r1;
    cta_rows(R, unit, c, G, r0, r1);
    const int n_tiles = (r1 - r0 + TR - 1) / TR;
    // (software-pipelining the tile registers over two buffers was tried: it spills at 168
    // registers and ran 2.15 ms/step even with setmaxnreg 232, so tiles are fetched in order)
    uint2 vA[TR * J];
    auto fetch = [&](int t, uint2 (&vv)[TR * J]) {
        const unsigned g = g0 + (unsigned)t;
        const unsigned stage = g % NSTAGE, ph = (g / NSTAGE) & 1u;
        if (MK_PROF) {
            const long long tw = clock64();
            mbar_wait(&sm.full[stage], ph, err, g);
            t_wait += clock64() - tw;
        } else {
            mbar_wait(&sm.full[stage], ph, err, g);
        }
        const int n = min(TR, r1 - (r0 + t * TR));
        const uint2 *tp = reinterpret_cast<const uint2 *>(sm.ring[stage]) + w * 32 + lane;
        #pragma unroll
        for (int r = 0; r < TR; r++)
            #pragma unroll
            for (int j = 0; j < J; j++)
                vv[r * J + j] = (r < n) ? tp[r * (K / 4) + j * 256] : make_uint2(0u, 0u);
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.empty[stage]);      // release orders the loads above before the refill
    };
    auto multiply = [&](const uint2 (&vv)[TR * J], float2 *out) {
        #pragma unroll
        for (int r = 0; r < TR; r++) {
            float2 acc = make_float2(0.f, 0.f);
            #pragma unroll
            for (int j = 0; j < J; j++) {
                const uint2 v = vv[r * J + j];
                acc = ffma2(make_float2(bf16_lo(v.x), bf16_hi(v.x)), make_float2(xr[4 * j], xr[4 * j + 1]), acc);
                acc = ffma2(make_float2(bf16_lo(v.y), bf16_hi(v.y)), make_float2(xr[4 * j + 2], xr[4 * j + 3]), acc);
            }
            out[r] = acc;                                  // rows beyond the tile were loaded as zeros
        }
    };
    int grp = 0;
    for (int tg = 0; tg < n_tiles; tg += GT, grp++) {
        float2 a2[GT * TR];
        #pragma unroll
        for (int i = 0; i < GT * TR; i++) a2[i] = make_float2(0.f, 0.f);
        #pragma unroll
        for (int ti = 0; ti < GT; ti++) {
            const int t = tg + ti;
            if (t < n_tiles) {
                fetch(t, vA);
                multiply(vA, &a2[ti * TR]);
            }
        }
        // transposed reduction of up to 16 row sums over the 32 lanes: lane l ends with row (l >> 1)
        float v[GROWS];
        #pragma unroll
        for (int i = 0; i < GROWS; i++) v[i] = (i < GT * TR) ? a2[i < GT * TR ? i : 0].x + a2[i < GT * TR ? i : 0].y : 0.f;
        #pragma unroll
        for (int h = GROWS / 2; h >= 1; h >>= 1) {
            const bool up = (lane & (2 * h)) != 0;
            #pragma unroll
            for (int k = 0; k < h; k++) {
                const float keep = up ? v[k + h] : v[k];
                const float send = up ? v[k] : v[k + h];
                v[k] = keep + __shfl_xor_sync(0xffffffffu, send, 2 * h);
            }
        }
        v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
        if ((lane & 1) == 0) sm.part[pbuf][lane >> 1][w] = v[0];
        consumer_sync();
        const int rbase = r0 + tg * TR;
        const int rows = min(r1 - rbase, GT * TR);
        if (w == (grp & (CW - 1))) {                        // epilogue duty rotates over the warps
            if (PAIR) {
                if (lane < (rows >> 1)) {
                    float v0 = 0.f, v1 = 0.f;
                    #pragma unroll
                    for (int k = 0; k < CW; k++) { v0 += sm.part[pbuf][2 * lane][k]; v1 += sm.part[pbuf][2 * lane + 1][k]; }
                    epi(rbase + 2 * lane, v0, v1);
                }
            } else {
                if (lane < rows) {
                    float v0 = 0.f;
                    #pragma unroll
                    for (int k = 0; k < CW; k++) v0 += sm.part[pbuf][lane][k];
                    epi(rbase + lane, v0, 0.f);
                }
            }
        }
        pbuf ^= 1;
    }
    g0 += (unsigned)n_tiles;
}

// ---- Q8 consumers: int8 weight tiles on the integer tensor cores --------------------------------------
// q8_matvec_fused (/root/reference/voxtral_kernels.c:277-310): y[n] = (sum_k q[n][k] x[k]) * scale[n].
// The activation vector is turned into 32-bit fixed point ONCE per phase (prep_digits): X[k] = rint(x[k] 2^(30-e)) with
// 2^e > max|x|, stored offset-binary (U = X + 2^31) as four planes of unsigned bytes.  A weight tile then needs no
// conversion at all: mma.sync.m16n8k32.s32.u8.s8 takes the digit planes as its A rows (4 of 16 used) and 8 weight
// rows straight from the shared-memory tile as its B columns, int32 accumulation is exact, and per row
//     sum_k q x = 2^(e-30) (D0 + 2^8 D1 + 2^16 D2 + 2^24 D3 - 2^31 sum_k q[k])       D_j = sum_k q[k] u_j[k]
// in 64-bit integers - the only rounding of the whole dot product is X itself (2^-30 of the vector's maximum; the
// reference's sequential fp32 sum carries more).  ~30 instructions per thread per 24 KB tile instead of ~180 for
// an unpack + FFMA loop over twice the weights of a bf16 tile.
template <int K> struct GroupTilesQ { static constexpr int v = QGROWS / TileRows<K, 1>::v; };   // 2 / 2 / 8 tiles at 16 rows

__device__ __forceinline__ void imma_u8s8(int (&d)[4], uint32_t a0, uint32_t a2, uint32_t b0, uint32_t b1) {
    const uint32_t z = 0u;               // A rows 8..15 are not used
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3]) : "r"(a0), "r"(z), "r"(a2), "r"(z), "r"(b0), "r"(b1));
}

// Digit planes of this phase's activation vector.  xr = the lane's slice in the layout of load_slice (element
// 4j + i of xr is x[j*1024 + col0 + i]); dg[sp] = the 16 digits of plane (lane >> 2) that this lane feeds to the two
// MMAs of step pair sp (k = (sp*8 + w)*64 + (lane & 3)*16 .. +15: the same bytes it reads of every weight row).
__device__ __forceinline__ void cp_async4(void *dst, const void *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
// start of a Q8 phase: row constants of the CTA's rows -> shared memory, asynchronously (consume_q8 waits for them)
__device__ __forceinline__ void stage_row_consts(SmemQ &sm, const float *__restrict__ scale, const int *__restrict__ rowsum,
                                                 int R, int unit, int c, int G) {
    int r0, r1;
    cta_rows(R, unit, c, G, r0, r1);
    for (int i = threadIdx.x; i < r1 - r0; i += CT) {
        cp_async4(&sm.esc[i], scale + r0 + i);
        cp_async4(&sm.ers[i], rowsum + r0 + i);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}

template <int K, bool WANT_DG>
__device__ __forceinline__ void prep_digits(SmemQ &sm, const float (&xr)[K / 256], int col0, uint4 (&dg)[K / 512],
                                            float &unit_scale, float mx_known = -1.f) {
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float mx = mx_known;
    if (mx_known < 0.f) {                  // (uniform) no maximum from the norm's reduction round: one of our own
        mx = 0.f;
        #pragma unroll
        for (int i = 0; i < K / 256; i++) mx = fmaxf(mx, fabsf(xr[i]));
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        consumer_sync();                   // previous readers of redm[] are done
        if (lane == 0) sm.redm[w] = mx;
        consumer_sync();
        #pragma unroll
        for (int k = 0; k < CW; k++) mx = fmaxf(mx, sm.redm[k]);
    }
    // 2^e > max|x| (e = biased exponent - 126), clamped so that both scale factors are normal numbers
    int e = (int)((__float_as_uint(mx) >> 23) & 255u) - 126;
    e = max(-90, min(e, 96));
    const float up = __uint_as_float((unsigned)(127 + 30 - e) << 23);       // 2^(30-e)
    unit_scale = __uint_as_float((unsigned)(127 + e - 30) << 23);           // 2^(e-30)
    #pragma unroll
    for (int j = 0; j < K / 1024; j++) {
        uint32_t u[4];
        #pragma unroll
        for (int i = 0; i < 4; i++) u[i] = (uint32_t)__float2int_rn(xr[4 * j + i] * up) ^ 0x80000000u;
        // 4 x 4 byte transpose: plane p gets byte p of the four words
        const uint32_t t0 = __byte_perm(u[0], u[1], 0x5140), t1 = __byte_perm(u[2], u[3], 0x5140);
        const uint32_t t2 = __byte_perm(u[0], u[1], 0x7362), t3 = __byte_perm(u[2], u[3], 0x7362);
        const int k = j * 1024 + col0;
        *reinterpret_cast<uint32_t *>(&sm.dig[0][k]) = __byte_perm(t0, t1, 0x5410);
        *reinterpret_cast<uint32_t *>(&sm.dig[1][k]) = __byte_perm(t0, t1, 0x7632);
        *reinterpret_cast<uint32_t *>(&sm.dig[2][k]) = __byte_perm(t2, t3, 0x5410);
        *reinterpret_cast<uint32_t *>(&sm.dig[3][k]) = __byte_perm(t2, t3, 0x7632);
    }
    if (WANT_DG) {
        // K-split consumer (consume_q8): the lane's digits stay in registers.  With the Q8 column map (q8_col0) the
        // 64-byte chunks (sp*8 + w) that warp w multiplies are exactly the ones its own lanes just converted: a warp-level
        // barrier is enough, no CTA barrier between conversion and use
        __syncwarp();
        const int m = lane >> 2, cq = lane & 3;
        #pragma unroll
        for (int sp = 0; sp < K / 512; sp++)
            dg[sp] = (m < QDIG) ? *reinterpret_cast<const uint4 *>(&sm.dig[m & (QDIG - 1)][(sp * 8 + w) * 64 + cq * 16])
                                : make_uint4(0u, 0u, 0u, 0u);
        __syncwarp();                      // the next phase's conversion overwrites the planes
    } else {
        consumer_sync();                   // tile-per-warp consumer: every warp reads all of the planes
    }
}
// Column map of the Q8 kernel: lane l of warp w owns x[j*1024 + q8_col0 .. +3], chosen so that the 16 lanes of a half
// warp convert one 64-byte chunk (sp*8 + w) of the digit planes, sp = 2j + (l >> 4) - the chunks warp w multiplies.
__device__ __forceinline__ int q8_col0(int w, int lane) { return (lane >> 4) * 512 + w * 64 + (lane & 15) * 4; }

// Loads that must be issued where they are written (the compiler would otherwise sink them to their first use,
// an L2 round trip in front of the epilogue)
__device__ __forceinline__ float ldg_f32_now(const float *p) {
    float v;
    asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ int ldg_s32_now(const int *p) {
    int v;
    asm volatile("ld.global.nc.s32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
// The int8 rows are stored with the two 64-byte halves of every 128 bytes swapped on odd rows (q8_swizzle_kernel,
// api_core.cu): the 8 rows of a tile, whose pitch K is a multiple of 128, then spread over all 32 banks
__device__ __forceinline__ int q8_swz(int row) { return (row & 1) << 6; }

// Tile-per-warp consumer for K = 3072 / 4096: a warp multiplies every NSTAGE-th tile of the CTA's list on its own - all K
// columns of the tile's <= 8 rows, digits read from the shared planes - so the GEMV phase has no CTA-wide barrier, no
// partial-sum exchange and no shared epilogue duty: the row sums end in the warp that owns the tile (digit planes
// combined with two 64-bit shuffles), lanes 0..3 finish rows (2c, 2c + 1).  The warp stands in for all CW arrivals
// on the stage's `empty` barrier.  Up to NSTAGE tiles are in work by different warps at the same time.
template <int K, bool PAIR, typename Epi>
__device__ __forceinline__ void consume_q8_w(SmemQ &sm, unsigned &g0, int R, int unit, int c, int G, float unit_scale,
                                             const float *__restrict__ scale, const int *__restrict__ rowsum,
                                             unsigned *err, Epi epi) {
    constexpr int TR = TileRows<K, 1>::v, NP = K / 64;
    static_assert(TR <= 8 && (TR & 1) == 0, "");
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31, rq = lane >> 2, cq = lane & 3;
    int r0, r1;
    cta_rows(R, unit, c, G, r0, r1);
    const int n_tiles = (r1 - r0 + TR - 1) / TR;
    const unsigned char *dp = &sm.dig[rq & (QDIG - 1)][cq * 16];
    // Warp w owns ring stage w: it takes the tiles whose GLOBAL index is w (mod NSTAGE), so it sees every phase of
    // that stage's mbarriers in turn.  (A parity wait may run at most one phase ahead of its barrier: a warp that
    // skipped tiles of its stage would take the completion of an older tile for the one it waits for.)  Six warps
    // multiply - more than the integer tensor pipe of one SM sustains - and warps 6, 7 have no tile in these phases.
    constexpr int NWK = NSTAGE < CW ? NSTAGE : CW;
    static_assert(NSTAGE <= CW, "a stage needs its own warp");
    const int first = (int)(((unsigned)w + (unsigned)NSTAGE - g0 % (unsigned)NSTAGE) % (unsigned)NSTAGE);
    for (int t = first; w < NWK && t < n_tiles; t += NSTAGE) {
        const int row0 = r0 + t * TR;
        const int n = min(TR, r1 - row0);
        // scale / row sum of the two rows this lane may finish, in flight while the tile is multiplied
        const int ea = min(row0 + 2 * cq, r1 - 1), eb = min(row0 + 2 * cq + 1, r1 - 1);
        const float sc0 = ldg_f32_now(scale + ea), sc1 = ldg_f32_now(scale + eb);
        const int rs0 = ldg_s32_now(rowsum + ea), rs1 = ldg_s32_now(rowsum + eb);
        const unsigned g = g0 + (unsigned)t;
        const unsigned stage = g % NSTAGE, ph = (g / NSTAGE) & 1u;
        mbar_wait(&sm.full[stage], ph, err, g);
        const int rr = min(rq, TR - 1);                     // lanes beyond the tile's rows re-read its last row (columns nobody uses)
        const unsigned char *tp = sm.ring[stage] + (size_t)rr * K + cq * 16;
        const int swz = q8_swz(row0 + rr);
        int dch[4][4];
        #pragma unroll
        for (int ch = 0; ch < 4; ch++) { dch[ch][0] = dch[ch][1] = dch[ch][2] = dch[ch][3] = 0; }
        #pragma unroll 8
        for (int p = 0; p < NP; p++) {
            const uint4 wv = *reinterpret_cast<const uint4 *>(tp + ((p * 64) ^ swz));
            uint4 dv = make_uint4(0u, 0u, 0u, 0u);
            if (rq < QDIG) dv = *reinterpret_cast<const uint4 *>(dp + p * 64);
            imma_u8s8(dch[(2 * p) & 3], dv.x, dv.y, wv.x, wv.y);
            imma_u8s8(dch[(2 * p + 1) & 3], dv.z, dv.w, wv.z, wv.w);
        }
        __syncwarp();
        if (lane < CW) mbar_arrive(&sm.empty[stage]);       // release orders the loads above before the refill
        // digit plane rq of rows 2cq, 2cq + 1 -> 64-bit sums over the planes (lanes 4m + cq, m < 4)
        long long y0 = 0, y1 = 0;
        if (rq < QDIG) {
            y0 = (long long)((dch[0][0] + dch[1][0]) + (dch[2][0] + dch[3][0])) << (8 * rq);
            y1 = (long long)((dch[0][1] + dch[1][1]) + (dch[2][1] + dch[3][1])) << (8 * rq);
        }
        y0 += __shfl_xor_sync(0xffffffffu, y0, 4); y1 += __shfl_xor_sync(0xffffffffu, y1, 4);
        y0 += __shfl_xor_sync(0xffffffffu, y0, 8); y1 += __shfl_xor_sync(0xffffffffu, y1, 8);
        if (rq == 0 && 2 * cq < n) {
            const float v0 = __ll2float_rn(y0 - ((long long)rs0 << 31)) * unit_scale * sc0;
            const float v1 = __ll2float_rn(y1 - ((long long)rs1 << 31)) * unit_scale * sc1;
            if (PAIR) epi(row0 + 2 * cq, v0, v1);
            else {
                epi(row0 + 2 * cq, v0, 0.f);
                if (2 * cq + 1 < n) epi(row0 + 2 * cq + 1, v1, 0.f);
            }
        }
    }
    g0 += (unsigned)n_tiles;
}

// Same contract as consume<>: epi(row, v0, v1) by one lane per output row (pair).  scale / rowsum are indexed by the
// matrix row; unit_scale = 2^(e-30) of prep_digits.
template <int K, bool PAIR, typename Epi>
__device__ __forceinline__ void consume_q8(SmemQ &sm, unsigned &g0, int &pbuf, int R, int unit, int c, int G,
                                           const uint4 (&dg)[K / 512], float unit_scale, const float *__restrict__ scale,
                                           const int *__restrict__ rowsum, unsigned *err, long long &t_wait, Epi epi) {
    constexpr int TR = TileRows<K, 1>::v, GT = GroupTilesQ<K>::v, SP = K / 512;
    static_assert(GT * TR <= QGROWS && QGROWS <= 32 && TR <= 8, "");
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31, rq = lane >> 2, cq = lane & 3;
    int r0, r1;
    cta_rows(R, unit, c, G, r0, r1);
    const int n_tiles = (r1 - r0 + TR - 1) / TR;
    int grp = 0;
    for (int tg = 0; tg < n_tiles; tg += GT, grp++) {
        const int rbase = r0 + tg * TR;
        const int rows = min(r1 - rbase, GT * TR);
        const int er = PAIR ? 2 * lane : lane;
        #pragma unroll
        for (int ti = 0; ti < GT; ti++) {
            const int t = tg + ti;
            if (t >= n_tiles) break;                       // uniform across the CTA
            const unsigned g = g0 + (unsigned)t;
            const unsigned stage = g % NSTAGE, ph = (g / NSTAGE) & 1u;
            if (MK_PROF) {
                const long long tw = clock64();
                mbar_wait(&sm.full[stage], ph, err, g);
                t_wait += clock64() - tw;
            } else {
                mbar_wait(&sm.full[stage], ph, err, g);
            }
            // lane (rq, cq) holds 16 consecutive weights of row rq per step pair
            const int rr = min(rq, TR - 1);                // lanes beyond the tile's rows re-read its last row
            const unsigned char *tp = sm.ring[stage] + (size_t)rr * K + ((w * 64) ^ q8_swz(r0 + t * TR + rr)) + cq * 16;
            // the tile's weights go to registers in batches of <= 9 step pairs (72 registers for K = 9216 would not
            // leave room for the digit registers); the stage is released once the last batch has been read
            constexpr int SB = SP > 9 ? 9 : SP;
            static_assert(SP % SB == 0, "");
            // NCH independent accumulator chains: a single chain of 12-36 dependent MMAs is bound by the MMA latency
            constexpr int NCH = 4;
            int dch[NCH][4];
            #pragma unroll
            for (int ch = 0; ch < NCH; ch++) { dch[ch][0] = dch[ch][1] = dch[ch][2] = dch[ch][3] = 0; }
            #pragma unroll
            for (int b0 = 0; b0 < SP; b0 += SB) {
                uint4 wv[SB];
                #pragma unroll
                for (int sp = 0; sp < SB; sp++)
                    wv[sp] = *reinterpret_cast<const uint4 *>(tp + (b0 + sp) * 512);
                if (b0 + SB == SP) {
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&sm.empty[stage]);  // release orders the loads above before the refill
                }
                #pragma unroll
                for (int sp = 0; sp < SB; sp++) {
                    imma_u8s8(dch[(2 * sp) % NCH], dg[b0 + sp].x, dg[b0 + sp].y, wv[sp].x, wv[sp].y);
                    imma_u8s8(dch[(2 * sp + 1) % NCH], dg[b0 + sp].z, dg[b0 + sp].w, wv[sp].z, wv[sp].w);
                }
            }
            int d[2];
            d[0] = (dch[0][0] + dch[1][0]) + (dch[2][0] + dch[3][0]);
            d[1] = (dch[0][1] + dch[1][1]) + (dch[2][1] + dch[3][1]);
            // d[0], d[1] = digit plane rq (< QDIG) x tile rows 2cq, 2cq + 1, this warp's K-slice
            if (rq < QDIG) {
                if (2 * cq < TR) sm.partq[pbuf][rq][w][ti * TR + 2 * cq] = d[0];
                if (2 * cq + 1 < TR) sm.partq[pbuf][rq][w][ti * TR + 2 * cq + 1] = d[1];
            }
        }
        if (tg == 0) asm volatile("cp.async.wait_all;" ::: "memory");       // this thread's share of stage_row_consts
        consumer_sync();
        auto row_value = [&](int lr) -> float {
            const float sc = sm.esc[tg * TR + lr];
            const int rs = sm.ers[tg * TR + lr];
            long long y = 0;
            #pragma unroll
            for (int j = QDIG - 1; j >= 0; j--) {
                int dj = 0;                                                       // |.| <= K * 128 * 255 < 2^31
                #pragma unroll
                for (int k = 0; k < CW; k++) dj += sm.partq[pbuf][j][k][lr];     // lanes = consecutive rows: no bank conflicts
                y = (y << 8) + (long long)dj;
            }
            y -= (long long)rs << 31;
            return __ll2float_rn(y) * unit_scale * sc;
        };
        if (w == (grp & (CW - 1))) {                        // epilogue duty rotates over the warps
            if (PAIR) {
                if (lane < (rows >> 1)) epi(rbase + er, row_value(er), row_value(er + 1));
            } else {
                if (lane < rows) epi(rbase + er, row_value(er), 0.f);
            }
        }
        pbuf ^= 1;
    }
    g0 += (unsigned)n_tiles;
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
//
// Latency matters more than throughput here (a 60 s clip never exceeds ~800 keys),
// so the unit of work is a GROUP of 8 keys handled by one warp with all 16 K and V
// row loads (coalesced 512 B each) in flight at once - one memory round trip per
// group.  Lanes own 4 of the 128 dims; the 8 keys x 4 GQA heads = 32 partial dot
// products per lane are summed with a 31-shuffle transposed reduction that leaves
// score (key i, head h) in lane 4i + h.  Groups are dealt to (split, warp) units;
// units are merged per CTA in smem, and the splits are merged in the wo prologue.
__device__ __forceinline__ int attn_splits(int n_keys, int G) {
    const int ng = (n_keys + 7) >> 3;
    return min(min(G >> 3, MAX_SPLITS), (ng + CW - 1) / CW);
}

// K/V ring element type: float (the reference's CPU / BLAS layout, voxtral_decoder.c:703-705) or __half (the
// reference's GPU default "fp16 KV", voxtral.c:189-190): rows are rounded once when appended and widened when read;
// scores, softmax and the output stay fp32.
__device__ __forceinline__ float4 ld_kv4(const float *p) { return __ldcg(reinterpret_cast<const float4 *>(p)); }
__device__ __forceinline__ float4 ld_kv4(const __half *p) {
    const uint2 u = __ldcg(reinterpret_cast<const uint2 *>(p));
    const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&u.x));
    const float2 b = __half22float2(*reinterpret_cast<const __half2 *>(&u.y));
    return make_float4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ void st_kv2(float *p, float a, float b) { *reinterpret_cast<float2 *>(p) = make_float2(a, b); }
__device__ __forceinline__ void st_kv2(__half *p, float a, float b) { *reinterpret_cast<__half2 *>(p) = __floats2half2_rn(a, b); }

template <typename KV>
__device__ __forceinline__ void attention_phase(Smem &sm, const Params &P, const KV *kc, const KV *vc,
                                                int pos, int n_keys, int c, int G, unsigned arrivals_expected,
                                                const unsigned *qkv_ready, unsigned qkv_target) {
    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
    const int kvh = c & 7, split = c >> 3;
    const int s_eff = attn_splits(n_keys, G);
    if (split >= s_eff) return;
    const int first = pos - n_keys + 1;
    const int ng = (n_keys + 7) >> 3;
    const int U = s_eff * CW, u = split * CW + w;
    const int g0 = (int)(((long long)u * ng) / U), g1 = (int)(((long long)(u + 1) * ng) / U);
    const float scale = 0.08838834764831845f;  // 1/sqrt(128)
    const size_t hoff = (size_t)kvh * DEC_HD + 4 * lane;
    // Keys of earlier tokens do not depend on this step: the first group's K / V rows are requested BEFORE waiting
    // for the owners of this kv head's q / k / v rows (the newest key, written in this step, sits in the last group).
    float4 kf[8], vf[8];
    const bool pre = g0 < g1 && g0 * 8 + 8 <= n_keys - 1;
    if (pre) {
        #pragma unroll
        for (int i = 0; i < 8; i++) {
            const int slot = (first + g0 * 8 + i) & (DEC_WINDOW - 1);
            kf[i] = ld_kv4(kc + (size_t)slot * DEC_KVDIM + hoff);
            vf[i] = ld_kv4(vc + (size_t)slot * DEC_KVDIM + hoff);
        }
    }
    wait_count(qkv_ready, qkv_target, P.err, 3u);
    float4 q[4];
    #pragma unroll
    for (int h = 0; h < 4; h++) q[h] = __ldcg(reinterpret_cast<const float4 *>(P.q + (kvh * 4 + h) * DEC_HD) + lane);
    float m = -INFINITY, l = 0.f;              // running max / sum of head (lane & 3)
    float4 acc[4];
    #pragma unroll
    for (int h = 0; h < 4; h++) acc[h] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int g = g0; g < g1; g++) {
        const int j0 = g * 8;
        const int nv = min(8, n_keys - j0);
        if (!(pre && g == g0)) {
            #pragma unroll
            for (int i = 0; i < 8; i++) {
                const int slot = (first + j0 + min(i, nv - 1)) & (DEC_WINDOW - 1);   // clamp: always a valid row
                kf[i] = ld_kv4(kc + (size_t)slot * DEC_KVDIM + hoff);
                vf[i] = ld_kv4(vc + (size_t)slot * DEC_KVDIM + hoff);
            }
        }
        float v[32];
        #pragma unroll
        for (int i = 0; i < 8; i++)
            #pragma unroll
            for (int h = 0; h < 4; h++)
                v[i * 4 + h] = kf[i].x * q[h].x + kf[i].y * q[h].y + kf[i].z * q[h].z + kf[i].w * q[h].w;
        // transposed reduction: after the step with offset o, lanes with bit o set hold the upper half
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
        const float mn = fmaxf(m, mx);                       // finite: every group has >= 1 valid key
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
    for (int h = 0; h < 4; h++) reinterpret_cast<float4 *>(&sm.u.at.wsm[w][h][0])[lane] = acc[h];
    if (lane < 4) { sm.u.at.wsm[w][lane][DEC_HD] = m; sm.u.at.wsm[w][lane][DEC_HD + 1] = l; }
    consumer_sync();
    {   // merge the 8 warps: thread -> (head, two dims)
        const int h = tid >> 6, d = (tid & 63) * 2;
        float M = -INFINITY;
        #pragma unroll
        for (int k = 0; k < CW; k++) M = fmaxf(M, sm.u.at.wsm[k][h][DEC_HD]);
        float n0 = 0.f, n1 = 0.f, den = 0.f;
        #pragma unroll
        for (int k = 0; k < CW; k++) {
            const float mk_ = sm.u.at.wsm[k][h][DEC_HD];
            const float e = (mk_ == -INFINITY) ? 0.f : expf(mk_ - M);
            n0 = fmaf(e, sm.u.at.wsm[k][h][d], n0);
            n1 = fmaf(e, sm.u.at.wsm[k][h][d + 1], n1);
            den = fmaf(e, sm.u.at.wsm[k][h][DEC_HD + 1], den);
        }
        float *o = P.attn_part + ((size_t)split * DEC_HEADS + kvh * 4 + h) * PART_LD;
        *reinterpret_cast<float2 *>(o + d) = make_float2(n0, n1);
        if (d == 0) { o[DEC_HD] = M; o[DEC_HD + 1] = den; }
    }
    // The split that finishes LAST for this kv head merges all of them and publishes the normalised rows
    // (4 heads x 128) once, instead of every CTA of the grid redoing the merge in its wo prologue (148 x the
    // loads and exps, all hitting the same L2 lines at the same moment).
    __threadfence();                                     // partial rows visible before the arrival is counted
    consumer_sync();
    if (tid == 0) {
        const unsigned old = atomicAdd(P.attn_cnt + kvh, 1u);
        sm.last_split = (old + 1u == arrivals_expected) ? 1 : 0;
    }
    consumer_sync();
    if (!sm.last_split) return;
    __threadfence();                                     // acquire side: the other splits' rows
    {
        const int h = tid >> 6, d = (tid & 63) * 2;
        const float *base = P.attn_part + ((size_t)(kvh * 4 + h)) * PART_LD;
        const size_t sstride = (size_t)DEC_HEADS * PART_LD;
        float M = -INFINITY, n0 = 0.f, n1 = 0.f, den = 0.f;
        for (int sp = 0; sp < s_eff; sp += 8) {              // 16 loads in flight per round trip; one round for <= 8 splits
            float2 nv[8], ml[8];
            #pragma unroll
            for (int k = 0; k < 8; k++) {
                const int si = min(sp + k, s_eff - 1);
                nv[k] = __ldcg(reinterpret_cast<const float2 *>(base + si * sstride + d));
                ml[k] = __ldcg(reinterpret_cast<const float2 *>(base + si * sstride + DEC_HD));
            }
            float Mn = M;
            #pragma unroll
            for (int k = 0; k < 8; k++) if (sp + k < s_eff) Mn = fmaxf(Mn, ml[k].x);
            const float c0 = (M == -INFINITY) ? 0.f : expf(M - Mn);
            n0 *= c0; n1 *= c0; den *= c0;
            #pragma unroll
            for (int k = 0; k < 8; k++) {
                if (sp + k < s_eff) {
                    const float e = (ml[k].x == -INFINITY) ? 0.f : expf(ml[k].x - Mn);
                    n0 = fmaf(e, nv[k].x, n0); n1 = fmaf(e, nv[k].y, n1); den = fmaf(e, ml[k].y, den);
                }
            }
            M = Mn;
        }
        const float inv = den > 0.f ? 1.0f / den : 0.f;
        *reinterpret_cast<float2 *>(P.attn + (kvh * 4 + h) * DEC_HD + d) = make_float2(n0 * inv, n1 * inv);
    }
    consumer_sync();
    if (tid == 0) red_release_add(P.attn_cnt + SYNC_ATTN_DONE, 1u);     // one of the 8 kv heads is ready for wo
}

// ---- the kernel ----------------------------------------------------------------------------
// load this lane's K-slice of a length-K fp32 vector (coherent at L2) and its sum of squares
template <int K>
__device__ __forceinline__ float load_slice(const float *v, int col0, float (&xr)[K / 256]) {
    float ss = 0.f;
    #pragma unroll
    for (int j = 0; j < K / 1024; j++) {
        const float4 t = __ldcg(reinterpret_cast<const float4 *>(v + j * 1024 + col0));
        xr[4 * j] = t.x; xr[4 * j + 1] = t.y; xr[4 * j + 2] = t.z; xr[4 * j + 3] = t.w;
        ss += t.x * t.x + t.y * t.y + t.z * t.z + t.w * t.w;
    }
    return ss;
}

// RMSNorm of the slice (voxtral_kernels.c:475-492), optional (1 + ada) (voxtral_decoder.c:742-745).
// The norm weights are fetched before the block reduction so their latency overlaps it.
__device__ __forceinline__ void norm_slice(Smem &sm, float (&xr)[DEC_DIM / 256], float ss, const float *nw,
                                           const float *ada, int col0) {
    float4 wv[DEC_DIM / 1024], av[DEC_DIM / 1024];
    #pragma unroll
    for (int j = 0; j < DEC_DIM / 1024; j++) {
        wv[j] = *reinterpret_cast<const float4 *>(nw + j * 1024 + col0);
        if (ada) av[j] = *reinterpret_cast<const float4 *>(ada + j * 1024 + col0);
    }
    const float tot = block_sum(sm, ss);
    const float rinv = 1.0f / sqrtf(tot / (float)DEC_DIM + NORM_EPS);
    #pragma unroll
    for (int j = 0; j < DEC_DIM / 1024; j++) {
        xr[4 * j] = xr[4 * j] * rinv * wv[j].x; xr[4 * j + 1] = xr[4 * j + 1] * rinv * wv[j].y;
        xr[4 * j + 2] = xr[4 * j + 2] * rinv * wv[j].z; xr[4 * j + 3] = xr[4 * j + 3] * rinv * wv[j].w;
        if (ada) {
            xr[4 * j] *= (1.0f + av[j].x); xr[4 * j + 1] *= (1.0f + av[j].y);
            xr[4 * j + 2] *= (1.0f + av[j].z); xr[4 * j + 3] *= (1.0f + av[j].w);
        }
    }
}

// Q8: norm_slice that also returns (an upper bound within 1e-6 of) max |normalised x| from the SAME reduction round as the
// sum of squares - prep_digits then needs no reduction of its own (2 CTA barriers per norm phase instead of 5)
__device__ __forceinline__ float norm_slice_max(SmemQ &sm, float (&xr)[DEC_DIM / 256], float ss, const float *nw,
                                                const float *ada, int col0) {
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float4 wv[DEC_DIM / 1024], av[DEC_DIM / 1024];
    float mraw = 0.f;
    #pragma unroll
    for (int j = 0; j < DEC_DIM / 1024; j++) {
        wv[j] = *reinterpret_cast<const float4 *>(nw + j * 1024 + col0);
        float4 pw = wv[j];                 // w (1 + ada): for the maximum only - the norm below multiplies in the reference's order
        if (ada) {
            av[j] = *reinterpret_cast<const float4 *>(ada + j * 1024 + col0);
            pw.x *= (1.0f + av[j].x); pw.y *= (1.0f + av[j].y); pw.z *= (1.0f + av[j].z); pw.w *= (1.0f + av[j].w);
        }
        mraw = fmaxf(fmaxf(mraw, fabsf(xr[4 * j] * pw.x)), fabsf(xr[4 * j + 1] * pw.y));
        mraw = fmaxf(fmaxf(mraw, fabsf(xr[4 * j + 2] * pw.z)), fabsf(xr[4 * j + 3] * pw.w));
    }
    ss = warp_sum(ss);
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) mraw = fmaxf(mraw, __shfl_xor_sync(0xffffffffu, mraw, o));
    consumer_sync();                       // previous readers of red[] / redm[] are done
    if (lane == 0) { sm.red[w] = ss; sm.redm[w] = mraw; }
    consumer_sync();
    float tot = 0.f;
    #pragma unroll
    for (int k = 0; k < CW; k++) { tot += sm.red[k]; mraw = fmaxf(mraw, sm.redm[k]); }
    const float rinv = 1.0f / sqrtf(tot / (float)DEC_DIM + NORM_EPS);
    #pragma unroll
    for (int j = 0; j < DEC_DIM / 1024; j++) {
        xr[4 * j] = xr[4 * j] * rinv * wv[j].x; xr[4 * j + 1] = xr[4 * j + 1] * rinv * wv[j].y;
        xr[4 * j + 2] = xr[4 * j + 2] * rinv * wv[j].z; xr[4 * j + 3] = xr[4 * j + 3] * rinv * wv[j].w;
        if (ada) {
            xr[4 * j] *= (1.0f + av[j].x); xr[4 * j + 1] *= (1.0f + av[j].y);
            xr[4 * j + 2] *= (1.0f + av[j].z); xr[4 * j + 3] *= (1.0f + av[j].w);
        }
    }
    return mraw * rinv * 1.000002f;        // the products above round differently from mraw's: a bound, a few ulps up
}

// KV = element type of the K/V rings; Q8 = the weights are int8 rows + row scales (vox_cuda_model_create_q8)
template <typename KV, bool Q8>
__global__ void __launch_bounds__(NT, 1) dec_mega_kernel(const Params P) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem &sm = *reinterpret_cast<Smem *>(smem_raw);
    SmemQ &smq = *reinterpret_cast<SmemQ *>(smem_raw);      // the digit planes exist in the Q8 instantiations only
    const int tid = threadIdx.x;
    const int c = blockIdx.x, G = gridDim.x;

    if (tid == 0) {
        sm.issued = 0;
        for (int i = 0; i < NSTAGE; i++) { mbar_init(&sm.full[i], 1); mbar_init(&sm.empty[i], CW); }  // every consumer warp releases a stage
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const int n_dec = P.n_dec;

    // Register reallocation between warpgroups: ptxas caps a 384-thread CTA at 168 registers per thread, which
    // spills in the consumers; the producer warpgroup keeps 40 and hands the rest to the two consumer warpgroups.
    if (tid >= CT) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        // =========================== PRODUCER ===========================
        if (!Q8 && tid == CT + 32 && P.l2_ahead > 0) {
            unsigned fg = 0;
            for (int step = 0; step < P.n_steps; step++) {
                for (int l = 0; l < n_dec; l++) {
                    const DecLayerDev L = P.layers[l];
                    prefetch_l2<DEC_DIM>(sm, fg, L.wqkv, DEC_QKV, 2, c, G, (unsigned)P.l2_ahead);
                    prefetch_l2<DEC_QDIM>(sm, fg, L.wo, DEC_DIM, 1, c, G, (unsigned)P.l2_ahead);
                    prefetch_l2<DEC_DIM>(sm, fg, L.w13, 2 * DEC_HIDDEN, 2, c, G, (unsigned)P.l2_ahead);
                    prefetch_l2<DEC_HIDDEN>(sm, fg, L.w2, DEC_DIM, 1, c, G, (unsigned)P.l2_ahead);
                }
                prefetch_l2<DEC_DIM>(sm, fg, P.tok_emb, VOCAB, 1, c, G, (unsigned)P.l2_ahead);
            }
            return;
        }
        if (tid != CT) return;
        const unsigned long long policy = policy_evict_first();
        unsigned pg = 0;
        for (int step = 0; step < P.n_steps; step++) {
            for (int l = 0; l < n_dec; l++) {
                const DecLayerDev L = P.layers[l];
                constexpr int EB = Q8 ? 1 : 2;
                produce<DEC_DIM, EB>(sm, pg, Q8 ? (const void *)L.q_wqkv : (const void *)L.wqkv, DEC_QKV, 2, c, G, policy, (unsigned)P.inflight, P.err);
                produce<DEC_QDIM, EB>(sm, pg, Q8 ? (const void *)L.q_wo : (const void *)L.wo, DEC_DIM, 1, c, G, policy, (unsigned)P.inflight, P.err);
                produce<DEC_DIM, EB>(sm, pg, Q8 ? (const void *)L.q_w13 : (const void *)L.w13, 2 * DEC_HIDDEN, 2, c, G, policy, (unsigned)P.inflight, P.err);
                produce<DEC_HIDDEN, EB>(sm, pg, Q8 ? (const void *)L.q_w2 : (const void *)L.w2, DEC_DIM, 1, c, G, policy, (unsigned)P.inflight, P.err);
            }
            produce<DEC_DIM, Q8 ? 1 : 2>(sm, pg, Q8 ? (const void *)P.q_tok_emb : (const void *)P.tok_emb, VOCAB, 1, c, G, policy, (unsigned)P.inflight, P.err);
        }
        return;
    }

    // =============================== CONSUMERS ===============================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    const int w = tid >> 5, lane = tid & 31;
    unsigned gt = 0;                              // global index of the next phase's first tile
    int pbuf = 0;
    unsigned bar_target = P.bar_base;
    long long t_wait = 0, t_grid = 0, t_attn = 0, t_pro = 0, t_own = 0, t_atom = 0, t_prep = 0;
    const long long t_start = clock64();
#define MK_GRID_SYNC()                                              \
    do {                                                            \
        const long long tg_ = clock64();                            \
        const bool rec_ = P.prof && tid == 0 && step == P.n_steps - 1 && bar_k < 12;   \
        if (rec_) P.prof[(size_t)c * 32 + 8 + 2 * bar_k] = globaltimer_ns();          \
        grid_sync(P.bar, bar_target += G, P.err, P.prof ? &t_own : nullptr, P.prof ? &t_atom : nullptr, (unsigned)P.spin_ns); \
        if (rec_) P.prof[(size_t)c * 32 + 9 + 2 * bar_k] = globaltimer_ns();          \
        bar_k++;                                                    \
        t_grid += clock64() - tg_;                                  \
    } while (0)
    // step scalars: every CTA tracks them in registers
    int pos = P.params->pos, n_valid = P.params->n_valid, prev_token = P.params->prev_token;
    int adapter_index = P.params->adapter_index, step_count = P.params->step_count;
    const int use_host_embed = P.params->use_host_embed;
    const size_t kv_layer = (size_t)DEC_WINDOW * DEC_KVDIM;
    // column base of this lane for unit j: (j*256 + w*32 + lane) * 4
    const int col0 = Q8 ? q8_col0(w, lane) : (w * 32 + lane) * 4;
    // rows of the [3072]-row matrices (wo, w2) this CTA owns: residual values are staged in smem
    int rr0, rr1;
    cta_rows(DEC_DIM, 1, c, G, rr0, rr1);
    // arrivals per kv head since kernel start (every CTA tracks the same number); the counters are zeroed here,
    // a whole qkv phase and a grid barrier before their first use
    unsigned attn_arrivals = 0, layer_phases = 0;
    // who feeds whom in the qkv -> attention hand-over
    int my_heads = 0;                           // bit g: this CTA owns rows of kv head g
    {
        int a, b;
        cta_rows(DEC_QKV, 2, c, G, a, b);
        #pragma unroll
        for (int g = 0; g < 8; g++) my_heads |= qkv_rows_touch(a, b, g) ? (1 << g) : 0;
        if (tid == 0) sm.owners = 0;
        consumer_sync();
        if (tid < G) {
            cta_rows(DEC_QKV, 2, tid, G, a, b);
            if (qkv_rows_touch(a, b, c & 7)) atomicAdd(&sm.owners, 1);
        }
        consumer_sync();
    }
    const unsigned owners = (unsigned)sm.owners;

    for (int step = 0; step < P.n_steps; step++) {
        const int n_keys = min(n_valid, DEC_WINDOW - 1) + 1;
        const int slot = pos & (DEC_WINDOW - 1);
        const int s_eff = attn_splits(n_keys, G);
        // RoPE table of this position (voxtral_kernels.c:617-629)
        if (tid < DEC_HD / 2) {
            const float angle = (float)pos * P.inv_freq[tid];
            sm.rope[2 * tid] = cosf(angle);
            sm.rope[2 * tid + 1] = sinf(angle);
        }
        int bar_k = 0;
        for (int l = 0; l < n_dec; l++) {
            if (l == 10) bar_k = 0; else if (l < 10) bar_k = 100;
            const DecLayerDev L = P.layers[l];
            KV *kc = reinterpret_cast<KV *>(P.kv_k) + l * kv_layer, *vc = reinterpret_cast<KV *>(P.kv_v) + l * kv_layer;
            // ---------------- phase 1: rmsnorm -> qkv -> rope -> q / KV ring ----------------
            {
                const long long tp = clock64();
                if constexpr (Q8) stage_row_consts(smq, L.s_wqkv, L.qs_wqkv, DEC_QKV, 2, c, G);
                float xr[DEC_DIM / 256];
                float ss = 0.f;
                if (l == 0) {
                    #pragma unroll
                    for (int j = 0; j < DEC_DIM / 1024; j++) {
                        const int col = j * 1024 + col0;
                        float4 v;
                        // x = adapter[idx] + bf16->f32(tok_emb[prev])  (voxtral.c:1106-1113)
                        if (use_host_embed) v = __ldcg(reinterpret_cast<const float4 *>(P.embed_in + col));
                        else {
                            v = __ldcg(reinterpret_cast<const float4 *>(P.adapter + (size_t)adapter_index * DEC_DIM + col));
                            if constexpr (Q8) {     // embedding row = q * scale (tok_embed_q8_to_f32, voxtral.c:443-452)
                                const char4 e = *reinterpret_cast<const char4 *>(P.q_tok_emb + (size_t)prev_token * DEC_DIM + (col ^ q8_swz(prev_token)));
                                const float sc = __ldg(P.s_tok_emb + prev_token);
                                v.x += (float)e.x * sc; v.y += (float)e.y * sc; v.z += (float)e.z * sc; v.w += (float)e.w * sc;
                            } else {
                                const uint2 e = *reinterpret_cast<const uint2 *>(P.tok_emb + (size_t)prev_token * DEC_DIM + col);
                                v.x += bf16_lo(e.x); v.y += bf16_hi(e.x); v.z += bf16_lo(e.y); v.w += bf16_hi(e.y);
                            }
                        }
                        if (c == 0) *reinterpret_cast<float4 *>(P.x + col) = v;
                        xr[4 * j] = v.x; xr[4 * j + 1] = v.y; xr[4 * j + 2] = v.z; xr[4 * j + 3] = v.w;
                        ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
                    }
                } else {
                    ss = load_slice<DEC_DIM>(P.x, col0, xr);
                }
                float mxn = -1.f;
                if constexpr (Q8) mxn = norm_slice_max(smq, xr, ss, L.attn_norm, nullptr, col0);
                else norm_slice(sm, xr, ss, L.attn_norm, nullptr, col0);
                auto epi_qkv = [&](int row, float v0, float v1) {
                        if (row < DEC_QDIM + DEC_KVDIM) {
                            const int d = (row & (DEC_HD - 1)) >> 1;
                            const float cs = sm.rope[2 * d], sn = sm.rope[2 * d + 1];
                            const float r0 = v0 * cs - v1 * sn, r1 = v0 * sn + v1 * cs;
                            v0 = r0; v1 = r1;
                        }
                        if (row < DEC_QDIM) *reinterpret_cast<float2 *>(P.q + row) = make_float2(v0, v1);
                        else if (row < DEC_QDIM + DEC_KVDIM) st_kv2(kc + (size_t)slot * DEC_KVDIM + (row - DEC_QDIM), v0, v1);
                        else st_kv2(vc + (size_t)slot * DEC_KVDIM + (row - DEC_QDIM - DEC_KVDIM), v0, v1);
                    };
                if constexpr (Q8) {
                    uint4 dg[DEC_DIM / 512];
                    float us;
                    { const long long tq_ = clock64(); prep_digits<DEC_DIM, !MK_Q8_TPW>(smq, xr, col0, dg, us, mxn); t_prep += clock64() - tq_; }
                    t_pro += clock64() - tp;
                    if constexpr (MK_Q8_TPW) consume_q8_w<DEC_DIM, true>(smq, gt, DEC_QKV, 2, c, G, us, L.s_wqkv, L.qs_wqkv, P.err, epi_qkv);
                    else consume_q8<DEC_DIM, true>(smq, gt, pbuf, DEC_QKV, 2, c, G, dg, us, L.s_wqkv, L.qs_wqkv, P.err, t_wait, epi_qkv);
                } else {
                    t_pro += clock64() - tp;
                    consume<DEC_DIM, true>(sm, gt, pbuf, DEC_QKV, 2, c, G, xr, P.err, t_wait, epi_qkv);
                }
            }
            // q / k / v rows of this CTA are written: tell the attention CTAs of the kv heads they belong to
            layer_phases += 1u;
            consumer_sync();
            if (tid < 8 && ((my_heads >> tid) & 1)) red_release_add(P.attn_cnt + SYNC_QKV + tid, 1u);
            // ---------------- phase 2: attention partials (the CTAs that own a key range of a kv head) ----------------
            {
                const long long ta = clock64();
                attn_arrivals += (unsigned)s_eff;
                if ((c >> 3) < s_eff)
                    attention_phase(sm, P, kc, vc, pos, n_keys, c, G, attn_arrivals,
                                    P.attn_cnt + SYNC_QKV + (c & 7), owners * layer_phases);
                t_attn += clock64() - ta;
            }
            // every CTA: all 8 kv heads merged?
            {
                const long long tg_ = clock64();
                wait_count(P.attn_cnt + SYNC_ATTN_DONE, 8u * layer_phases, P.err, 4u);
                t_grid += clock64() - tg_;
            }
            // ---------------- phase 3: merge partials -> wo -> residual ----------------
            {
                const long long tp = clock64();
                if constexpr (Q8) stage_row_consts(smq, L.s_wo, L.qs_wo, DEC_DIM, 1, c, G);
                if (tid < rr1 - rr0) sm.resid[tid] = __ldcg(P.x + rr0 + tid);
                float xr[DEC_QDIM / 256];
                (void)load_slice<DEC_QDIM>(P.attn, col0, xr);     // rows merged by the last split of each kv head
                auto epi_res = [&](int row, float v0, float) { P.x[row] = sm.resid[row - rr0] + v0; };
                if constexpr (Q8) {
                    uint4 dg[DEC_QDIM / 512];
                    float us;
                    { const long long tq_ = clock64(); prep_digits<DEC_QDIM, !MK_Q8_TPW>(smq, xr, col0, dg, us); t_prep += clock64() - tq_; }
                    t_pro += clock64() - tp;
                    if constexpr (MK_Q8_TPW) consume_q8_w<DEC_QDIM, false>(smq, gt, DEC_DIM, 1, c, G, us, L.s_wo, L.qs_wo, P.err, epi_res);
                    else consume_q8<DEC_QDIM, false>(smq, gt, pbuf, DEC_DIM, 1, c, G, dg, us, L.s_wo, L.qs_wo, P.err, t_wait, epi_res);
                } else {
                    t_pro += clock64() - tp;
                    consume<DEC_QDIM, false>(sm, gt, pbuf, DEC_DIM, 1, c, G, xr, P.err, t_wait, epi_res);
                }
            }
            MK_GRID_SYNC();
            // ---------------- phase 4: rmsnorm * (1 + ada) -> w1|w3 -> SwiGLU ----------------
            {
                const long long tp = clock64();
                float xr[DEC_DIM / 256];
                const float ss = load_slice<DEC_DIM>(P.x, col0, xr);
                if constexpr (Q8) stage_row_consts(smq, L.s_w13, L.qs_w13, 2 * DEC_HIDDEN, 2, c, G);
                float mxn = -1.f;
                if constexpr (Q8) mxn = norm_slice_max(smq, xr, ss, L.ffn_norm, P.ada_scale + (size_t)l * DEC_DIM, col0);
                else norm_slice(sm, xr, ss, L.ffn_norm, P.ada_scale + (size_t)l * DEC_DIM, col0);
                auto epi_glu = [&](int row, float v0, float v1) { P.h[row >> 1] = silu(v0) * v1; };
                if constexpr (Q8) {
                    uint4 dg[DEC_DIM / 512];
                    float us;
                    { const long long tq_ = clock64(); prep_digits<DEC_DIM, !MK_Q8_TPW>(smq, xr, col0, dg, us, mxn); t_prep += clock64() - tq_; }
                    t_pro += clock64() - tp;
                    if constexpr (MK_Q8_TPW) consume_q8_w<DEC_DIM, true>(smq, gt, 2 * DEC_HIDDEN, 2, c, G, us, L.s_w13, L.qs_w13, P.err, epi_glu);
                    else consume_q8<DEC_DIM, true>(smq, gt, pbuf, 2 * DEC_HIDDEN, 2, c, G, dg, us, L.s_w13, L.qs_w13, P.err, t_wait, epi_glu);
                } else {
                    t_pro += clock64() - tp;
                    consume<DEC_DIM, true>(sm, gt, pbuf, 2 * DEC_HIDDEN, 2, c, G, xr, P.err, t_wait, epi_glu);
                }
            }
            MK_GRID_SYNC();
            // ---------------- phase 5: w2 -> residual ----------------
            {
                const long long tp = clock64();
                if constexpr (Q8) stage_row_consts(smq, L.s_w2, L.qs_w2, DEC_DIM, 1, c, G);
                if (tid < rr1 - rr0) sm.resid[tid] = __ldcg(P.x + rr0 + tid);
                float xr[DEC_HIDDEN / 256];
                (void)load_slice<DEC_HIDDEN>(P.h, col0, xr);
                auto epi_res = [&](int row, float v0, float) { P.x[row] = sm.resid[row - rr0] + v0; };
                if constexpr (Q8) {
                    uint4 dg[DEC_HIDDEN / 512];
                    float us;
                    { const long long tq_ = clock64(); prep_digits<DEC_HIDDEN, true>(smq, xr, col0, dg, us); t_prep += clock64() - tq_; }
                    t_pro += clock64() - tp;
                    consume_q8<DEC_HIDDEN, false>(smq, gt, pbuf, DEC_DIM, 1, c, G, dg, us, L.s_w2, L.qs_w2, P.err, t_wait, epi_res);
                } else {
                    t_pro += clock64() - tp;
                    consume<DEC_HIDDEN, false>(sm, gt, pbuf, DEC_DIM, 1, c, G, xr, P.err, t_wait, epi_res);
                }
            }
            MK_GRID_SYNC();
        }
        // ---------------- final norm -> logits -> argmax (voxtral_decoder.c:765-779) ----------------
        {
            const long long tp = clock64();
            float xr[DEC_DIM / 256];
            const float ss = load_slice<DEC_DIM>(P.x, col0, xr);
            if constexpr (Q8) stage_row_consts(smq, P.s_tok_emb, P.qs_tok_emb, VOCAB, 1, c, G);
            float mxn = -1.f;
            if constexpr (Q8) mxn = norm_slice_max(smq, xr, ss, P.dec_norm, nullptr, col0);
            else norm_slice(sm, xr, ss, P.dec_norm, nullptr, col0);
            float bv = -INFINITY;
            int bi = 0x7fffffff;
            Top4 top;
            top.clear();
            auto epi_alt = [&](int row, float v0, float) {
                        P.logits[row] = v0; argmax_merge(bv, bi, v0, row);
                        if (row >= 1000) top.insert(v0, row);            // TOKEN_TEXT_MIN, voxtral.c:399,986
                    };
            auto epi_max = [&](int row, float v0, float) { P.logits[row] = v0; argmax_merge(bv, bi, v0, row); };
            if constexpr (Q8) {
                uint4 dg[DEC_DIM / 512];
                float us;
                { const long long tq_ = clock64(); prep_digits<DEC_DIM, !MK_Q8_TPW>(smq, xr, col0, dg, us, mxn); t_prep += clock64() - tq_; }
                t_pro += clock64() - tp;
                if constexpr (MK_Q8_TPW) {
                    if (P.alt_on) consume_q8_w<DEC_DIM, false>(smq, gt, VOCAB, 1, c, G, us, P.s_tok_emb, P.qs_tok_emb, P.err, epi_alt);
                    else consume_q8_w<DEC_DIM, false>(smq, gt, VOCAB, 1, c, G, us, P.s_tok_emb, P.qs_tok_emb, P.err, epi_max);
                } else {
                    if (P.alt_on) consume_q8<DEC_DIM, false>(smq, gt, pbuf, VOCAB, 1, c, G, dg, us, P.s_tok_emb, P.qs_tok_emb, P.err, t_wait, epi_alt);
                    else consume_q8<DEC_DIM, false>(smq, gt, pbuf, VOCAB, 1, c, G, dg, us, P.s_tok_emb, P.qs_tok_emb, P.err, t_wait, epi_max);
                }
            } else {
                t_pro += clock64() - tp;
                if (P.alt_on) consume<DEC_DIM, false>(sm, gt, pbuf, VOCAB, 1, c, G, xr, P.err, t_wait, epi_alt);
                else consume<DEC_DIM, false>(sm, gt, pbuf, VOCAB, 1, c, G, xr, P.err, t_wait, epi_max);
            }
            if (P.alt_on) {
                top.merge_warp();
                if (lane == 0) {
                    #pragma unroll
                    for (int k = 0; k < 4; k++) { sm.altv[w][k] = top.v[k]; sm.alti[w][k] = top.i[k]; }
                }
            }
            // epilogue lanes of every warp hold candidates: block-wide merge
            #pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                argmax_merge(bv, bi, ov, oi);
            }
            if (lane == 0) { sm.bv[w] = bv; sm.bi[w] = bi; }
            consumer_sync();
            if (tid == 0) {
                #pragma unroll
                for (int k = 1; k < CW; k++) argmax_merge(bv, bi, sm.bv[k], sm.bi[k]);
                P.blockmax[2 * c] = bv;
                P.blockmax[2 * c + 1] = __int_as_float(bi);
                if (P.alt_on) {
                    for (int k = 1; k < CW; k++)
                        for (int j = 0; j < 4; j++) top.insert(sm.altv[k][j], sm.alti[k][j]);
                    #pragma unroll
                    for (int j = 0; j < 4; j++) P.alt_cand[4 * c + j] = make_float2(top.v[j], __int_as_float(top.i[j]));
                }
            }
        }
        MK_GRID_SYNC();
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
            if (lane == 0) { sm.bv[w] = bv; sm.bi[w] = bi; }   // (the grid barrier separates this from the reads above)
            consumer_sync();
            if (tid == 0) {
                for (int k = 1; k < CW; k++) argmax_merge(bv, bi, sm.bv[k], sm.bi[k]);
                if (bi == 0x7fffffff) bi = 0;   // all-NaN logits: the reference returns index 0
                sm.token = bi;
            }
            consumer_sync();
            prev_token = sm.token;
            if (c == 0 && tid == 0) P.tokens[step_count & (TOKENS_RING - 1)] = prev_token;
            if (P.alt_on && c == 0 && w == 1) {
                // merge the per-CTA candidates; the greedy token itself is skipped (voxtral.c:987)
                Top4 top;
                top.clear();
                for (int k = lane; k < 4 * G; k += 32) {
                    const float2 cd = __ldcg(P.alt_cand + k);
                    if (__float_as_int(cd.y) != prev_token) top.insert(cd.x, __float_as_int(cd.y));
                }
                top.merge_warp();
                if (lane == 0) {
                    AltRecord r;
                    r.best = __ldcg(P.logits + prev_token);
                    #pragma unroll
                    for (int k = 0; k < 3; k++) { r.v[k] = top.v[k]; r.id[k] = top.i[k] == 0x7fffffff ? -1 : top.i[k]; }
                    r.pad = 0;
                    P.alt_out[step_count & (TOKENS_RING - 1)] = r;
                }
            }
        }
        step_count += 1;
        pos += 1;
        n_valid = min(n_valid + 1, DEC_WINDOW);
        adapter_index += 1;
    }
#undef MK_GRID_SYNC
    if (c == 0 && tid == 0) {
        DecStepParams *p = P.params;
        p->out_token = prev_token; p->prev_token = prev_token; p->step_count = step_count;
        p->pos = pos; p->n_valid = n_valid; p->adapter_index = adapter_index;
    }
    if (P.prof && tid == 0) {
        unsigned long long *o = P.prof + (size_t)c * 32;
        o[0] = (unsigned long long)(clock64() - t_start);
        o[1] = (unsigned long long)t_wait; o[2] = (unsigned long long)t_grid;
        o[3] = (unsigned long long)t_attn; o[4] = (unsigned long long)t_pro; o[5] = (unsigned long long)t_own; o[6] = (unsigned long long)t_atom; o[7] = (unsigned long long)t_prep;   // Q8: digit conversion share of the prologues
    }
}

}  // namespace mk

// ---- host ---------------------------------------------------------------------------------------
struct MegaHost {
    DecLayerDev *d_layers;
    unsigned *d_bar, *d_err;
    unsigned long long *d_prof;   // [grid][8] when VOX_CUDA_MEGA_PROF=1
    int inflight, l2_ahead, spin_ns;
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
    CUDA_TRY(cudaFuncSetAttribute(mk::dec_mega_kernel<float, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)sizeof(mk::Smem)));
    CUDA_TRY(cudaFuncSetAttribute(mk::dec_mega_kernel<__half, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)sizeof(mk::Smem)));
    if (s->m->q8) {
        CUDA_TRY(cudaFuncSetAttribute(mk::dec_mega_kernel<float, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)sizeof(mk::SmemQ)));
        CUDA_TRY(cudaFuncSetAttribute(mk::dec_mega_kernel<__half, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)sizeof(mk::SmemQ)));
        CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, mk::dec_mega_kernel<float, true>, mk::NT, sizeof(mk::SmemQ)));
    } else
        CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, mk::dec_mega_kernel<float, false>, mk::NT, sizeof(mk::Smem)));
    if (occ < 1 || sms < 8) { vox_set_error("dec_mega: kernel does not fit (occupancy %d, %d SMs)", occ, sms); return -1; }
    h->grid = sms;
    h->barriers_per_step = 3 * s->m->n_dec + 1;       // after wo, w1|w3, w2 of every layer + after the logits
    CUDA_TRY(cudaMalloc((void **)&h->d_layers, sizeof(DecLayerDev) * VOX_CUDA_DEC_LAYERS));
    CUDA_TRY(cudaMemcpy(h->d_layers, s->m->dec, sizeof(DecLayerDev) * VOX_CUDA_DEC_LAYERS, cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMalloc((void **)&h->d_bar, 256));                  // [0] grid barrier, [8 ...] hand-over counters (SYNC_*)
    CUDA_TRY(cudaMemset(h->d_bar, 0, 256));
    // watchdog record in mapped host memory: still readable after a device-side trap
    CUDA_TRY(cudaHostAlloc((void **)&h->d_err, 16, cudaHostAllocMapped));
    memset(h->d_err, 0, 16);
    h->bar_count = 0;
    h->inflight = mk::NSTAGE;
    h->l2_ahead = 0;
    h->spin_ns = 0;
    {
        const char *env = getenv("VOX_CUDA_MEGA_SPIN_NS");
        if (env && atoi(env) >= 0) h->spin_ns = atoi(env);
    }
    {
        const char *env = getenv("VOX_CUDA_MEGA_L2AHEAD");
        if (env && atoi(env) >= 0 && atoi(env) <= 256) h->l2_ahead = atoi(env);
    }
    {
        const char *env = getenv("VOX_CUDA_MEGA_INFLIGHT");
        if (env && atoi(env) >= 1 && atoi(env) <= mk::NSTAGE) h->inflight = atoi(env);
    }
    {
        const char *env = getenv("VOX_CUDA_MEGA_PROF");
        if (env && atoi(env)) {
            CUDA_TRY(cudaMalloc((void **)&h->d_prof, (size_t)h->grid * 32 * sizeof(unsigned long long)));
            CUDA_TRY(cudaMemset(h->d_prof, 0, (size_t)h->grid * 32 * sizeof(unsigned long long)));
        }
    }
    if (h->grid > 8 * 64) { vox_set_error("dec_mega: grid too large"); return -1; }
    if (s->m->q8 && (VOCAB + h->grid - 1) / h->grid + 1 > mk::QROWS) { vox_set_error("dec_mega: Q8 needs >= %d SMs", VOCAB / (mk::QROWS - 2) + 1); return -1; }
    return 0;
}

void dec_mega_free(vox_cuda_state *s) {
    MegaHost *h = (MegaHost *)s->mega;
    if (!h) return;
    cudaFree(h->d_layers);
    cudaFree(h->d_bar);
    if (h->d_err) cudaFreeHost(h->d_err);
    if (h->d_prof) cudaFree(h->d_prof);
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
    P.tok_emb = m->tok_emb; P.q_tok_emb = m->q_tok_emb; P.s_tok_emb = m->s_tok_emb; P.qs_tok_emb = m->qs_tok_emb;
    P.dec_norm = m->dec_norm; P.ada_scale = m->ada_scale; P.inv_freq = m->inv_freq_dec;
    P.kv_k = s->kv16 ? (void *)s->kv_k16 : (void *)s->kv_k; P.kv_v = s->kv16 ? (void *)s->kv_v16 : (void *)s->kv_v;
    P.attn = s->d_attn; P.attn_cnt = h->d_bar + 8;
    P.x = s->d_x; P.q = s->d_q; P.attn_part = s->d_attn_part; P.h = s->d_h; P.logits = s->d_logits;
    P.blockmax = s->d_blockmax;
    P.embed_in = s->d_embed_in; P.adapter = s->adapter_buf;
    P.params = s->d_params; P.tokens = s->d_tokens;
    P.alt_cand = s->d_alt_cand; P.alt_out = s->d_alt; P.alt_on = (s->alt_n > 1 && s->d_alt) ? 1 : 0;
    P.bar = h->d_bar; P.bar_base = h->bar_count; P.n_steps = n_steps; P.inflight = h->inflight; P.l2_ahead = h->l2_ahead; P.spin_ns = h->spin_ns; P.err = h->d_err; P.prof = h->d_prof;
    CUDA_TRY(cudaMemsetAsync(h->d_bar + 8, 0, mk::SYNC_WORDS * sizeof(unsigned), s->stream));   // hand-over counters start at 0
    void *args[] = {&P};
    const void *fn = m->q8 ? (s->kv16 ? (const void *)mk::dec_mega_kernel<__half, true> : (const void *)mk::dec_mega_kernel<float, true>)
                           : (s->kv16 ? (const void *)mk::dec_mega_kernel<__half, false> : (const void *)mk::dec_mega_kernel<float, false>);
    CUDA_TRY(cudaLaunchCooperativeKernel(fn, dim3(h->grid), dim3(mk::NT), args,
                                         m->q8 ? sizeof(mk::SmemQ) : sizeof(mk::Smem), s->stream));
    h->bar_count += (unsigned)h->grid * (unsigned)h->barriers_per_step * (unsigned)n_steps;
    LAUNCHED();
    return 0;
}

int dec_mega_attn_part_floats(void) { return mk::MAX_SPLITS * DEC_HEADS * mk::PART_LD; }

// Debug: cycle counters of the last launch, averaged over CTAs:
// out[0..7] = total, waiting for weight tiles, grid barriers, attention, prologues, barrier spin, atomic round trip, Q8 digit prep.
extern "C" int vox_cuda_debug_mega_profile(vox_cuda_state_t *s, double *out) {
    MegaHost *h = s ? (MegaHost *)s->mega : nullptr;
    if (!h || !h->d_prof) return -1;
    const int n = h->grid * 32;
    unsigned long long *buf = (unsigned long long *)malloc((size_t)n * 8);
    if (cudaMemcpy(buf, h->d_prof, (size_t)n * 8, cudaMemcpyDeviceToHost) != cudaSuccess) { free(buf); return -1; }
    for (int k = 0; k < 8; k++) {
        double a = 0;
        for (int c = 0; c < h->grid; c++) a += (double)buf[c * 32 + k];
        out[k] = a / h->grid;
    }
    free(buf);
    return 0;
}

// Watchdog record of the megakernel (code 1 = tile mbarrier wait timed out, 2 = grid barrier timed out).
const unsigned *dec_mega_error_record(vox_cuda_state *s) {
    MegaHost *h = s ? (MegaHost *)s->mega : nullptr;
    return h ? h->d_err : nullptr;
}

// Debug: arrival / exit timestamps (ns) of barriers 0..11 counted from layer 10 of the last step; out[grid][24]
extern "C" int vox_cuda_debug_mega_barriers(vox_cuda_state_t *s, unsigned long long *out, int *grid) {
    MegaHost *h = s ? (MegaHost *)s->mega : nullptr;
    if (!h || !h->d_prof) return -1;
    const int n = h->grid * 32;
    unsigned long long *buf = (unsigned long long *)malloc((size_t)n * 8);
    if (cudaMemcpy(buf, h->d_prof, (size_t)n * 8, cudaMemcpyDeviceToHost) != cudaSuccess) { free(buf); return -1; }
    for (int c = 0; c < h->grid; c++)
        for (int k = 0; k < 24; k++) out[c * 24 + k] = buf[c * 32 + 8 + k];
    *grid = h->grid;
    free(buf);
    return 0;
}
