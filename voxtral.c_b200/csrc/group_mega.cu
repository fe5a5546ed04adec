// group_mega.cu - the stream group's lock-step decode as ONE persistent cooperative kernel.
//
// Same arithmetic as the kernel-per-op path of api_group.cu (the device bodies are shared through
// group_dev.cuh and the GEMM is the swap-AB tcgen05 product of gemm_wt.cu), but the 185 dependent
// launches of a step become 185 grid barriers inside one kernel with one CTA per SM, so that
//   * the TMA producer keeps requesting WEIGHT tiles of the next product while the CTAs sit in a
//     barrier or in a norm / attention phase (activation tiles are requested once the barrier that
//     publishes them has been passed), and
//   * launch latency, per-kernel prologues (TMEM allocation, barrier init, descriptor fetch) and
//     kernel-completion flushes disappear.
//
// Roles per CTA (320 threads): warp 0 = TMA producer (one lane), warp 1 = MMA issuer (one lane),
// warps 2-9 = 8 worker warps: GEMM epilogues (two warps per TMEM lane quarter, half of the columns
// each), RMSNorm rows, attention tasks, argmax rows, and the grid barrier after every phase.
// Phases of one step (global phase index q; barrier q is passed when the counter reaches (q+1)*G):
//   per layer: N1 | G wqkv | A | G wo | N2 | G w1|w3 (SwiGLU) | G w2      then N3 | G logits | ARGMAX
#include "group.h"
#include "tc_ptx.cuh"
#include <cudaTypedefs.h>
#include <string.h>
#include <stdlib.h>

namespace gm {
using namespace tcx;
using namespace grp;

constexpr int BMW = 128, BK = 64, ACC = 2;
constexpr int NT = 320, NWORK = 256;
constexpr int A_BYTES = BMW * BK * 2;
constexpr int N_AMAP = 3;
constexpr int MAX_OPS = 4 * VOX_CUDA_DEC_LAYERS + 1;

template <int NB> struct Cfg {
    static constexpr int B_BYTES = NB * BK * 2;
    static constexpr int STAGE = A_BYTES + 2 * B_BYTES;
    static constexpr int STAGES = (196608 / STAGE) > 8 ? 8 : (196608 / STAGE);
    static constexpr int ACC_COLS = 2 * NB;
    static constexpr int TMEM_COLS = ACC * ACC_COLS;
};

struct alignas(64) OpDev {
    CUtensorMap wmap;
    int n_tiles, nkb, ksplit, mode, amap, pad0, pad1, pad2;
    float *out_f32; long ldo; long split_stride;
    uint16_t *out_bf16; long ldb; long lo_off;
};
struct alignas(64) Tables {
    CUtensorMap amap[N_AMAP];               // xn_b, att_b, hh_b
    OpDev ops[MAX_OPS];                     // 4 per layer (wqkv, wo, w13, w2) + logits
    const float *attn_norm[VOX_CUDA_DEC_LAYERS], *ffn_norm[VOX_CUDA_DEC_LAYERS];
};

struct MParams {
    const Tables *tab;
    int n_dec, B, n_steps, ring_mask;
    const float *dec_norm, *ada_scale, *inv_freq;
    const uint16_t *tok_emb;
    float *x, *part, *logits;
    uint16_t *xn_b, *att_b, *hh_b;
    float *kv_k, *kv_v; size_t kv_layer, kv_stream;
    const float *adapter; long adapter_stride;
    Slot *slots; int *tokens;
    unsigned *bar, *err;
    unsigned long long *prof;               // [grid][8] cycles: N, G epilogue wait+work, A, ARGMAX, barriers (VOX_CUDA_GROUP_PROF=1)
};

struct Misc {
    uint64_t full[8], empty[8], tfull[ACC], tempty[ACC];
    uint32_t tmem_slot;
    int pend_kb[8];
    float red[8];
    float sv[32];
    int si[32];
};

__device__ __forceinline__ void worker_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }
struct WorkerSync { __device__ __forceinline__ void operator()() const { worker_sync(); } };

__device__ __forceinline__ unsigned ld_acquire(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release(unsigned *p, unsigned v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void proxy_fence() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __noinline__ void die(unsigned *err, unsigned code, unsigned arg) {
    if (err) { err[0] = code; err[1] = arg; err[2] = blockIdx.x; err[3] = threadIdx.x; __threadfence_system(); }
    __trap();
}
__device__ __forceinline__ void wait_count(const unsigned *bar, unsigned target, unsigned *err, unsigned code) {
    if ((int)(ld_acquire(bar) - target) >= 0) return;
    const long long t0 = clock64();
    while ((int)(ld_acquire(bar) - target) < 0)
        if (clock64() - t0 > 3000000000LL) die(err, code, target);
}
// grid barrier of the worker warps: everything the CTA wrote in this phase is published
__device__ __forceinline__ void grid_bar(unsigned *bar, unsigned target, unsigned *err, int wtid) {
    proxy_fence();                          // generic-proxy writes vs the TMA reads of other CTAs' producers
    worker_sync();
    if (wtid == 0) {
        __threadfence();
        red_release(bar, 1u);
        wait_count(bar, target, err, 3u);
    }
    worker_sync();
}

template <int NB>
__global__ void __launch_bounds__(NT, 1) group_mega_kernel(const MParams P) {
    using C = Cfg<NB>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t *base = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    AttnSmem &at = *reinterpret_cast<AttnSmem *>(base + C::STAGES * C::STAGE);
    Misc &mi = *reinterpret_cast<Misc *>(base + C::STAGES * C::STAGE + ((sizeof(AttnSmem) + 127) & ~(size_t)127));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned G = gridDim.x;
    const Tables &T = *P.tab;
    if (threadIdx.x == 0) {
        for (int i = 0; i < C::STAGES; i++) { mb_init(&mi.full[i], 1); mb_init(&mi.empty[i], 1); }
        for (int i = 0; i < ACC; i++) { mb_init(&mi.tfull[i], 1); mb_init(&mi.tempty[i], 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&mi.tmem_slot)), "n"(C::TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = mi.tmem_slot;

    const int n_dec = P.n_dec;
    const unsigned pps = 7u * (unsigned)n_dec + 3u;         // phases (= barriers) per step
    const int n_gemm = 4 * n_dec + 1;                       // GEMM phases per step

    // GEMM phase g (0 .. n_gemm-1) of a step: op index and its phase index inside the step
    auto gemm_op = [&](int g) { return g < 4 * n_dec ? g : MAX_OPS - 1; };
    auto gemm_phase = [&](int g) -> unsigned {
        if (g == 4 * n_dec) return 7u * n_dec + 1u;
        const int l = g >> 2, j = g & 3;                    // j: 0 wqkv, 1 wo, 2 w13, 3 w2
        return 7u * l + (j == 0 ? 1u : j == 1 ? 3u : j == 2 ? 5u : 6u);
    };

    if (warp == 0) {
        if (lane == 0) {
            int s = 0; unsigned ph = 0;
            const unsigned tx = A_BYTES + 2 * C::B_BYTES;
            for (int step = 0; step < P.n_steps; step++) {
                for (int g = 0; g < n_gemm; g++) {
                    const OpDev &op = T.ops[gemm_op(g)];
                    const CUtensorMap *am = &T.amap[op.amap];
                    const unsigned need = ((unsigned)step * pps + gemm_phase(g)) * G;   // barriers before this phase
                    const int total = op.n_tiles * op.ksplit, nkb = op.nkb, ks = op.ksplit;
                    const int Kw = nkb * BK;
                    int pend = 0;
                    bool ready = false;
                    auto flush = [&]() {
                        for (int i = pend; i > 0; i--) {
                            const int st_i = (s - i + 2 * C::STAGES) % C::STAGES;
                            uint8_t *st = base + st_i * C::STAGE;
                            const int kb = mi.pend_kb[st_i];
                            tma_2d(st + A_BYTES, am, kb * BK, 0, &mi.full[st_i]);
                            tma_2d(st + A_BYTES + C::B_BYTES, am, Kw + kb * BK, 0, &mi.full[st_i]);
                        }
                        pend = 0;
                    };
                    for (int it = blockIdx.x; it < total; it += G) {
                        const int tile = it / ks, sp = it % ks;
                        const int kb0 = (int)((long)sp * nkb / ks), kb1 = (int)((long)(sp + 1) * nkb / ks);
                        for (int kb = kb0; kb < kb1; kb++) {
                            if (pend == C::STAGES) {            // the whole ring waits for activations
                                wait_count(P.bar, need, P.err, 4u);
                                ready = true;
                                proxy_fence();
                                flush();
                            }
                            mb_wait(&mi.empty[s], ph ^ 1u);
                            mb_expect_tx(&mi.full[s], tx);
                            tma_2d(base + s * C::STAGE, &op.wmap, kb * BK, tile * BMW, &mi.full[s]);
                            mi.pend_kb[s] = kb;
                            pend++;
                            if (++s == C::STAGES) { s = 0; ph ^= 1u; }
                            if (!ready && (int)(ld_acquire(P.bar) - need) >= 0) { ready = true; proxy_fence(); }
                            if (ready) flush();
                        }
                    }
                    if (pend) {
                        wait_count(P.bar, need, P.err, 5u);
                        proxy_fence();
                        flush();
                    }
                }
            }
        }
    } else if (warp == 1) {
        constexpr uint32_t IDESC2 = umma_idesc(BMW, 2 * NB);
        int s = 0; unsigned ph = 0;
        int n = 0;
        for (int step = 0; step < P.n_steps; step++) {
            for (int g = 0; g < n_gemm; g++) {
                const OpDev &op = T.ops[gemm_op(g)];
                const int total = op.n_tiles * op.ksplit, nkb = op.nkb, ks = op.ksplit;
                for (int it = blockIdx.x; it < total; it += G, n++) {
                    const int sp = it % ks;
                    const int kb0 = (int)((long)sp * nkb / ks), kb1 = (int)((long)(sp + 1) * nkb / ks);
                    const int as = n & 1;
                    const unsigned aph = (unsigned)(n >> 1) & 1u;
                    mb_wait(&mi.tempty[as], aph ^ 1u);
                    tc_fence_after();
                    const uint32_t d = tmem_base + (uint32_t)(as * C::ACC_COLS);
                    for (int kb = kb0; kb < kb1; kb++) {
                        mb_wait(&mi.full[s], ph);
                        tc_fence_after();
                        if (lane == 0) {
                            const uint32_t st = s32(base + s * C::STAGE);
                            const uint64_t da = umma_desc(st), dh = umma_desc(st + A_BYTES);
                            #pragma unroll
                            for (int k = 0; k < BK / 16; k++)
                                umma_f16(d, da + (uint64_t)(2 * k), dh + (uint64_t)(2 * k), IDESC2, (kb > kb0 || k) ? 1u : 0u);
                            umma_commit(&mi.empty[s]);
                            if (kb == kb1 - 1) umma_commit(&mi.tfull[as]);
                        }
                        __syncwarp();
                        if (++s == C::STAGES) { s = 0; ph ^= 1u; }
                    }
                }
            }
        }
    } else {
        const int wtid = threadIdx.x - 64;                  // 0..255
        const int q = warp & 3, half = (warp - 2) >> 2;     // TMEM lane quarter, column half
        const WorkerSync wsync;
        int n = 0;
        unsigned bars = 0;                                  // barriers passed so far
        // one GEMM phase: epilogues of this CTA's items
        auto gemm_epilogue = [&](const OpDev &op) {
            const int total = op.n_tiles * op.ksplit, ks = op.ksplit;
            for (int it = blockIdx.x; it < total; it += G, n++) {
                const int tile = it / ks, sp = it % ks;
                const int as = n & 1;
                const unsigned aph = (unsigned)(n >> 1) & 1u;
                const int row = tile * BMW + q * 32 + lane;
                mb_wait(&mi.tfull[as], aph);
                tc_fence_after();
                #pragma unroll 1
                for (int cc = half; cc < NB / 32; cc += 2) {
                    uint32_t r[32], r2[32];
                    tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * C::ACC_COLS + cc * 32), r);
                    tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * C::ACC_COLS + NB + cc * 32), r2);
                    #pragma unroll
                    for (int c = 0; c < 32; c++) r[c] = __float_as_uint(__uint_as_float(r[c]) + __uint_as_float(r2[c]));
                    if (op.mode == WT_OUT_F32) {
                        float *o = op.out_f32 + (size_t)sp * op.split_stride + row;
                        #pragma unroll
                        for (int c = 0; c < 32; c++)
                            if (cc * 32 + c < P.B) o[(size_t)(cc * 32 + c) * op.ldo] = __uint_as_float(r[c]);
                    } else {
                        const int odd = lane & 1;
                        #pragma unroll
                        for (int c = 0; c < 32; c += 2) {
                            const float v0 = __uint_as_float(r[c]), v1 = __uint_as_float(r[c + 1]);
                            const float recv = __shfl_xor_sync(0xffffffffu, odd ? v0 : v1, 1);
                            const float gte = odd ? recv : v0, up = odd ? v1 : recv;
                            const int col = cc * 32 + c + odd;
                            if (col < P.B) {
                                const float h = silu(gte) * up;
                                const __nv_bfloat16 hi = __float2bfloat16_rn(h);
                                const __nv_bfloat16 lo = __float2bfloat16_rn(h - __bfloat162float(hi));
                                __nv_bfloat16 *o = reinterpret_cast<__nv_bfloat16 *>(op.out_bf16) + (size_t)col * op.ldb + (row >> 1);
                                o[0] = hi;
                                o[op.lo_off] = lo;
                            }
                        }
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mb_arrive(&mi.tempty[as]);
            }
        };
        long long t_n = 0, t_g = 0, t_a = 0, t_x = 0, t_b = 0, t0_;
        #define GM_T0() (t0_ = clock64())
        #define GM_ACC(v) ((v) += clock64() - t0_)
        #define GM_BAR() do { GM_T0(); grid_bar(P.bar, (++bars) * G, P.err, wtid); GM_ACC(t_b); } while (0)
        const int b_own = blockIdx.x;                       // row handled by this CTA in N / ARGMAX phases
        for (int step = 0; step < P.n_steps; step++) {
            for (int l = 0; l < n_dec; l++) {
                // N1
                GM_T0();
                if (b_own < P.B) {
                    if (l == 0)
                        norm_row(b_own, wtid, P.x, 0, nullptr, 0, 0L, P.adapter, P.adapter_stride, P.tok_emb, nullptr, P.slots,
                                 T.attn_norm[l], nullptr, P.xn_b, mi.red, wsync);
                    else
                        norm_row(b_own, wtid, P.x, 1, P.part, KS_W2, (long)P.B * DEC_DIM, nullptr, 0L, nullptr, nullptr, P.slots,
                                 T.attn_norm[l], nullptr, P.xn_b, mi.red, wsync);
                }
                GM_ACC(t_n);
                GM_BAR();
                GM_T0(); gemm_epilogue(T.ops[4 * l + 0]); GM_ACC(t_g);
                GM_BAR();
                // A
                GM_T0();
                for (int t = blockIdx.x; t < P.B * DEC_KVH; t += G)
                    attn_task(t >> 3, t & 7, wtid, P.part, KS_QKV, (long)P.B * DEC_QKV, P.slots, P.inv_freq,
                              P.kv_k + l * P.kv_layer, P.kv_v + l * P.kv_layer, P.kv_stream, P.ring_mask, P.att_b, at, wsync);
                GM_ACC(t_a);
                GM_BAR();
                GM_T0(); gemm_epilogue(T.ops[4 * l + 1]); GM_ACC(t_g);
                GM_BAR();
                // N2
                GM_T0();
                if (b_own < P.B)
                    norm_row(b_own, wtid, P.x, 1, P.part, KS_WO, (long)P.B * DEC_DIM, nullptr, 0L, nullptr, nullptr, P.slots,
                             T.ffn_norm[l], P.ada_scale + (size_t)l * DEC_DIM, P.xn_b, mi.red, wsync);
                GM_ACC(t_n);
                GM_BAR();
                GM_T0(); gemm_epilogue(T.ops[4 * l + 2]); GM_ACC(t_g);
                GM_BAR();
                GM_T0(); gemm_epilogue(T.ops[4 * l + 3]); GM_ACC(t_g);
                GM_BAR();
            }
            GM_T0();
            if (b_own < P.B)
                norm_row(b_own, wtid, P.x, 1, P.part, KS_W2, (long)P.B * DEC_DIM, nullptr, 0L, nullptr, nullptr, P.slots,
                         P.dec_norm, nullptr, P.xn_b, mi.red, wsync);
            GM_ACC(t_n);
            GM_BAR();
            GM_T0(); gemm_epilogue(T.ops[MAX_OPS - 1]); GM_ACC(t_x);
            GM_BAR();
            GM_T0();
            if (b_own < P.B)
                argmax_row<NWORK>(b_own, wtid, P.logits, P.slots, P.tokens, step, P.B, mi.sv, mi.si, wsync);
            GM_ACC(t_x);
            GM_BAR();
        }
        if (P.prof && wtid == 0) {
            unsigned long long *o = P.prof + (size_t)blockIdx.x * 8;
            o[0] = t_n; o[1] = t_g; o[2] = t_a; o[3] = t_x; o[4] = t_b;
        }
        #undef GM_BAR
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(C::TMEM_COLS));
    }
}

template <int NB> constexpr size_t smem_bytes() {
    return (size_t)Cfg<NB>::STAGES * Cfg<NB>::STAGE + ((sizeof(AttnSmem) + 127) & ~(size_t)127) + sizeof(Misc) + 1024 + 128;
}

}  // namespace gm

struct GroupMega {
    gm::Tables *d_tab;
    unsigned *d_bar, *d_err;     // d_err: mapped host memory (readable after a device trap)
    unsigned long long *d_prof;
    int grid;
};

void group_mega_free(vox_cuda_group *g) {
    GroupMega *h = (GroupMega *)g->mega;
    if (!h) return;
    cudaFree(h->d_tab); cudaFree(h->d_bar); cudaFree(h->d_prof);
    if (h->d_err) cudaFreeHost(h->d_err);
    free(h);
    g->mega = nullptr;
}

template <int NB> static int mega_setup(int *occ) {
    CUDA_TRY(cudaFuncSetAttribute(gm::group_mega_kernel<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gm::smem_bytes<NB>()));
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, gm::group_mega_kernel<NB>, gm::NT, gm::smem_bytes<NB>()));
    return 0;
}

int group_mega_init(vox_cuda_group *g) {
    // Opt-in (VOX_CUDA_GROUP_MEGA=1): measured 4.33 ms per 64-stream step against 3.82 ms for the kernel-per-op
    // path with programmatic dependent launch (one CTA per SM halves the attention phase's loads in flight and
    // the 185 grid barriers cost what the PDL launches cost), so the per-op path is the default.
    const char *env = getenv("VOX_CUDA_GROUP_MEGA");
    if (!env || !atoi(env)) return -1;
    if (g->m->q8) return -1;                     // the persistent step kernel has no row-scale epilogue: Q8 groups run per op
    int dev = 0, sms = 0, coop = 0, occ = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    CUDA_TRY(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
    if (!coop || g->B > sms) return -1;
    if (g->NB == 32) { if (mega_setup<32>(&occ)) return -1; }
    else if (g->NB == 64) { if (mega_setup<64>(&occ)) return -1; }
    else { if (mega_setup<128>(&occ)) return -1; }
    if (occ < 1) return -1;
    GroupMega *h = (GroupMega *)calloc(1, sizeof(GroupMega));
    if (!h) return -1;
    g->mega = h;
    h->grid = sms;
    gm::Tables *t = (gm::Tables *)aligned_alloc(64, sizeof(gm::Tables));
    if (!t) { group_mega_free(g); return -1; }
    memset(t, 0, sizeof(gm::Tables));
    memcpy(&t->amap[0], g->op_qkv[0].map_a, sizeof(CUtensorMap));
    memcpy(&t->amap[1], g->op_wo[0].map_a, sizeof(CUtensorMap));
    memcpy(&t->amap[2], g->op_w2[0].map_a, sizeof(CUtensorMap));
    auto fill = [&](gm::OpDev &o, const WtOp &w, int amap) {
        memcpy(&o.wmap, w.map_w, sizeof(CUtensorMap));
        o.n_tiles = w.N / 128; o.nkb = w.K / 64; o.ksplit = w.ksplit; o.mode = w.mode; o.amap = amap;
        o.out_f32 = w.out_f32; o.ldo = w.ldo; o.split_stride = w.split_stride;
        o.out_bf16 = w.out_bf16; o.ldb = w.ldb; o.lo_off = w.lo_off;
    };
    for (int l = 0; l < g->m->n_dec; l++) {
        fill(t->ops[4 * l + 0], g->op_qkv[l], 0);
        fill(t->ops[4 * l + 1], g->op_wo[l], 1);
        fill(t->ops[4 * l + 2], g->op_w13[l], 0);
        fill(t->ops[4 * l + 3], g->op_w2[l], 2);
        t->attn_norm[l] = g->m->dec[l].attn_norm;
        t->ffn_norm[l] = g->m->dec[l].ffn_norm;
    }
    fill(t->ops[gm::MAX_OPS - 1], g->op_logits, 0);
    bool ok = cudaMalloc((void **)&h->d_tab, sizeof(gm::Tables)) == cudaSuccess &&
              cudaMemcpy(h->d_tab, t, sizeof(gm::Tables), cudaMemcpyHostToDevice) == cudaSuccess &&
              cudaMalloc((void **)&h->d_bar, 8) == cudaSuccess &&
              cudaHostAlloc((void **)&h->d_err, 16, cudaHostAllocMapped) == cudaSuccess;
    free(t);
    if (!ok) { vox_set_error("group_mega_init: %s", cudaGetErrorString(cudaGetLastError())); group_mega_free(g); return -1; }
    memset(h->d_err, 0, 16);
    const char *pe = getenv("VOX_CUDA_GROUP_PROF");
    if (pe && atoi(pe)) {
        if (cudaMalloc((void **)&h->d_prof, (size_t)h->grid * 64) != cudaSuccess) h->d_prof = nullptr;
        else cudaMemset(h->d_prof, 0, (size_t)h->grid * 64);
    }
    return 0;
}

int group_mega_run(vox_cuda_group *g, int n_steps) {
    GroupMega *h = (GroupMega *)g->mega;
    if (!h) return -1;
    cudaStream_t st = g->eng->stream;
    vox_cuda_model *m = g->m;
    gm::MParams P;
    P.tab = h->d_tab; P.n_dec = m->n_dec; P.B = g->B; P.n_steps = n_steps; P.ring_mask = g->mask;
    P.dec_norm = m->dec_norm; P.ada_scale = m->ada_scale; P.inv_freq = m->inv_freq_dec; P.tok_emb = m->tok_emb;
    P.x = g->x; P.part = g->part; P.logits = g->logits; P.xn_b = g->xn_b; P.att_b = g->att_b; P.hh_b = g->hh_b;
    P.kv_k = g->kv_k; P.kv_v = g->kv_v; P.kv_layer = g->kv_layer; P.kv_stream = g->kv_stream;
    P.adapter = g->adapter; P.adapter_stride = (long)g->max_pos * DEC_DIM;
    P.slots = g->d_slots; P.tokens = g->d_tokens;
    P.bar = h->d_bar;
    unsigned *err_dev = nullptr;
    CUDA_TRY(cudaHostGetDevicePointer((void **)&err_dev, h->d_err, 0));
    P.err = err_dev;
    P.prof = h->d_prof;
    CUDA_TRY(cudaMemsetAsync(h->d_bar, 0, 8, st));
    void *args[] = {&P};
    cudaError_t e;
    if (g->NB == 32) e = cudaLaunchCooperativeKernel((void *)gm::group_mega_kernel<32>, dim3(h->grid), dim3(gm::NT), args, gm::smem_bytes<32>(), st);
    else if (g->NB == 64) e = cudaLaunchCooperativeKernel((void *)gm::group_mega_kernel<64>, dim3(h->grid), dim3(gm::NT), args, gm::smem_bytes<64>(), st);
    else e = cudaLaunchCooperativeKernel((void *)gm::group_mega_kernel<128>, dim3(h->grid), dim3(gm::NT), args, gm::smem_bytes<128>(), st);
    if (e != cudaSuccess) { vox_set_error("group_mega_run: %s", cudaGetErrorString(e)); return -1; }
    LAUNCHED();
    return 0;
}

// watchdog record of the last run (code, arg, cta, thread), all zero when nothing fired
const unsigned *group_mega_error_record(vox_cuda_group *g) {
    GroupMega *h = (GroupMega *)g->mega;
    return h ? h->d_err : nullptr;
}

// VOX_CUDA_GROUP_PROF=1: per-phase worker cycles of the last run (mean over CTAs), after the stream has been synchronised
void group_mega_dump_prof(vox_cuda_group *g, int n_steps) {
    GroupMega *h = (GroupMega *)g->mega;
    if (!h || !h->d_prof) return;
    unsigned long long *p = (unsigned long long *)malloc((size_t)h->grid * 64);
    if (!p) return;
    if (cudaMemcpy(p, h->d_prof, (size_t)h->grid * 64, cudaMemcpyDeviceToHost) == cudaSuccess) {
        double a[5] = {0, 0, 0, 0, 0}, mx[5] = {0, 0, 0, 0, 0};
        for (int c = 0; c < h->grid; c++)
            for (int k = 0; k < 5; k++) { a[k] += (double)p[c * 8 + k]; if ((double)p[c * 8 + k] > mx[k]) mx[k] = (double)p[c * 8 + k]; }
        const char *nm[5] = {"norm", "gemm-epilogue(wait+store)", "attention", "logits+argmax", "barriers"};
        fprintf(stderr, "group_mega: worker cycles per step, mean over %d CTAs (max CTA):", h->grid);
        for (int k = 0; k < 5; k++) fprintf(stderr, "  %s %.0f (%.0f)", nm[k], a[k] / h->grid / n_steps, mx[k] / n_steps);
        fprintf(stderr, "\n");
    }
    free(p);
}
