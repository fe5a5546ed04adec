// gemm_wt.cu - "weights as the M operand" skinny GEMM for lock-step multi-stream decoding.
//
//   out[b][n] = sum_k act[b][k] * W[n][k]      b < B <= NB streams, W bf16 [N,K] row-major
//
// This is vox_linear_nobias_bf16 (/root/reference/voxtral_kernels.c:215-236) applied to the B
// activation rows of B independent streams at once, so that the weights - the 6.86 GB per token
// that bound the single-stream decoder (voxtral_kernels.c:154-195) - cross HBM once per step for
// all streams.  The operand roles are swapped with respect to gemm_tc.cu: the 128 x 64 weight tile
// is the UMMA A operand (M = 128 TMEM lanes = 128 output features) and the activations are the
// narrow B operand (N = NB streams), so a CTA moves 16 KB of weights per 2 * NB * 128 B of
// L2-resident activations.  Activations are split bf16 pairs [hi(K) | lo(K)] (a = hi + lo to ~16
// mantissa bits): every weight tile is multiplied by both halves into the same fp32 TMEM
// accumulator, which keeps fp32-class accuracy (greedy tokens stay identical to the fp32 path).
//
// HBM-bound by construction, so the work is cut for bandwidth, not for the tensor pipe: every
// matrix is split into (N/128 tiles) x (ksplit K-ranges) >= ~144 items so that all 148 SMs pull
// weights; K-range partial sums go to out[split][b][n] and are added in fixed order by the
// consumer kernel (batch_dec.cu), which keeps the result deterministic.
//
//   warp 0   TMA producer (one lane): weight tile + hi/lo activation tiles per stage, mbarrier tx
//   warp 1   MMA issuer (one lane): 2 x 4 tcgen05.mma (M=128, N=NB, K=16) per stage, commit
//   warps 2-5 epilogue: tcgen05.ld of one TMEM lane quarter each -> partial fp32 / SwiGLU bf16x2
#include "vox_internal.h"
#include "gemm_wt.h"
#include "tc_ptx.cuh"
#include <cudaTypedefs.h>

namespace wt {
using namespace tcx;

constexpr int BMW = 128, BK = 64, ACC = 2, NT = 192;
constexpr int A_BYTES = BMW * BK * 2;                       // 16 KB of weights per stage

__device__ __forceinline__ void bulk_1d(void *dst, const void *src, unsigned bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(s32(dst)), "l"(src), "r"(bytes), "r"(s32(bar)) : "memory");
}

template <int NB> struct Cfg {
    static constexpr int B_BYTES = NB * BK * 2;             // one activation half (hi or lo)
    static constexpr int STAGE = A_BYTES + 2 * B_BYTES;
    static constexpr int STAGES = (196608 / STAGE) > 8 ? 8 : (196608 / STAGE);
    static constexpr int ACC_COLS = 2 * NB;                 // [hi products | lo products], added in the epilogue
    static constexpr int TMEM_COLS = ACC * ACC_COLS;        // 128 / 256 / 512
    static constexpr int SMEM = STAGES * STAGE + 512 + 1024;
};

struct KParams {
    int n_tiles, nkb, ksplit, B, split, mode;
    const uint8_t *wt, *at;                                 // tiled operands (see gemm_wt.h); null: tensor maps
    float *out_f32; long ldo; long split_stride;            // mode 0: out_f32[s*split_stride + b*ldo + n]
    uint16_t *out_bf16; long ldb; long lo_off;              // mode 1: rows (2i, 2i+1) = (gate, up)
    const float *row_scale;                                 // Q8: scale of output feature n, or null
};

template <int NB>
__global__ void __launch_bounds__(NT, 1) gemm_wt_kernel(const __grid_constant__ CUtensorMap tmW,
                                                        const __grid_constant__ CUtensorMap tmA, const KParams p) {
    using C = Cfg<NB>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t *base = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t *full = reinterpret_cast<uint64_t *>(base + C::STAGES * C::STAGE);
    uint64_t *empty = full + C::STAGES, *tfull = empty + C::STAGES, *tempty = tfull + ACC;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + ACC);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int i = 0; i < C::STAGES; i++) { mb_init(&full[i], 1); mb_init(&empty[i], 1); }
        for (int i = 0; i < ACC; i++) { mb_init(&tfull[i], 1); mb_init(&tempty[i], 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmW) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(tmem_slot)), "n"(C::TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const int total = p.n_tiles * p.ksplit;
    const int Kw = p.nkb * BK;                              // activation lo half starts at column Kw

    pdl_launch_dependents();                                // the next kernel may start its own prologue
    if (warp == 0) {
        if (lane == 0) {
            int s = 0; unsigned ph = 0;
            const unsigned tx = A_BYTES + (p.split ? 2 : 1) * C::B_BYTES;
            // Weights do not depend on the previous kernel: the first ring of weight tiles is requested
            // before the dependency wait, so HBM keeps streaming across the kernel boundary.
            int pre = 0;
            if (blockIdx.x < total) {
                const int tile = blockIdx.x / p.ksplit, sp = blockIdx.x % p.ksplit;
                const int kb0 = (int)((long)sp * p.nkb / p.ksplit), kb1 = (int)((long)(sp + 1) * p.nkb / p.ksplit);
                pre = min(C::STAGES, kb1 - kb0);
                for (int i = 0; i < pre; i++) {
                    uint8_t *st = base + i * C::STAGE;
                    mb_expect_tx(&full[i], tx);
                    if (p.wt) bulk_1d(st, p.wt + ((size_t)tile * p.nkb + kb0 + i) * A_BYTES, A_BYTES, &full[i]);
                    else tma_2d(st, &tmW, (kb0 + i) * BK, tile * BMW, &full[i]);
                }
            }
            pdl_wait();
            for (int it = blockIdx.x; it < total; it += gridDim.x) {
                const int tile = it / p.ksplit, sp = it % p.ksplit;
                const int kb0 = (int)((long)sp * p.nkb / p.ksplit), kb1 = (int)((long)(sp + 1) * p.nkb / p.ksplit);
                for (int kb = kb0; kb < kb1; kb++) {
                    uint8_t *st = base + s * C::STAGE;
                    const bool prefetched = pre > 0;        // weight tile and expect_tx already issued above
                    if (prefetched) pre--;
                    else {
                        mb_wait(&empty[s], ph ^ 1u);
                        mb_expect_tx(&full[s], tx);
                    }
                    if (p.wt) {
                        // tiled operands: one contiguous 16 KB weight tile + one contiguous [hi | lo] activation tile
                        if (!prefetched) bulk_1d(st, p.wt + ((size_t)tile * p.nkb + kb) * A_BYTES, A_BYTES, &full[s]);
                        bulk_1d(st + A_BYTES, p.at + (size_t)kb * (2 * C::B_BYTES), (p.split ? 2 : 1) * C::B_BYTES, &full[s]);
                    } else {
                        if (!prefetched) tma_2d(st, &tmW, kb * BK, tile * BMW, &full[s]);
                        tma_2d(st + A_BYTES, &tmA, kb * BK, 0, &full[s]);
                        if (p.split) tma_2d(st + A_BYTES + C::B_BYTES, &tmA, Kw + kb * BK, 0, &full[s]);
                    }
                    if (++s == C::STAGES) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // hi and lo activation tiles are adjacent in the stage, so one MMA with N = 2 NB multiplies the
        // weight tile by both halves (the A-operand read, not N, sets the cost of a narrow MMA)
        constexpr uint32_t IDESC2 = umma_idesc(BMW, 2 * NB), IDESC1 = umma_idesc(BMW, NB);
        int s = 0; unsigned ph = 0;
        int n = 0;
        for (int it = blockIdx.x; it < total; it += gridDim.x, n++) {
            const int sp = it % p.ksplit;
            const int kb0 = (int)((long)sp * p.nkb / p.ksplit), kb1 = (int)((long)(sp + 1) * p.nkb / p.ksplit);
            const int as = n & 1;
            const unsigned aph = (unsigned)(n >> 1) & 1u;
            mb_wait(&tempty[as], aph ^ 1u);                 // epilogue has drained this accumulator
            tc_fence_after();
            const uint32_t d = tmem_base + (uint32_t)(as * C::ACC_COLS);
            const uint32_t idesc = p.split ? IDESC2 : IDESC1;
            for (int kb = kb0; kb < kb1; kb++) {
                mb_wait(&full[s], ph);
                tc_fence_after();
                if (lane == 0) {
                    const uint32_t st = s32(base + s * C::STAGE);
                    const uint64_t da = umma_desc(st), dh = umma_desc(st + A_BYTES);
                    #pragma unroll
                    for (int k = 0; k < BK / 16; k++)       // +32 B per K=16 step inside the 128 B swizzle row
                        umma_f16(d, da + (uint64_t)(2 * k), dh + (uint64_t)(2 * k), idesc, (kb > kb0 || k) ? 1u : 0u);
                    umma_commit(&empty[s]);                 // stage reusable once these MMAs retire
                    if (kb == kb1 - 1) umma_commit(&tfull[as]);
                }
                __syncwarp();
                if (++s == C::STAGES) { s = 0; ph ^= 1u; }
            }
        }
    } else {
        const int q = warp & 3;                             // TMEM lane quarter this warp may access
        int n = 0;
        pdl_wait();                                         // outputs may still be read by an earlier kernel
        for (int it = blockIdx.x; it < total; it += gridDim.x, n++) {
            const int tile = it / p.ksplit, sp = it % p.ksplit;
            const int as = n & 1;
            const unsigned aph = (unsigned)(n >> 1) & 1u;
            const int row = tile * BMW + q * 32 + lane;     // output feature of this thread
            const float rsc = p.row_scale ? __ldg(p.row_scale + row) : 1.0f;
            mb_wait(&tfull[as], aph);
            tc_fence_after();
            #pragma unroll 1
            for (int cc = 0; cc < NB / 32; cc++) {
                uint32_t r[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * C::ACC_COLS + cc * 32), r);
                if (p.split) {
                    uint32_t r2[32];
                    tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * C::ACC_COLS + NB + cc * 32), r2);
                    #pragma unroll
                    for (int c = 0; c < 32; c++) r[c] = __float_as_uint(__uint_as_float(r[c]) + __uint_as_float(r2[c]));
                }
                if (p.row_scale) {
                    #pragma unroll
                    for (int c = 0; c < 32; c++) r[c] = __float_as_uint(__uint_as_float(r[c]) * rsc);
                }
                if (p.mode == 0) {
                    float *o = p.out_f32 + (size_t)sp * p.split_stride + row;
                    #pragma unroll
                    for (int c = 0; c < 32; c++)
                        if (cc * 32 + c < p.B) o[(size_t)(cc * 32 + c) * p.ldo] = __uint_as_float(r[c]);
                } else {
                    // SwiGLU (voxtral_decoder.c:747-757): even lane = gate row 2i, odd lane = up row 2i+1 of the
                    // same feature.  The pair shares the work: the even lane finishes column c, the odd lane c + 1.
                    const int odd = lane & 1;
                    #pragma unroll
                    for (int c = 0; c < 32; c += 2) {
                        const float v0 = __uint_as_float(r[c]), v1 = __uint_as_float(r[c + 1]);
                        const float recv = __shfl_xor_sync(0xffffffffu, odd ? v0 : v1, 1);
                        const float gte = odd ? recv : v0, up = odd ? v1 : recv;
                        const int col = cc * 32 + c + odd;
                        if (col < p.B) {
                            const float h = silu(gte) * up;
                            const __nv_bfloat16 hi = __float2bfloat16_rn(h);
                            const __nv_bfloat16 lo = __float2bfloat16_rn(h - __bfloat162float(hi));
                            __nv_bfloat16 *o = reinterpret_cast<__nv_bfloat16 *>(p.out_bf16) + (size_t)col * p.ldb + (row >> 1);
                            o[0] = hi;
                            o[p.lo_off] = lo;
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mb_arrive(&tempty[as]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(C::TMEM_COLS));
    }
}

}  // namespace wt

// ---- host ----------------------------------------------------------------------------------------
static PFN_cuTensorMapEncodeTiled_v12000 g_wt_encode = nullptr;
static int g_wt_sms = 0;

static int wt_init(void) {
    if (g_wt_encode) return 0;
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess || !fn) {
        vox_set_error("gemm_wt: cuTensorMapEncodeTiled not available from the driver");
        return -1;
    }
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    CUDA_TRY(cudaDeviceGetAttribute(&g_wt_sms, cudaDevAttrMultiProcessorCount, dev));
    CUDA_TRY(cudaFuncSetAttribute(wt::gemm_wt_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, wt::Cfg<32>::SMEM));
    CUDA_TRY(cudaFuncSetAttribute(wt::gemm_wt_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, wt::Cfg<64>::SMEM));
    CUDA_TRY(cudaFuncSetAttribute(wt::gemm_wt_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, wt::Cfg<128>::SMEM));
    g_wt_encode = (PFN_cuTensorMapEncodeTiled_v12000)fn;
    return 0;
}

static int wt_map(CUtensorMap *m, const uint16_t *ptr, int rows, long cols, long ld, int box_rows) {
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)wt::BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_wt_encode(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void *)ptr, dims, strides, box, estr,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                             CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        vox_set_error("gemm_wt: cuTensorMapEncodeTiled failed (%d) rows=%d cols=%ld ld=%ld", (int)r, rows, cols, ld);
        return -1;
    }
    return 0;
}

int gemm_wt_prepare(WtOp *op, int NB, int B, int N, int K, int split, const uint16_t *act, long lda,
                    const uint16_t *W, int ksplit) {
    if (wt_init()) return -1;
    if ((NB != 32 && NB != 64 && NB != 128) || B < 1 || B > NB || N % wt::BMW || K % wt::BK || ksplit < 1 ||
        ksplit > K / wt::BK || (lda * 2) % 16 || ((uintptr_t)act & 15) || ((uintptr_t)W & 15)) {
        vox_set_error("gemm_wt: unsupported shape NB=%d B=%d N=%d K=%d ksplit=%d lda=%ld", NB, B, N, K, ksplit, lda);
        return -1;
    }
    static_assert(sizeof(op->map_w) >= sizeof(CUtensorMap) && sizeof(op->map_a) >= sizeof(CUtensorMap), "map storage");
    if (wt_map(reinterpret_cast<CUtensorMap *>(op->map_w), W, N, K, K, wt::BMW)) return -1;
    if (wt_map(reinterpret_cast<CUtensorMap *>(op->map_a), act, NB, split ? 2L * K : (long)K, lda, NB)) return -1;
    op->NB = NB; op->B = B; op->N = N; op->K = K; op->split = split; op->ksplit = ksplit;
    op->mode = WT_OUT_F32; op->out_f32 = nullptr; op->ldo = 0; op->split_stride = 0;
    op->out_bf16 = nullptr; op->ldb = 0; op->lo_off = 0;
    op->w_tiled = nullptr; op->act_tiled = nullptr;
    op->row_scale = nullptr;
    return 0;
}

int gemm_wt_launch(cudaStream_t st, const WtOp *op) {
    wt::KParams p;
    p.n_tiles = op->N / wt::BMW; p.nkb = op->K / wt::BK; p.ksplit = op->ksplit; p.B = op->B; p.split = op->split;
    p.mode = op->mode; p.out_f32 = op->out_f32; p.ldo = op->ldo; p.split_stride = op->split_stride;
    p.out_bf16 = op->out_bf16; p.ldb = op->ldb; p.lo_off = op->lo_off;
    p.wt = (const uint8_t *)op->w_tiled; p.at = (const uint8_t *)op->act_tiled;
    p.row_scale = op->row_scale;
    if ((p.wt == nullptr) != (p.at == nullptr)) { vox_set_error("gemm_wt: tiled weights need tiled activations"); return -1; }
    if ((p.mode == WT_OUT_F32 && !p.out_f32) || (p.mode == WT_OUT_SWIGLU_BF16X2 && (!p.out_bf16 || p.ksplit != 1))) {
        vox_set_error("gemm_wt: epilogue not set");
        return -1;
    }
    const int total = p.n_tiles * p.ksplit;
    const int grid = total < g_wt_sms ? total : g_wt_sms;
    const CUtensorMap &mw = *reinterpret_cast<const CUtensorMap *>(op->map_w);
    const CUtensorMap &ma = *reinterpret_cast<const CUtensorMap *>(op->map_a);
    if (op->NB == 32) CUDA_TRY(launch_pdl(wt::gemm_wt_kernel<32>, grid, wt::NT, wt::Cfg<32>::SMEM, st, mw, ma, p));
    else if (op->NB == 64) CUDA_TRY(launch_pdl(wt::gemm_wt_kernel<64>, grid, wt::NT, wt::Cfg<64>::SMEM, st, mw, ma, p));
    else CUDA_TRY(launch_pdl(wt::gemm_wt_kernel<128>, grid, wt::NT, wt::Cfg<128>::SMEM, st, mw, ma, p));
    LAUNCHED();
    return 0;
}

// fp32 rows -> split bf16 rows [hi | lo] (also used by the op-level hook below)
__global__ void wt_split_rows_kernel(const float *__restrict__ in, uint16_t *__restrict__ out, int cols) {
    const float *ir = in + (size_t)blockIdx.x * cols;
    uint16_t *orow = out + (size_t)blockIdx.x * 2 * cols;
    for (int i = threadIdx.x * 2; i < cols; i += blockDim.x * 2) {
        const float2 v = *reinterpret_cast<const float2 *>(ir + i);
        const __nv_bfloat162 hi = __floats2bfloat162_rn(v.x, v.y);
        *reinterpret_cast<__nv_bfloat162 *>(orow + i) = hi;
        *reinterpret_cast<__nv_bfloat162 *>(orow + cols + i) =
            __floats2bfloat162_rn(v.x - __bfloat162float(hi.x), v.y - __bfloat162float(hi.y));
    }
}
__global__ void wt_sum_splits_kernel(const float *__restrict__ part, int n_split, size_t stride, float *__restrict__ out, size_t n) {
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    float a = 0.f;
    for (int s = 0; s < n_split; s++) a += part[s * stride + i];
    out[i] = a;
}

// C ABI (include/vox_cuda.h): the multi-stream product with host buffers, for parity tests of the
// kernel alone.  A [B,K] fp32, W [N,K] bf16, C [B,N] fp32 (or [B,N/2] with swiglu).
extern "C" int vox_cuda_gemm_bf16_streams(int B, int N, int K, const float *A, const uint16_t *W_bf16, int ksplit,
                                          int swiglu, float *C) {
    if (!A || !W_bf16 || !C || B <= 0 || B > 128) return -1;
    if (!vox_cuda_available() && !vox_cuda_init()) return -1;
    const int NB = B <= 32 ? 32 : (B <= 64 ? 64 : 128);
    const int nout = swiglu ? N / 2 : N;
    float *dA = nullptr, *dP = nullptr, *dC = nullptr;
    uint16_t *dAb = nullptr, *dW = nullptr, *dH = nullptr;
    uint16_t *hH = nullptr;
    int rc = -1;
    WtOp op;
    if (swiglu && ksplit != 1) { vox_set_error("gemm_bf16_streams: swiglu needs ksplit 1"); return -1; }
    if (cudaMalloc((void **)&dA, (size_t)B * K * 4) != cudaSuccess) goto done;
    if (cudaMalloc((void **)&dAb, (size_t)NB * 2 * K * 2) != cudaSuccess) goto done;
    if (cudaMemset(dAb, 0, (size_t)NB * 2 * K * 2) != cudaSuccess) goto done;
    if (cudaMalloc((void **)&dW, (size_t)N * K * 2) != cudaSuccess) goto done;
    if (cudaMalloc((void **)&dP, (size_t)ksplit * B * N * 4) != cudaSuccess) goto done;
    if (cudaMalloc((void **)&dC, (size_t)B * N * 4) != cudaSuccess) goto done;
    if (cudaMalloc((void **)&dH, (size_t)B * 2 * nout * 2) != cudaSuccess) goto done;
    if (cudaMemcpy(dA, A, (size_t)B * K * 4, cudaMemcpyHostToDevice) != cudaSuccess) goto done;
    if (cudaMemcpy(dW, W_bf16, (size_t)N * K * 2, cudaMemcpyHostToDevice) != cudaSuccess) goto done;
    wt_split_rows_kernel<<<B, 256>>>(dA, dAb, K);
    LAUNCHED();
    if (gemm_wt_prepare(&op, NB, B, N, K, 1, dAb, 2L * K, dW, ksplit)) goto done;
    if (swiglu) {
        op.mode = WT_OUT_SWIGLU_BF16X2; op.out_bf16 = dH; op.ldb = 2L * nout; op.lo_off = nout;
        if (gemm_wt_launch(0, &op)) goto done;
        hH = (uint16_t *)malloc((size_t)B * 2 * nout * 2);
        if (!hH || cudaMemcpy(hH, dH, (size_t)B * 2 * nout * 2, cudaMemcpyDeviceToHost) != cudaSuccess) goto done;
        for (int b = 0; b < B; b++)
            for (int i = 0; i < nout; i++) {
                uint32_t hi = (uint32_t)hH[(size_t)b * 2 * nout + i] << 16, lo = (uint32_t)hH[(size_t)b * 2 * nout + nout + i] << 16;
                float fh, fl;
                memcpy(&fh, &hi, 4); memcpy(&fl, &lo, 4);
                C[(size_t)b * nout + i] = fh + fl;
            }
    } else {
        op.out_f32 = dP; op.ldo = N; op.split_stride = (long)B * N;
        if (gemm_wt_launch(0, &op)) goto done;
        wt_sum_splits_kernel<<<(unsigned)(((size_t)B * N + 255) / 256), 256>>>(dP, ksplit, (size_t)B * N, dC, (size_t)B * N);
        LAUNCHED();
        if (cudaMemcpy(C, dC, (size_t)B * N * 4, cudaMemcpyDeviceToHost) != cudaSuccess) goto done;
    }
    rc = 0;
done:
    if (rc && !*vox_cuda_last_error())
        vox_set_error("vox_cuda_gemm_bf16_streams(%d,%d,%d) failed: %s", B, N, K, cudaGetErrorString(cudaGetLastError()));
    free(hH);
    cudaFree(dA); cudaFree(dAb); cudaFree(dW); cudaFree(dP); cudaFree(dC); cudaFree(dH);
    return rc;
}

// C ABI (include/vox_cuda.h): device-resident timing of the multi-stream product alone.  Weights
// rotate over enough copies (> 400 MB) that no launch finds its matrix in L2.  Returns the mean
// microseconds per launch over iters launches (CUDA events), or a negative value on error.
extern "C" float vox_cuda_bench_gemm_streams(int B, int N, int K, int ksplit, int swiglu, int iters) {
    if (B <= 0 || B > 128 || iters < 1) return -1.f;
    if (!vox_cuda_available() && !vox_cuda_init()) return -1.f;
    const int NB = B <= 32 ? 32 : (B <= 64 ? 64 : 128);
    const size_t wbytes = (size_t)N * K * 2;
    int copies = (int)((400u << 20) / wbytes) + 1;
    if (copies < 2) copies = 2;
    if (copies > 64) copies = 64;
    uint16_t *dW = nullptr, *dA = nullptr, *dH = nullptr;
    float *dP = nullptr;
    WtOp *ops = (WtOp *)aligned_alloc(64, sizeof(WtOp) * copies);
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    float us = -1.f, ms = 0.f;
    const int nout = swiglu ? N / 2 : N;
    const char *tenv = getenv("VOX_WT_BENCH_TILED");
    const bool tiled = tenv && atoi(tenv);
    if (!ops) return -1.f;
    if (cudaMalloc((void **)&dW, wbytes * copies) != cudaSuccess) goto done;
    if (cudaMemset(dW, 0x3c, wbytes * copies) != cudaSuccess) goto done;
    if (cudaMalloc((void **)&dA, (size_t)NB * 2 * K * 2) != cudaSuccess) goto done;
    if (cudaMemset(dA, 0x3c, (size_t)NB * 2 * K * 2) != cudaSuccess) goto done;
    if (cudaMalloc((void **)&dP, (size_t)ksplit * B * N * 4) != cudaSuccess) goto done;
    if (cudaMalloc((void **)&dH, (size_t)B * 2 * nout * 2) != cudaSuccess) goto done;
    for (int c = 0; c < copies; c++) {
        if (gemm_wt_prepare(&ops[c], NB, B, N, K, 1, dA, 2L * K, dW + (size_t)c * N * K, ksplit)) goto done;
        if (swiglu) { ops[c].mode = WT_OUT_SWIGLU_BF16X2; ops[c].out_bf16 = dH; ops[c].ldb = 2L * nout; ops[c].lo_off = nout; }
        else { ops[c].out_f32 = dP; ops[c].ldo = N; ops[c].split_stride = (long)B * N; }
        if (tiled) { ops[c].w_tiled = dW + (size_t)c * N * K; ops[c].act_tiled = dA; }   // same bytes read as tiles (timing only)
    }
    if (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess) goto done;
    for (int i = 0; i < copies; i++) if (gemm_wt_launch(0, &ops[i])) goto done;
    if (cudaEventRecord(e0, 0) != cudaSuccess) goto done;
    for (int i = 0; i < iters; i++) if (gemm_wt_launch(0, &ops[i % copies])) goto done;
    if (cudaEventRecord(e1, 0) != cudaSuccess || cudaEventSynchronize(e1) != cudaSuccess) goto done;
    cudaEventElapsedTime(&ms, e0, e1);
    us = ms * 1000.f / iters;
done:
    if (us < 0.f && !*vox_cuda_last_error()) vox_set_error("bench_gemm_streams: %s", cudaGetErrorString(cudaGetLastError()));
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    cudaFree(dW); cudaFree(dA); cudaFree(dP); cudaFree(dH);
    free(ops);
    return us;
}

// one thread per 16-byte chunk: (tile, kb, row, chunk j) <- W[tile*128 + row][kb*64 + 8j .. +8]
__global__ void wt_tile_weights_kernel(const uint4 *__restrict__ W, uint4 *__restrict__ dst, int N, int K) {
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;          // chunk index in the row-major source
    const size_t n_chunks = (size_t)N * K / 8;
    if (i >= n_chunks) return;
    const int cpr = K / 8;                                            // chunks per row
    const int row = (int)(i / cpr), ck = (int)(i % cpr);
    const int tile = row >> 7, r = row & 127, kb = ck >> 3, j = ck & 7;
    const size_t o = (((size_t)tile * (K / 64) + kb) * 128 + r) * 8 + (j ^ (r & 7));
    dst[o] = W[i];
}
int wt_tile_weights(cudaStream_t st, const uint16_t *W, uint16_t *dst, int N, int K) {
    if (N % 128 || K % 64) { vox_set_error("wt_tile_weights: N %% 128 / K %% 64"); return -1; }
    const size_t n_chunks = (size_t)N * K / 8;
    wt_tile_weights_kernel<<<(unsigned)((n_chunks + 255) / 256), 256, 0, st>>>((const uint4 *)W, (uint4 *)dst, N, K);
    LAUNCHED();
    CUDA_TRY(cudaGetLastError());
    return 0;
}
