// gemm_tc.cu - bf16 x bf16 -> fp32 GEMM on the 5th-generation tensor cores.
//
//   C[M,N] = epi( A[M,K] (bf16, row stride lda) x W[N,K]^T (bf16) )
//
// This is the throughput matmul of the encoder (voxtral_encoder.c:562-686 is a
// chain of vox_linear*_bf16 calls, i.e. sgemm on widened weights,
// voxtral_kernels.c:197-264): QKV, wo, w1|w3, w2, the conv stem as an
// im2col-free GEMM (overlapping A rows through the TMA row stride) and the
// adapter.  Activations are rounded to bf16 (the documented 2e-2 budget);
// accumulation is fp32 in TMEM.
//
// Persistent, warp-specialised, one CTA per SM:
//   warp 0   TMA producer: cp.async.bulk.tensor 2-D tiles (128 x 64 bf16, SWIZZLE_128B)
//            of A and W into a 6-stage shared-memory ring, mbarrier complete_tx
//   warp 1   MMA issuer: one elected lane issues tcgen05.mma.cta_group::1.kind::f16
//            (M=128, N=128, K=16) x4 per stage from UMMA smem descriptors into one
//            of two 128-column TMEM accumulators; tcgen05.commit frees the stage /
//            publishes the accumulator
//   warps 2-5 epilogue: tcgen05.ld 32x32b.x32 (one TMEM lane quarter each), then
//            bias / GELU / residual add / SwiGLU / bf16 cast straight to global,
//            overlapping the next tile's main loop (double-buffered TMEM).
#include "vox_internal.h"
#include "gemm_tc.h"
#include "tc_ptx.cuh"
#include <cuda.h>
#include <cudaTypedefs.h>
#include <string.h>

namespace tc {

constexpr int BM = 128, BK = 64;
constexpr int ACC = 2;
constexpr int A_BYTES = BM * BK * 2;                              // 16 KB per activation half (hi or lo)
constexpr int NT = 192;
// One stage = [A hi | A lo | W]: the weight k-block is fetched once and multiplied by both activation
// halves.  What bounds this kernel is the bytes an SM can take in per FLOP (measured ~60 GB/s per SM
// whatever the source), so the tile is as wide as TMEM allows: BN = 256 -> 64 KB per 8.4 MFLOP.
template <int BN> struct Cfg {
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE = 2 * A_BYTES + B_BYTES;
    static constexpr int STAGES = 196608 / STAGE;                 // 3 (BN 256) or 4 (BN 128)
    static constexpr int SMEM_BYTES = STAGES * STAGE + 256 + 1024; // + barriers + alignment slack
    static constexpr uint32_t IDESC = (1u << 4) /*D=f32*/ | (1u << 7) /*A=bf16*/ | (1u << 10) /*B=bf16*/ |
                                      ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);   // K-major A and B
};

// mbarrier / TMA / UMMA-descriptor / tcgen05 wrappers: tc_ptx.cuh (shared with gemm_wt.cu)
using namespace tcx;

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&v);
}

struct KParams {
    int lo_sep;             // 1: the lo half of A is a separate array with its own tensor map (overlapping rows)
    int M, N, K, Kw;        // K = walked A columns (Kw or 2*Kw), Kw = weight columns
    long split_off;
    float *out_f32;
    long ldo_f32;
    uint16_t *out_bf16;
    long ldo_bf16;
    const float *bias;
    int flags;
    const float *col_scale;
};

// bf16 store of NV consecutive values; TC_OUT_SPLIT adds the rounding residual lo = bf16(v - hi)
template <int NV>
__device__ __forceinline__ void store_bf16(const KParams &p, int row, int col, const float (&v)[NV]) {
    uint32_t hi[NV / 2], lo[NV / 2];
    #pragma unroll
    for (int i = 0; i < NV / 2; i++) {
        hi[i] = pack_bf16(v[2 * i], v[2 * i + 1]);
        lo[i] = pack_bf16(v[2 * i] - bf16_lo(hi[i]), v[2 * i + 1] - bf16_hi(hi[i]));
    }
    uint4 *o = reinterpret_cast<uint4 *>(p.out_bf16 + (size_t)row * p.ldo_bf16 + col);
    #pragma unroll
    for (int i = 0; i < NV / 8; i++) o[i] = make_uint4(hi[4 * i], hi[4 * i + 1], hi[4 * i + 2], hi[4 * i + 3]);
    if (p.flags & TC_OUT_SPLIT) {
        uint4 *ol = reinterpret_cast<uint4 *>(p.out_bf16 + (size_t)row * p.ldo_bf16 + p.split_off + col);
        #pragma unroll
        for (int i = 0; i < NV / 8; i++) ol[i] = make_uint4(lo[4 * i], lo[4 * i + 1], lo[4 * i + 2], lo[4 * i + 3]);
    }
}

// epilogue of one thread: row `row`, 32 consecutive columns starting at `col`
__device__ __forceinline__ void epilogue_row(const KParams &p, int row, int col, const uint32_t (&r)[32]) {
    float v[32];
    #pragma unroll
    for (int i = 0; i < 32; i++) v[i] = __uint_as_float(r[i]);
    if (p.col_scale) {
        #pragma unroll
        for (int i = 0; i < 32; i += 4) {
            const float4 s = *reinterpret_cast<const float4 *>(p.col_scale + col + i);
            v[i] *= s.x; v[i + 1] *= s.y; v[i + 2] *= s.z; v[i + 3] *= s.w;
        }
    }
    if (p.flags & TC_BIAS) {
        #pragma unroll
        for (int i = 0; i < 32; i += 4) {
            const float4 b = *reinterpret_cast<const float4 *>(p.bias + col + i);
            v[i] += b.x; v[i + 1] += b.y; v[i + 2] += b.z; v[i + 3] += b.w;
        }
    }
    if (p.flags & TC_GELU) {
        #pragma unroll
        for (int i = 0; i < 32; i++) v[i] = gelu_tanh(v[i]);
    }
    if (p.flags & TC_SWIGLU) {
        // 16 outputs: h[col/2 + i] = silu(v[2i]) * v[2i+1]   (voxtral_encoder.c:670-676)
        float h[16];
        #pragma unroll
        for (int i = 0; i < 16; i++) h[i] = silu(v[2 * i]) * v[2 * i + 1];
        if (p.flags & TC_OUT_BF16) store_bf16<16>(p, row, col >> 1, h);
        if (p.flags & TC_OUT_F32) {
            float4 *o = reinterpret_cast<float4 *>(p.out_f32 + (size_t)row * p.ldo_f32 + (col >> 1));
            #pragma unroll
            for (int i = 0; i < 4; i++) o[i] = make_float4(h[4 * i], h[4 * i + 1], h[4 * i + 2], h[4 * i + 3]);
        }
        return;
    }
    if (p.flags & (TC_OUT_F32 | TC_RESIDUAL)) {
        float4 *o = reinterpret_cast<float4 *>(p.out_f32 + (size_t)row * p.ldo_f32 + col);
        #pragma unroll
        for (int i = 0; i < 8; i++) {
            float4 t = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
            if (p.flags & TC_RESIDUAL) {
                const float4 x = o[i];
                t.x += x.x; t.y += x.y; t.z += x.z; t.w += x.w;
            }
            o[i] = t;
        }
    }
    if (p.flags & TC_OUT_BF16) store_bf16<32>(p, row, col, v);
}

template <int BN>
__global__ void __launch_bounds__(NT, 1) gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA,
                                                        const __grid_constant__ CUtensorMap tmAlo,
                                                        const __grid_constant__ CUtensorMap tmB, const KParams p) {
    using C = Cfg<BN>;
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
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(tmem_slot)), "n"(ACC * BN));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const int tiles_n = p.N / BN, tiles_m = (p.M + BM - 1) / BM;
    const int total = tiles_n * tiles_m, nkb = p.Kw / BK;
    const bool split = p.K != p.Kw;                         // A rows are [hi(Kw) | lo(Kw)]

    tcx::pdl_launch_dependents();                           // the next kernel may run its prologue now
    if (warp == 0) {
        if (lane == 0) {
            int s = 0; unsigned ph = 0;
            const unsigned tx = (split ? 2 : 1) * A_BYTES + C::B_BYTES;
            tcx::pdl_wait();                                // A is produced by the previous kernel
            for (int t = blockIdx.x; t < total; t += gridDim.x) {
                const int m0 = (t / tiles_n) * BM, n0 = (t % tiles_n) * BN;
                for (int kb = 0; kb < nkb; kb++) {
                    uint8_t *st = base + s * C::STAGE;
                    mb_wait(&empty[s], ph ^ 1u);
                    mb_expect_tx(&full[s], tx);
                    tma_2d(st, &tmA, kb * BK, m0, &full[s]);
                    if (split) {
                        if (p.lo_sep) tma_2d(st + A_BYTES, &tmAlo, kb * BK, m0, &full[s]);
                        else tma_2d(st + A_BYTES, &tmA, p.Kw + kb * BK, m0, &full[s]);
                    }
                    tma_2d(st + 2 * A_BYTES, &tmB, kb * BK, n0, &full[s]);
                    if (++s == C::STAGES) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        int s = 0; unsigned ph = 0;
        int it = 0;
        for (int t = blockIdx.x; t < total; t += gridDim.x, it++) {
            const int as = it & 1;
            const unsigned aph = (unsigned)(it >> 1) & 1u;
            mb_wait(&tempty[as], aph ^ 1u);                 // epilogue has drained this accumulator
            tc_fence_after();
            const uint32_t d = tmem_base + (uint32_t)(as * BN);
            for (int kb = 0; kb < nkb; kb++) {
                mb_wait(&full[s], ph);
                tc_fence_after();
                if (lane == 0) {
                    const uint32_t st = s32(base + s * C::STAGE);
                    const uint64_t dh = umma_desc(st), dl = umma_desc(st + A_BYTES), db = umma_desc(st + 2 * A_BYTES);
                    #pragma unroll
                    for (int k = 0; k < BK / 16; k++)       // +32 B per K=16 step inside the 128 B swizzle row
                        umma_f16(d, dh + (uint64_t)(2 * k), db + (uint64_t)(2 * k), C::IDESC, (kb | k) ? 1u : 0u);
                    if (split) {
                        #pragma unroll
                        for (int k = 0; k < BK / 16; k++)
                            umma_f16(d, dl + (uint64_t)(2 * k), db + (uint64_t)(2 * k), C::IDESC, 1u);
                    }
                    umma_commit(&empty[s]);                 // stage reusable once these MMAs retire
                    if (kb == nkb - 1) umma_commit(&tfull[as]);
                }
                __syncwarp();
                if (++s == C::STAGES) { s = 0; ph ^= 1u; }
            }
        }
    } else {
        const int q = warp & 3;                             // TMEM lane quarter this warp may access
        int it = 0;
        tcx::pdl_wait();                                    // outputs (and the residual read) order after earlier kernels
        for (int t = blockIdx.x; t < total; t += gridDim.x, it++) {
            const int as = it & 1;
            const unsigned aph = (unsigned)(it >> 1) & 1u;
            const int m0 = (t / tiles_n) * BM, n0 = (t % tiles_n) * BN;
            const int row = m0 + q * 32 + lane;
            mb_wait(&tfull[as], aph);
            tc_fence_after();
            #pragma unroll 1
            for (int cc = 0; cc < BN / 32; cc++) {
                uint32_t r[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN + cc * 32), r);
                if (row < p.M) epilogue_row(p, row, n0 + cc * 32, r);
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
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(ACC * BN));
    }
}

}  // namespace tc

// ---- host ----------------------------------------------------------------------------------------
static PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;
static int g_sms = 0;

static int tc_init(void) {
    if (g_encode) return 0;
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    // resolved at run time: libvoxcuda.so keeps no link-time dependency on libcuda
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess || !fn) {
        vox_set_error("gemm_tc: cuTensorMapEncodeTiled not available from the driver");
        return -1;
    }
    g_encode = (PFN_cuTensorMapEncodeTiled_v12000)fn;
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    CUDA_TRY(cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, dev));
    CUDA_TRY(cudaFuncSetAttribute(tc::gemm_tc_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::Cfg<128>::SMEM_BYTES));
    CUDA_TRY(cudaFuncSetAttribute(tc::gemm_tc_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::Cfg<256>::SMEM_BYTES));
    return 0;
}

// 2-D bf16 tensor map: dim0 = K (contiguous), dim1 = rows with byte stride ld*2; box 64 x 128, 128 B swizzle
static int make_map(CUtensorMap *m, const uint16_t *ptr, int rows, int K, long ld, int box_rows) {
    cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)tc::BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void *)ptr, dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        vox_set_error("gemm_tc: cuTensorMapEncodeTiled failed (%d) rows=%d K=%d ld=%ld ptr=%p", (int)r, rows, K, ld, ptr);
        return -1;
    }
    return 0;
}

static int gemm_tc_impl(cudaStream_t st, int M, int N, int K, int KA, const uint16_t *A, const uint16_t *A_lo, long lda,
                        const uint16_t *W, const TcEpilogue &epi) {
    if (M <= 0) return 0;
    if (tc_init()) return -1;
    if ((KA != K && KA != 2 * K) || N % 128 || K % tc::BK || (lda * 2) % 16 || ((uintptr_t)A & 15) || ((uintptr_t)W & 15) ||
        ((uintptr_t)A_lo & 15)) {
        vox_set_error("gemm_tc: unsupported shape M=%d N=%d K=%d lda=%ld", M, N, K, lda);
        return -1;
    }
    static int force128 = -1;
    if (force128 < 0) { const char *e = getenv("VOX_CUDA_TC_BN128"); force128 = (e && atoi(e)) ? 1 : 0; }
    // 256-wide tiles carry twice the work of 128-wide ones at better operand reuse: take them unless the
    // machine then needs more rounds of tiles than it saves (small-M calls stream weights: more CTAs matter more)
    const int mt = (M + tc::BM - 1) / tc::BM;
    const int t256 = (N / 256) * mt, t128 = (N / 128) * mt;
    const int r256 = (t256 + g_sms - 1) / g_sms, r128 = (t128 + g_sms - 1) / g_sms;
    const int BN = (N % 256 == 0 && !force128 && t256 >= 64 && 2 * r256 <= r128) ? 256 : 128;
    CUtensorMap ma, mlo, mb;
    if (make_map(&ma, A, M, A_lo ? K : KA, lda, tc::BM) || make_map(&mb, W, N, K, K, BN)) return -1;
    if (A_lo) { if (make_map(&mlo, A_lo, M, K, lda, tc::BM)) return -1; }
    else mlo = ma;
    tc::KParams p;
    p.lo_sep = A_lo ? 1 : 0;
    p.M = M; p.N = N; p.K = KA; p.Kw = K; p.split_off = epi.split_off;
    p.out_f32 = epi.out_f32; p.ldo_f32 = epi.ldo_f32; p.out_bf16 = epi.out_bf16; p.ldo_bf16 = epi.ldo_bf16;
    p.bias = epi.bias; p.flags = epi.flags; p.col_scale = epi.col_scale;
    const int total = (N / BN) * mt;
    const int grid = total < g_sms ? total : g_sms;
    if (BN == 256) CUDA_TRY(launch_pdl(tc::gemm_tc_kernel<256>, grid, tc::NT, tc::Cfg<256>::SMEM_BYTES, st, ma, mlo, mb, p));
    else CUDA_TRY(launch_pdl(tc::gemm_tc_kernel<128>, grid, tc::NT, tc::Cfg<128>::SMEM_BYTES, st, ma, mlo, mb, p));
    LAUNCHED();
    return 0;
}

int gemm_tc(cudaStream_t st, int M, int N, int K, int KA, const uint16_t *A, long lda, const uint16_t *W,
            const TcEpilogue &epi) {
    return gemm_tc_impl(st, M, N, K, KA, A, nullptr, lda, W, epi);
}

// split activations as two planes (hi and lo arrays of identical shape): rows may overlap (lda < K), which is how
// the causal convolutions read their [3 x channels] windows without an im2col copy
int gemm_tc_planes(cudaStream_t st, int M, int N, int K, const uint16_t *A_hi, const uint16_t *A_lo, long lda,
                   const uint16_t *W, const TcEpilogue &epi) {
    if (!A_lo) return -1;
    return gemm_tc_impl(st, M, N, K, 2 * K, A_hi, A_lo, lda, W, epi);
}

// fp32 -> two bf16 planes (hi = rn(x), lo = rn(x - hi)), elementwise
__global__ void f32_to_bf16_planes_kernel(const float *__restrict__ in, uint16_t *__restrict__ hi, uint16_t *__restrict__ lo, size_t n2) {
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= n2) return;
    const float2 v = reinterpret_cast<const float2 *>(in)[i];
    const uint32_t h = tc::pack_bf16(v.x, v.y);
    reinterpret_cast<uint32_t *>(hi)[i] = h;
    reinterpret_cast<uint32_t *>(lo)[i] = tc::pack_bf16(v.x - bf16_lo(h), v.y - bf16_hi(h));
}
int f32_to_bf16_planes(cudaStream_t st, const float *in, uint16_t *hi, uint16_t *lo, size_t n) {
    if (!n) return 0;
    if (n & 1) { vox_set_error("f32_to_bf16_planes: odd element count"); return -1; }
    f32_to_bf16_planes_kernel<<<(unsigned)((n / 2 + 255) / 256), 256, 0, st>>>(in, hi, lo, n / 2);
    LAUNCHED();
    CUDA_TRY(cudaGetLastError());
    return 0;
}

// fp32 -> bf16 (round to nearest even) rows for the A operand; split writes [hi | lo]
__global__ void f32_to_bf16_rows_kernel(const float *__restrict__ in, long ldi, uint16_t *__restrict__ out, long ldo,
                                        int cols, int split) {
    const float *ir = in + (size_t)blockIdx.x * ldi;
    uint16_t *orow = out + (size_t)blockIdx.x * ldo;
    for (int i = threadIdx.x * 2; i < cols; i += blockDim.x * 2) {
        const float2 v = *reinterpret_cast<const float2 *>(ir + i);
        const uint32_t hi = tc::pack_bf16(v.x, v.y);
        *reinterpret_cast<uint32_t *>(orow + i) = hi;
        if (split)
            *reinterpret_cast<uint32_t *>(orow + cols + i) = tc::pack_bf16(v.x - bf16_lo(hi), v.y - bf16_hi(hi));
    }
}
int f32_to_bf16_rows(cudaStream_t st, const float *in, long ldi, uint16_t *out, long ldo, int rows, int cols, int split) {
    if (rows <= 0) return 0;
    f32_to_bf16_rows_kernel<<<rows, 256, 0, st>>>(in, ldi, out, ldo, cols, split);
    LAUNCHED();
    CUDA_TRY(cudaGetLastError());
    return 0;
}

// C ABI (include/vox_cuda.h): tensor-core twin of vox_cuda_sgemm_bf16 with host buffers.
extern "C" int vox_cuda_gemm_bf16_tc(int M, int N, int K, const float *A, const uint16_t *B_bf16, const float *bias,
                                     int flags, float *C) {
    if (!A || !B_bf16 || !C || M <= 0) return -1;
    if (!vox_cuda_available() && !vox_cuda_init()) return -1;
    float *dA = nullptr, *dC = nullptr, *dbias = nullptr;
    uint16_t *dAb = nullptr, *dW = nullptr;
    const int nout = (flags & TC_SWIGLU) ? N / 2 : N;
    int rc = -1;
    if (cudaMalloc((void **)&dA, (size_t)M * K * 4) != cudaSuccess) goto done;
    if (cudaMalloc((void **)&dAb, (size_t)M * K * 4) != cudaSuccess) goto done;
    if (cudaMalloc((void **)&dW, (size_t)N * K * 2) != cudaSuccess) goto done;
    if (cudaMalloc((void **)&dC, (size_t)M * nout * 4) != cudaSuccess) goto done;
    if (cudaMemcpy(dA, A, (size_t)M * K * 4, cudaMemcpyHostToDevice) != cudaSuccess) goto done;
    if (cudaMemcpy(dW, B_bf16, (size_t)N * K * 2, cudaMemcpyHostToDevice) != cudaSuccess) goto done;
    if (flags & TC_RESIDUAL)
        if (cudaMemcpy(dC, C, (size_t)M * nout * 4, cudaMemcpyHostToDevice) != cudaSuccess) goto done;
    if (bias) {
        if (cudaMalloc((void **)&dbias, (size_t)N * 4) != cudaSuccess) goto done;
        if (cudaMemcpy(dbias, bias, (size_t)N * 4, cudaMemcpyHostToDevice) != cudaSuccess) goto done;
    }
    {
        const int split = (flags & 128) ? 1 : 0;           // VOX_CUDA_TC_SPLIT_A
        if (f32_to_bf16_rows(0, dA, K, dAb, split ? 2 * K : K, M, K, split)) goto done;
        TcEpilogue e;
        e.out_f32 = dC; e.ldo_f32 = nout; e.out_bf16 = nullptr; e.ldo_bf16 = 0; e.bias = dbias; e.split_off = 0;
        e.flags = (flags & (TC_BIAS | TC_GELU | TC_RESIDUAL | TC_SWIGLU)) | TC_OUT_F32;
        if (!bias) e.flags &= ~TC_BIAS;
        if (gemm_tc(0, M, N, K, split ? 2 * K : K, dAb, split ? 2 * K : K, dW, e)) goto done;
    }
    if (cudaMemcpy(C, dC, (size_t)M * nout * 4, cudaMemcpyDeviceToHost) != cudaSuccess) goto done;
    rc = 0;
done:
    if (rc && !*vox_cuda_last_error())
        vox_set_error("vox_cuda_gemm_bf16_tc(%d,%d,%d) failed: %s", M, N, K, cudaGetErrorString(cudaGetLastError()));
    else if (rc)
        fprintf(stderr, "vox_cuda_gemm_bf16_tc(%d,%d,%d): cuda status %s\n", M, N, K, cudaGetErrorString(cudaGetLastError()));
    cudaFree(dA); cudaFree(dAb); cudaFree(dW); cudaFree(dC); cudaFree(dbias);
    return rc;
}
