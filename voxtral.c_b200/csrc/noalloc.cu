// noalloc.cu - an entry point that allocates NOTHING (no cudaMalloc / cudaFree / allocating library call), for callers
// that cannot allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid it.
//
// vox_cuda_alternatives_noalloc = the alternatives rule of stream_fill_alts (/root/reference/voxtral.c:955-1010; the
// existing entry point is vox_cuda_decoder_alternatives, api_dec.cu) on a logits row the CALLER passes: the best
// (max_alt - 1) ids >= 1000 other than the greedy token, by logit, lowest id on ties, kept while
// 1 - p_i / p_best <= cutoff.  The kernel's scratch - the top-4 candidates of every 1024-logit slice plus an arrival
// counter - lives in the file-scope __device__ array below instead of a cudaMalloc'd buffer: NA_SCRATCH_BYTES (8448 B)
// on every GPU that loads the library, until exit.  The array bounds the input: n_logits <= VOX_CUDA_NOALLOC_MAX_LOGITS.
//
// Every call shares the array, so two calls that overlap on the GPU (different streams, no dependency) race; calls on
// one stream, or otherwise ordered, are fine.  The array keeps the last call's contents: the candidates of a call are
// all rewritten before they are read, the arrival counter is cleared on the stream in front of every launch.
//
// No counterpart in the reference.
#include "vox_common.cuh"
#include <math.h>

namespace na {

constexpr int SLICE = 1024;                                        // logits per CTA
constexpr int MAX_SLICES = VOX_CUDA_NOALLOC_MAX_LOGITS / SLICE;    // 256
constexpr int NT = 256;
constexpr int TEXT_MIN = 1000;                                     // TOKEN_TEXT_MIN, voxtral.c:399
constexpr int NONE = 0x7fffffff;

struct Scratch {
    float2 cand[MAX_SLICES][4];      // (logit, id bits) - the four best text logits of a slice, best first
    unsigned arrived;                // slices finished in the current call
    unsigned pad[63];
};
constexpr size_t NA_SCRATCH_BYTES = sizeof(Scratch);
static_assert(NA_SCRATCH_BYTES == 8448, "stated size of the static scratch array");

}  // namespace na

// In host code this name is a host-side stand-in for the device array, NOT a device address: the address the kernel
// gets comes from cudaGetSymbolAddress (na_setup below).  Device code may name it directly.
static __device__ __align__(256) unsigned char g_na_scratch[na::NA_SCRATCH_BYTES];

namespace na {

struct Top4 {
    float v[4]; int i[4];
    __device__ __forceinline__ void clear() {
        #pragma unroll
        for (int k = 0; k < 4; k++) { v[k] = -INFINITY; i[k] = NONE; }
    }
    // larger logit first, lower id on ties; NaN never enters (the rule of alt_topk_kernel, api_dec.cu)
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

// block-wide merge through shared memory; the result is valid in thread 0
__device__ __forceinline__ void merge_block(Top4 &t, float (*sv)[4], int (*si)[4]) {
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    t.merge_warp();
    if (lane == 0) {
        #pragma unroll
        for (int k = 0; k < 4; k++) { sv[w][k] = t.v[k]; si[w][k] = t.i[k]; }
    }
    __syncthreads();
    if (threadIdx.x == 0)
        for (int ww = 1; ww < NT / 32; ww++)
            for (int k = 0; k < 4; k++) t.insert(sv[ww][k], si[ww][k]);
    __syncthreads();
}

// One CTA per 1024-logit slice writes its four best text candidates to the scratch; the CTA that arrives last merges
// all slices, applies the cutoff and writes ids_out[0 .. *count_out).
__global__ void __launch_bounds__(NT) alternatives_kernel(const float *__restrict__ logits, int n_logits, int best, int k_alt,
                                                          float cutoff, Scratch *__restrict__ sc, int *__restrict__ ids_out,
                                                          int *__restrict__ count_out) {
    __shared__ float sv[NT / 32][4];
    __shared__ int si[NT / 32][4];
    __shared__ int s_last;
    const int slice = blockIdx.x, n_slices = gridDim.x;
    Top4 t;
    t.clear();
    for (int i = slice * SLICE + threadIdx.x; i < min(n_logits, (slice + 1) * SLICE); i += NT)
        if (i >= TEXT_MIN && i != best) t.insert(logits[i], i);
    merge_block(t, sv, si);
    if (threadIdx.x == 0) {
        #pragma unroll
        for (int k = 0; k < 4; k++) sc->cand[slice][k] = make_float2(t.v[k], __int_as_float(t.i[k]));
        __threadfence();                                          // candidates visible before the arrival is counted
        s_last = (atomicAdd(&sc->arrived, 1u) + 1u == (unsigned)n_slices) ? 1 : 0;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    t.clear();
    for (int k = threadIdx.x; k < 4 * n_slices; k += NT) {
        const float2 c = __ldcg(&sc->cand[0][0] + k);
        t.insert(c.x, __float_as_int(c.y));
    }
    merge_block(t, sv, si);
    if (threadIdx.x == 0) {
        const float lb = logits[best];
        int n = 0;
        for (int k = 0; k < k_alt && k < 4; k++) {
            if (t.i[k] == NONE) break;
            const float r = 1.0f - expf(t.v[k] - lb);             // p_i / p_best = exp(l_i - l_best)
            if (r > cutoff) break;
            ids_out[n++] = t.i[k];
        }
        *count_out = n;
        sc->arrived = 0;                                          // leave the counter clean for the next call
    }
}

// device address of the scratch array, per device; resolved once per device with that device current
constexpr int MAX_DEVICES = 64;
static Scratch *g_ptr[MAX_DEVICES];

static Scratch *setup(void) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) {
        vox_set_error("alternatives_noalloc: no current CUDA device");
        return nullptr;
    }
    if (g_ptr[dev]) return g_ptr[dev];
    void *p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_na_scratch);
    if (e != cudaSuccess || !p) {
        vox_set_error("alternatives_noalloc: cudaGetSymbolAddress(scratch) on device %d failed: %s", dev, cudaGetErrorString(e));
        return nullptr;
    }
    g_ptr[dev] = (Scratch *)p;
    return g_ptr[dev];
}

}  // namespace na

extern "C" int vox_cuda_alternatives_noalloc(const float *logits, int n_logits, int best_token, int max_alt, float cutoff,
                                             int *ids_out, int *count_out, void *stream, int sync) {
    if (!logits || !ids_out || !count_out) { vox_set_error("alternatives_noalloc: null pointer"); return -1; }
    if (n_logits <= 0 || n_logits > VOX_CUDA_NOALLOC_MAX_LOGITS) {
        vox_set_error("alternatives_noalloc: %d logits, the static scratch holds at most %d (VOX_CUDA_NOALLOC_MAX_LOGITS)",
                      n_logits, VOX_CUDA_NOALLOC_MAX_LOGITS);
        return -1;
    }
    if (best_token < 0 || best_token >= n_logits) { vox_set_error("alternatives_noalloc: best_token %d outside [0,%d)", best_token, n_logits); return -1; }
    if (max_alt > VOX_CUDA_MAX_ALT) max_alt = VOX_CUDA_MAX_ALT;
    const int k_alt = max_alt - 1 > 0 ? max_alt - 1 : 0;
    na::Scratch *sc = na::setup();
    if (!sc) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    const int n_slices = (n_logits + na::SLICE - 1) / na::SLICE;
    CUDA_TRY(cudaMemsetAsync(&sc->arrived, 0, sizeof(unsigned), st));   // what must start at zero, whatever the last call left
    na::alternatives_kernel<<<n_slices, na::NT, 0, st>>>(logits, n_logits, best_token, k_alt, cutoff, sc, ids_out, count_out);
    CUDA_TRY(cudaGetLastError());
    LAUNCHED();
    if (sync) CUDA_TRY(cudaStreamSynchronize(st));
    return 0;
}
