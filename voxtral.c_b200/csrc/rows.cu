// rows.cu - row-wise fp32 kernels shared by decoder prefill and the encoder:
// RMSNorm, RoPE + KV-ring append, sliding-window causal attention, SwiGLU.
#include "vox_internal.h"
#include "rows.h"
#include "tc_ptx.cuh"

// ---- RMSNorm over rows (voxtral_kernels.c:475-492), optional (1+ada) ----------
__global__ void __launch_bounds__(256) rows_rmsnorm_kernel(
    const float *__restrict__ x, long ldx, const float *__restrict__ w,
    const float *__restrict__ ada, float *__restrict__ out, long ldo, int dim) {
    const float *xr = x + (size_t)blockIdx.x * ldx;
    float *orow = out + (size_t)blockIdx.x * ldo;
    __shared__ float red[8];
    float ss = 0.f;
    for (int i = threadIdx.x; i < dim; i += 256) { const float v = xr[i]; ss += v * v; }
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
    #pragma unroll
    for (int i = 0; i < 8; i++) tot += red[i];
    const float rinv = 1.0f / sqrtf(tot / (float)dim + NORM_EPS);
    for (int i = threadIdx.x; i < dim; i += 256) {
        float y = xr[i] * rinv * w[i];
        if (ada) y *= (1.0f + ada[i]);
        orow[i] = y;
    }
}

// same, output rounded to bf16 (A operand of the tensor-core GEMMs)
__global__ void __launch_bounds__(256) rows_rmsnorm_bf16_kernel(
    const float *__restrict__ x, long ldx, const float *__restrict__ w, const float *__restrict__ ada,
    uint16_t *__restrict__ out, long ldo, int dim, int split) {
    tcx::pdl_launch_dependents();
    tcx::pdl_wait();
    const float *xr = x + (size_t)blockIdx.x * ldx;
    uint16_t *orow = out + (size_t)blockIdx.x * ldo;
    __shared__ float red[8];
    float ss = 0.f;
    for (int i = threadIdx.x; i < dim; i += 256) { const float v = xr[i]; ss += v * v; }
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
    #pragma unroll
    for (int i = 0; i < 8; i++) tot += red[i];
    const float rinv = 1.0f / sqrtf(tot / (float)dim + NORM_EPS);
    for (int i = threadIdx.x * 2; i < dim; i += 512) {
        float y0 = xr[i] * rinv * w[i], y1 = xr[i + 1] * rinv * w[i + 1];
        if (ada) { y0 *= (1.0f + ada[i]); y1 *= (1.0f + ada[i + 1]); }
        const __nv_bfloat162 v = __floats2bfloat162_rn(y0, y1);
        *reinterpret_cast<__nv_bfloat162 *>(orow + i) = v;
        if (split)   // second bf16 term: y = hi + lo to ~16 mantissa bits
            *reinterpret_cast<__nv_bfloat162 *>(orow + dim + i) =
                __floats2bfloat162_rn(y0 - __bfloat162float(v.x), y1 - __bfloat162float(v.y));
    }
}

int rows_rmsnorm_bf16(cudaStream_t st, const float *x, long ldx, const float *w, const float *ada, uint16_t *out,
                      long ldo, int rows, int dim, int split) {
    if (rows <= 0) return 0;
    CUDA_TRY(launch_pdl(rows_rmsnorm_bf16_kernel, rows, 256, 0, st, x, ldx, w, ada, out, ldo, dim, split));
    LAUNCHED();
    return 0;
}

int rows_rmsnorm(cudaStream_t st, const float *x, long ldx, const float *w, const float *ada,
                 float *out, long ldo, int rows, int dim) {
    if (rows <= 0) return 0;
    rows_rmsnorm_kernel<<<rows, 256, 0, st>>>(x, ldx, w, ada, out, ldo, dim);
    LAUNCHED();
    CUDA_TRY(cudaGetLastError());
    return 0;
}

// ---- RoPE on q,k + append k,v to the ring -----------------------------------------
// qkv: [rows, q_dim + 2*kv_dim] (q | k | v).  Interleaved-pair rotation with the
// logical position pos0 + row (voxtral_kernels.c:617-655); q is rotated in place,
// rotated k and plain v are written to ring slot (pos0 + row) & ring_mask.
__global__ void __launch_bounds__(256) rows_rope_kv_kernel(
    float *__restrict__ qkv, int q_dim, int kv_dim, int head_dim,
    const float *__restrict__ inv_freq, int pos0, int ring_mask,
    float *__restrict__ kc, float *__restrict__ vc, const int *__restrict__ pos_dev) {
    tcx::pdl_launch_dependents();
    tcx::pdl_wait();
    if (pos_dev) pos0 = *pos_dev;
    const int row = blockIdx.x;
    const int ld = q_dim + 2 * kv_dim;
    float *r = qkv + (size_t)row * ld;
    const int pos = pos0 + row;
    const size_t slot = (size_t)(pos & ring_mask) * kv_dim;
    const int half = head_dim >> 1;
    // thread -> pair index d inside a head (one cosf/sinf per thread), then a stride over the q and k heads
    const int d = threadIdx.x % half, hg = threadIdx.x / half, hstep = 256 / half;
    const float angle = (float)pos * inv_freq[d];
    const float c = cosf(angle), s = sinf(angle);
    const int n_heads = (q_dim + kv_dim) / head_dim;
    for (int h = hg; h < n_heads; h += hstep) {
        const int col = h * head_dim + 2 * d;
        const float2 x = *reinterpret_cast<const float2 *>(r + col);
        const float y0 = x.x * c - x.y * s, y1 = x.x * s + x.y * c;
        if (col < q_dim) *reinterpret_cast<float2 *>(r + col) = make_float2(y0, y1);
        else *reinterpret_cast<float2 *>(kc + slot + col - q_dim) = make_float2(y0, y1);
    }
    for (int i = threadIdx.x * 4; i < kv_dim; i += 1024)
        *reinterpret_cast<float4 *>(vc + slot + i) = *reinterpret_cast<const float4 *>(r + q_dim + kv_dim + i);
}

int rows_rope_kv(cudaStream_t st, float *qkv, int rows, int q_dim, int kv_dim, int head_dim,
                 const float *inv_freq, int pos0, int ring_mask, float *kc, float *vc) {
    return rows_rope_kv_dev(st, qkv, rows, q_dim, kv_dim, head_dim, inv_freq, pos0, ring_mask, kc, vc, nullptr);
}
int rows_rope_kv_dev(cudaStream_t st, float *qkv, int rows, int q_dim, int kv_dim, int head_dim,
                     const float *inv_freq, int pos0, int ring_mask, float *kc, float *vc, const int *pos_dev) {
    if (rows <= 0) return 0;
    CUDA_TRY(launch_pdl(rows_rope_kv_kernel, rows, 256, 0, st, qkv, q_dim, kv_dim, head_dim, inv_freq, pos0, ring_mask, kc, vc, pos_dev));
    LAUNCHED();
    return 0;
}

// ---- sliding-window causal attention over a ring (fp32) -----------------------------
// One warp per (query row, head).  Query i has logical position q_pos0 + i and
// attends to logical keys [max(first_valid, g - window + 1), g]
// (voxtral_kernels.c:541-611).  Lane-per-key scores on 32-key tiles, then
// lanes-over-dims accumulation with the probabilities broadcast by shuffle.
template <int HD>
__global__ void __launch_bounds__(128) rows_attention_kernel(
    const float *__restrict__ q, long ldq, const float *__restrict__ kc,
    const float *__restrict__ vc, int kv_dim, int ring_mask, int heads_per_kv,
    int q_pos0, int first_valid, int window, float scale, float *__restrict__ out, long ldo,
    uint16_t *__restrict__ out_b, long ldb, long lo_off) {
    tcx::pdl_launch_dependents();   // no-ops under a plain launch
    tcx::pdl_wait();
    constexpr int DPL = HD / 32;  // dims per lane in the PV phase
    const int row = blockIdx.x;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int head = blockIdx.y * 4 + warp;
    const int kvh = head / heads_per_kv;
    __shared__ float4 qs[4][HD / 4];
    {
        const float *qr = q + (size_t)row * ldq + head * HD;
        for (int i = lane; i < HD / 4; i += 32) qs[warp][i] = reinterpret_cast<const float4 *>(qr)[i];
    }
    __syncwarp();
    const int g = q_pos0 + row;
    int ks = g - window + 1;
    if (ks < first_valid) ks = first_valid;
    float m = -INFINITY, l = 0.f;
    float acc[DPL];
    #pragma unroll
    for (int d = 0; d < DPL; d++) acc[d] = 0.f;
    for (int base = ks; base <= g; base += 32) {
        const int j = base + lane;
        float s = -INFINITY;
        if (j <= g) {
            const float4 *kr = reinterpret_cast<const float4 *>(
                kc + (size_t)(j & ring_mask) * kv_dim + kvh * HD);
            float d0 = 0.f;
            #pragma unroll 8
            for (int i = 0; i < HD / 4; i++) {
                const float4 kv = kr[i], qv = qs[warp][i];
                d0 += kv.x * qv.x + kv.y * qv.y + kv.z * qv.z + kv.w * qv.w;
            }
            s = d0 * scale;
        }
        const float mt = warp_max(s);
        const float mn = fmaxf(m, mt);
        const float corr = (m == -INFINITY) ? 0.f : expf(m - mn);
        const float p = (j <= g) ? expf(s - mn) : 0.f;
        l = l * corr + warp_sum(p);
        #pragma unroll
        for (int d = 0; d < DPL; d++) acc[d] *= corr;
        const int nk = min(32, g - base + 1);
        for (int i = 0; i < nk; i++) {
            const float pi = __shfl_sync(0xffffffffu, p, i);
            const float *vr = vc + (size_t)((base + i) & ring_mask) * kv_dim + kvh * HD + lane * DPL;
            #pragma unroll
            for (int d = 0; d < DPL; d++) acc[d] = fmaf(pi, vr[d], acc[d]);
        }
        m = mn;
    }
    const float inv = l > 0.f ? 1.0f / l : 0.f;
    if (out) {
        float *o = out + (size_t)row * ldo + head * HD + lane * DPL;
        #pragma unroll
        for (int d = 0; d < DPL; d++) o[d] = acc[d] * inv;
    }
    if (out_b) {   // split-bf16 row (hi at column c, rounding residual at lo_off + c): the next product's A operand
        uint16_t *o = out_b + (size_t)row * ldb + head * HD + lane * DPL;
        #pragma unroll
        for (int d = 0; d < DPL; d += 2) {
            const float y0 = acc[d] * inv, y1 = acc[d + 1] * inv;
            const __nv_bfloat162 hi = __floats2bfloat162_rn(y0, y1);
            *reinterpret_cast<__nv_bfloat162 *>(o + d) = hi;
            *reinterpret_cast<__nv_bfloat162 *>(o + lo_off + d) =
                __floats2bfloat162_rn(y0 - __bfloat162float(hi.x), y1 - __bfloat162float(hi.y));
        }
    }
}

int rows_attention(cudaStream_t st, int head_dim, const float *q, long ldq, const float *kc,
                   const float *vc, int kv_dim, int ring_mask, int n_heads, int heads_per_kv,
                   int rows, int q_pos0, int first_valid, int window, float *out, long ldo) {
    if (rows <= 0) return 0;
    const float scale = 1.0f / sqrtf((float)head_dim);
    dim3 grid(rows, n_heads / 4);
    if (head_dim == 64)
        rows_attention_kernel<64><<<grid, 128, 0, st>>>(q, ldq, kc, vc, kv_dim, ring_mask, heads_per_kv,
                                                        q_pos0, first_valid, window, scale, out, ldo, nullptr, 0L, 0L);
    else if (head_dim == 128)
        rows_attention_kernel<128><<<grid, 128, 0, st>>>(q, ldq, kc, vc, kv_dim, ring_mask, heads_per_kv,
                                                         q_pos0, first_valid, window, scale, out, ldo, nullptr, 0L, 0L);
    else { vox_set_error("rows_attention: head_dim %d unsupported", head_dim); return -1; }
    LAUNCHED();
    CUDA_TRY(cudaGetLastError());
    return 0;
}

// head_dim 128, output as a split-bf16 row, launched with programmatic dependent launch (decoder prefill, dec_prefill.cu)
int rows_attention_split128(cudaStream_t st, const float *q, long ldq, const float *kc, const float *vc, int kv_dim,
                            int ring_mask, int n_heads, int heads_per_kv, int rows, int q_pos0, int first_valid, int window,
                            uint16_t *out_b, long ldb, long lo_off) {
    if (rows <= 0) return 0;
    const float scale = 1.0f / sqrtf(128.0f);
    CUDA_TRY(launch_pdl(rows_attention_kernel<128>, dim3(rows, n_heads / 4), 128, 0, st, q, ldq, kc, vc, kv_dim, ring_mask,
                        heads_per_kv, q_pos0, first_valid, window, scale, (float *)nullptr, 0L, out_b, ldb, lo_off));
    LAUNCHED();
    return 0;
}

// ---- SwiGLU on interleaved (gate, up) columns: h[i] = silu(gu[2i]) * gu[2i+1] ---------
__global__ void __launch_bounds__(256) rows_swiglu_kernel(const float *__restrict__ gu,
                                                          float *__restrict__ h, size_t n) {
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i < n) {
        const float2 v = reinterpret_cast<const float2 *>(gu)[i];
        h[i] = silu(v.x) * v.y;
    }
}

int rows_swiglu(cudaStream_t st, const float *gu, float *h, int rows, int hidden) {
    const size_t n = (size_t)rows * hidden;
    if (!n) return 0;
    rows_swiglu_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(gu, h, n);
    LAUNCHED();
    CUDA_TRY(cudaGetLastError());
    return 0;
}
