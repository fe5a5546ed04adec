// group_dev.cuh - device bodies of the stream-group decode step, shared by the kernel-per-op path
// (api_group.cu) and the persistent step kernel (group_mega.cu).  Every body runs on 256 threads
// (tid 0..255, argmax: NTH threads) and synchronises them through `sync()` (__syncthreads in a plain
// kernel, a named barrier among the worker warps of the persistent kernel).  Data another CTA may have
// written during the same kernel is read with ld.global.cg.
#pragma once
#include <cuda_fp16.h>
#include "vox_internal.h"

namespace grp {

constexpr int MAX_STREAMS = 128;
constexpr int KS_QKV = 3, KS_WO = 6, KS_W2 = 6;     // K-range splits: 48x3, 24x6, 24x6 work items for 148 SMs
constexpr int KS_MAX = 6;
constexpr int AT_WARPS = 8;

struct Slot {
    int pos;        // logical position of the token being processed (= valid K/V rows before it)
    int prev;       // token whose embedding is added to the adapter row (voxtral.c:1106-1113)
    int aidx;       // adapter row of this position
    int remaining;  // steps this stream still has to take; 0 = finished (rows are computed but ignored)
};

struct SyncThreads { __device__ __forceinline__ void operator()() const { __syncthreads(); } };

// ---- x (+= partials | = embedding) -> RMSNorm -> split bf16 rows --------------------------------
// mode 0: x[b] = adapter[b][aidx] + tok_emb[prev];  mode 1: x[b] += sum_s part[s][b]
template <class Sync>
__device__ __forceinline__ void norm_row(
    int b, int tid, float *__restrict__ x, int mode, const float *__restrict__ part, int n_split, long split_stride,
    const float *__restrict__ adapter, long adapter_stride, const uint16_t *__restrict__ tok_emb,
    const float *__restrict__ tok_scale /* Q8 model: row scales of the int-valued embedding rows, else null */,
    const Slot *__restrict__ slots, const float *__restrict__ w, const float *__restrict__ ada,
    uint16_t *__restrict__ out, float *red /* smem [8] */, Sync sync) {
    float *xr = x + (size_t)b * DEC_DIM;
    float v[12];
    if (mode == 0) {
        Slot sl;
        *reinterpret_cast<int4 *>(&sl) = __ldcg(reinterpret_cast<const int4 *>(slots + b));
        const float *a = adapter + (size_t)b * adapter_stride + (size_t)sl.aidx * DEC_DIM;
        const uint16_t *e = tok_emb + (size_t)sl.prev * DEC_DIM;
        const float es = tok_scale ? tok_scale[sl.prev] : 1.0f;      // q * scale (tok_embed_q8_to_f32, voxtral.c:443-452)
        #pragma unroll
        for (int j = 0; j < 3; j++) {
            const int col = j * 1024 + tid * 4;
            const float4 av = *reinterpret_cast<const float4 *>(a + col);
            const uint2 ev = *reinterpret_cast<const uint2 *>(e + col);
            if (tok_scale) {
                v[4 * j] = av.x + bf16_lo(ev.x) * es; v[4 * j + 1] = av.y + bf16_hi(ev.x) * es;
                v[4 * j + 2] = av.z + bf16_lo(ev.y) * es; v[4 * j + 3] = av.w + bf16_hi(ev.y) * es;
            } else {
                v[4 * j] = av.x + bf16_lo(ev.x); v[4 * j + 1] = av.y + bf16_hi(ev.x);
                v[4 * j + 2] = av.z + bf16_lo(ev.y); v[4 * j + 3] = av.w + bf16_hi(ev.y);
            }
        }
    } else {
        #pragma unroll
        for (int j = 0; j < 3; j++) {
            const int col = j * 1024 + tid * 4;
            float4 t = *reinterpret_cast<const float4 *>(xr + col);
            for (int s = 0; s < n_split; s++) {
                const float4 pv = __ldcg(reinterpret_cast<const float4 *>(part + (size_t)s * split_stride + (size_t)b * DEC_DIM + col));
                t.x += pv.x; t.y += pv.y; t.z += pv.z; t.w += pv.w;
            }
            v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
        }
    }
    float ss = 0.f;
    #pragma unroll
    for (int j = 0; j < 3; j++) {
        *reinterpret_cast<float4 *>(xr + j * 1024 + tid * 4) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        ss += v[4 * j] * v[4 * j] + v[4 * j + 1] * v[4 * j + 1] + v[4 * j + 2] * v[4 * j + 2] + v[4 * j + 3] * v[4 * j + 3];
    }
    ss = warp_sum(ss);
    if ((tid & 31) == 0) red[tid >> 5] = ss;
    sync();
    float tot = 0.f;
    #pragma unroll
    for (int i = 0; i < 8; i++) tot += red[i];
    const float rinv = 1.0f / sqrtf(tot / (float)DEC_DIM + NORM_EPS);
    uint16_t *orow = out + (size_t)b * 2 * DEC_DIM;
    #pragma unroll
    for (int j = 0; j < 3; j++) {
        const int col = j * 1024 + tid * 4;
        const float4 wv = *reinterpret_cast<const float4 *>(w + col);
        float y[4] = {v[4 * j] * rinv * wv.x, v[4 * j + 1] * rinv * wv.y, v[4 * j + 2] * rinv * wv.z, v[4 * j + 3] * rinv * wv.w};
        if (ada) {
            const float4 av = *reinterpret_cast<const float4 *>(ada + col);
            y[0] *= (1.0f + av.x); y[1] *= (1.0f + av.y); y[2] *= (1.0f + av.z); y[3] *= (1.0f + av.w);
        }
        const __nv_bfloat162 h0 = __floats2bfloat162_rn(y[0], y[1]), h1 = __floats2bfloat162_rn(y[2], y[3]);
        const __nv_bfloat162 l0 = __floats2bfloat162_rn(y[0] - __bfloat162float(h0.x), y[1] - __bfloat162float(h0.y));
        const __nv_bfloat162 l1 = __floats2bfloat162_rn(y[2] - __bfloat162float(h1.x), y[3] - __bfloat162float(h1.y));
        uint2 hv, lv;
        hv.x = *reinterpret_cast<const uint32_t *>(&h0); hv.y = *reinterpret_cast<const uint32_t *>(&h1);
        lv.x = *reinterpret_cast<const uint32_t *>(&l0); lv.y = *reinterpret_cast<const uint32_t *>(&l1);
        *reinterpret_cast<uint2 *>(orow + col) = hv;
        *reinterpret_cast<uint2 *>(orow + DEC_DIM + col) = lv;
    }
}

// ---- RoPE + ring append + GQA attention of one (stream, kv head) -------------------------------
// Key groups of 8 per warp with all 16 row loads in flight (the single-stream kernel's scheme,
// dec_mega.cu); the 8 warps' online-softmax states are merged in shared memory.
struct AttnSmem {
    float qs[4][DEC_HD];
    float wsm[AT_WARPS][4][DEC_HD + 4];
};
// K/V element type: float (the reference's CPU / BLAS layout, voxtral_decoder.c:703-705) or __half (the
// reference's GPU default, "fp16 KV", voxtral.c:189-190): rows are rounded once when appended and widened
// in registers when read; scores, softmax and the output stay fp32.
__device__ __forceinline__ float4 ld_kv4(const float *p) { return __ldcg(reinterpret_cast<const float4 *>(p)); }
__device__ __forceinline__ float4 ld_kv4(const __half *p) {
    const uint2 u = __ldcg(reinterpret_cast<const uint2 *>(p));
    const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&u.x));
    const float2 b = __half22float2(*reinterpret_cast<const __half2 *>(&u.y));
    return make_float4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ void st_kv(float *p, float v) { *p = v; }
__device__ __forceinline__ void st_kv(__half *p, float v) { *p = __float2half_rn(v); }

template <class Sync, class KV>
__device__ __forceinline__ void attn_task(
    int b, int kvh, int tid, const float *__restrict__ part, int n_split, long split_stride, const Slot *__restrict__ slots,
    const float *__restrict__ inv_freq, KV *__restrict__ kv_k, KV *__restrict__ kv_v, size_t stream_stride,
    int ring_mask, uint16_t *__restrict__ out, AttnSmem &sm, Sync sync) {
    const int w = tid >> 5, lane = tid & 31;
    float (*qs)[DEC_HD] = sm.qs;
    float (*wsm)[4][DEC_HD + 4] = sm.wsm;
    const int pos = __ldcg(&slots[b].pos);
    KV *kc = kv_k + (size_t)b * stream_stride, *vc = kv_v + (size_t)b * stream_stride;
    const size_t slot_new = (size_t)(pos & ring_mask) * DEC_KVDIM + (size_t)kvh * DEC_HD;
    {
        // 256 q pairs (4 heads), 64 k pairs, 128 v values of this kv head; partial sums in split order
        const float *pr = part + (size_t)b * DEC_QKV;
        for (int i = tid; i < 320 + 128; i += 256) {
            if (i < 320) {
                const int col = i < 256 ? kvh * 4 * DEC_HD + 2 * i : DEC_QDIM + kvh * DEC_HD + 2 * (i - 256);
                float x0 = 0.f, x1 = 0.f;
                for (int s = 0; s < n_split; s++) {
                    const float2 t = __ldcg(reinterpret_cast<const float2 *>(pr + (size_t)s * split_stride + col));
                    x0 += t.x; x1 += t.y;
                }
                const int d = (col % DEC_HD) >> 1;
                const float angle = (float)pos * inv_freq[d];
                const float c = cosf(angle), sn = sinf(angle);
                const float y0 = x0 * c - x1 * sn, y1 = x0 * sn + x1 * c;
                if (i < 256) { qs[i >> 6][2 * (i & 63)] = y0; qs[i >> 6][2 * (i & 63) + 1] = y1; }
                else { st_kv(kc + slot_new + 2 * (i - 256), y0); st_kv(kc + slot_new + 2 * (i - 256) + 1, y1); }
            } else {
                const int d = i - 320;
                float x0 = 0.f;
                for (int s = 0; s < n_split; s++) x0 += __ldcg(pr + (size_t)s * split_stride + DEC_QDIM + DEC_KVDIM + kvh * DEC_HD + d);
                st_kv(vc + slot_new + d, x0);
            }
        }
    }
    sync();                                              // q in smem, new K/V row visible to the CTA
    const int n_keys = min(pos + 1, DEC_WINDOW);
    const int first = pos - n_keys + 1;
    const int ng = (n_keys + 7) >> 3;
    const float scale = 0.08838834764831845f;            // 1/sqrt(128)
    float4 q[4];
    #pragma unroll
    for (int h = 0; h < 4; h++) q[h] = reinterpret_cast<const float4 *>(&qs[h][0])[lane];
    float m = -INFINITY, l = 0.f;                        // running max / sum of head (lane & 3)
    float4 acc[4];
    #pragma unroll
    for (int h = 0; h < 4; h++) acc[h] = make_float4(0.f, 0.f, 0.f, 0.f);
    const size_t hoff = (size_t)kvh * DEC_HD + 4 * lane;
    for (int g = w; g < ng; g += AT_WARPS) {
        const int j0 = g * 8;
        const int nv = min(8, n_keys - j0);
        float4 kf[8], vf[8];
        #pragma unroll
        for (int i = 0; i < 8; i++) {
            const int slot = (first + j0 + min(i, nv - 1)) & ring_mask;     // clamp: always a valid row
            kf[i] = ld_kv4(kc + (size_t)slot * DEC_KVDIM + hoff);
            vf[i] = ld_kv4(vc + (size_t)slot * DEC_KVDIM + hoff);
        }
        float v[32];
        #pragma unroll
        for (int i = 0; i < 8; i++)
            #pragma unroll
            for (int h = 0; h < 4; h++)
                v[i * 4 + h] = kf[i].x * q[h].x + kf[i].y * q[h].y + kf[i].z * q[h].z + kf[i].w * q[h].w;
        // transposed reduction: score (key i, head h) ends in lane 4i + h
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
        const float mn = fmaxf(m, mx);                   // finite: every group has >= 1 valid key
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
    for (int h = 0; h < 4; h++) reinterpret_cast<float4 *>(&wsm[w][h][0])[lane] = acc[h];
    if (lane < 4) { wsm[w][lane][DEC_HD] = m; wsm[w][lane][DEC_HD + 1] = l; }
    sync();
    {   // merge the 8 warps: thread -> (head, two dims); normalise; split bf16 out
        const int h = tid >> 6, d = (tid & 63) * 2;
        float M = -INFINITY;
        #pragma unroll
        for (int k = 0; k < AT_WARPS; k++) M = fmaxf(M, wsm[k][h][DEC_HD]);
        float n0 = 0.f, n1 = 0.f, den = 0.f;
        #pragma unroll
        for (int k = 0; k < AT_WARPS; k++) {
            const float mk_ = wsm[k][h][DEC_HD];
            const float e = (mk_ == -INFINITY) ? 0.f : expf(mk_ - M);
            n0 = fmaf(e, wsm[k][h][d], n0);
            n1 = fmaf(e, wsm[k][h][d + 1], n1);
            den = fmaf(e, wsm[k][h][DEC_HD + 1], den);
        }
        const float inv = den > 0.f ? 1.0f / den : 0.f;
        const float y0 = n0 * inv, y1 = n1 * inv;
        const __nv_bfloat162 hi = __floats2bfloat162_rn(y0, y1);
        const __nv_bfloat162 lo = __floats2bfloat162_rn(y0 - __bfloat162float(hi.x), y1 - __bfloat162float(hi.y));
        uint16_t *o = out + (size_t)b * 2 * DEC_QDIM + (kvh * 4 + h) * DEC_HD + d;
        *reinterpret_cast<__nv_bfloat162 *>(o) = hi;
        *reinterpret_cast<__nv_bfloat162 *>(o + DEC_QDIM) = lo;
    }
    sync();                                              // smem is free for the next task
}

// ---- first-maximum argmax per stream + advance the stream ---------------------------------------
template <int NTH, class Sync>
__device__ __forceinline__ void argmax_row(int b, int tid, const float *__restrict__ logits, Slot *__restrict__ slots,
                                           int *__restrict__ tokens, int step, int n_streams, float *sv, int *si /* smem [32] */,
                                           Sync sync) {
    const float4 *row = reinterpret_cast<const float4 *>(logits + (size_t)b * VOCAB);
    float best = -INFINITY; int bi = 0x7fffffff;
    for (int i = tid; i < VOCAB / 4; i += NTH) {        // ascending indices per thread: strict > keeps the first
        const float4 t = __ldcg(row + i);
        if (t.x > best) { best = t.x; bi = 4 * i; }
        if (t.y > best) { best = t.y; bi = 4 * i + 1; }
        if (t.z > best) { best = t.z; bi = 4 * i + 2; }
        if (t.w > best) { best = t.w; bi = 4 * i + 3; }
    }
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    if ((tid & 31) == 0) { sv[tid >> 5] = best; si[tid >> 5] = bi; }
    sync();
    if (tid < 32) {
        best = tid < NTH / 32 ? sv[tid] : -INFINITY; bi = tid < NTH / 32 ? si[tid] : 0x7fffffff;
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
        }
        if (tid == 0) {
            Slot sl;
            *reinterpret_cast<int4 *>(&sl) = __ldcg(reinterpret_cast<const int4 *>(slots + b));
            if (bi == 0x7fffffff) bi = 0;                // all -inf / NaN row: the reference's loop leaves index 0
            tokens[(size_t)step * n_streams + b] = sl.remaining > 0 ? bi : -1;
            if (sl.remaining > 0) {
                sl.remaining--;
                if (sl.remaining > 0) { sl.pos++; sl.aidx++; sl.prev = bi; }   // a finished stream idles on its last row
                slots[b] = sl;
            }
        }
    }
}

}  // namespace grp
