// pcm.cu - device-side audio ingestion: the arithmetic of the reference's WAV / stdin
// front end without its parsing - interleaved s16 frames -> mono f32 in [-1, 1) -> 16 kHz by linear
// interpolation (vox_parse_wav_buffer, /root/reference/voxtral_audio.c:94-137; raw s16le stdin :199-202) -
// and the microphone silence gate's window energy (main.c:262-292).  Container parsing (RIFF chunks, stdin
// buffering) stays on the host.  Every operation is an explicit round-to-nearest intrinsic in the order the
// C source evaluates it (no FMA contraction), so results equal the reference's strict-IEEE build bit for bit.
#include "vox_internal.h"

// mono[i] = channels == 1 ? s / 32768 : (sum_c (float)s_c / channels) / 32768      voxtral_audio.c:98-110
__global__ void __launch_bounds__(256) pcm_downmix_kernel(const int16_t *__restrict__ src, int n_frames, int channels,
                                                          float *__restrict__ mono) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n_frames) return;
    if (channels == 1) {
        mono[i] = __fdiv_rn((float)src[i], 32768.0f);
    } else {
        float sum = 0.f;
        for (int c = 0; c < channels; c++) sum = __fadd_rn(sum, (float)src[(size_t)i * channels + c]);
        mono[i] = __fdiv_rn(__fdiv_rn(sum, (float)channels), 32768.0f);
    }
}

// out[i] = mono[idx] * (1 - frac) + mono[idx + 1] * frac, src_pos = (float)i * rate / 16000      voxtral_audio.c:113-131
__global__ void __launch_bounds__(256) pcm_resample_kernel(const float *__restrict__ mono, int n_frames, int sample_rate,
                                                           float *__restrict__ out, int new_n) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= new_n) return;
    const float src_pos = __fdiv_rn(__fmul_rn((float)i, (float)sample_rate), 16000.0f);
    const int idx = (int)src_pos;
    const float frac = __fsub_rn(src_pos, (float)idx);
    float v;
    if (idx + 1 < n_frames)
        v = __fadd_rn(__fmul_rn(mono[idx], __fsub_rn(1.0f, frac)), __fmul_rn(mono[idx + 1], frac));
    else
        v = idx < n_frames ? mono[idx] : 0.0f;
    out[i] = v;
}

// rms[w] = sqrt(sum_{i in window w} v^2 / window), summed in index order as main.c:266-272 does
__global__ void __launch_bounds__(128) pcm_window_rms_kernel(const float *__restrict__ x, int n_windows, int window,
                                                             float *__restrict__ rms) {
    const int w = blockIdx.x * 128 + threadIdx.x;
    if (w >= n_windows) return;
    const float *p = x + (size_t)w * window;
    float e = 0.f;
    for (int i = 0; i < window; i++) e = __fadd_rn(e, __fmul_rn(p[i], p[i]));
    rms[w] = __fsqrt_rn(__fdiv_rn(e, (float)window));
}

extern "C" int vox_cuda_pcm_resampled_len(int n_frames, int sample_rate) {
    if (n_frames < 0 || sample_rate <= 0) return -1;
    return sample_rate == 16000 ? n_frames : (int)((long long)n_frames * 16000 / sample_rate);   // voxtral_audio.c:114
}

extern "C" int vox_cuda_pcm_from_s16(const int16_t *pcm, int n_frames, int channels, int sample_rate, float *out, int out_cap) {
    if (!vox_cuda_available()) { vox_set_error("pcm_from_s16: backend not initialised (no CPU fallback)"); return -1; }
    if (!pcm || !out || n_frames <= 0 || channels < 1 || sample_rate <= 0) return -1;
    const int new_n = vox_cuda_pcm_resampled_len(n_frames, sample_rate);
    if (new_n > out_cap) { vox_set_error("pcm_from_s16: %d output samples > capacity %d", new_n, out_cap); return -1; }
    int16_t *d_src = nullptr; float *d_mono = nullptr, *d_out = nullptr;
    int rc = -1;
    do {
        if (cudaMalloc((void **)&d_src, (size_t)n_frames * channels * 2) != cudaSuccess) break;
        if (cudaMalloc((void **)&d_mono, (size_t)n_frames * 4) != cudaSuccess) break;
        if (cudaMemcpy(d_src, pcm, (size_t)n_frames * channels * 2, cudaMemcpyHostToDevice) != cudaSuccess) break;
        pcm_downmix_kernel<<<(n_frames + 255) / 256, 256>>>(d_src, n_frames, channels, d_mono);
        LAUNCHED();
        const float *res = d_mono;
        if (sample_rate != 16000 && new_n > 0) {
            if (cudaMalloc((void **)&d_out, (size_t)new_n * 4) != cudaSuccess) break;
            pcm_resample_kernel<<<(new_n + 255) / 256, 256>>>(d_mono, n_frames, sample_rate, d_out, new_n);
            LAUNCHED();
            res = d_out;
        }
        if (new_n > 0 && cudaMemcpy(out, res, (size_t)new_n * 4, cudaMemcpyDeviceToHost) != cudaSuccess) break;
        rc = new_n;
    } while (0);
    if (rc < 0) vox_set_error("pcm_from_s16: %s", cudaGetErrorString(cudaGetLastError()));
    cudaFree(d_src); cudaFree(d_mono); cudaFree(d_out);
    return rc;
}

extern "C" int vox_cuda_window_rms(const float *samples, int n, int window, float *rms_out) {
    if (!vox_cuda_available()) { vox_set_error("window_rms: backend not initialised (no CPU fallback)"); return -1; }
    if (!samples || !rms_out || window <= 0 || n < 0) return -1;
    const int nw = n / window;
    if (nw == 0) return 0;
    float *d_x = nullptr, *d_r = nullptr;
    int rc = -1;
    do {
        if (cudaMalloc((void **)&d_x, (size_t)nw * window * 4) != cudaSuccess) break;
        if (cudaMalloc((void **)&d_r, (size_t)nw * 4) != cudaSuccess) break;
        if (cudaMemcpy(d_x, samples, (size_t)nw * window * 4, cudaMemcpyHostToDevice) != cudaSuccess) break;
        pcm_window_rms_kernel<<<(nw + 127) / 128, 128>>>(d_x, nw, window, d_r);
        LAUNCHED();
        if (cudaMemcpy(rms_out, d_r, (size_t)nw * 4, cudaMemcpyDeviceToHost) != cudaSuccess) break;
        rc = nw;
    } while (0);
    if (rc < 0) vox_set_error("window_rms: %s", cudaGetErrorString(cudaGetLastError()));
    cudaFree(d_x); cudaFree(d_r);
    return rc;
}
