// mel.cu - batched log-mel front end (hot path #3).
//
// Restates mel_compute_available (/root/reference/voxtral_audio.c:454-513) for
// many frames at once: frame t = samples[160 t .. 160 t + 399] x periodic Hann
// -> direct 201-bin real DFT from cos/sin tables -> power -> 128 x 201 Slaney
// filterbank -> log10(max(.,1e-10)), clamp to >= 1.5 - 8, (v + 4) / 4.
//
// The tables are built on the HOST with the reference's own float expressions
// (voxtral_audio.c:531-542 and build_mel_filters :248-285) - the float rounding
// of the angle 2*pi*k*n/400 is worth up to 1e-4 rad, so the table bits are part
// of the reference's definition of "mel", not an implementation detail.
#include "vox_internal.h"
#include <math.h>
#include <stdlib.h>
#include <vector>

constexpr int N_FFT = 400, N_FREQ = 201, N_MEL = VOX_CUDA_MEL_BINS, HOP = 160;
constexpr int MEL_FR = 8;     // frames per CTA
constexpr int MEL_T = 256;

static float *g_dft = nullptr;    // [400][2][208]: (n, cos|sin, k) padded to 208
static float *g_filt = nullptr;   // [201][128]  (k, m)
static float *g_win = nullptr;    // [400]
constexpr int KP = 208;

static float hz_to_mel(float f) {
    const float min_log_hertz = 1000.0f, min_log_mel = 15.0f;
    const float logstep = 27.0f / logf(6.4f);
    float mels = 3.0f * f / 200.0f;
    if (f >= min_log_hertz) mels = min_log_mel + logf(f / min_log_hertz) * logstep;
    return mels;
}
static float mel_to_hz(float mels) {
    const float min_log_hertz = 1000.0f, min_log_mel = 15.0f;
    const float logstep = logf(6.4f) / 27.0f;
    float f = 200.0f * mels / 3.0f;
    if (mels >= min_log_mel) f = min_log_hertz * expf(logstep * (mels - min_log_mel));
    return f;
}

int mel_tables_init(void) {
    if (g_dft) return 0;
    std::vector<float> dft((size_t)N_FFT * 2 * KP, 0.f), filt((size_t)N_FREQ * N_MEL, 0.f), win(N_FFT);
    for (int k = 0; k < N_FREQ; k++)
        for (int n = 0; n < N_FFT; n++) {
            // voxtral_audio.c:533-535, evaluated left to right in float
            volatile float angle = 2.0f * (float)M_PI * (float)k * (float)n / (float)N_FFT;
            dft[((size_t)n * 2 + 0) * KP + k] = cosf(angle);
            dft[((size_t)n * 2 + 1) * KP + k] = sinf(angle);
        }
    for (int i = 0; i < N_FFT; i++)  // voxtral_audio.c:540-542
        win[i] = 0.5f * (1.0f - cosf(2.0f * (float)M_PI * (float)i / (float)N_FFT));
    {   // voxtral_audio.c:248-285
        float fft_freqs[N_FREQ], ff[N_MEL + 2], fd[N_MEL + 1];
        for (int i = 0; i < N_FREQ; i++) fft_freqs[i] = (float)i * (16000.0f / 2.0f) / (float)(N_FREQ - 1);
        const float mel_min = hz_to_mel(0.0f), mel_max = hz_to_mel(8000.0f);
        for (int i = 0; i < N_MEL + 2; i++) ff[i] = mel_to_hz(mel_min + (mel_max - mel_min) * (float)i / (float)(N_MEL + 1));
        for (int i = 0; i < N_MEL + 1; i++) { fd[i] = ff[i + 1] - ff[i]; if (fd[i] == 0.0f) fd[i] = 1e-6f; }
        for (int m = 0; m < N_MEL; m++) {
            const float enorm = 2.0f / (ff[m + 2] - ff[m]);
            for (int f = 0; f < N_FREQ; f++) {
                const float down = (fft_freqs[f] - ff[m]) / fd[m];
                const float up = (ff[m + 2] - fft_freqs[f]) / fd[m + 1];
                float v = fminf(down, up);
                if (v < 0.0f) v = 0.0f;
                filt[(size_t)f * N_MEL + m] = v * enorm;
            }
        }
    }
    CUDA_TRY(cudaMalloc((void **)&g_dft, dft.size() * 4));
    CUDA_TRY(cudaMalloc((void **)&g_filt, filt.size() * 4));
    CUDA_TRY(cudaMalloc((void **)&g_win, win.size() * 4));
    CUDA_TRY(cudaMemcpy(g_dft, dft.data(), dft.size() * 4, cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(g_filt, filt.data(), filt.size() * 4, cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(g_win, win.data(), win.size() * 4, cudaMemcpyHostToDevice));
    return 0;
}

__global__ void __launch_bounds__(MEL_T) mel_frames_kernel(
    const float *__restrict__ samples, int n_frames, const float *__restrict__ dft,
    const float *__restrict__ filt, const float *__restrict__ win, float *__restrict__ mel, long ldm) {
    __shared__ float fr[MEL_FR][N_FFT];
    __shared__ float pw[MEL_FR][N_FREQ + 3];
    const int t0 = blockIdx.x * MEL_FR;
    const int nf = min(MEL_FR, n_frames - t0);
    const int t = threadIdx.x;
    for (int i = t; i < MEL_FR * N_FFT; i += MEL_T) {
        const int f = i / N_FFT, n = i % N_FFT;
        fr[f][n] = (f < nf) ? samples[(size_t)(t0 + f) * HOP + n] * win[n] : 0.f;
    }
    __syncthreads();
    if (t < N_FREQ) {
        float re[MEL_FR], im[MEL_FR];
        #pragma unroll
        for (int f = 0; f < MEL_FR; f++) { re[f] = 0.f; im[f] = 0.f; }
        const float *c = dft + t;
        #pragma unroll 4
        for (int n = 0; n < N_FFT; n++) {
            const float cv = c[(size_t)n * 2 * KP], sv = c[(size_t)n * 2 * KP + KP];
            #pragma unroll
            for (int f = 0; f < MEL_FR; f++) {
                const float x = fr[f][n];
                re[f] = fmaf(x, cv, re[f]);
                im[f] = fmaf(x, sv, im[f]);
            }
        }
        #pragma unroll
        for (int f = 0; f < MEL_FR; f++) pw[f][t] = re[f] * re[f] + im[f] * im[f];
    }
    __syncthreads();
    // 128 mel bins x MEL_FR frames over 256 threads: thread -> (bin, frame half)
    const int m = t & (N_MEL - 1), fh = t >> 7;
    float acc[MEL_FR / 2];
    #pragma unroll
    for (int f = 0; f < MEL_FR / 2; f++) acc[f] = 0.f;
    for (int k = 0; k < N_FREQ; k++) {
        const float w = filt[(size_t)k * N_MEL + m];
        #pragma unroll
        for (int f = 0; f < MEL_FR / 2; f++) acc[f] = fmaf(w, pw[fh * (MEL_FR / 2) + f][k], acc[f]);
    }
    #pragma unroll
    for (int f = 0; f < MEL_FR / 2; f++) {
        const int fi = fh * (MEL_FR / 2) + f;
        if (fi < nf) {
            float s = acc[f];
            if (s < 1e-10f) s = 1e-10f;
            float v = log10f(s);
            const float min_val = 1.5f - 8.0f;
            if (v < min_val) v = min_val;
            mel[(size_t)(t0 + fi) * ldm + m] = (v + 4.0f) / 4.0f;
        }
    }
}

// samples: device, (n_frames-1)*160+400 values; mel: device [n_frames][ldm]
int mel_frames_device(cudaStream_t st, const float *samples, int n_frames, float *mel, long ldm) {
    if (n_frames <= 0) return 0;
    if (mel_tables_init()) return -1;
    mel_frames_kernel<<<(n_frames + MEL_FR - 1) / MEL_FR, MEL_T, 0, st>>>(samples, n_frames, g_dft, g_filt,
                                                                         g_win, mel, ldm);
    LAUNCHED();
    CUDA_TRY(cudaGetLastError());
    return 0;
}

extern "C" int vox_cuda_mel_frames(const float *samples, int n_frames, float *mel_out) {
    if (!samples || !mel_out || n_frames < 0) return -1;
    if (!n_frames) return 0;
    if (!vox_cuda_available() && !vox_cuda_init()) return -1;
    const size_t ns = (size_t)(n_frames - 1) * HOP + N_FFT;
    float *ds = nullptr, *dm = nullptr;
    int rc = -1;
    if (cudaMalloc((void **)&ds, ns * 4) != cudaSuccess) goto done;
    if (cudaMalloc((void **)&dm, (size_t)n_frames * N_MEL * 4) != cudaSuccess) goto done;
    if (cudaMemcpy(ds, samples, ns * 4, cudaMemcpyHostToDevice) != cudaSuccess) goto done;
    if (mel_frames_device(0, ds, n_frames, dm, N_MEL)) goto done;
    if (cudaMemcpy(mel_out, dm, (size_t)n_frames * N_MEL * 4, cudaMemcpyDeviceToHost) != cudaSuccess) goto done;
    rc = 0;
done:
    if (rc) vox_set_error("vox_cuda_mel_frames(%d) failed: %s", n_frames, cudaGetErrorString(cudaGetLastError()));
    cudaFree(ds); cudaFree(dm);
    return rc;
}
