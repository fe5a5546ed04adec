// api_batch.cu - whole-clip ("batch mode") entry point: the hot path of
// vox_transcribe_audio (/root/reference/voxtral.c:1390-1396 -> vox_stream_feed +
// vox_stream_finish) with every stage on the device:
//   padded PCM -> log-mel (mel.cu) -> conv stem + 32 encoder layers + adapter
//   (api_enc.cu / enc_fast.cu) -> prompt prefill (api_dec.cu) -> greedy decode of
//   every remaining adapter position in one persistent kernel (dec_mega.cu).
// Only the PCM goes in and the token ids come out.
#include "vox_internal.h"
#include "enc_state.h"
#include <string.h>

int mel_frames_device(cudaStream_t st, const float *samples, int n_frames, float *mel, long ldm);
int encode_chunk_device_mel(vox_cuda_state *s, int n_mel);

struct BatchBuf {
    float *d_pcm; size_t pcm_cap;     // floats
    float *d_mel; size_t mel_cap;     // frames
    int pcm_n;                        // samples currently resident
};

static BatchBuf *bb_get(vox_cuda_state *s) {
    if (!s->batch) s->batch = calloc(1, sizeof(BatchBuf));
    return (BatchBuf *)s->batch;
}

void batch_free(vox_cuda_state *s) {
    BatchBuf *b = (BatchBuf *)s->batch;
    if (!b) return;
    cudaFree(b->d_pcm); cudaFree(b->d_mel);
    free(b);
    s->batch = nullptr;
}

static int bb_reserve(BatchBuf *b, size_t n_samples, size_t n_frames) {
    if (n_samples > b->pcm_cap) {
        cudaFree(b->d_pcm); b->d_pcm = nullptr; b->pcm_cap = 0;
        CUDA_TRY(cudaMalloc((void **)&b->d_pcm, n_samples * 4));
        b->pcm_cap = n_samples;
    }
    if (n_frames > b->mel_cap) {
        cudaFree(b->d_mel); b->d_mel = nullptr; b->mel_cap = 0;
        CUDA_TRY(cudaMalloc((void **)&b->d_mel, n_frames * VOX_CUDA_MEL_BINS * 4));
        b->mel_cap = n_frames;
    }
    return 0;
}

extern "C" void *vox_cuda_host_alloc(size_t bytes) {
    void *p = nullptr;
    if (cudaMallocHost(&p, bytes) != cudaSuccess) { vox_set_error("cudaMallocHost(%zu) failed", bytes); return nullptr; }
    return p;
}
extern "C" void vox_cuda_host_free(void *p) { if (p) cudaFreeHost(p); }

extern "C" int vox_cuda_pcm_upload(vox_cuda_state_t *s, const float *pcm, int n_samples) {
    if (!s || !pcm || n_samples < 400) return -1;
    BatchBuf *b = bb_get(s);
    if (bb_reserve(b, (size_t)n_samples, (size_t)((n_samples - 400) / 160 + 1))) return -1;
    CUDA_TRY(cudaMemcpyAsync(b->d_pcm, pcm, (size_t)n_samples * 4, cudaMemcpyHostToDevice, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    b->pcm_n = n_samples;
    return 0;
}

// front half of the whole-clip call: PCM (host, or resident when pcm == NULL) -> adapter rows on the device
int batch_encode_clip(vox_cuda_state *s, const float *pcm, int n_samples, int n_frames) {
    BatchBuf *b = bb_get(s);
    EncState *e = (EncState *)s->enc_priv;
    cudaStream_t st = s->stream;
    if (pcm) {
        if (n_samples < (n_frames - 1) * 160 + 400) { vox_set_error("transcribe: %d samples < %d frames", n_samples, n_frames); return -1; }
        if (bb_reserve(b, (size_t)n_samples, (size_t)n_frames)) return -1;
    } else {
        if (!b->d_pcm || b->pcm_n < (n_frames - 1) * 160 + 400) { vox_set_error("transcribe: no resident PCM (vox_cuda_pcm_upload)"); return -1; }
        if (bb_reserve(b, 0, (size_t)n_frames)) return -1;
    }
    if (pcm) CUDA_TRY(cudaMemcpyAsync(b->d_pcm, pcm, (size_t)n_samples * 4, cudaMemcpyHostToDevice, st));
    if (mel_frames_device(st, b->d_pcm, n_frames, b->d_mel, VOX_CUDA_MEL_BINS)) return -1;
    vox_cuda_state_reset_encoder(s);
    vox_cuda_state_reset_decoder(s);
    for (int off = 0; off < n_frames; off += ENC_MAXMEL) {
        const int n = n_frames - off < ENC_MAXMEL ? n_frames - off : ENC_MAXMEL;
        CUDA_TRY(cudaMemcpyAsync(e->mel_in + 2 * VOX_CUDA_MEL_BINS, b->d_mel + (size_t)off * VOX_CUDA_MEL_BINS,
                                 (size_t)n * VOX_CUDA_MEL_BINS * 4, cudaMemcpyDeviceToDevice, st));
        if (encode_chunk_device_mel(s, n)) return -1;
    }
    return 0;
}

// CUDA-event time (ms) of the encoder side alone on the resident PCM (vox_cuda_pcm_upload): log-mel -> conv stem ->
// 32 layers -> adapter, i.e. what stream_run_encoder does for a whole clip (voxtral.c:826-951).  bench.py's tensor roofline.
extern "C" float vox_cuda_bench_encoder(vox_cuda_state_t *s, int n_samples, int n_frames) {
    if (!s || n_frames <= 0) return -1.f;
    cudaStream_t st = s->stream;
    if (cudaEventRecord(s->ev0, st) != cudaSuccess) return -1.f;
    if (batch_encode_clip(s, nullptr, n_samples, n_frames)) return -1.f;
    if (cudaEventRecord(s->ev1, st) != cudaSuccess || cudaEventSynchronize(s->ev1) != cudaSuccess) return -1.f;
    float ms = 0.f;
    cudaEventElapsedTime(&ms, s->ev0, s->ev1);
    return ms;
}

extern "C" int vox_cuda_transcribe(vox_cuda_state_t *s, const float *pcm, int n_samples, int n_frames,
                                   int prompt_len, int *tokens_out, int max_tokens, float *device_ms) {
    if (!s || !tokens_out || n_frames <= 0 || prompt_len < 2 || prompt_len > 128) return -1;
    cudaStream_t st = s->stream;
    CUDA_TRY(cudaEventRecord(s->ev0, st));
    if (batch_encode_clip(s, pcm, n_samples, n_frames)) return -1;
    const int T = s->adapter_count;
    const int n_pre = prompt_len - 1;            // rows fed to prefill; row n_pre produces the first token
    if (T < prompt_len) { vox_set_error("transcribe: only %d adapter tokens (< prompt %d)", T, prompt_len); return -1; }
    int prompt[128];
    prompt[0] = 1;                               // BOS, then [STREAMING_PAD] (voxtral.c:396-399,1039-1048)
    for (int i = 1; i < prompt_len; i++) prompt[i] = 32;
    if (vox_cuda_decoder_prefill_fused(s, prompt, 0, n_pre)) return -1;
    int n = T - n_pre;
    if (n > max_tokens) n = max_tokens;
    if (vox_cuda_decoder_run(s, prompt[n_pre], n_pre, n, tokens_out) != n) return -1;
    CUDA_TRY(cudaEventRecord(s->ev1, st));
    CUDA_TRY(cudaEventSynchronize(s->ev1));
    if (device_ms) cudaEventElapsedTime(device_ms, s->ev0, s->ev1);
    return n;
}
