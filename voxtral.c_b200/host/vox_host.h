/*
 * vox_host.h - C host over the CUDA backend (libvoxcuda.so): the slice of the reference's
 * public API that sits on the hot path, with the same names' meaning and error conventions
 * (NULL / -1 on failure plus a line on stderr, /root/reference/voxtral.h:252-337):
 *
 *   reference                               here
 *   vox_load(model_dir)                  -> vox_b200_load(model_dir)        mmap + weight table + upload
 *   vox_free(ctx)                        -> vox_b200_free(ctx)
 *   vox_set_delay(ctx, ms)               -> vox_b200_set_delay_tokens(ctx, n)
 *   vox_stream_init / feed / finish      -> vox_b200_stream_init / feed / finish
 *   vox_stream_get (token text)          -> vox_b200_stream_get_ids (token ids; the tokenizer is
 *                                           outside the hot path)
 *   vox_set_processing_interval          -> vox_b200_set_processing_interval
 *   vox_transcribe_audio                 -> vox_b200_transcribe_ids (whole clip in one device call)
 *
 * Plain C, no CPU fallback: every call fails if the CUDA backend is unavailable.
 */
#ifndef VOX_HOST_H
#define VOX_HOST_H
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct vox_b200_ctx vox_b200_ctx;
typedef struct vox_b200_stream vox_b200_stream;

#define VOX_TOKEN_BOS 1
#define VOX_TOKEN_EOS 2
#define VOX_TOKEN_STREAMING_PAD 32

vox_b200_ctx *vox_b200_load(const char *model_dir);
void vox_b200_free(vox_b200_ctx *ctx);
int vox_b200_set_delay_tokens(vox_b200_ctx *ctx, int delay_tokens);   /* default 6 (480 ms) */

/* Whole clip: samples f32 mono 16 kHz in [-1, 1].  Writes the greedy token ids up to (not
 * including) EOS; returns their count or -1.  device_ms (optional) = CUDA-event time. */
int vox_b200_transcribe_ids(vox_b200_ctx *ctx, const float *samples, int n_samples,
                            int *ids_out, int max_ids, float *device_ms);

vox_b200_stream *vox_b200_stream_init(vox_b200_ctx *ctx);
void vox_b200_set_processing_interval(vox_b200_stream *s, float seconds);   /* -I, default 2.0 */
int vox_b200_stream_feed(vox_b200_stream *s, const float *samples, int n_samples);
int vox_b200_stream_finish(vox_b200_stream *s);
/* Pop up to max pending token ids (all generated tokens, control tokens included). */
int vox_b200_stream_get_ids(vox_b200_stream *s, int *ids_out, int max);
/* Wall-clock split like the reference's summary lines (voxtral.c:1359-1370). */
void vox_b200_stream_stats(const vox_b200_stream *s, double *mel_ms, double *encoder_ms, double *decoder_ms,
                           int *encoder_calls, int *decoder_steps);
/* "Encoder: %d mel -> %d tokens (%.0f ms)" and "Decoder: %d text tokens (%d steps) in ..." exactly as the
 * reference prints them (voxtral.c:1359-1370; parsed by its benchmark.py:25-30); verbose adds the row count
 * of every encoder call. */
void vox_b200_stream_report(const vox_b200_stream *s, FILE *f, int verbose);
void vox_b200_stream_free(vox_b200_stream *s);

/* 16-bit PCM mono 16 kHz WAV -> f32 samples (malloc'd, caller frees); NULL on anything else. */
float *vox_b200_load_wav(const char *path, int *n_samples);

#ifdef __cplusplus
}
#endif
#endif
