/*
 * vox_cuda_dropin.h - the hooks the reference's host sources call when they are built with -DUSE_CUDA
 * (integration/use_cuda.patch).  Each one is the CUDA twin of a `#ifdef USE_METAL` site of the
 * reference and is implemented in vox_cuda_dropin.c on top of the C ABI of include/vox_cuda.h.
 * Include after voxtral.h (uses vox_ctx_t).
 *
 * The host keeps owning its integers (kv_cache_len, kv_pos_offset, enc_kv_cache_len,
 * enc_kv_pos_offset, voxtral.h:197-224) and its stream state machine; every tensor lives on the device.
 * There is no CPU fallback: a failed hook prints vox_cuda_last_error() and the caller stops.
 */
#ifndef VOX_CUDA_DROPIN_H
#define VOX_CUDA_DROPIN_H
#include "vox_cuda.h"

/* vox_load, after the weights are mapped and ada_scale is computed (twin of the warm-up block
 * voxtral.c:186-285): upload, create the device state.  0 on success. */
int vox_cuda_dropin_load(vox_ctx_t *ctx);
/* vox_free (twin of voxtral.c:331-351) */
void vox_cuda_dropin_free(vox_ctx_t *ctx);
/* vox_set_delay, after vox_update_time_conditioning (voxtral.c:1681-1687) */
void vox_cuda_dropin_set_delay(vox_ctx_t *ctx);
/* vox_decoder_forward at the vox_metal_decoder_full_step site (voxtral_decoder.c:686-699): one token on
 * the device at logical position kv_pos_offset + kv_cache_len; kv_cache_len++ on success.  Returns the
 * greedy id, or 2 (EOS, the reference's own out-of-memory convention, :675) after printing the error. */
int vox_cuda_dropin_decoder_forward(vox_ctx_t *ctx, const float *embed, float *logits);
/* vox_decoder_prefill at the vox_metal_decoder_prefill_step site (voxtral_decoder.c:485-493);
 * kv_cache_len += seq_len on success */
void vox_cuda_dropin_decoder_prefill(vox_ctx_t *ctx, const float *x, int seq_len);
/* vox_encoder_forward_incremental at the vox_metal_encoder_full_step site (voxtral_encoder.c:551-560):
 * x [new_len,1280] in/out; the caller sets enc_kv_cache_len = cache_len + new_len.  0 on success. */
int vox_cuda_dropin_encoder_step(vox_ctx_t *ctx, float *x, int new_len, int cache_len);
/* vox_adapter_forward (voxtral_encoder.c:699-737): malloc'd [enc_len/4,3072] */
float *vox_cuda_dropin_adapter(vox_ctx_t *ctx, const float *enc_out, int enc_len, int *out_len);
/* stream_conv_stem (voxtral.c:581-759): malloc'd [*out_len,1280] or NULL; first = the stream's
 * conv_stem_initialized is still 0 (device tails are cleared) */
float *vox_cuda_dropin_conv_stem(vox_ctx_t *ctx, int first, const float *mel, int n_mel, int *out_len);
/* vox_metal_sgemm_bf16 sites (voxtral_kernels.c:199-204,218-229,249-254): C = A x B^T.  1 = done */
int vox_cuda_dropin_sgemm_bf16(int M, int N, int K, const float *A, const uint16_t *B, float *C);
#endif
