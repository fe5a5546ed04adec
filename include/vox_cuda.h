/*
 * vox_cuda.h - C ABI of the B200 (sm_100a) CUDA backend for voxtral.c
 *
 * This header is the drop-in boundary.  It plays the role voxtral_metal.h plays
 * for the reference's Apple backend (/root/reference/voxtral_metal.h:20-282):
 * plain C, pointers and sizes only, callable from one host thread per GPU.
 * Every entry point names the reference interface it stands in for.
 *
 * Ownership model (differs from Metal's unified memory on purpose): weights are
 * uploaded ONCE from the mmap'd safetensors pointers into HBM
 * (vox_cuda_model_create), and every per-stream tensor - decoder KV ring,
 * encoder KV ring, conv-stem tails, adapter-token buffer - lives in a device
 * resident vox_cuda_state_t.  The host keeps only integers (lengths, offsets).
 *
 * Error convention follows the reference (/root/reference/voxtral.c:149,158):
 * a negative / NULL return plus one line on stderr; vox_cuda_last_error()
 * returns the same text.  There is no CPU fallback anywhere behind this ABI.
 */
#ifndef VOX_CUDA_H
#define VOX_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Model constants - identical to /root/reference/voxtral.h:19-50. */
#define VOX_CUDA_MEL_BINS      128
#define VOX_CUDA_ENC_DIM       1280
#define VOX_CUDA_ENC_LAYERS    32
#define VOX_CUDA_ENC_HEADS     32
#define VOX_CUDA_ENC_HEAD_DIM  64
#define VOX_CUDA_ENC_HIDDEN    5120
#define VOX_CUDA_ENC_WINDOW    750
#define VOX_CUDA_DEC_DIM       3072
#define VOX_CUDA_DEC_LAYERS    26
#define VOX_CUDA_DEC_HEADS     32
#define VOX_CUDA_DEC_KV_HEADS  8
#define VOX_CUDA_DEC_HEAD_DIM  128
#define VOX_CUDA_DEC_HIDDEN    9216
#define VOX_CUDA_DEC_WINDOW    8192
#define VOX_CUDA_VOCAB         131072
#define VOX_CUDA_MAX_ALT       4

/* ------------------------------------------------------------------------
 * Weight table: host pointers exactly as the reference loaders leave them
 * (bf16 matrices = pointers into the mmap, small tensors = malloc'd f32);
 * /root/reference/voxtral.h:56-182, loaders voxtral_encoder.c:58-146,
 * voxtral_decoder.c:57-145, voxtral.c:102-125.
 * ------------------------------------------------------------------------ */
typedef struct {
    const uint16_t *wq, *wk, *wv, *wo;        /* [2048,1280]x3, [1280,2048] bf16 */
    const uint16_t *w1, *w2, *w3;             /* [5120,1280], [1280,5120], [5120,1280] */
    const float *wq_bias, *wv_bias;           /* [2048] (wk has no bias) */
    const float *wo_bias, *w2_bias;           /* [1280] */
    const float *attention_norm, *ffn_norm;   /* [1280] */
} vox_cuda_enc_layer_w_t;

typedef struct {
    const uint16_t *wq, *wk, *wv, *wo;        /* [4096,3072],[1024,3072]x2,[3072,4096] */
    const uint16_t *w1, *w2, *w3;             /* [9216,3072],[3072,9216],[9216,3072] */
    const float *attention_norm, *ffn_norm;   /* [3072] */
} vox_cuda_dec_layer_w_t;

typedef struct {
    int n_enc_layers;                         /* 32 (fewer allowed: test models) */
    int n_dec_layers;                         /* 26 */
    const float *conv0_weight, *conv0_bias;   /* [1280,128,3], [1280] f32 */
    const float *conv1_weight, *conv1_bias;   /* [1280,1280,3], [1280] f32 */
    vox_cuda_enc_layer_w_t enc[VOX_CUDA_ENC_LAYERS];
    const float *enc_norm;                    /* [1280] */
    const uint16_t *adapter_w0;               /* [3072,5120] bf16 */
    const uint16_t *adapter_w1;               /* [3072,3072] bf16 */
    const uint16_t *tok_embeddings;           /* [131072,3072] bf16 (tied LM head) */
    vox_cuda_dec_layer_w_t dec[VOX_CUDA_DEC_LAYERS];
    const float *dec_norm;                    /* [3072] */
    const float *ada_scale;                   /* [n_dec_layers*3072] f32, voxtral.c:47-80 */
} vox_cuda_weights_t;

/* Q8 container: what the reference's loader hands out for a dtype "Q8" tensor -
 * safetensors_get_q8_data_direct / _scales_direct (voxtral_safetensors.c:457-468), fields *_weight_q8 /
 * *_scale_q8 of vox_enc_layer_t / vox_dec_layer_t / vox_adapter_t (voxtral.h:56-182): int8 rows, one
 * fp32 scale per row, w[r][c] = q[r][c] * scale[r] (quantize.py:34-43). */
typedef struct {
    const int8_t *q;                          /* [rows, cols] row-major */
    const float *scale;                       /* [rows] */
} vox_cuda_q8_t;

typedef struct { vox_cuda_q8_t wq, wk, wv, wo, w1, w2, w3; } vox_cuda_layer_q8_t;   /* shapes as the bf16 tables above */

typedef struct {
    vox_cuda_layer_q8_t enc[VOX_CUDA_ENC_LAYERS];
    vox_cuda_q8_t adapter_w0, adapter_w1;     /* [3072,5120], [3072,3072] */
    vox_cuda_q8_t tok_embeddings;             /* [131072,3072] (tied LM head; embedding row = q * scale, voxtral.c:443-452) */
    vox_cuda_layer_q8_t dec[VOX_CUDA_DEC_LAYERS];
} vox_cuda_weights_q8_t;

typedef struct vox_cuda_model vox_cuda_model_t;   /* weights in HBM, shared by streams */
typedef struct vox_cuda_state vox_cuda_state_t;   /* one stream's device state */
typedef struct vox_cuda_group vox_cuda_group_t;   /* B streams decoded in lock step (one weight pass per step) */
typedef struct vox_cuda_tp vox_cuda_tp_t;         /* one stream decoded by G ranks (tensor-parallel decoder step) */

/* ---- lifetime (replaces vox_metal_init/available/shutdown,
 *      voxtral_metal.h:20-26, call sites main.c:181-183,407-409) ---------- */
int  vox_cuda_init(void);            /* 1 = ready, 0 = no usable sm_100 device */
int  vox_cuda_available(void);
void vox_cuda_shutdown(void);
const char *vox_cuda_last_error(void);
size_t vox_cuda_memory_used(void);   /* vox_metal_memory_used, voxtral_metal.h:282 */

/* ---- weights (replaces the warm-up block in vox_load, voxtral.c:186-285:
 *      vox_metal_warmup_bf16 / _merged_2 / _merged_3 / _decoder_ops) ------- */
vox_cuda_model_t *vox_cuda_model_create(const vox_cuda_weights_t *w);
/* Q8 checkpoint (ctx->use_q8, voxtral.c:175-181): `w` supplies the fp32 tensors (conv stem, biases, norms,
 * ada_scale; its bf16 pointers are ignored), `q` every 2-D matrix.  Arithmetic follows the reference's Q8 CPU
 * path: y = (sum_k q[n][k] * x[k]) * scale[n] with the sum exact to fp32 class (voxtral_kernels.c:277-310 for
 * M = 1; for M > 1 the reference widens q * scale to fp32 and calls sgemm, :322-336 - the same product up to
 * one fp32 rounding per weight).  The single-stream decoder streams the int8 rows (half the bytes per token);
 * encoder, adapter, prompt prefill, the stream groups' lock-step decode and the tensor-parallel decoder run on
 * the int8 values held as bf16 with the scale applied to each product (no byte saving there).  Only the legacy
 * graph engine and the fp32 debug modes refuse a Q8 model (an error - not a fallback). */
vox_cuda_model_t *vox_cuda_model_create_q8(const vox_cuda_weights_t *w, const vox_cuda_weights_q8_t *q);
int  vox_cuda_model_is_q8(const vox_cuda_model_t *m);
/* vox_set_delay recomputes ada_scale on the host (voxtral.c:1681-1687). */
int  vox_cuda_model_set_ada_scale(vox_cuda_model_t *m, const float *ada_scale);
void vox_cuda_model_free(vox_cuda_model_t *m);

/* ---- per-stream state (replaces vox_metal_shared_alloc'd KV caches,
 *      voxtral_metal.h:206-207, voxtral_decoder.c:208-243,
 *      voxtral_encoder.c:371-391, and the stream's conv tails
 *      voxtral.c:469-483) ------------------------------------------------- */
vox_cuda_state_t *vox_cuda_state_create(vox_cuda_model_t *m);
/* kv_bits 32 = the call above (fp32 rings: the reference's CPU / BLAS layout, ids equal the reference);
 * 16 = fp16 decoder K/V rings, the reference's GPU default ("fp16 KV", voxtral.c:189-190, SPEED.md:24-43):
 * rows are rounded once when appended, scores / softmax / output stay fp32; half the K/V bytes per token
 * (0.87 instead of 1.74 GB at the full 8192-key window).  Not bit-comparable with the fp32 BLAS oracle: the
 * tests state the logit tolerance (tests/test_gpu_kv16.py).  Prompt prefill: at most 128 rows at position 0.
 * vox_cuda_decoder_kv_read / _write keep fp32 host rows and convert.  Default engine only (not the legacy
 * graph decoder); the tensor-parallel decoder takes its own kv_bits (vox_cuda_tp_create_ex). */
vox_cuda_state_t *vox_cuda_state_create_ex(vox_cuda_model_t *m, int kv_bits);
int  vox_cuda_state_kv_bits(const vox_cuda_state_t *s);
void vox_cuda_state_free(vox_cuda_state_t *s);
/* kv_cache_len = kv_pos_offset = 0 (voxtral.c:781-782,1050-1051). */
void vox_cuda_state_reset_decoder(vox_cuda_state_t *s);
/* enc_kv_cache_len = enc_kv_pos_offset = 0, conv tails / residuals cleared,
 * adapter buffer emptied (voxtral.c:797-824,1279-1280). */
void vox_cuda_state_reset_encoder(vox_cuda_state_t *s);
/* Encoder twin of the two calls below: logical position of the next encoder row
 * (enc_kv_pos_offset + enc_kv_cache_len, voxtral.h:218-224).  A host that resets or
 * compacts its own integers (voxtral.c:819-820, voxtral_encoder.c:431-449) passes the
 * sum on before the next vox_cuda_encoder_step. */
int  vox_cuda_encoder_position(const vox_cuda_state_t *s);
int  vox_cuda_encoder_set_position(vox_cuda_state_t *s, int logical_pos);
/* conv_stem_initialized = 0 (voxtral.c:807-817): tails and parity residual only. */
void vox_cuda_state_reset_stem(vox_cuda_state_t *s);
/* Logical decoder position = kv_pos_offset + kv_cache_len of the reference. */
int  vox_cuda_decoder_position(const vox_cuda_state_t *s);
/* Override the decoder position bookkeeping (test hook for the 8192-window
 * wrap: lets a harness start right before a compaction boundary the way the
 * reference allows through its public ctx fields, voxtral.h:197-205). */
int  vox_cuda_decoder_set_position(vox_cuda_state_t *s, int logical_pos, int n_valid);
/* Copy n rows [n,1024] of fp32 K and V of one decoder layer between the host and
 * the device ring, addressed by LOGICAL position (slot = pos & 8191); pos + n may
 * cross the ring end.  The host layout is the reference's kv_cache_k/v row layout
 * (voxtral_decoder.c:188-196), so a harness - or a host that switches a running
 * stream from the CPU path to the GPU - can move ctx->kv_cache_k + (layer *
 * kv_cache_max + physical) * 1024 straight across (voxtral.h:197-205). */
int  vox_cuda_decoder_kv_write(vox_cuda_state_t *s, int layer, int logical_pos, int n,
                               const float *k, const float *v);
int  vox_cuda_decoder_kv_read(vox_cuda_state_t *s, int layer, int logical_pos, int n,
                              float *k, float *v);

/* ---- decoder (hot path #1) -------------------------------------------- */
/* vox_decoder_prefill / vox_metal_decoder_prefill_step
 * (voxtral_decoder.c:447-612, voxtral_metal.h:254). x: host [seq_len,3072]. */
int vox_cuda_decoder_prefill(vox_cuda_state_t *s, const float *x, int seq_len);
/* vox_decoder_forward / vox_metal_decoder_full_step (voxtral_decoder.c:640-780,
 * voxtral_metal.h:219).  embed: host [3072].  logits: host [131072] or NULL.
 * Returns the greedy token id (lowest index wins ties), <0 on error. */
int vox_cuda_decoder_step(vox_cuda_state_t *s, const float *embed, float *logits);
/* Same step with the embedding built on the device:
 * embed = adapter_buf[adapter_index] + bf16->f32(tok_embeddings[prev_token])
 * (voxtral.c:1106-1113).  Only the 4-byte token id crosses PCIe. */
int vox_cuda_decoder_step_fused(vox_cuda_state_t *s, int prev_token,
                                int adapter_index, float *logits);
/* n greedy steps back to back on the device: step i consumes adapter token
 * adapter_index+i and the token produced by step i-1 (the while loop of
 * stream_run_decoder, voxtral.c:1105-1145, without a host round trip per
 * token).  The caller truncates at EOS.  Returns n or <0. */
int vox_cuda_decoder_run(vox_cuda_state_t *s, int prev_token, int adapter_index,
                         int n, int *tokens_out);
/* Logits [131072] of the last step (device -> host). */
int vox_cuda_decoder_read_logits(vox_cuda_state_t *s, float *logits);
/* Prompt prefill with device-built embeddings (voxtral.c:1036-1055). */
int vox_cuda_decoder_prefill_fused(vox_cuda_state_t *s, const int *tokens,
                                   int adapter_start, int seq_len);
/* Alternatives of the LAST step (stream_fill_alts, voxtral.c:955-1010):
 * candidate ids >= 1000 and != best, ordered by probability, kept while
 * 1 - p_i/p_best <= cutoff.  Returns the count written (<= max_alt-1). */
int vox_cuda_decoder_alternatives(vox_cuda_state_t *s, int max_alt, float cutoff,
                                  int *ids);

/* The same rule fused into the decoder kernel (vox_stream_set_alt, voxtral.c:1330-1339): with n_alt > 1 every
 * following step - single, vox_cuda_decoder_run, vox_cuda_transcribe - also records the best three other
 * text logits while the logits row is produced (no second pass over the 131072 logits, no logits copy).
 * vox_cuda_decoder_get_alts returns, for the LAST n steps, n rows of VOX_CUDA_MAX_ALT ids: [0] = the greedy
 * id, then the alternatives that pass the cutoff, -1 = none (what vox_stream_get_alt hands out per text
 * token, voxtral.c:1341-1357; non-text ids get no alternatives).  n_alt = 1 switches recording off. */
int vox_cuda_decoder_set_alt(vox_cuda_state_t *s, int n_alt, float cutoff);
int vox_cuda_decoder_get_alts(vox_cuda_state_t *s, int n, int *alts_out);

/* ---- encoder (hot path #2) -------------------------------------------- */
/* vox_encoder_forward_incremental / vox_metal_encoder_full_step
 * (voxtral_encoder.c:495-693, voxtral_metal.h:245): x host [new_len,1280]
 * in/out, 32 layers + final norm against the device encoder KV ring. */
int vox_cuda_encoder_step(vox_cuda_state_t *s, float *x, int new_len);
/* stream_conv_stem (voxtral.c:581-759): mel host [n_mel,128] ->
 * out host [*out_len,1280]; tails and the parity residual stay on device.
 * out must hold (n_mel/2 + 2) rows. */
int vox_cuda_conv_stem(vox_cuda_state_t *s, const float *mel, int n_mel,
                       float *out, int *out_len);
/* vox_adapter_forward (voxtral_encoder.c:699-737): enc_out host
 * [enc_len,1280] -> out host [enc_len/4,3072]. */
int vox_cuda_adapter(vox_cuda_state_t *s, const float *enc_out, int enc_len,
                     float *out);
/* Whole encoder side of stream_run_encoder (voxtral.c:845-934) on the device:
 * conv stem -> 32 layers -> 4-alignment residual -> adapter -> append to the
 * device adapter buffer.  mel: host [n_mel,128].  Returns the total number of
 * adapter tokens now buffered (>=0) or <0 on error. */
int vox_cuda_encode_chunk(vox_cuda_state_t *s, const float *mel, int n_mel);
int vox_cuda_adapter_count(const vox_cuda_state_t *s);
/* Copy adapter tokens [start, start+n) to host [n,3072] (tests, fallback API). */
int vox_cuda_adapter_read(vox_cuda_state_t *s, int start, int n, float *out);
/* Host-provided adapter tokens (used when the encoder ran elsewhere). */
int vox_cuda_adapter_append(vox_cuda_state_t *s, const float *tok, int n);
/* stream_adapter_compact (voxtral.c:762-775): forget tokens < keep_from. */
int vox_cuda_adapter_discard_before(vox_cuda_state_t *s, int keep_from);

/* ---- log-mel front end (hot path #3) ---------------------------------- */
/* mel_compute_available (voxtral_audio.c:454-513), batched over frames:
 * frame t reads samples[t*160 .. t*160+399]; out host [n_frames,128].
 * samples must hold (n_frames-1)*160+400 values. */
int vox_cuda_mel_frames(const float *samples, int n_frames, float *mel_out);

/* ---- whole clip (batch mode) --------------------------------------------
 * The hot path of vox_transcribe_audio (voxtral.c:1390-1396: vox_stream_feed of
 * the whole clip + vox_stream_finish) in one call with every stage on the
 * device.  pcm: the PADDED sample buffer exactly as the reference's mel context
 * sees it - 200 + 32*1280 zeros, the samples, the flush padding
 * (voxtral.c:1645-1658) and the 200 reflected samples of vox_mel_finish
 * (voxtral_audio.c:602-629); n_frames = mel frames to compute.  pcm may be a
 * host pointer (pinned memory from vox_cuda_host_alloc avoids a staging copy)
 * or NULL to use the buffer last uploaded with vox_cuda_pcm_upload (device
 * resident input).  prompt_len = 1 + 32 + delay_tokens (39).  Writes the greedy
 * token ids of all decode steps (the caller truncates at EOS = 2) and returns
 * their count; *device_ms (optional) = CUDA-event time of the whole call. */
void *vox_cuda_host_alloc(size_t bytes);
void  vox_cuda_host_free(void *p);
int   vox_cuda_pcm_upload(vox_cuda_state_t *s, const float *pcm, int n_samples);
int   vox_cuda_transcribe(vox_cuda_state_t *s, const float *pcm, int n_samples, int n_frames,
                          int prompt_len, int *tokens_out, int max_tokens, float *device_ms);

/* ---- stream groups: B independent streams, one pass over the weights per step ----------------
 * The reference holds one stream per vox_ctx_t (voxtral.h:188-239, voxtral.c:1242-1286) and
 * vox_decoder_forward (voxtral_decoder.c:640-780) reads all 6.86 GB of decoder weights for each
 * token of each stream.  A group keeps the adapter rows and K/V rings of n_streams (<= 128)
 * streams on the device and advances all of them by one token per step, so the weights cross HBM
 * once per step for every stream.  Per-stream arithmetic is that of vox_decoder_prefill /
 * vox_decoder_forward on the stream's own rows: the greedy ids of stream i equal a single-stream
 * run on the same audio.  max_positions bounds the adapter tokens of one stream (its K/V ring
 * holds min(8192, next power of two) rows, 212,992 B per position).
 *   vox_cuda_group_encode      whole-clip front end of one stream (same pcm / n_frames contract as
 *                              vox_cuda_transcribe); returns its adapter-token count
 *   vox_cuda_group_set_adapter adapter rows computed elsewhere (host fp32 [n,3072])
 *   vox_cuda_group_decode      prompt prefill (prompt_len = 39) of every encoded stream, then
 *                              lock-step greedy decode of all remaining positions; tokens_out is
 *                              [n_streams][max_tokens], counts_out[i] = ids written for stream i
 *                              (the caller truncates at EOS = 2); returns the number of steps */
vox_cuda_group_t *vox_cuda_group_create(vox_cuda_model_t *m, int n_streams, int max_positions);
/* kv_bits 32 = the call above; 16 = K/V rings in fp16, the reference's GPU default ("fp16 KV",
 * voxtral.c:189-190): rows are rounded once when appended, scores / softmax / output
 * stay fp32.  Half the K/V bytes per lock step.  Not bit-comparable with the fp32 BLAS oracle:
 * tests assert equal greedy ids on the test clips (top-2 margins ~1e-1 vs ~1e-3 of rounding). */
vox_cuda_group_t *vox_cuda_group_create_ex(vox_cuda_model_t *m, int n_streams, int max_positions, int kv_bits);
void vox_cuda_group_free(vox_cuda_group_t *g);
int  vox_cuda_group_streams(const vox_cuda_group_t *g);
int  vox_cuda_group_encode(vox_cuda_group_t *g, int stream, const float *pcm, int n_samples, int n_frames);
int  vox_cuda_group_set_adapter(vox_cuda_group_t *g, int stream, const float *rows, int n);
int  vox_cuda_group_decode(vox_cuda_group_t *g, int prompt_len, int *tokens_out, int max_tokens,
                           int *counts_out, float *device_ms);

/* The group's matrix product alone, host buffers (parity tests of the kernel): C[b][n] =
 * sum_k A[b][k] W[n][k] for B <= 128 activation rows at once - vox_linear_nobias_bf16
 * (voxtral_kernels.c:215-236) for B streams in one pass over W.  A is split into two bf16 terms on
 * the device (fp32-class accuracy), products accumulate in fp32 in TMEM; the K range is cut into
 * ksplit parts summed in order.  swiglu != 0: rows (2i, 2i+1) of W are (gate, up) and C is
 * [B, N/2] = silu(g) * u (needs ksplit 1).  N % 128 == 0, K % 64 == 0.  0 on success. */
int vox_cuda_gemm_bf16_streams(int B, int N, int K, const float *A, const uint16_t *W_bf16, int ksplit,
                               int swiglu, float *C);

/* Device-resident timing of that product alone (weights rotate over > 400 MB of copies so no launch
 * finds its matrix in L2): mean microseconds per launch over iters launches, < 0 on error. */
float vox_cuda_bench_gemm_streams(int B, int N, int K, int ksplit, int swiglu, int iters);

/* ---- op-level hook (replaces vox_metal_sgemm_bf16, voxtral_metal.h:38,
 *      call sites voxtral_kernels.c:199-204,218-229,249-254) -------------- */
/* C[M,N] = A[M,K] (f32 host) x B[N,K]^T.  B_bf16 is a host pointer: if it is
 * one of the matrices uploaded by vox_cuda_model_create (looked up by pointer
 * identity, as Metal's weight cache does, voxtral_metal.m:165-201) the HBM
 * copy is used, otherwise B is uploaded for the call.  m may be NULL.
 * N % 64 == 0 and K % 32 == 0.  fp32 accumulate.  0 on success. */
int vox_cuda_sgemm_bf16(vox_cuda_model_t *m, int M, int N, int K, const float *A,
                        const uint16_t *B_bf16, float *C);

/* Tensor-core twin of the hook above (the encoder's throughput path): A is
 * rounded to bf16 on the device, products accumulate in fp32 on the tcgen05
 * tensor cores (TMEM).  Differs from vox_cuda_sgemm_bf16 by the bf16 rounding of
 * A only (relative 2^-9 per element); the encoder tolerance of 2e-2 max-abs
 * covers it.  flags: VOX_CUDA_TC_* ; bias [N] may be NULL; with
 * VOX_CUDA_TC_RESIDUAL C is read and accumulated into; with VOX_CUDA_TC_SWIGLU
 * columns are (gate, up) pairs and C is [M, N/2].  N % 128 == 0, K % 64 == 0. */
#define VOX_CUDA_TC_BIAS      1
#define VOX_CUDA_TC_GELU      2
#define VOX_CUDA_TC_RESIDUAL  4
#define VOX_CUDA_TC_SWIGLU    8
#define VOX_CUDA_TC_SPLIT_A   128  /* A = hi + lo (two bf16 terms): fp32-class accuracy, 2x tensor work */
int vox_cuda_gemm_bf16_tc(int M, int N, int K, const float *A, const uint16_t *B_bf16,
                          const float *bias, int flags, float *C);

/* ---- audio ingestion on the device --------------------------------------
 * The arithmetic of vox_parse_wav_buffer (voxtral_audio.c:94-137) and of the raw s16le stdin reader
 * (:199-202) without the container parsing: interleaved s16 frames -> mono (mean of the channels)
 * / 32768 -> 16 kHz by linear interpolation when sample_rate != 16000.  Bit-identical to the
 * reference compiled in strict IEEE mode.  Host pointers in and out; returns the number of output
 * samples (vox_cuda_pcm_resampled_len) or -1. */
int vox_cuda_pcm_resampled_len(int n_frames, int sample_rate);
int vox_cuda_pcm_from_s16(const int16_t *pcm, int n_frames, int channels, int sample_rate, float *out, int out_cap);
/* the microphone silence gate's energy (main.c:262-292): rms_out[w] = sqrt(mean(x^2)) of consecutive
 * windows of `window` samples (160 = 10 ms there); returns n / window */
int vox_cuda_window_rms(const float *samples, int n, int window, float *rms_out);

/* ---- tensor-parallel single-stream decoder (csrc/dec_tp.cu) ----------
 * vox_decoder_forward (voxtral_decoder.c:640-780) cut across G = 1/2/4/8 GPUs, one
 * process per GPU: rank r owns kv heads [r*8/G, ...) with their query heads (rows of
 * wq|wk|wv), rows of w1|w3, the matching column range of wo and w2, and a vocabulary
 * slice of the logits.  Weights stay replicated (the shard is a view of the model this
 * process already holds).  The sums after wo and w2 and the greedy (max, index) pairs are
 * exchanged by the GEMV kernels themselves, which store into every peer's exchange window
 * over NVLink (CUDA IPC peer memory) and publish a sequence number; the next kernel's
 * prologue waits and adds the G partials in rank order, so all ranks hold identical rows.
 * The reference has no multi-device path; this is the north star's optional TP decoder.
 *
 *   tp = vox_cuda_tp_create(model, rank, world);  vox_cuda_tp_export(tp, my_handle);
 *   ... all-gather the handles between the processes (any transport) ...
 *   vox_cuda_tp_connect(tp, all_handles);
 *   s = vox_cuda_tp_state(tp);   encode + prefill s on EVERY rank with the usual calls
 *   vox_cuda_tp_decode(tp, prev_token, adapter_index, n, ids, &ms);   (all ranks together)
 * Before vox_cuda_tp_free every rank must have returned from its last decode (the peers
 * write into the window being freed): put a barrier between them.
 * After a TP decode the state's K/V ring holds only this rank's kv heads for the decoded
 * positions: keep decoding it through vox_cuda_tp_decode. */
#define VOX_CUDA_TP_MAX 8
#define VOX_CUDA_TP_HANDLE_BYTES 80
vox_cuda_tp_t *vox_cuda_tp_create(vox_cuda_model_t *m, int rank, int world);
/* all `world` ranks on THIS device, run kernel by kernel in lock step: checks the shard
 * arithmetic and the exchange protocol on one GPU (not a fast path) */
vox_cuda_tp_t *vox_cuda_tp_create_local(vox_cuda_model_t *m, int world);
/* kv_bits 32 = the two calls above; 16 = this rank's K/V rings in fp16 (the reference's GPU default
 * "fp16 KV", voxtral.c:189-190; same rounding rule and the same prompt-prefill limits as vox_cuda_state_create_ex).
 * No counterpart call in the reference (it has no multi-device code). */
vox_cuda_tp_t *vox_cuda_tp_create_ex(vox_cuda_model_t *m, int rank, int world, int kv_bits);
vox_cuda_tp_t *vox_cuda_tp_create_local_ex(vox_cuda_model_t *m, int world, int kv_bits);
int  vox_cuda_tp_export(vox_cuda_tp_t *tp, void *handle_out /* VOX_CUDA_TP_HANDLE_BYTES */);
int  vox_cuda_tp_connect(vox_cuda_tp_t *tp, const void *handles /* world x VOX_CUDA_TP_HANDLE_BYTES, rank order */);
vox_cuda_state_t *vox_cuda_tp_state(vox_cuda_tp_t *tp);
/* same contract as vox_cuda_decoder_run; ms_per_step (may be NULL) = CUDA-event time / n */
int  vox_cuda_tp_decode(vox_cuda_tp_t *tp, int prev_token, int adapter_index, int n, int *tokens_out,
                        float *ms_per_step);
/* vox_cuda_transcribe with the greedy steps tensor-parallel (encoder + prefill replicated per rank) */
int  vox_cuda_tp_transcribe(vox_cuda_tp_t *tp, const float *pcm, int n_samples, int n_frames,
                            int prompt_len, int *tokens_out, int max_tokens, float *ms_per_step);
void vox_cuda_tp_free(vox_cuda_tp_t *tp);

/* ---- an entry point that allocates nothing (csrc/noalloc.cu) ---------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The rule of vox_cuda_decoder_alternatives on a logits row the caller passes: the best (max_alt - 1) ids >= 1000
 * other than best_token, by logit (lowest id on ties), kept while 1 - p_i / p_best <= cutoff; ids_out gets up to
 * VOX_CUDA_MAX_ALT - 1 ids, *count_out their number.  logits, ids_out and count_out must be DEVICE-ACCESSIBLE
 * (device memory, or pinned host memory such as vox_cuda_host_alloc returns).  The kernel's scratch (per-slice
 * candidates + an arrival counter, 8448 bytes) is a file-scope static __device__ array that every GPU loading the
 * library holds until exit; it bounds the input: n_logits <= VOX_CUDA_NOALLOC_MAX_LOGITS, larger rows are rejected
 * (-1 + vox_cuda_last_error()).  The call performs no cudaMalloc / cudaFree / allocating library call: it enqueues
 * one 4-byte memset and one kernel on `stream` (a cudaStream_t, NULL = the default stream) and, if sync != 0, waits
 * for them; with sync == 0 the outputs are valid once the caller has synchronised the stream.  The first call on a
 * device resolves the array's address (cudaGetSymbolAddress) - make it outside a capture.  All calls share the
 * array: calls that overlap on the GPU (different streams without a dependency) race; ordered calls are fine. */
#define VOX_CUDA_NOALLOC_MAX_LOGITS 262144
int vox_cuda_alternatives_noalloc(const float *logits, int n_logits, int best_token, int max_alt, float cutoff,
                                  int *ids_out, int *count_out, void *stream, int sync);

/* ---- measurement helpers (bench.py / profiles) ------------------------ */
/* vox_cuda_decoder_run timed with CUDA events on the launch stream: returns
 * the average device time per step in milliseconds (<0 on error). */
float vox_cuda_bench_decoder_steps(vox_cuda_state_t *s, int prev_token,
                                   int adapter_index, int n, int *tokens_out);
/* CUDA-event time (ms) of the encoder side of a whole clip on the resident PCM (vox_cuda_pcm_upload):
 * log-mel + conv stem + 32 layers + adapter (stream_run_encoder, voxtral.c:826-951); <0 on error. */
float vox_cuda_bench_encoder(vox_cuda_state_t *s, int n_samples, int n_frames);
/* Number of kernel launches issued by this library since init. */
unsigned long long vox_cuda_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* VOX_CUDA_H */
