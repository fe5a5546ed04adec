// vox_internal.h - host-side structures shared by the .cu files of the backend.
#pragma once
#include "vox_common.cuh"

// Device-resident scalars of one decoder stream.  Kernels read them from HBM so
// that a captured CUDA graph can be replayed for every token without the host
// patching parameters; the last kernel of a step advances them.
struct DecStepParams {
    int pos;            // logical position of the token being processed (kv_pos_offset + kv_cache_len)
    int n_valid;        // KV entries valid BEFORE this step (<= DEC_WINDOW-1 after clamping)
    int prev_token;     // token whose embedding is added (voxtral.c:1106-1113)
    int adapter_index;  // physical row in the device adapter buffer
    int use_host_embed; // 1: x comes from d_embed_in, 0: built on device
    int out_token;      // greedy token of the last step
    int step_count;     // steps executed since the counter was last reset
    int pad;
};

// --alt: best logit and the three best OTHER text tokens (ids >= 1000, by logit, lowest id on ties) of one step
struct AltRecord { float best; float v[3]; int id[3]; int pad; };

struct DecLayerDev {
    const uint16_t *wqkv;   // [6144,3072] rows: wq(4096) | wk(1024) | wv(1024)
    const uint16_t *wo;     // [3072,4096]
    const uint16_t *w13;    // [18432,3072] row 2i = w1[i], row 2i+1 = w3[i]
    const uint16_t *w2;     // [3072,9216]
    const float *attn_norm; // [3072]
    const float *ffn_norm;  // [3072]
    // Q8 model (vox_cuda_model_create_q8): the bf16 matrices above hold the int8 values exactly (|q| <= 128) for the
    // M > 1 tensor-core products, s_* are the per-row scales in the merged row order (applied in the epilogue),
    // q_* the int8 rows themselves for the single-token kernel, qs_* = sum_k q[row][k] (offset-binary correction,
    // dec_mega.cu).  All null for a bf16 model.
    const float *s_wqkv, *s_wo, *s_w13, *s_w2;
    const int8_t *q_wqkv, *q_wo, *q_w13, *q_w2;
    const int *qs_wqkv, *qs_wo, *qs_w13, *qs_w2;
};

struct EncLayerDev {
    const uint16_t *wqkv;   // [6144,1280] rows: wq | wk | wv
    const uint16_t *wo;     // [1280,2048]
    const uint16_t *w13;    // [10240,1280] row 2i = w1[i], row 2i+1 = w3[i]
    const uint16_t *w2;     // [1280,5120]
    const float *bqkv;      // [6144] (wk part zero)
    const float *bo;        // [1280]
    const float *b2;        // [1280]
    const float *attn_norm; // [1280]
    const float *ffn_norm;  // [1280]
    const float *s_wqkv, *s_wo, *s_w13, *s_w2;   // Q8 model: per-row scales (merged row order), else null
};

struct vox_cuda_model {
    int n_enc, n_dec;
    // one arena, laid out in execution order (decoder first: its walk per token
    // is then a single linear pass through HBM)
    uint8_t *arena;
    size_t arena_bytes;
    DecLayerDev dec[VOX_CUDA_DEC_LAYERS];
    EncLayerDev enc[VOX_CUDA_ENC_LAYERS];
    const uint16_t *tok_emb;   // [131072,3072] (Q8 model: the int8 values as bf16, to be scaled by s_tok_emb)
    int q8;                    // 1: created by vox_cuda_model_create_q8
    const int8_t *q_tok_emb;   // [131072,3072] int8
    const float *s_tok_emb;    // [131072]
    const int *qs_tok_emb;     // [131072] row sums of q
    const float *s_adapter0, *s_adapter1;   // [3072] each
    const float *dec_norm;     // [3072]
    float *ada_scale;          // [n_dec,3072]
    const float *enc_norm;     // [1280]
    const uint16_t *conv0_w;   // [1280, 3*128]  tap-major: col = tap*128 + ic  (bf16, exact: source is bf16)
    const uint16_t *conv1_w;   // [1280, 3*1280] tap-major
    const float *conv0_b, *conv1_b;
    const uint16_t *adapter_w0; // [3072,5120]
    const uint16_t *adapter_w1; // [3072,3072]
    float *inv_freq_dec;       // [64]  1/theta^(2d/128)
    float *inv_freq_enc;       // [32]  1/theta^(2d/64)
    // host pointer -> device pointer registry for vox_cuda_sgemm_bf16
    struct Reg { const void *host; const void *dev; size_t bytes; } reg[512];
    int n_reg;
};

struct vox_cuda_state {
    vox_cuda_model *m;
    cudaStream_t stream;
    // ---- decoder ----
    float *kv_k, *kv_v;          // [n_dec][DEC_WINDOW][1024] fp32 ring, slot = logical & 8191 (kv16 == 0)
    // kv16 == 1 (vox_cuda_state_create_ex(m, 16)): the rings are fp16 (void pointers: this header stays free of
    // cuda_fp16.h), kv_k / kv_v shrink to [n_dec][PF_KV_ROWS][1024] fp32 rows that the prompt prefill writes before
    // they are rounded into the ring (the arrangement of the stream groups, api_group.cu)
    void *kv_k16, *kv_v16;
    int kv16;
    DecStepParams *d_params;     // device
    DecStepParams h_params;      // host mirror (authoritative between calls)
    float *d_x, *d_q, *d_attn, *d_h, *d_logits, *d_embed_in, *d_rope;
    float *d_attn_part;          // [NSPLIT][32][DEC_HD+2]
    float *d_blockmax;           // per-CTA (val, idx) pairs of the logits kernel
    int *d_tokens;               // ring of produced tokens (device), [TOKENS_RING]
    int *h_tokens;               // pinned host copy
    float *h_pinned;             // pinned staging (embed / logits)
    // --alt (vox_cuda_decoder_set_alt): per-step records written by the decoder kernel, ring indexed like d_tokens
    AltRecord *d_alt, *h_alt;    // device ring [TOKENS_RING] + pinned host mirror (allocated on first use)
    float2 *d_alt_cand;          // [grid][4] per-CTA candidates
    int alt_n;                   // 1 = off
    float alt_cutoff;
    cudaGraphExec_t step_graph;
    int graph_ready;
    void *mega;                  // MegaHost*, dec_mega.cu (persistent cooperative step kernel)
    int use_graph_steps;         // 1: legacy kernel-per-op CUDA graph (VOX_CUDA_DEC_GRAPH=1), 0: megakernel
    // prefill scratch (allocated on demand)
    float *pf_buf; size_t pf_bytes;
    void *pf_wt;                 // PfWt*, dec_prefill.cu (skinny-GEMM prefill: prepared products + row buffers, on first use)
    // ---- encoder ----
    void *enc_priv;              // EncState*, see enc_host.cu
    void *batch;                 // BatchBuf*, api_batch.cu (resident PCM / mel of the whole-clip entry point)
    // ---- adapter tokens (device) ----
    float *adapter_buf;          // [adapter_cap][3072]
    int adapter_cap, adapter_count, adapter_offset; // count = total logical, offset = first kept
    // ---- timing ----
    cudaEvent_t ev0, ev1;
};

constexpr int PF_KV_ROWS = 128;   // prompt rows an fp16-ring state can prefill (the prompt has 38)
constexpr int ATTN_NSPLIT = 16;
constexpr int TOKENS_RING = 65536;
constexpr int LOGITS_ROWS_PER_CTA = 128;

// dec_mega.cu
int dec_mega_init(vox_cuda_state *s);
void dec_mega_free(vox_cuda_state *s);
int dec_mega_launch(vox_cuda_state *s, int n_steps);
int dec_mega_attn_part_floats(void);

// api_dec.cu: prompt prefill into an arbitrary K/V ring (used by the state itself and by stream groups)
int dec_prefill_into(vox_cuda_state *s, float *x_dev, int seq_len, float *kbase, float *vbase, size_t kv_layer,
                     int mask, int *pos, int *n_valid);
// dec_prefill.cu: one block of <= 64 prompt rows on the skinny swap-AB GEMM (x = residual stream, in place)
int dec_prefill_block_wt(vox_cuda_state *s, float *x, int M, float *kbase, float *vbase, size_t kv_layer, int mask, int pos0,
                         int first_valid);
void dec_prefill_wt_free(vox_cuda_state *s);
// api_batch.cu: PCM -> mel -> conv stem -> encoder -> adapter rows in s->adapter_buf (resets the state)
int batch_encode_clip(vox_cuda_state *s, const float *pcm, int n_samples, int n_frames);

// dec_kernels.cu
int dec_launch_step(vox_cuda_state *s, cudaStream_t st);
int dec_prefill(vox_cuda_state *s, const float *d_x_in, int seq_len);
void dec_launch_embed(cudaStream_t st, const DecStepParams *p, const float *embed_in, const float *adapter,
                      const uint16_t *tok_emb, const float *tok_scale, const float *inv_freq, float *x, float *rope);
void dec_launch_attn(cudaStream_t st, const DecStepParams *p, const float *q, const float *kc, const float *vc,
                     float *part, float *attn, int kvh0, int n_kvh);

// sgemm.cu : C[M,N] = A[M,K] fp32 (lda) x W[N,K]^T bf16, fp32 accumulate, CUDA cores
enum { EPI_NONE = 0, EPI_BIAS = 1, EPI_GELU = 2, EPI_RESIDUAL = 4 };
int sgemm_bf16w(cudaStream_t st, int M, int N, int K, const float *A, long lda,
                const uint16_t *W, float *C, long ldc, const float *bias, int epi);
