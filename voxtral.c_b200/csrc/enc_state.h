// enc_state.h - device state of one stream's encoder side (owned by vox_cuda_state)
#pragma once
#include "vox_internal.h"

constexpr int ENC_CHUNK = 3328;    // max encoder positions per internal pass (26 row tiles: a 60 s clip = 3196 positions is ONE pass,
                                   // so its N = 1280 products are 125 tiles in one wave instead of 80 + 45 in two)
constexpr int ENC_MAXMEL = 6656;   // max mel frames per internal pass (= 2 * ENC_CHUNK)
constexpr int ATTN_KSPLITS = 14;   // key splits of the streaming attention: ceil((750 + 128) / 64) tiles, one per CTA
constexpr int ENC_RING = 4096;     // KV ring slots (>= ENC_CHUNK + ENC_WINDOW - 1 = 4077), power of two
static_assert(ENC_RING >= ENC_CHUNK + 750 - 1 && ENC_MAXMEL == 2 * ENC_CHUNK, "encoder pass limits");

struct EncState {
    // KV ring, slot = logical position & (ENC_RING-1).  Replaces the memmove
    // compaction of voxtral_encoder.c:431-449: the 750-key window never reaches
    // further back than ENC_RING - ENC_CHUNK slots.
    float *kv_k, *kv_v;        // precise mode: [n_enc][ENC_RING][2048] fp32
    size_t kv_bytes;
    int pos;                   // logical position of the next row (enc_kv_pos_offset + enc_kv_cache_len)
    int pos_min;               // logical position of the oldest valid key (a host-side position jump empties the cache)
    // incremental conv stem (voxtral.c:469-473)
    float *mel_in;             // [2 + ENC_MAXMEL][128]; rows 0-1 = last two mel frames
    float *c0;                 // [2 + ENC_MAXMEL][1280]; row 0 = last conv0 row fed to conv1,
                               // row 1 = parity residual (if c0_resid) else first new row
    int stem_init, c0_resid;
    // scratch for one pass
    float *x, *xn, *qkv, *att, *gu, *hh;
    float *encout;             // [4 + ENC_CHUNK][1280]; rows 0..enc_resid-1 = 4-alignment residual
    int enc_resid;
    float *mid;                // adapter hidden [ENC_CHUNK/4 + 1][3072]
    float *h_stage;            // pinned host staging
    void *fast;                // EncFast*, enc_fast.cu
    void *small;               // EncSmall*, enc_fast.cu (M <= 128 streaming path)
    float *attn_part;          // key-split attention partials [32 heads][ATTN_KSPLITS][66][128] (attn_tc.cu, M <= 128), lazily
    unsigned *attn_cnt;        // [32] arrival counters of the key splits (self-resetting)
    uint16_t *mel_hi, *mel_lo; // bf16 planes of mel_in (tensor-core conv stem)
    uint16_t *c0_hi, *c0_lo;   // bf16 planes of c0
};
