// group.h - host-side state of a stream group (api_group.cu, group_mega.cu)
#pragma once
#include <cuda_fp16.h>
#include "vox_internal.h"
#include "gemm_wt.h"
#include "group_dev.cuh"

struct vox_cuda_group {
    vox_cuda_model *m;
    vox_cuda_state *eng;         // encoder engine + prefill scratch; its stream orders all group work
    int B, NB, max_pos, cap, mask;
    float *kv_k, *kv_v;          // [B][n_dec][cap][1024] fp32 rings (kv16 == 0)
    __half *kv_k16, *kv_v16;     // the same rings as fp16 (kv16 == 1: vox_cuda_group_create_ex(..., 16))
    float *pf_k, *pf_v;          // kv16: fp32 rows [n_dec][128][1024] the prompt prefill writes before they are rounded into the ring
    int kv16;
    size_t kv_layer, kv_stream;  // elements
    float *adapter;              // [B][max_pos][3072]
    int count[grp::MAX_STREAMS];      // adapter rows per stream (0 = not encoded)
    grp::Slot *d_slots;
    float *x;                    // [B][3072] residual rows
    uint16_t *xn_b, *att_b, *hh_b;   // split bf16 rows [NB][2*3072], [NB][2*4096], [NB][2*9216]
    float *part;                 // [KS_MAX][B][6144] K-range partials of the current product
    float *logits;               // [B][131072]
    int *d_tokens; size_t tok_cap;   // [steps][B]
    float *pf_x; int *pf_tok;    // prefill rows / prompt ids
    WtOp op_qkv[VOX_CUDA_DEC_LAYERS], op_wo[VOX_CUDA_DEC_LAYERS], op_w13[VOX_CUDA_DEC_LAYERS], op_w2[VOX_CUDA_DEC_LAYERS], op_logits;
    cudaEvent_t ev0, ev1;
    void *mega;                  // GroupMega*, group_mega.cu (persistent step kernel), or null
};


// group_mega.cu: all n_steps lock steps in one persistent cooperative kernel.  group_mega_init returns 0 when the
// kernel is usable for this group; group_mega_run enqueues on the group's stream (tokens land in g->d_tokens).
int group_mega_init(vox_cuda_group *g);
void group_mega_free(vox_cuda_group *g);
int group_mega_run(vox_cuda_group *g, int n_steps);
const unsigned *group_mega_error_record(vox_cuda_group *g);
void group_mega_dump_prof(vox_cuda_group *g, int n_steps);
