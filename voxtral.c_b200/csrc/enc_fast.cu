// enc_fast.cu - throughput mode of the encoder (tcgen05 GEMMs).  Until the
// tensor-core kernels land this forwards to the precise path.
#include "vox_internal.h"
#include "enc_state.h"

int enc_layers_precise(vox_cuda_state *s, float *x, int M, float *out);

int enc_layers_fast(vox_cuda_state *s, float *x, int M, float *out) {
    return enc_layers_precise(s, x, M, out);
}
