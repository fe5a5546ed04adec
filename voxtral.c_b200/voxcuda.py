"""ctypes binding of include/vox_cuda.h (libvoxcuda.so).

Plumbing only: every call goes straight to the C ABI.  Importing this module
never touches the test oracle; it raises if the CUDA library is missing (there is no
CPU fallback in the product path).
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VOX_CUDA_LIB") or os.path.join(HERE, "libvoxcuda.so")

ENC_LAYERS, DEC_LAYERS = 32, 26
ENC_DIM, DEC_DIM, VOCAB, MEL_BINS = 1280, 3072, 131072, 128
TP_HANDLE_BYTES = 80

f32p = C.POINTER(C.c_float)
u16p = C.POINTER(C.c_uint16)
i32p = C.POINTER(C.c_int)


class EncLayerW(C.Structure):
    _fields_ = [(n, u16p) for n in ("wq", "wk", "wv", "wo", "w1", "w2", "w3")] + \
               [(n, f32p) for n in ("wq_bias", "wv_bias", "wo_bias", "w2_bias", "attention_norm", "ffn_norm")]


class DecLayerW(C.Structure):
    _fields_ = [(n, u16p) for n in ("wq", "wk", "wv", "wo", "w1", "w2", "w3")] + \
               [(n, f32p) for n in ("attention_norm", "ffn_norm")]


class Weights(C.Structure):
    _fields_ = [("n_enc_layers", C.c_int), ("n_dec_layers", C.c_int),
                ("conv0_weight", f32p), ("conv0_bias", f32p),
                ("conv1_weight", f32p), ("conv1_bias", f32p),
                ("enc", EncLayerW * ENC_LAYERS), ("enc_norm", f32p),
                ("adapter_w0", u16p), ("adapter_w1", u16p), ("tok_embeddings", u16p),
                ("dec", DecLayerW * DEC_LAYERS), ("dec_norm", f32p), ("ada_scale", f32p)]


i8p = C.POINTER(C.c_int8)


class Q8(C.Structure):                      # vox_cuda_q8_t
    _fields_ = [("q", i8p), ("scale", f32p)]


class LayerQ8(C.Structure):                 # vox_cuda_layer_q8_t
    _fields_ = [(n, Q8) for n in ("wq", "wk", "wv", "wo", "w1", "w2", "w3")]


class WeightsQ8(C.Structure):               # vox_cuda_weights_q8_t
    _fields_ = [("enc", LayerQ8 * ENC_LAYERS), ("adapter_w0", Q8), ("adapter_w1", Q8), ("tok_embeddings", Q8),
                ("dec", LayerQ8 * DEC_LAYERS)]


_lib = None

# name -> (restype, argtypes); also the list the CPU test checks against the header
SIGNATURES = {
    "vox_cuda_init": (C.c_int, []),
    "vox_cuda_available": (C.c_int, []),
    "vox_cuda_shutdown": (None, []),
    "vox_cuda_last_error": (C.c_char_p, []),
    "vox_cuda_memory_used": (C.c_size_t, []),
    "vox_cuda_launch_count": (C.c_ulonglong, []),
    "vox_cuda_bench_encoder": (C.c_float, [C.c_void_p, C.c_int, C.c_int]),
    "vox_cuda_model_create": (C.c_void_p, [C.POINTER(Weights)]),
    "vox_cuda_model_create_q8": (C.c_void_p, [C.POINTER(Weights), C.POINTER(WeightsQ8)]),
    "vox_cuda_model_is_q8": (C.c_int, [C.c_void_p]),
    "vox_cuda_model_set_ada_scale": (C.c_int, [C.c_void_p, f32p]),
    "vox_cuda_model_free": (None, [C.c_void_p]),
    "vox_cuda_state_create": (C.c_void_p, [C.c_void_p]),
    "vox_cuda_state_create_ex": (C.c_void_p, [C.c_void_p, C.c_int]),
    "vox_cuda_state_kv_bits": (C.c_int, [C.c_void_p]),
    "vox_cuda_state_free": (None, [C.c_void_p]),
    "vox_cuda_state_reset_decoder": (None, [C.c_void_p]),
    "vox_cuda_state_reset_encoder": (None, [C.c_void_p]),
    "vox_cuda_encoder_position": (C.c_int, [C.c_void_p]),
    "vox_cuda_encoder_set_position": (C.c_int, [C.c_void_p, C.c_int]),
    "vox_cuda_state_reset_stem": (None, [C.c_void_p]),
    "vox_cuda_decoder_position": (C.c_int, [C.c_void_p]),
    "vox_cuda_decoder_set_position": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "vox_cuda_decoder_kv_write": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, f32p, f32p]),
    "vox_cuda_decoder_kv_read": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, f32p, f32p]),
    "vox_cuda_decoder_prefill": (C.c_int, [C.c_void_p, f32p, C.c_int]),
    "vox_cuda_decoder_step": (C.c_int, [C.c_void_p, f32p, f32p]),
    "vox_cuda_decoder_step_fused": (C.c_int, [C.c_void_p, C.c_int, C.c_int, f32p]),
    "vox_cuda_decoder_run": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, i32p]),
    "vox_cuda_decoder_read_logits": (C.c_int, [C.c_void_p, f32p]),
    "vox_cuda_decoder_prefill_fused": (C.c_int, [C.c_void_p, i32p, C.c_int, C.c_int]),
    "vox_cuda_decoder_alternatives": (C.c_int, [C.c_void_p, C.c_int, C.c_float, i32p]),
    "vox_cuda_alternatives_noalloc": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p,
                                              C.c_void_p, C.c_int]),
    "vox_cuda_decoder_set_alt": (C.c_int, [C.c_void_p, C.c_int, C.c_float]),
    "vox_cuda_decoder_get_alts": (C.c_int, [C.c_void_p, C.c_int, i32p]),
    "vox_cuda_gemm_bf16_streams": (C.c_int, [C.c_int, C.c_int, C.c_int, f32p, u16p, C.c_int, C.c_int, f32p]),
    "vox_cuda_bench_gemm_streams": (C.c_float, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "vox_cuda_group_create": (C.c_void_p, [C.c_void_p, C.c_int, C.c_int]),
    "vox_cuda_group_create_ex": (C.c_void_p, [C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "vox_cuda_group_free": (None, [C.c_void_p]),
    "vox_cuda_group_streams": (C.c_int, [C.c_void_p]),
    "vox_cuda_group_encode": (C.c_int, [C.c_void_p, C.c_int, f32p, C.c_int, C.c_int]),
    "vox_cuda_group_set_adapter": (C.c_int, [C.c_void_p, C.c_int, f32p, C.c_int]),
    "vox_cuda_group_decode": (C.c_int, [C.c_void_p, C.c_int, i32p, C.c_int, i32p, f32p]),
    "vox_cuda_encoder_step": (C.c_int, [C.c_void_p, f32p, C.c_int]),
    "vox_cuda_conv_stem": (C.c_int, [C.c_void_p, f32p, C.c_int, f32p, i32p]),
    "vox_cuda_adapter": (C.c_int, [C.c_void_p, f32p, C.c_int, f32p]),
    "vox_cuda_encode_chunk": (C.c_int, [C.c_void_p, f32p, C.c_int]),
    "vox_cuda_adapter_count": (C.c_int, [C.c_void_p]),
    "vox_cuda_adapter_read": (C.c_int, [C.c_void_p, C.c_int, C.c_int, f32p]),
    "vox_cuda_adapter_append": (C.c_int, [C.c_void_p, f32p, C.c_int]),
    "vox_cuda_adapter_discard_before": (C.c_int, [C.c_void_p, C.c_int]),
    "vox_cuda_mel_frames": (C.c_int, [f32p, C.c_int, f32p]),
    "vox_cuda_sgemm_bf16": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, f32p, u16p, f32p]),
    "vox_cuda_host_alloc": (C.c_void_p, [C.c_size_t]),
    "vox_cuda_host_free": (None, [C.c_void_p]),
    "vox_cuda_pcm_upload": (C.c_int, [C.c_void_p, f32p, C.c_int]),
    "vox_cuda_transcribe": (C.c_int, [C.c_void_p, f32p, C.c_int, C.c_int, C.c_int, i32p, C.c_int, C.POINTER(C.c_float)]),
    "vox_cuda_gemm_bf16_tc": (C.c_int, [C.c_int, C.c_int, C.c_int, f32p, u16p, f32p, C.c_int, f32p]),
    "vox_cuda_bench_decoder_steps": (C.c_float, [C.c_void_p, C.c_int, C.c_int, C.c_int, i32p]),
    "vox_cuda_pcm_resampled_len": (C.c_int, [C.c_int, C.c_int]),
    "vox_cuda_pcm_from_s16": (C.c_int, [C.POINTER(C.c_int16), C.c_int, C.c_int, C.c_int, f32p, C.c_int]),
    "vox_cuda_window_rms": (C.c_int, [f32p, C.c_int, C.c_int, f32p]),
    "vox_cuda_tp_create": (C.c_void_p, [C.c_void_p, C.c_int, C.c_int]),
    "vox_cuda_tp_create_local": (C.c_void_p, [C.c_void_p, C.c_int]),
    "vox_cuda_tp_create_ex": (C.c_void_p, [C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "vox_cuda_tp_create_local_ex": (C.c_void_p, [C.c_void_p, C.c_int, C.c_int]),
    "vox_cuda_tp_export": (C.c_int, [C.c_void_p, C.c_void_p]),
    "vox_cuda_tp_connect": (C.c_int, [C.c_void_p, C.c_void_p]),
    "vox_cuda_tp_state": (C.c_void_p, [C.c_void_p]),
    "vox_cuda_tp_decode": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, i32p, C.POINTER(C.c_float)]),
    "vox_cuda_tp_transcribe": (C.c_int, [C.c_void_p, f32p, C.c_int, C.c_int, C.c_int, i32p, C.c_int, C.POINTER(C.c_float)]),
    "vox_cuda_tp_free": (None, [C.c_void_p]),
}


def lib():
    """Load libvoxcuda.so (built by `make cuda` / __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} missing: run `make -C voxtral.c_b200 cuda` "
                               "(the product has no CPU fallback)")
        _lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(_lib, name)
            fn.restype, fn.argtypes = res, args
    return _lib


def fptr(a):
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(f32p)


def iptr(a):
    assert a.dtype == np.int32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(i32p)


def last_error():
    return lib().vox_cuda_last_error().decode()


def check(rc, what):
    if rc is None or (isinstance(rc, int) and rc < 0):
        raise RuntimeError(f"{what} failed: {last_error()}")
    return rc
