"""Test-only harness around oracle/_ref/libvoxref.so (the UNMODIFIED reference
BLAS path compiled by oracle/build_ref.sh) - the checker, never the product.

Nothing here reads /root/reference at run time: the shared object is prebuilt
and travels with the repo snapshot.
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libvoxref.so")
REF_STRICT_SO = os.path.join(ROOT, "oracle", "_ref", "libvoxref_strict.so")
sys.path.insert(0, os.path.join(ROOT, "voxtral.c_b200"))

f32p = C.POINTER(C.c_float)


def have_ref():
    return os.path.exists(REF_SO)


def ensure_gen_model():
    exe = os.path.join(ROOT, "build", "gen_model")
    src = os.path.join(ROOT, "tools", "gen_model.c")
    if not os.path.exists(exe) or os.path.getmtime(exe) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(exe), exist_ok=True)
        subprocess.check_call(["gcc", "-O3", "-fopenmp", "-w", "-o", exe, src, "-lm"])
    return exe


def model_dir(seed=1234, enc_layers=32, dec_layers=26, emb_sigma=0.02, extra=()):
    """Generate (once per box) the deterministic random-init checkpoint."""
    base = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else "/tmp"
    tag = f"voxm2_s{seed}_e{enc_layers}_d{dec_layers}_g{emb_sigma}" + "".join(extra).replace("-", "").replace(" ", "")
    d = os.path.join(base, tag)
    if not os.path.exists(os.path.join(d, "tekken.json")):
        os.makedirs(d, exist_ok=True)
        subprocess.check_call([ensure_gen_model(), d, "--seed", str(seed), "--enc-layers", str(enc_layers),
                               "--dec-layers", str(dec_layers), "--emb-sigma", str(emb_sigma), *extra])
    return d


class Ref:
    """The reference engine through its own exported C functions
    (/root/reference/voxtral.h:252-363, voxtral_audio.h:34-69)."""

    def __init__(self, mdir, verbose=0):
        self.L = L = C.CDLL(REF_SO)
        L.vox_load.restype = C.c_void_p
        L.vox_load.argtypes = [C.c_char_p]
        L.vox_free.argtypes = [C.c_void_p]
        L.vox_mel_spectrogram.restype = f32p
        L.vox_mel_spectrogram.argtypes = [f32p, C.c_int, C.POINTER(C.c_int)]
        L.refshim_stream_mel.restype = f32p
        L.refshim_stream_mel.argtypes = [f32p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)]
        L.vox_encoder_forward.restype = f32p
        L.vox_encoder_forward.argtypes = [C.c_void_p, f32p, C.c_int, C.POINTER(C.c_int)]
        L.vox_encoder_forward_incremental.restype = f32p
        L.vox_encoder_forward_incremental.argtypes = [C.c_void_p, f32p, C.c_int, C.POINTER(C.c_int)]
        L.vox_adapter_forward.restype = f32p
        L.vox_adapter_forward.argtypes = [C.c_void_p, f32p, C.c_int, C.POINTER(C.c_int)]
        L.vox_decoder_prefill.argtypes = [C.c_void_p, f32p, C.c_int]
        L.vox_decoder_forward.restype = C.c_int
        L.vox_decoder_forward.argtypes = [C.c_void_p, f32p, f32p]
        L.refshim_get_int.argtypes = [C.c_void_p, C.c_char_p]
        L.refshim_set_int.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        L.refshim_fill_weights.argtypes = [C.c_void_p, C.c_void_p]
        L.refshim_free.argtypes = [C.c_void_p]
        L.refshim_set_verbose.argtypes = [C.c_int]
        L.refshim_tok_embeddings.restype = C.POINTER(C.c_uint16)
        L.refshim_tok_embeddings.argtypes = [C.c_void_p]
        L.refshim_ada_scale.restype = f32p
        L.refshim_ada_scale.argtypes = [C.c_void_p]
        L.vox_set_delay.argtypes = [C.c_void_p, C.c_int]
        L.refshim_set_verbose(verbose)
        self.ctx = L.vox_load(mdir.encode())
        if not self.ctx:
            raise RuntimeError(f"reference vox_load({mdir}) failed")

    def close(self):
        if self.ctx:
            self.L.vox_free(self.ctx)
            self.ctx = None

    def _take(self, p, shape):
        a = np.ctypeslib.as_array(p, shape=shape).copy()
        self.L.refshim_free(p)
        return a

    def get(self, field):
        return self.L.refshim_get_int(self.ctx, field.encode())

    def set(self, field, v):
        assert self.L.refshim_set_int(self.ctx, field.encode(), int(v)) == 0

    def mel_spectrogram(self, samples):
        x = np.ascontiguousarray(samples, dtype=np.float32)
        n = C.c_int()
        p = self.L.vox_mel_spectrogram(x.ctypes.data_as(f32p), len(x), C.byref(n))
        return self._take(p, (n.value, 128))

    def stream_mel(self, samples, delay_tokens=6, chunk=0):
        x = np.ascontiguousarray(samples, dtype=np.float32)
        n = C.c_int()
        p = self.L.refshim_stream_mel(x.ctypes.data_as(f32p), len(x), delay_tokens, chunk, C.byref(n))
        return self._take(p, (n.value, 128))

    def encoder_forward(self, mel):
        mel = np.ascontiguousarray(mel, dtype=np.float32)
        n = C.c_int()
        p = self.L.vox_encoder_forward(self.ctx, mel.ctypes.data_as(f32p), mel.shape[0], C.byref(n))
        return self._take(p, (n.value, 1280))

    def encoder_incremental(self, x_new):
        x = np.ascontiguousarray(x_new, dtype=np.float32)
        n = C.c_int()
        p = self.L.vox_encoder_forward_incremental(self.ctx, x.ctypes.data_as(f32p), x.shape[0], C.byref(n))
        return self._take(p, (n.value, 1280))

    def adapter(self, enc_out):
        x = np.ascontiguousarray(enc_out, dtype=np.float32)
        n = C.c_int()
        p = self.L.vox_adapter_forward(self.ctx, x.ctypes.data_as(f32p), x.shape[0], C.byref(n))
        return self._take(p, (n.value, 3072))

    def prefill(self, embeds):
        x = np.ascontiguousarray(embeds, dtype=np.float32)
        self.L.vox_decoder_prefill(self.ctx, x.ctypes.data_as(f32p), x.shape[0])

    def decoder_forward(self, embed):
        x = np.ascontiguousarray(embed, dtype=np.float32)
        logits = np.empty(131072, dtype=np.float32)
        tok = self.L.vox_decoder_forward(self.ctx, x.ctypes.data_as(f32p), logits.ctypes.data_as(f32p))
        return tok, logits

    def fill_weights_q8(self, q):
        """binds the reference's Q8 tables; returns ctx->use_q8"""
        return self.L.refshim_fill_weights_q8(self.ctx, C.byref(q))

    def tok_embedding(self, token):
        p = self.L.refshim_tok_embeddings(self.ctx)
        if not p:                             # Q8 checkpoint: row = q * scale (tok_embed_q8_to_f32, voxtral.c:443-452)
            self.L.refshim_tok_embeddings_q8.restype = C.POINTER(C.c_int8)
            self.L.refshim_tok_embeddings_q8.argtypes = [C.c_void_p]
            self.L.refshim_tok_embeddings_scale_q8.restype = f32p
            self.L.refshim_tok_embeddings_scale_q8.argtypes = [C.c_void_p]
            q = np.ctypeslib.as_array(self.L.refshim_tok_embeddings_q8(self.ctx), shape=(131072, 3072))[token]
            sc = np.ctypeslib.as_array(self.L.refshim_tok_embeddings_scale_q8(self.ctx), shape=(131072,))[token]
            return q.astype(np.float32) * np.float32(sc)
        row = np.ctypeslib.as_array(p, shape=(131072, 3072))[token]
        return (row.astype(np.uint32) << 16).view(np.float32)

    def kv_views(self):
        """numpy views [26, kv_cache_max, 1024] of the reference's fp32 K and V caches
        (voxtral.h:197-205; allocated by the first prefill, voxtral_decoder.c:457-462)."""
        kmax = self.get("kv_cache_max")
        self.L.refshim_kv_k.restype = f32p
        self.L.refshim_kv_v.restype = f32p
        self.L.refshim_kv_k.argtypes = [C.c_void_p]
        self.L.refshim_kv_v.argtypes = [C.c_void_p]
        k = np.ctypeslib.as_array(self.L.refshim_kv_k(self.ctx), shape=(26, kmax, 1024))
        v = np.ctypeslib.as_array(self.L.refshim_kv_v(self.ctx), shape=(26, kmax, 1024))
        return k, v

    def fill_weights(self, w):
        self.L.refshim_fill_weights(self.ctx, C.byref(w))


def strict_stream_mel(samples, delay_tokens=6, chunk=0):
    """Streaming mel of the strict-IEEE reference build (oracle/build_ref.sh):
    the pin for the 1e-5 mel tolerance (the -ffast-math build moves DFT table
    entries by ~1e-4 and its mel by ~2e-4)."""
    L = C.CDLL(REF_STRICT_SO)
    L.refshim_stream_mel.restype = f32p
    L.refshim_stream_mel.argtypes = [f32p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)]
    L.refshim_free.argtypes = [C.c_void_p]
    x = np.ascontiguousarray(samples, dtype=np.float32)
    n = C.c_int()
    p = L.refshim_stream_mel(x.ctypes.data_as(f32p), len(x), delay_tokens, chunk, C.byref(n))
    a = np.ctypeslib.as_array(p, shape=(n.value, 128)).copy()
    L.refshim_free(p)
    return a


def make_wav(frames_s16, channels, sample_rate):
    """A minimal RIFF/WAVE file image (16-bit PCM) around interleaved s16 frames."""
    import struct
    data = np.asarray(frames_s16, np.int16).tobytes()
    hdr = b"RIFF" + struct.pack("<I", 36 + len(data)) + b"WAVE" + b"fmt " + struct.pack(
        "<IHHIIHH", 16, 1, channels, sample_rate, sample_rate * channels * 2, channels * 2, 16)
    return hdr + b"data" + struct.pack("<I", len(data)) + data


def ref_parse_wav(wav_bytes, strict=True):
    """vox_parse_wav_buffer of the compiled reference (voxtral_audio.c:49-141) -> f32 samples."""
    L = C.CDLL(REF_STRICT_SO if strict else REF_SO)
    L.vox_parse_wav_buffer.restype = f32p
    L.vox_parse_wav_buffer.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(C.c_int)]
    n = C.c_int(0)
    p = L.vox_parse_wav_buffer(wav_bytes, len(wav_bytes), C.byref(n))
    assert p, "reference rejected the WAV image"
    out = np.ctypeslib.as_array(p, shape=(n.value,)).copy()
    C.CDLL(None).free(C.cast(p, C.c_void_p))
    return out
