#!/usr/bin/env python3
"""Generate tests/golden/ref_golden.npz by running the UNMODIFIED reference
(oracle/_ref/libvoxref.so and libvoxref_strict.so, compiled from /root/reference by
oracle/build_ref.sh) on the seeded random-init checkpoint (tools/gen_model.c, seed 1234).

Run in the build container (needs oracle/_ref):   python tests/golden/make_golden.py
The committed .npz is what the CPU tests pin oracle/vox_oracle.py against and what the
GPU tests compare the CUDA path with when oracle/_ref is not on the box.

Inputs are either stored in the file (small) or regenerated from the seeds below.
Large outputs are stored subsampled (strides named in the keys) to keep the fixture small.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import refharness as rh  # noqa: E402

AUDIO_SEED, ENC_SEED, DEC_SEED = 0, 11, 5
MEL_ROWS = slice(240, 380)      # stream-mel rows kept (the 1 s of audio sits at frames 256..356)
N_STEPS = 3


def audio():
    return (0.1 * np.random.default_rng(AUDIO_SEED).standard_normal(16000)).astype(np.float32)


def enc_inputs():
    rng = np.random.default_rng(ENC_SEED)
    return [(0.5 * rng.standard_normal((n, 1280))).astype(np.float32) for n in (24, 16)]


def dec_adapter_rows():
    return (0.5 * np.random.default_rng(DEC_SEED).standard_normal((38 + N_STEPS, 3072))).astype(np.float32)


def main():
    ref = rh.Ref(rh.model_dir())
    out = {}
    x = audio()
    # mel: strict-IEEE build (see oracle/build_ref.sh for why) - streaming and batch
    out["mel_stream_rows240_380"] = rh.strict_stream_mel(x)[MEL_ROWS]
    out["mel_stream_frames"] = np.int32(rh.strict_stream_mel(x).shape[0])
    import ctypes as C
    Ls = C.CDLL(rh.REF_STRICT_SO)
    Ls.vox_mel_spectrogram.restype = rh.f32p
    Ls.vox_mel_spectrogram.argtypes = [rh.f32p, C.c_int, C.POINTER(C.c_int)]
    n = C.c_int()
    p = Ls.vox_mel_spectrogram(x.ctypes.data_as(rh.f32p), len(x), C.byref(n))
    mel_batch = np.ctypeslib.as_array(p, shape=(n.value, 128)).copy()
    out["mel_batch"] = mel_batch

    # encoder, one pass: first 64 batch-mel frames -> 32 positions
    enc_out = ref.encoder_forward(mel_batch[:64])
    out["enc_forward_cols8"] = enc_out[:, ::8]
    ad = ref.adapter(enc_out)
    out["adapter_of_enc_forward"] = ad
    # encoder, incremental: two chunks of conv-stem rows through the K/V cache
    ref.set("enc_kv_cache_len", 0)
    ref.set("enc_kv_pos_offset", 0)
    for i, xn in enumerate(enc_inputs()):
        out[f"enc_incremental{i}_cols8"] = ref.encoder_incremental(xn)[:, ::8]

    # decoder: prefill(38) + greedy steps, teacher-forced with the reference's own tokens
    rows = dec_adapter_rows()
    prompt = [1] + [32] * 37
    emb = np.stack([rows[i] + ref.tok_embedding(t) for i, t in enumerate(prompt)]).astype(np.float32)
    ref.set("kv_cache_len", 0)
    ref.set("kv_pos_offset", 0)
    ref.prefill(emb)
    prev, toks, lg = 32, [], []
    for i in range(N_STEPS):
        tok, logits = ref.decoder_forward((rows[38 + i] + ref.tok_embedding(prev)).astype(np.float32))
        toks.append(tok)
        lg.append(logits[::32].copy())
        prev = tok
    out["dec_tokens"] = np.array(toks, np.int32)
    out["dec_logits_cols32"] = np.stack(lg)
    out["ada_scale_cols8"] = np.ctypeslib.as_array(ref.L.refshim_ada_scale(ref.ctx), shape=(26, 3072))[:, ::8].copy()
    ref.close()
    path = os.path.join(HERE, "ref_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", {k: getattr(v, "shape", ()) for k, v in out.items()})


if __name__ == "__main__":
    main()
