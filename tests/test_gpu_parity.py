"""GPU parity: every stage of the hot path through the C ABI (libvoxcuda.so)
against the reference BLAS path (oracle/_ref) on the same seeded inputs.

Tolerances (north_star): mel 1e-5 relative; encoder activations and decoder
logits 2e-2 max-abs; greedy token ids bit-exact.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def synth_audio(n, seed=1):
    """chirps + noise at -20 dBFS + silent gaps, quantised to s16 like a WAV."""
    rng = np.random.default_rng(seed)
    t = np.arange(n) / 16000.0
    x = np.zeros(n)
    for f0, f1 in ((100, 7000), (3000, 300), (500, 5000)):
        x += 0.1 * np.sin(2 * np.pi * (f0 * t + (f1 - f0) * t * t / (2 * max(t[-1], 1e-3))))
    x += 0.05 * rng.standard_normal(n)
    for g in range(4):
        a = int((0.15 + 0.2 * g) * n)
        x[a:a + min(8000, n // 20)] = 0
    return (np.clip(x, -1, 1) * 32767).astype(np.int16).astype(np.float32) / 32768.0


def relerr(a, b):
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-3)))


def test_mel_frames(full_env):
    e = full_env
    x = synth_audio(16000 * 4)
    ref_mel = e.ref.stream_mel(x)  # left pad + feed + flush + finish, like vox_stream_*
    # host framing exactly as vox_mel_ctx does: 200 + 32*1280 zeros, samples, right pad, 200 reflected
    n = len(x)
    right = (1280 - n % 1280) % 1280 + 17 * 1280
    buf = np.concatenate([np.zeros(200 + 32 * 1280, np.float32), x, np.zeros(right, np.float32)])
    refl = np.array([buf[len(buf) - 2 - i] for i in range(200)], np.float32)
    buf = np.concatenate([buf, refl])
    nf = (len(buf) - 400) // 160 + 1 - 1  # last frame dropped
    assert nf == ref_mel.shape[0]
    out = np.empty((nf, 128), np.float32)
    assert e.L.vox_cuda_mel_frames(e.vc.fptr(buf), nf, e.vc.fptr(out)) == 0, e.vc.last_error()
    import refharness as rh
    strict = rh.strict_stream_mel(x)
    d_strict = np.abs(out - strict)
    err = float(np.max(d_strict / np.maximum(np.abs(strict), 1.0)))
    print("mel vs strict-IEEE reference build: max abs", d_strict.max(), "rel", err)
    print("mel vs -ffast-math reference build: max abs", np.abs(out - ref_mel).max(),
          "(the two reference builds differ by", np.abs(strict - ref_mel).max(), ")")
    assert err < 1e-5
    assert np.abs(out - ref_mel).max() < 1e-3


def _embeds(e, adapter, tokens):
    return np.stack([adapter[i] + e.ref.tok_embedding(t) for i, t in enumerate(tokens)]).astype(np.float32)


def test_decoder_prefill_and_steps(full_env):
    e = full_env
    rng = np.random.default_rng(7)
    n_steps = 6
    adapter = (0.5 * rng.standard_normal((38 + n_steps, 3072))).astype(np.float32)
    prompt = [1] + [32] * 37
    emb = _embeds(e, adapter, prompt)
    # reference
    e.ref.set("kv_cache_len", 0); e.ref.set("kv_pos_offset", 0)
    e.ref.prefill(emb)
    # CUDA
    st = e.L.vox_cuda_state_create(e.model)
    assert st, e.vc.last_error()
    try:
        assert e.L.vox_cuda_decoder_prefill(st, e.vc.fptr(emb), 38) == 0, e.vc.last_error()
        assert e.L.vox_cuda_decoder_position(st) == 38
        prev_r = prev_c = 32
        logits_c = np.empty(131072, np.float32)
        for i in range(n_steps):
            x_r = (adapter[38 + i] + e.ref.tok_embedding(prev_r)).astype(np.float32)
            tok_r, logits_r = e.ref.decoder_forward(x_r)
            x_c = (adapter[38 + i] + e.ref.tok_embedding(prev_c)).astype(np.float32)
            tok_c = e.L.vox_cuda_decoder_step(st, e.vc.fptr(x_c), e.vc.fptr(logits_c))
            assert tok_c >= 0, e.vc.last_error()
            srt = np.sort(logits_r)
            print(f"step {i}: ref {tok_r} cuda {tok_c} max|dlogit| {np.abs(logits_c - logits_r).max():.3e} "
                  f"margin {srt[-1] - srt[-2]:.3e}")
            assert np.abs(logits_c - logits_r).max() < 2e-2
            assert tok_c == tok_r
            prev_r, prev_c = tok_r, tok_c
    finally:
        e.L.vox_cuda_state_free(st)


def test_decoder_prefill_two_blocks(full_env):
    """a 100-row prefill = two blocks (64 + 36 rows) of the skinny-GEMM path (csrc/dec_prefill.cu): the second block's
    causal attention reads the first block's K/V rows from the ring.  K/V rows of the first and last layer and the
    logits of the following step against vox_decoder_prefill / vox_decoder_forward of the reference."""
    e = full_env
    rng = np.random.default_rng(11)
    n = 100
    emb = (0.5 * rng.standard_normal((n, 3072))).astype(np.float32)
    e.ref.set("kv_cache_len", 0); e.ref.set("kv_pos_offset", 0)
    e.ref.prefill(emb)
    kref, vref = e.ref.kv_views()
    st = e.L.vox_cuda_state_create(e.model)
    assert st, e.vc.last_error()
    try:
        assert e.L.vox_cuda_decoder_prefill(st, e.vc.fptr(emb), n) == 0, e.vc.last_error()
        assert e.L.vox_cuda_decoder_position(st) == n
        k = np.empty((n, 1024), np.float32); v = np.empty((n, 1024), np.float32)
        for layer in (0, 25):
            assert e.L.vox_cuda_decoder_kv_read(st, layer, 0, n, e.vc.fptr(k), e.vc.fptr(v)) == 0
            dk, dv = np.abs(k - kref[layer, :n]).max(), np.abs(v - vref[layer, :n]).max()
            print(f"layer {layer}: 100-row prefill max|dK| {dk:.2e} max|dV| {dv:.2e}")
            assert dk < 2e-2 and dv < 2e-2
        x = (0.5 * rng.standard_normal(3072)).astype(np.float32)
        tok_r, logits_r = e.ref.decoder_forward(x)
        logits_c = np.empty(131072, np.float32)
        tok_c = e.L.vox_cuda_decoder_step(st, e.vc.fptr(x), e.vc.fptr(logits_c))
        d = float(np.abs(logits_c - logits_r).max())
        print(f"step after the 100-row prompt: ref {tok_r} cuda {tok_c} max|dlogit| {d:.3e}")
        assert d < 2e-2 and tok_c == tok_r
    finally:
        e.L.vox_cuda_state_free(st)


def test_encoder_stem_layers_adapter(full_env):
    e = full_env
    x = synth_audio(16000 * 2, seed=3)
    mel = e.ref.stream_mel(x)[:312]          # first streaming chunk size (voxtral.c:405)
    ref_enc = e.ref.encoder_forward(mel)     # conv stem + 32 layers + final norm
    ref_ad = e.ref.adapter(ref_enc)
    st = e.L.vox_cuda_state_create(e.model)
    assert st, e.vc.last_error()
    try:
        stem = np.empty((mel.shape[0] // 2 + 2, 1280), np.float32)
        n = C.c_int()
        assert e.L.vox_cuda_conv_stem(st, e.vc.fptr(mel), mel.shape[0], e.vc.fptr(stem), C.byref(n)) == 0
        assert n.value == ref_enc.shape[0] == 156
        xs = stem[:n.value].copy()
        assert e.L.vox_cuda_encoder_step(st, e.vc.fptr(xs), n.value) == 0, e.vc.last_error()
        d = np.abs(xs - ref_enc).max()
        print("encoder max|d|", d, "ref rms", np.sqrt((ref_enc ** 2).mean()))
        assert d < 2e-2
        ad = np.empty((n.value // 4, 3072), np.float32)
        assert e.L.vox_cuda_adapter(st, e.vc.fptr(xs), n.value, e.vc.fptr(ad)) == 0
        d2 = np.abs(ad - ref_ad).max()
        print("adapter max|d|", d2, "ref rms", np.sqrt((ref_ad ** 2).mean()))
        assert d2 < 2e-2
        # fused device path must agree with the staged one
        e.L.vox_cuda_state_reset_encoder(st)
        assert e.L.vox_cuda_encode_chunk(st, e.vc.fptr(mel), mel.shape[0]) == 39
        ad2 = np.empty((39, 3072), np.float32)
        assert e.L.vox_cuda_adapter_read(st, 0, 39, e.vc.fptr(ad2)) == 39
        assert np.abs(ad2 - ad).max() < 1e-5
    finally:
        e.L.vox_cuda_state_free(st)
