"""GPU parity against (a) the committed golden outputs of the reference
(tests/golden/ref_golden.npz) and (b) the numpy restatement oracle/vox_oracle.py,
all through the C ABI; plus the whole-clip call against the reference pipeline
and the size-independent properties used at the full bench size."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

MEL_REL, ACT_ABS = 1e-5, 2e-2          # north-star tolerances (BASELINE.json)


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "ref_golden.npz"))


def test_mel_against_golden_and_oracle(full_env, golden):
    import make_golden as mg
    import vox_oracle as vo
    e = full_env
    x = mg.audio()
    buf, nf = vo.stream_buffer(x)
    assert nf == int(golden["mel_stream_frames"])
    out = np.empty((nf, 128), np.float32)
    assert e.L.vox_cuda_mel_frames(e.vc.fptr(buf), nf, e.vc.fptr(out)) == 0, e.vc.last_error()
    g = golden["mel_stream_rows240_380"]
    assert float(np.max(np.abs(out[mg.MEL_ROWS] - g) / np.maximum(np.abs(g), 1.0))) < MEL_REL
    o = vo.mel_frames(buf, nf)
    assert float(np.max(np.abs(out - o) / np.maximum(np.abs(o), 1.0))) < MEL_REL


def test_mel_edge_cases(full_env):
    """one frame; an all-zero buffer (clamp floor); a full-scale square wave (no overflow)."""
    import vox_oracle as vo
    e = full_env
    for buf in (np.zeros(400, np.float32), np.ones(400, np.float32),
                np.sign(np.sin(np.arange(400 + 160 * 9) * 0.3)).astype(np.float32)):
        nf = (len(buf) - 400) // 160 + 1
        out = np.empty((nf, 128), np.float32)
        assert e.L.vox_cuda_mel_frames(e.vc.fptr(buf), nf, e.vc.fptr(out)) == 0
        o = vo.mel_frames(buf, nf)
        assert float(np.max(np.abs(out - o) / np.maximum(np.abs(o), 1.0))) < MEL_REL
    assert e.L.vox_cuda_mel_frames(e.vc.fptr(np.zeros(400, np.float32)), 0, e.vc.fptr(out)) == 0   # empty input


def test_encoder_against_golden(full_env, golden):
    import make_golden as mg
    e = full_env
    st = e.L.vox_cuda_state_create(e.model)
    assert st, e.vc.last_error()
    try:
        mel = np.ascontiguousarray(golden["mel_batch"][:64])
        stem = np.empty((40, 1280), np.float32)
        n = C.c_int()
        assert e.L.vox_cuda_conv_stem(st, e.vc.fptr(mel), 64, e.vc.fptr(stem), C.byref(n)) == 0
        assert n.value == 32
        xs = stem[:32].copy()
        assert e.L.vox_cuda_encoder_step(st, e.vc.fptr(xs), 32) == 0, e.vc.last_error()
        assert float(np.abs(xs[:, ::8] - golden["enc_forward_cols8"]).max()) < ACT_ABS
        ad = np.empty((8, 3072), np.float32)
        assert e.L.vox_cuda_adapter(st, e.vc.fptr(xs), 32, e.vc.fptr(ad)) == 0
        assert float(np.abs(ad - golden["adapter_of_enc_forward"]).max()) < ACT_ABS
        # incremental: two ragged chunks through the device K/V cache
        e.L.vox_cuda_state_reset_encoder(st)
        for i, xn in enumerate(mg.enc_inputs()):
            xc = xn.copy()
            assert e.L.vox_cuda_encoder_step(st, e.vc.fptr(xc), xc.shape[0]) == 0, e.vc.last_error()
            assert float(np.abs(xc[:, ::8] - golden[f"enc_incremental{i}_cols8"]).max()) < ACT_ABS
    finally:
        e.L.vox_cuda_state_free(st)


def test_decoder_against_golden(full_env, golden):
    import make_golden as mg
    e = full_env
    rows = mg.dec_adapter_rows()
    prompt = [1] + [32] * 37
    emb = np.stack([rows[i] + e.ref.tok_embedding(t) for i, t in enumerate(prompt)]).astype(np.float32)
    st = e.L.vox_cuda_state_create(e.model)
    assert st, e.vc.last_error()
    try:
        assert e.L.vox_cuda_decoder_prefill(st, e.vc.fptr(emb), 38) == 0, e.vc.last_error()
        prev = 32
        logits = np.empty(131072, np.float32)
        for i in range(mg.N_STEPS):
            xin = (rows[38 + i] + e.ref.tok_embedding(prev)).astype(np.float32)
            tok = e.L.vox_cuda_decoder_step(st, e.vc.fptr(xin), e.vc.fptr(logits))
            assert tok == int(golden["dec_tokens"][i])
            assert float(np.abs(logits[::32] - golden["dec_logits_cols32"][i]).max()) < ACT_ABS
            prev = tok
    finally:
        e.L.vox_cuda_state_free(st)


def _frame_buffer(x):
    import vox_oracle as vo
    return vo.stream_buffer(x)


def test_whole_clip_tokens_match_reference(full_env):
    """BASELINE config 2 at a CPU-affordable length: a 2 s clip through vox_cuda_transcribe
    (mel -> encoder -> adapter -> prefill -> greedy decode on the device) against the
    reference pipeline driven stage by stage (voxtral.c:1390-1396 order).  Token ids bit-exact."""
    import test_gpu_parity as tp
    e = full_env
    x = tp.synth_audio(16000 * 2, seed=5)
    buf, nf = _frame_buffer(x)
    # reference
    mel = e.ref.stream_mel(x)
    assert mel.shape[0] == nf
    ad = e.ref.adapter(e.ref.encoder_forward(mel))
    T = ad.shape[0]
    prompt = [1] + [32] * 38                                  # BOS + (32 + delay 6) streaming pads
    emb = np.stack([ad[i] + e.ref.tok_embedding(t) for i, t in enumerate(prompt[:-1])]).astype(np.float32)
    e.ref.set("kv_cache_len", 0); e.ref.set("kv_pos_offset", 0)
    e.ref.prefill(emb)
    ref_toks, prev = [], prompt[-1]
    for pos in range(38, T):
        tok, _ = e.ref.decoder_forward((ad[pos] + e.ref.tok_embedding(prev)).astype(np.float32))
        ref_toks.append(tok)
        prev = tok
    # CUDA, host PCM in
    st = e.L.vox_cuda_state_create(e.model)
    assert st, e.vc.last_error()
    try:
        toks = np.empty(T + 8, np.int32)
        ms = C.c_float()
        n = e.L.vox_cuda_transcribe(st, e.vc.fptr(buf), len(buf), nf, 39, e.vc.iptr(toks), len(toks), C.byref(ms))
        assert n == len(ref_toks), (n, len(ref_toks), e.vc.last_error())
        assert toks[:n].tolist() == ref_toks
        # idempotence: a second call on the same state starts from a clean slate
        toks2 = np.empty_like(toks)
        assert e.L.vox_cuda_transcribe(st, e.vc.fptr(buf), len(buf), nf, 39, e.vc.iptr(toks2), len(toks2), None) == n
        assert np.array_equal(toks[:n], toks2[:n])
    finally:
        e.L.vox_cuda_state_free(st)


def test_full_size_properties_60s(full_env):
    """At the bench size (60 s, 6392 mel frames, 799 adapter tokens, 761 decode steps) the CPU
    reference is too slow to diff, so check size-independent properties: (1) determinism -
    host-PCM and device-resident-PCM runs give identical token ids; (2) causality - the
    adapter tokens of the first 20 s are unchanged by the audio that follows; (3) the chunked
    encoder equals the one-pass encoder on the same mel."""
    import bench
    e = full_env
    x = bench.synth_clip(60.0, 1)
    buf, nf = bench.frame_buffer(x)
    st = e.L.vox_cuda_state_create(e.model)
    assert st, e.vc.last_error()
    try:
        toks = np.empty(nf // 8 + 8, np.int32)
        n1 = e.L.vox_cuda_transcribe(st, e.vc.fptr(buf), len(buf), nf, 39, e.vc.iptr(toks), len(toks), None)
        assert n1 == nf // 8 - 38, (n1, e.vc.last_error())
        a = toks[:n1].copy()
        assert e.L.vox_cuda_pcm_upload(st, e.vc.fptr(buf), len(buf)) == 0
        n2 = e.L.vox_cuda_transcribe(st, None, len(buf), nf, 39, e.vc.iptr(toks), len(toks), None)
        assert n2 == n1 and np.array_equal(a, toks[:n2])
        T = e.L.vox_cuda_adapter_count(st)
        full = np.empty((T, 3072), np.float32)
        assert e.L.vox_cuda_adapter_read(st, 0, T, e.vc.fptr(full)) == T
        # causality / chunking: encode only the first 2048 mel frames, in 3 ragged chunks
        mel = np.empty((nf, 128), np.float32)
        assert e.L.vox_cuda_mel_frames(e.vc.fptr(buf), nf, e.vc.fptr(mel)) == 0
        e.L.vox_cuda_state_reset_encoder(st)
        got = 0
        for a0, a1 in ((0, 312), (312, 1000), (1000, 2048)):
            got = e.L.vox_cuda_encode_chunk(st, e.vc.fptr(np.ascontiguousarray(mel[a0:a1])), a1 - a0)
        assert got == 256
        part = np.empty((256, 3072), np.float32)
        assert e.L.vox_cuda_adapter_read(st, 0, 256, e.vc.fptr(part)) == 256
        d = float(np.abs(part - full[:256]).max())
        print("chunked vs one-pass adapter tokens max|d|", d)
        assert d < 2e-3
    finally:
        e.L.vox_cuda_state_free(st)


def test_alternatives_match_oracle(full_env):
    """--alt (stream_fill_alts, voxtral.c:955-1010): ids >= 1000, by probability, cut by 1 - p_i/p_best."""
    import vox_oracle as vo
    e = full_env
    rng = np.random.default_rng(17)
    st = e.L.vox_cuda_state_create(e.model)
    assert st, e.vc.last_error()
    try:
        emb = (0.5 * rng.standard_normal((38, 3072))).astype(np.float32)
        assert e.L.vox_cuda_decoder_prefill(st, e.vc.fptr(emb), 38) == 0
        logits = np.empty(131072, np.float32)
        ids = np.empty(8, np.int32)
        for step in range(3):
            x = (0.5 * rng.standard_normal(3072)).astype(np.float32)
            tok = e.L.vox_cuda_decoder_step(st, e.vc.fptr(x), e.vc.fptr(logits))
            assert tok == int(np.argmax(logits))
            for n_alt, cutoff in ((4, 1.0), (4, 0.5), (3, 0.2), (2, 0.05), (1, 1.0), (4, 0.0)):
                n = e.L.vox_cuda_decoder_alternatives(st, n_alt, C.c_float(cutoff), e.vc.iptr(ids))
                want = vo.fill_alts(logits, tok, n_alt, cutoff)
                assert ids[:n].tolist() == want, (step, n_alt, cutoff, ids[:n].tolist(), want)
    finally:
        e.L.vox_cuda_state_free(st)
