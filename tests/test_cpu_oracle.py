"""Pin oracle/vox_oracle.py (the numpy restatement) against golden outputs of the
UNMODIFIED reference (tests/golden/ref_golden.npz, made by tests/golden/make_golden.py
from oracle/_ref).  CPU only; the checkpoint is regenerated from its seed (~15 s)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

import refharness as rh  # noqa: E402
import vox_oracle as vo  # noqa: E402
import make_golden as mg  # noqa: E402

# tolerances of the north star (BASELINE.json): mel 1e-5 relative, activations / logits 2e-2 max-abs.
# The restatement computes in fp32 like the reference, so it is held to much tighter bounds.
MEL_REL = 1e-5
ACT_ABS = 2e-3


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "ref_golden.npz"))


@pytest.fixture(scope="module")
def ck():
    return vo.Checkpoint(rh.model_dir())


def rel(a, b):
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1.0)))


def test_mel_filters_are_slaney_normalised():
    bank = vo.mel_filters()
    assert bank.shape == (128, 201) and bank.min() >= 0
    assert (bank > 0).sum(axis=1).min() >= 1          # every filter touches at least one bin
    assert np.all(bank[:, 0] == 0)                    # DC bin is outside every triangle (f_min = 0)


def test_stream_mel_matches_reference(golden):
    m = vo.stream_mel(mg.audio())
    assert m.shape[0] == int(golden["mel_stream_frames"])
    assert rel(m[mg.MEL_ROWS], golden["mel_stream_rows240_380"]) < MEL_REL
    # frames that only see the zero left padding are the clamp floor (1.5 - 8 + 4) / 4
    assert np.all(m[:200] == np.float32(-0.625))


def test_batch_mel_matches_reference(golden):
    m = vo.batch_mel(mg.audio())
    assert m.shape == golden["mel_batch"].shape
    assert rel(m, golden["mel_batch"]) < MEL_REL


def test_mel_edge_cases():
    # shorter than one window: reflect padding reads zeros past the ends (voxtral_audio.c:305-314)
    assert vo.batch_mel(np.ones(161, np.float32)).shape == (1, 128)
    buf, nf = vo.stream_buffer(np.zeros(0, np.float32))
    assert nf == (len(buf) - 400) // 160 and len(buf) == 200 + 32 * 1280 + 17 * 1280 + 200


def test_ada_scale_matches_reference(golden, ck):
    a = vo.ada_scale(ck, 6)
    assert float(np.abs(a[:, ::8] - golden["ada_scale_cols8"]).max()) < 1e-5


def test_encoder_one_pass_and_adapter_match_reference(golden, ck):
    enc = vo.Encoder(ck)
    out = enc.forward(golden["mel_batch"][:64])
    assert out.shape == (32, 1280)
    assert float(np.abs(out[:, ::8] - golden["enc_forward_cols8"]).max()) < ACT_ABS
    ad = vo.adapter(ck, out)
    assert ad.shape == (8, 3072)
    assert float(np.abs(ad - golden["adapter_of_enc_forward"]).max()) < ACT_ABS


def test_encoder_incremental_matches_reference(golden, ck):
    enc = vo.Encoder(ck)
    for i, xn in enumerate(mg.enc_inputs()):
        out = enc.forward_incremental(xn)
        assert float(np.abs(out[:, ::8] - golden[f"enc_incremental{i}_cols8"]).max()) < ACT_ABS
    assert enc.cache_len == 40 and enc.pos_offset == 0


def test_encoder_chunking_is_equivalent_and_window_compacts(ck):
    """Size-independent property: feeding the conv-stem rows in chunks gives the rows of the
    one-pass result (causal attention + K/V cache), and the cache never keeps more than the
    750-row window once a chunk boundary passes it (2-layer checkpoint to stay fast)."""
    small = vo.Checkpoint(rh.model_dir(enc_layers=2, dec_layers=1))
    rng = np.random.default_rng(3)
    x = (0.5 * rng.standard_normal((900, 1280))).astype(np.float32)
    e1 = vo.Encoder(small)
    one = e1.forward_incremental(x[:700])
    e2 = vo.Encoder(small)
    parts = np.concatenate([e2.forward_incremental(x[a:b]) for a, b in ((0, 256), (256, 300), (300, 700))])
    assert float(np.abs(one - parts).max()) < 1e-4
    e2.forward_incremental(x[700:800])      # 700 + 100 > 750 but cache_len 700 <= 750: no compaction yet
    assert e2.cache_len == 800 and e2.pos_offset == 0
    e2.forward_incremental(x[800:900])      # now 800 > 750: compacted to 750, offset 50
    assert e2.cache_len == 850 and e2.pos_offset == 50


def test_decoder_prefill_and_steps_match_reference(golden, ck):
    dec = vo.Decoder(ck)
    rows = mg.dec_adapter_rows()
    prompt = [1] + [32] * 37
    emb = np.stack([rows[i] + dec.tok_embedding(t) for i, t in enumerate(prompt)]).astype(np.float32)
    dec.prefill(emb)
    prev = 32
    for i in range(mg.N_STEPS):
        tok, logits = dec.forward(rows[38 + i] + dec.tok_embedding(prev))
        assert tok == int(golden["dec_tokens"][i])
        assert float(np.abs(logits[::32] - golden["dec_logits_cols32"][i]).max()) < ACT_ABS
        prev = tok
    assert dec.cache_len == 38 + mg.N_STEPS and dec.cache_max == 8192 + 38 + 1024


def test_decoder_kv_compaction_keeps_window_and_positions():
    """kv_cache_compact semantics (voxtral_decoder.c:354-385) on a 1-layer checkpoint with a
    synthetic cache: after compaction the last 8192 rows remain and RoPE positions continue."""
    small = vo.Checkpoint(rh.model_dir(enc_layers=2, dec_layers=1))
    dec = vo.Decoder(small)
    rng = np.random.default_rng(9)
    n = 8192 + 38 + 1024
    dec.cache_max = n
    dec.k = [(0.1 * rng.standard_normal((n, 1024))).astype(np.float32)]
    dec.v = [(0.1 * rng.standard_normal((n, 1024))).astype(np.float32)]
    dec.cache_len = n
    k_tail = dec.k[0][-8192:].copy()
    tok, logits = dec.forward((0.5 * rng.standard_normal(3072)).astype(np.float32))
    assert dec.pos_offset == 38 + 1024 and dec.cache_len == 8193
    assert np.array_equal(dec.k[0][:8192], k_tail)
    assert np.isfinite(logits).all() and 0 <= tok < 131072


def test_fill_alts_rules():
    """stream_fill_alts (voxtral.c:955-1010): control ids < 1000 never appear, the best token is
    skipped, order is by probability, the cutoff is on 1 - p_i / p_best."""
    lg = np.full(131072, -5.0, np.float32)
    lg[7] = 9.0                      # control token with the highest logit: never an alternative
    lg[5000], lg[6000], lg[7000], lg[8000] = 4.0, 3.9, 3.0, 1.0
    assert vo.fill_alts(lg, 5000, 4, 1.0) == [6000, 7000, 8000]
    assert vo.fill_alts(lg, 5000, 3, 1.0) == [6000, 7000]
    assert vo.fill_alts(lg, 5000, 4, 0.2) == [6000]          # 1 - e^-0.1 = 0.095 kept, 1 - e^-1 = 0.63 cut
    assert vo.fill_alts(lg, 5000, 1, 1.0) == [] and vo.fill_alts(lg, 5000, 4, 0.0) == []
