"""The encoder's 750-key sliding window, the reference's compaction and the device ring's wrap, against
the REFERENCE (vox_encoder_forward_incremental of oracle/_ref/libvoxref.so, voxtral_encoder.c:495-693;
compaction :431-449): the same ragged chunks of stem rows go through both, outputs within 2e-2 max-abs
(BASELINE tolerance for encoder activations) and the host-visible position bookkeeping must agree."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ACT_ABS = 2e-2


def run_both(e, chunks, start=0, seed=3):
    rng = np.random.default_rng(seed)
    st = e.L.vox_cuda_state_create(e.model)
    assert st, e.vc.last_error()
    worst = 0.0
    try:
        e.ref.set("enc_kv_cache_len", 0)
        e.ref.set("enc_kv_pos_offset", start)
        assert e.L.vox_cuda_encoder_set_position(st, start) == 0
        for n in chunks:
            x = (0.5 * rng.standard_normal((n, 1280))).astype(np.float32)
            want = e.ref.encoder_incremental(x)
            got = x.copy()
            assert e.L.vox_cuda_encoder_step(st, e.vc.fptr(got), n) == 0, e.vc.last_error()
            d = float(np.abs(got - want).max())
            worst = max(worst, d)
            ref_pos = e.ref.get("enc_kv_pos_offset") + e.ref.get("enc_kv_cache_len")
            assert e.L.vox_cuda_encoder_position(st) == ref_pos, (e.L.vox_cuda_encoder_position(st), ref_pos)
            assert d < ACT_ABS, (n, ref_pos, d)
        return worst, e.ref.get("enc_kv_pos_offset")
    finally:
        e.L.vox_cuda_state_free(st)
        e.ref.set("enc_kv_cache_len", 0)
        e.ref.set("enc_kv_pos_offset", 0)


def test_window_and_compaction_beyond_750_positions(full_env):
    """1115 positions in ragged chunks (large-tile path, M > 128, and the skinny streaming path, M <= 128):
    the reference compacts its cache twice (enc_kv_pos_offset > 0), queries past position 750 must not see the
    keys the window dropped"""
    e = full_env
    worst, offset = run_both(e, [300, 250, 260, 100, 128, 77])
    assert offset > 0, "the reference never compacted: the test does not reach the window"
    print("encoder vs reference over 1115 positions (window 750): max|d| =", worst)


def test_device_ring_wrap_at_4096(full_env):
    """The device K/V ring has 4096 slots; a stream reaches them after ~82 s of audio.  Both engines start with
    an empty cache at logical position 3950 (the reference through its public enc_kv_pos_offset field,
    voxtral.h:218-224) and run across 4096: RoPE at the logical position, ring slots wrapping, window intact."""
    worst, _ = run_both(full_env, [100, 60, 200, 25], start=3950, seed=4)
    print("encoder vs reference across the 4096-slot ring wrap: max|d| =", worst)
