"""the oracle's restatement of the WAV front-end arithmetic (oracle/vox_oracle.py::pcm_from_s16)
pinned bit for bit against the compiled reference's own vox_parse_wav_buffer (strict-IEEE build; the
-ffast-math build within 1e-2 on full-scale white noise), on mono / stereo / 3-channel input at 8, 16, 22.05, 44.1 and 48 kHz."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

import refharness as rh  # noqa: E402
import vox_oracle as vo  # noqa: E402

CASES = [(1, 16000, 4000), (1, 8000, 3001), (2, 44100, 9000), (2, 16000, 777), (3, 48000, 5000), (1, 22050, 1), (2, 48000, 2)]


def frames(channels, n, seed):
    rng = np.random.default_rng(seed)
    a = rng.integers(-32768, 32768, size=(n, channels), dtype=np.int32).astype(np.int16)
    a[0] = 32767
    a[-1] = -32768
    return a


@pytest.mark.parametrize("channels,rate,n", CASES)
def test_oracle_matches_reference_parse_wav(channels, rate, n):
    if not rh.have_ref():
        pytest.skip("oracle/_ref not built")
    a = frames(channels, n, channels * 100000 + rate + n)
    wav = rh.make_wav(a, channels, rate)
    want = rh.ref_parse_wav(wav, strict=True)
    got = vo.pcm_from_s16(a, channels, rate)
    assert got.shape == want.shape
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    fast = rh.ref_parse_wav(wav, strict=False)
    # -ffast-math turns /16000 into a reciprocal multiply: src_pos moves by an ulp, frac by ~1e-4 at position ~1e3,
    # and on full-scale white noise (neighbouring samples differ by up to 2) the sample by up to ~1e-3
    assert fast.shape == want.shape and (len(want) == 0 or float(np.abs(fast - want).max()) < 1e-2)


def test_window_rms_oracle():
    rng = np.random.default_rng(3)
    x = (0.05 * rng.standard_normal(1000)).astype(np.float32)
    r = vo.window_rms(x, 160)
    assert r.shape == (6,)
    want = np.sqrt((x[:960].astype(np.float64).reshape(6, 160) ** 2).mean(axis=1))
    assert float(np.abs(r - want).max()) < 1e-6
