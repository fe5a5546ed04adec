"""device-side audio ingestion (csrc/pcm.cu) through the C ABI against the compiled
reference's vox_parse_wav_buffer (strict-IEEE build): bit-exact for mono / multi-channel input with and
without resampling; the silence gate's window RMS against the oracle; then the decoded clip: a 44.1 kHz
stereo s16 clip ingested on the device transcribes to the ids of the reference-ingested samples."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

from test_cpu_pcm import CASES, frames  # noqa: E402


@pytest.mark.parametrize("channels,rate,n", CASES + [(2, 44100, 441000)])
def test_pcm_from_s16_bit_exact(full_env, channels, rate, n):
    import refharness as rh
    e = full_env
    a = frames(channels, n, channels * 100000 + rate + n)
    want = rh.ref_parse_wav(rh.make_wav(a, channels, rate), strict=True)
    assert e.L.vox_cuda_pcm_resampled_len(n, rate) == len(want)
    out = np.empty(len(want) + 3, np.float32)
    got_n = e.L.vox_cuda_pcm_from_s16(a.ctypes.data_as(C.POINTER(C.c_int16)), n, channels, rate, e.vc.fptr(out), len(out))
    assert got_n == len(want), e.vc.last_error()
    assert np.array_equal(out[:got_n].view(np.uint32), want.view(np.uint32))
    # capacity too small: refused, nothing written past it
    assert e.L.vox_cuda_pcm_from_s16(a.ctypes.data_as(C.POINTER(C.c_int16)), n, channels, rate, e.vc.fptr(out), max(0, len(want) - 1)) == -1 or len(want) == 0


def test_window_rms(full_env):
    import vox_oracle as vo
    e = full_env
    rng = np.random.default_rng(9)
    x = (0.02 * rng.standard_normal(4800 + 37)).astype(np.float32)
    x[1600:3200] = 0
    out = np.empty(len(x) // 160, np.float32)
    assert e.L.vox_cuda_window_rms(e.vc.fptr(x), len(x), 160, e.vc.fptr(out)) == 30, e.vc.last_error()
    want = vo.window_rms(x, 160)
    assert np.array_equal(out.view(np.uint32), want.view(np.uint32))
    assert np.all(out[10:20] == 0)


def test_ingested_clip_transcribes_like_reference_ingest(full_env):
    import refharness as rh
    import vox_oracle as vo
    import test_gpu_parity as tpar
    e = full_env
    x16 = tpar.synth_audio(2 * 16000, seed=5)
    n44 = 2 * 44100
    t = np.arange(n44) * (len(x16) - 1) / (n44 - 1)
    mono44 = np.interp(t, np.arange(len(x16)), x16)
    st = np.stack([mono44 * 0.9, mono44 * 0.7], axis=1)
    a = np.clip(st * 32767, -32768, 32767).astype(np.int16)
    want_pcm = rh.ref_parse_wav(rh.make_wav(a, 2, 44100), strict=True)
    pcm = np.empty(len(want_pcm), np.float32)
    assert e.L.vox_cuda_pcm_from_s16(a.ctypes.data_as(C.POINTER(C.c_int16)), n44, 2, 44100, e.vc.fptr(pcm), len(pcm)) == len(want_pcm)
    s = e.L.vox_cuda_state_create(e.model)
    assert s, e.vc.last_error()
    try:
        ids = []
        for samples in (pcm, want_pcm):
            buf, nf = vo.stream_buffer(samples)
            toks = np.empty(nf // 8 + 8, np.int32)
            n = e.L.vox_cuda_transcribe(s, e.vc.fptr(buf), len(buf), nf, 39, e.vc.iptr(toks), len(toks), None)
            assert n > 0, e.vc.last_error()
            ids.append(toks[:n].tolist())
        assert ids[0] == ids[1]
    finally:
        e.L.vox_cuda_state_free(s)
