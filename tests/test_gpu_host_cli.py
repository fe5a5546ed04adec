"""The C host (voxtral.c_b200/host, binary voxtral_b200) against the ctypes path: same WAV,
batch mode and streaming mode (-I 0.5, fed in 0.3 s pieces; s16le on stdin) must print the same
greedy token ids that vox_cuda_transcribe returns - which test_gpu_golden.py ties to the reference."""
import os
import struct
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "voxtral.c_b200", "voxtral_b200")


def write_wav(path, x):
    pcm = (np.clip(x, -1, 1) * 32767).astype("<i2").tobytes()
    with open(path, "wb") as f:
        f.write(b"RIFF" + struct.pack("<I", 36 + len(pcm)) + b"WAVEfmt " +
                struct.pack("<IHHIIHH", 16, 1, 1, 16000, 32000, 2, 16) + b"data" + struct.pack("<I", len(pcm)) + pcm)
    return np.frombuffer(pcm, "<i2").astype(np.float32) / 32768.0


def run_cli(args, stdin=None):
    p = subprocess.run([EXE] + args, input=stdin, capture_output=True, timeout=600)
    assert p.returncode == 0, p.stderr.decode()[-2000:]
    print(p.stderr.decode()[-600:])
    return [int(t) for t in p.stdout.decode().split()]


def test_cli_batch_and_stream_match_ctypes(full_env, tmp_path):
    import ctypes as C
    import test_gpu_parity as tp
    import vox_oracle_path  # noqa: F401
    import vox_oracle as vo
    e = full_env
    assert os.path.exists(EXE), "run make -C voxtral.c_b200 cuda"
    wav = str(tmp_path / "clip.wav")
    x = write_wav(wav, tp.synth_audio(16000 * 3 + 333, seed=9))       # not a multiple of 1280: exercises the align pad
    buf, nf = vo.stream_buffer(x)
    st = e.L.vox_cuda_state_create(e.model)
    assert st, e.vc.last_error()
    try:
        toks = np.empty(nf // 8 + 8, np.int32)
        n = e.L.vox_cuda_transcribe(st, e.vc.fptr(buf), len(buf), nf, 39, e.vc.iptr(toks), len(toks), None)
        assert n > 0, e.vc.last_error()
        want = toks[:n].tolist()
        if 2 in want:
            want = want[:want.index(2)]
    finally:
        e.L.vox_cuda_state_free(st)
    assert run_cli(["-d", e.mdir, "-i", wav]) == want
    assert run_cli(["-d", e.mdir, "-i", wav, "-I", "0.5", "--stream", "0.3"]) == want
    pcm = (x * 32768.0).astype("<i2").tobytes()
    assert run_cli(["-d", e.mdir, "--stdin", "-I", "2.0"], stdin=pcm) == want


def test_cli_ingests_stereo_44k_like_the_reference(full_env, tmp_path):
    """A 44.1 kHz stereo WAV through the C CLI (RIFF parsing on the host, downmix + resampling on the device,
    csrc/pcm.cu) gives the ids of the samples the reference's own vox_parse_wav_buffer produces from the file."""
    import refharness as rh
    import test_gpu_parity as tp
    import vox_oracle_path  # noqa: F401
    import vox_oracle as vo
    e = full_env
    x16 = tp.synth_audio(2 * 16000 + 500, seed=14)
    n44 = int(len(x16) * 44100 / 16000)
    mono44 = np.interp(np.arange(n44) * (len(x16) - 1) / (n44 - 1), np.arange(len(x16)), x16)
    a = np.clip(np.stack([0.9 * mono44, 0.6 * mono44], axis=1) * 32767, -32768, 32767).astype(np.int16)
    image = rh.make_wav(a, 2, 44100)
    wav = str(tmp_path / "stereo44.wav")
    with open(wav, "wb") as f:
        f.write(image)
    samples = rh.ref_parse_wav(image, strict=True)
    buf, nf = vo.stream_buffer(samples)
    st = e.L.vox_cuda_state_create(e.model)
    assert st, e.vc.last_error()
    try:
        toks = np.empty(nf // 8 + 8, np.int32)
        n = e.L.vox_cuda_transcribe(st, e.vc.fptr(buf), len(buf), nf, 39, e.vc.iptr(toks), len(toks), None)
        assert n > 0, e.vc.last_error()
        want = toks[:n].tolist()
        if 2 in want:
            want = want[:want.index(2)]
    finally:
        e.L.vox_cuda_state_free(st)
    assert run_cli(["-d", e.mdir, "-i", wav]) == want
