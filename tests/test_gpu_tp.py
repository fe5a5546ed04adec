"""Tensor-parallel single-stream decoder (csrc/dec_tp.cu) through the C ABI.

One GPU: all G ranks on the device in lock step (vox_cuda_tp_create_local) - the shard arithmetic, the
exchange windows and the sequence numbers are those of the multi-GPU run.  Greedy ids must equal the
single-stream path, whose ids on this 2 s clip tests/test_gpu_golden.py pins to the reference BLAS path;
logits within the north star's 2e-2 max-abs (summation order is the only difference).
Two or more GPUs: tools/tp_probe.py under torchrun (one process per GPU, CUDA IPC windows over NVLink)."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def _clip(seconds, seed):
    import test_gpu_parity as tp
    import vox_oracle as vo
    x = tp.synth_audio(int(16000 * seconds), seed=seed)
    return vo.stream_buffer(x)


def _single(e, buf, nf):
    st = e.L.vox_cuda_state_create(e.model)
    assert st, e.vc.last_error()
    try:
        toks = np.empty(nf // 8 + 8, np.int32)
        n = e.L.vox_cuda_transcribe(st, e.vc.fptr(buf), len(buf), nf, 39, e.vc.iptr(toks), len(toks), None)
        assert n > 0, e.vc.last_error()
        logits = np.empty(131072, np.float32)
        assert e.L.vox_cuda_decoder_read_logits(st, e.vc.fptr(logits)) == 0
        return toks[:n].tolist(), logits
    finally:
        e.L.vox_cuda_state_free(st)


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_tp_lockstep_ids_equal_single_stream(full_env, world):
    e = full_env
    buf, nf = _clip(2.0, 5)
    want, want_logits = _single(e, buf, nf)
    tp = e.L.vox_cuda_tp_create_local(e.model, world)
    assert tp, e.vc.last_error()
    try:
        toks = np.empty(nf // 8 + 8, np.int32)
        ms = C.c_float(0)
        n = e.L.vox_cuda_tp_transcribe(tp, e.vc.fptr(buf), len(buf), nf, 39, e.vc.iptr(toks), len(toks), C.byref(ms))
        assert n == len(want), e.vc.last_error()
        assert toks[:n].tolist() == want
        # in lock step the ranks share one logits buffer, each writing its vocabulary slice
        logits = np.empty(131072, np.float32)
        assert e.L.vox_cuda_decoder_read_logits(e.L.vox_cuda_tp_state(tp), e.vc.fptr(logits)) == 0
        assert float(np.abs(logits - want_logits).max()) < 2e-2
        # a second clip on the same handle: sequence numbers keep counting, state resets cleanly
        buf2, nf2 = _clip(2.4, 12)
        want2, _ = _single(e, buf2, nf2)
        n2 = e.L.vox_cuda_tp_transcribe(tp, e.vc.fptr(buf2), len(buf2), nf2, 39, e.vc.iptr(toks), len(toks), None)
        assert n2 == len(want2), e.vc.last_error()
        assert toks[:n2].tolist() == want2
    finally:
        e.L.vox_cuda_tp_free(tp)


@pytest.mark.parametrize("world", [2, 4])
def test_tp_fp16_kv_rings(full_env, world):
    """SURVEY 8(f2) for the tensor-parallel decoder: vox_cuda_tp_create_local_ex(model, world, 16) - every rank's K/V rows
    rounded to fp16 when appended.  Stated as a tolerance against the fp32-ring tensor-parallel run on the same clip (the
    fp32 BLAS oracle cannot pin a 16-bit cache bit for bit): last-step logits within KV16_LOGIT_TOL, ids equal up to the
    first step whose margin is within reach of that difference (free-running afterwards)."""
    from test_gpu_kv16 import KV16_LOGIT_TOL
    e = full_env
    buf, nf = _clip(2.0, 5)
    out = []
    for bits in (32, 16):
        tp = e.L.vox_cuda_tp_create_local_ex(e.model, world, bits)
        assert tp, e.vc.last_error()
        try:
            assert e.L.vox_cuda_state_kv_bits(e.L.vox_cuda_tp_state(tp)) == bits
            toks = np.empty(nf // 8 + 8, np.int32)
            n = e.L.vox_cuda_tp_transcribe(tp, e.vc.fptr(buf), len(buf), nf, 39, e.vc.iptr(toks), len(toks), None)
            assert n > 0, e.vc.last_error()
            logits = np.empty(131072, np.float32)
            assert e.L.vox_cuda_decoder_read_logits(e.L.vox_cuda_tp_state(tp), e.vc.fptr(logits)) == 0
            out.append((toks[:n].copy(), logits))
        finally:
            e.L.vox_cuda_tp_free(tp)
    (t32, l32), (t16, l16) = out
    assert len(t32) == len(t16)
    nz = np.nonzero(t32 != t16)[0]
    d = float(np.abs(l32 - l16).max())
    print(f"TP x{world}: fp16 vs fp32 K/V rings: {len(t32) - len(nz)} of {len(t32)} ids equal, last-step max|dlogit| {d:.2e}")
    if len(nz) == 0:
        assert d < KV16_LOGIT_TOL
    else:
        assert nz[0] >= 10


def test_tp_argument_errors(full_env):
    e = full_env
    assert not e.L.vox_cuda_tp_create_local(e.model, 3)
    assert not e.L.vox_cuda_tp_create(e.model, 2, 2)
    tp = e.L.vox_cuda_tp_create(e.model, 0, 2)
    assert tp, e.vc.last_error()
    try:
        toks = np.empty(8, np.int32)
        # peers never attached: decode must refuse instead of waiting for them
        assert e.L.vox_cuda_tp_decode(tp, 32, 38, 1, e.vc.iptr(toks), None) == -1
        h = (C.c_ubyte * e.vc.TP_HANDLE_BYTES)()
        assert e.L.vox_cuda_tp_export(tp, h) == 0, e.vc.last_error()
        bad = (C.c_ubyte * (2 * e.vc.TP_HANDLE_BYTES))()
        assert e.L.vox_cuda_tp_connect(tp, bad) == -1
    finally:
        e.L.vox_cuda_tp_free(tp)


def test_tp_two_processes_two_gpus(full_env):
    """One process per GPU, windows mapped through CUDA IPC: ids equal on both ranks and equal to the
    single-stream run (the probe checks and prints them)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29533",
                          os.path.join(ROOT, "tools", "tp_probe.py"), "--seconds", "2", "--model", full_env.mdir],
                         capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "ids_equal_single_stream=True" in out.stdout


@pytest.mark.parametrize("mega", ["1", "0"])
def test_tp_single_rank_world_paths(full_env, monkeypatch, mega):
    """world = 1 through vox_cuda_tp_create: the persistent cooperative step kernel (dec_tp_mega.cu, default) and
    the kernel-per-op graph (VOX_CUDA_TP_MEGA=0) both give the single-stream ids; a second clip on the same
    handle continues the exchange sequence numbers and the barrier generation."""
    e = full_env
    monkeypatch.setenv("VOX_CUDA_TP_MEGA", mega)
    tp = e.L.vox_cuda_tp_create(e.model, 0, 1)
    assert tp, e.vc.last_error()
    try:
        for seconds, seed in ((2.0, 5), (2.4, 12), (6.0, 3)):
            buf, nf = _clip(seconds, seed)
            want, want_logits = _single(e, buf, nf)
            toks = np.empty(nf // 8 + 8, np.int32)
            n = e.L.vox_cuda_tp_transcribe(tp, e.vc.fptr(buf), len(buf), nf, 39, e.vc.iptr(toks), len(toks), None)
            assert n == len(want), e.vc.last_error()
            assert toks[:n].tolist() == want
            logits = np.empty(131072, np.float32)
            assert e.L.vox_cuda_decoder_read_logits(e.L.vox_cuda_tp_state(tp), e.vc.fptr(logits)) == 0
            assert float(np.abs(logits - want_logits).max()) < 2e-2
    finally:
        e.L.vox_cuda_tp_free(tp)


@pytest.mark.parametrize("mode", ["lockstep2", "lockstep8", "persistent1"])
def test_tp_window_wrap_matches_single_stream(full_env, monkeypatch, mode):
    """BASELINE config 4 for the tensor-parallel step: start right before the 8192-position ring wraps with the
    SAME injected K/V state and adapter rows as a single-stream state (whose wrap behaviour
    tests/test_gpu_kvwrap.py pins to the reference's cache compaction) and step across the wrap: ids equal,
    logits within 2e-2.  Exercises the sharded attention's ring arithmetic (full window, key splits, first
    logical key) in the kernel-per-op engine (2 and 8 ranks in lock step) and in the persistent kernel."""
    e = full_env
    rng = np.random.default_rng(33)
    start, n_steps = 8192 + 1060, 6
    K = (0.3 * rng.standard_normal((26, 8192, 1024), dtype=np.float32))
    V = (0.3 * rng.standard_normal((26, 8192, 1024), dtype=np.float32))
    ad = (0.5 * rng.standard_normal((n_steps, 3072))).astype(np.float32)

    def inject(s):
        for l in range(26):
            assert e.L.vox_cuda_decoder_kv_write(s, l, start - 8192, 8192, e.vc.fptr(K[l]), e.vc.fptr(V[l])) == 0
        assert e.L.vox_cuda_decoder_set_position(s, start, 8192) == 0
        assert e.L.vox_cuda_adapter_append(s, e.vc.fptr(ad), n_steps) >= 0, e.vc.last_error()

    st = e.L.vox_cuda_state_create(e.model)
    assert st, e.vc.last_error()
    try:
        inject(st)
        want = np.empty(n_steps, np.int32)
        assert e.L.vox_cuda_decoder_run(st, 32, 0, n_steps, e.vc.iptr(want)) == n_steps, e.vc.last_error()
        want_logits = np.empty(131072, np.float32)
        assert e.L.vox_cuda_decoder_read_logits(st, e.vc.fptr(want_logits)) == 0
    finally:
        e.L.vox_cuda_state_free(st)
    if mode == "persistent1":
        monkeypatch.setenv("VOX_CUDA_TP_MEGA", "1")
        tp = e.L.vox_cuda_tp_create(e.model, 0, 1)
    else:
        tp = e.L.vox_cuda_tp_create_local(e.model, int(mode[len("lockstep"):]))
    assert tp, e.vc.last_error()
    try:
        ts = e.L.vox_cuda_tp_state(tp)
        inject(ts)
        got = np.empty(n_steps, np.int32)
        assert e.L.vox_cuda_tp_decode(tp, 32, 0, n_steps, e.vc.iptr(got), None) == n_steps, e.vc.last_error()
        assert got.tolist() == want.tolist()
        assert e.L.vox_cuda_decoder_position(ts) == start + n_steps
        logits = np.empty(131072, np.float32)
        assert e.L.vox_cuda_decoder_read_logits(ts, e.vc.fptr(logits)) == 0
        assert float(np.abs(logits - want_logits).max()) < 2e-2
    finally:
        e.L.vox_cuda_tp_free(tp)
