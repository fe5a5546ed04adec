#!/usr/bin/env python3
"""BASELINE config 4 (GPU): a 30-minute synthetic clip through the streaming calls - the mel is
fed to vox_cuda_encode_chunk in chunks of 6656 frames (the library's largest one-pass chunk; argv[2] changes it), the decoder catches up after every chunk
(vox_cuda_decoder_run) and consumed adapter rows are dropped (vox_cuda_adapter_discard_before), the
way stream_run_encoder / stream_run_decoder do (/root/reference/voxtral.c:826-1145).  22 511 decode
steps; the K/V ring (8192 positions) wraps at step 8154 and the reference's compaction points
(9216, then every 1062 steps, voxtral_decoder.c:354-385) are passed 13 times.
Reports total time, ms/step per 2048-step bucket (context growth then the full-window plateau)
and checks causality: the ids of the first 60 s equal those of a run that only saw the first 90 s.
usage: long_run.py [seconds] [mel frames per encoder call]"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "voxtral.c_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
import refharness as rh  # noqa: E402
import voxcuda as vc  # noqa: E402
import hostmodel  # noqa: E402
import bench  # noqa: E402

secs = float(sys.argv[1]) if len(sys.argv) > 1 else 1800.0
CHUNK = int(sys.argv[2]) if len(sys.argv) > 2 else 6656
L = vc.lib()
assert L.vox_cuda_init() == 1, vc.last_error()
hm = hostmodel.HostModel(rh.model_dir())
model = vc.check(L.vox_cuda_model_create(C.byref(hm.weights)), "model")
st = vc.check(L.vox_cuda_state_create(model), "state")


def clip(seconds):
    # 60 s segments of the bench generator with different seeds, concatenated
    parts = [bench.synth_clip(min(60.0, seconds - 60.0 * i), 2 + i) for i in range(int(np.ceil(seconds / 60.0)))]
    return np.concatenate(parts)[: int(16000 * seconds)]


def run(x, chunk=None):
    chunk = chunk or CHUNK
    buf, nf = bench.frame_buffer(x)
    mel = np.empty((nf, 128), np.float32)
    t0 = time.perf_counter()
    vc.check(L.vox_cuda_mel_frames(vc.fptr(buf), nf, vc.fptr(mel)), "mel")
    t_mel = time.perf_counter() - t0
    L.vox_cuda_state_reset_encoder(st)
    L.vox_cuda_state_reset_decoder(st)
    prompt = np.array([1] + [32] * 38, np.int32)
    toks, stamps = [], []
    prefilled, next_pos, prev = False, 38, 32
    t_enc = t_dec = 0.0
    for off in range(0, nf, chunk):
        n = min(chunk, nf - off)
        part = np.ascontiguousarray(mel[off:off + n])
        t0 = time.perf_counter()
        total = vc.check(L.vox_cuda_encode_chunk(st, vc.fptr(part), n), "encode")
        t1 = time.perf_counter()
        t_enc += t1 - t0
        if not prefilled and total >= 39:
            vc.check(L.vox_cuda_decoder_prefill_fused(st, vc.iptr(prompt), 0, 38), "prefill")
            prefilled = True
        if prefilled and total > next_pos:
            k = total - next_pos
            out = np.empty(k, np.int32)
            t2 = time.perf_counter()
            vc.check(L.vox_cuda_decoder_run(st, prev, next_pos, k, vc.iptr(out)), "run")
            dt = time.perf_counter() - t2
            stamps.append((next_pos, k, dt))
            toks += out.tolist()
            prev, next_pos = int(out[-1]), total
            vc.check(L.vox_cuda_adapter_discard_before(st, next_pos), "discard")
        t_dec += time.perf_counter() - t1
    return toks, t_mel, t_enc, t_dec, stamps, nf


x = clip(secs)
toks, t_mel, t_enc, t_dec, stamps, nf = run(x)
total = t_mel + t_enc + t_dec
print(f"{secs:.0f} s clip: {nf} mel frames, {nf // 8} adapter tokens, {len(toks)} decode steps; mel {t_mel * 1e3:.0f} ms, "
      f"encoder {t_enc * 1e3:.0f} ms, decoder {t_dec * 1e3:.0f} ms; total {total:.2f} s = {secs / total:.1f} audio-s/s; "
      f"final position {L.vox_cuda_decoder_position(st)}")
bucket = {}
for pos, k, dt in stamps:
    b = pos // 2048
    a = bucket.setdefault(b, [0, 0.0])
    a[0] += k
    a[1] += dt
for b in sorted(bucket):
    n, dt = bucket[b]
    print(f"  positions {b * 2048:6d}-{b * 2048 + 2047:6d}: {n:5d} steps, {1e3 * dt / n:.3f} ms/step")
if secs > 100:
    short, *_ = run(x[: 16000 * 90])
    same = short[:700] == toks[:700]
    print(f"causality: first 700 ids of the {secs:.0f} s run == ids of a run on its first 90 s: {same}")
    assert same
print("distinct ids:", len(set(toks)), " ids >= 1000:", sum(t >= 1000 for t in toks))
