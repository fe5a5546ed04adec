#!/usr/bin/env python3
"""BASELINE config 5 on one GPU: n_streams independent clips through a stream group
(vox_cuda_group_*: per-stream encoder passes, then lock-step decoding with one pass over the
weights per step).  Reports encoder and decoder time, ms per lock step, aggregate tok/s and
audio-s/s, and the same clips through the single-stream path for comparison.
usage: group_probe.py [n_streams] [seconds] [--no-single]"""
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

args = [a for a in sys.argv[1:] if not a.startswith("--")]
B = int(args[0]) if len(args) > 0 else 64
secs = float(args[1]) if len(args) > 1 else 60.0
single = "--no-single" not in sys.argv

L = vc.lib()
assert L.vox_cuda_init() == 1, vc.last_error()
hm = hostmodel.HostModel(rh.model_dir())
model = vc.check(L.vox_cuda_model_create(C.byref(hm.weights)), "model")
clips = [bench.frame_buffer(bench.synth_clip(secs, 100 + i)) for i in range(min(B, 4))]     # 4 distinct clips, cycled
T = clips[0][1] // 8
g = vc.check(L.vox_cuda_group_create(model, B, T + 1), "group")
toks = np.empty((B, T), np.int32)
counts = np.zeros(B, np.int32)
ms = C.c_float()
for rep in range(2):
    t0 = time.perf_counter()
    for i in range(B):
        buf, nf = clips[i % len(clips)]
        vc.check(L.vox_cuda_group_encode(g, i, vc.fptr(buf), len(buf), nf), "encode")
    t1 = time.perf_counter()
    steps = vc.check(L.vox_cuda_group_decode(g, 39, vc.iptr(toks), T, vc.iptr(counts), C.byref(ms)), "decode")
    t2 = time.perf_counter()
    enc_s, dec_s = t1 - t0, t2 - t1
    n_tok = int(counts.sum())
    print(f"rep {rep}: {B} streams x {secs:.0f} s: encode {enc_s * 1e3:.0f} ms ({enc_s * 1e3 / B:.1f} ms/clip), "
          f"decode {dec_s * 1e3:.0f} ms wall / {ms.value:.0f} ms device incl. {B} prefills = {ms.value / steps:.3f} ms per lock step, "
          f"{n_tok / dec_s:.0f} tok/s; total {B * secs / (enc_s + dec_s):.1f} audio-s/s")
L.vox_cuda_group_free(g)
if single:
    st = vc.check(L.vox_cuda_state_create(model), "state")
    buf, nf = clips[0]
    t1 = np.empty(T + 8, np.int32)
    L.vox_cuda_transcribe(st, vc.fptr(buf), len(buf), nf, 39, vc.iptr(t1), len(t1), None)
    t0 = time.perf_counter()
    n = L.vox_cuda_transcribe(st, vc.fptr(buf), len(buf), nf, 39, vc.iptr(t1), len(t1), None)
    dt = time.perf_counter() - t0
    print(f"single stream: {dt * 1e3:.0f} ms per clip = {secs / dt:.1f} audio-s/s; group/single = "
          f"{(B * secs / (enc_s + dec_s)) / (secs / dt):.1f}x; ids equal: {np.array_equal(t1[:n], toks[0, :n])}")
    L.vox_cuda_state_free(st)
