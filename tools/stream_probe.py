#!/usr/bin/env python3
"""BASELINE config 3 (GPU): streaming with -I 0.5 and -I 2.0.

Feeds the mel of a 60 s clip to vox_cuda_encode_chunk the way stream_run_encoder
(/root/reference/voxtral.c:826-934) is driven by vox_stream_feed: a first chunk of >= 312 mel
frames, then fixed chunks of interval * 100 frames (50 for -I 0.5, 200 for -I 2.0), and after
every encoder call the decoder catches up on the new adapter tokens (vox_cuda_decoder_run).
Reports the per-chunk encoder latency distribution, the fixed cost per encoder call (intercept
of a latency-vs-rows fit over extra chunk sizes), the decoder catch-up time, and checks that the
streamed adapter tokens / token ids equal the one-pass run.
usage: stream_probe.py [seconds]"""
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

secs = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
L = vc.lib()
assert L.vox_cuda_init() == 1, vc.last_error()
hm = hostmodel.HostModel(rh.model_dir())
model = vc.check(L.vox_cuda_model_create(C.byref(hm.weights)), "model")
st = vc.check(L.vox_cuda_state_create(model), "state")
x = bench.synth_clip(secs, 1)
buf, nf = bench.frame_buffer(x)
mel = np.empty((nf, 128), np.float32)
vc.check(L.vox_cuda_mel_frames(vc.fptr(buf), nf, vc.fptr(mel)), "mel")

# one-pass reference run of this backend (tokens + adapter rows)
max_tok = nf // 8 + 8
toks_ref = np.empty(max_tok, np.int32)
n_ref = vc.check(L.vox_cuda_transcribe(st, vc.fptr(buf), len(buf), nf, 39, vc.iptr(toks_ref), max_tok, None), "transcribe")
T = L.vox_cuda_adapter_count(st)
ad_ref = np.empty((T, 3072), np.float32)
vc.check(L.vox_cuda_adapter_read(st, 0, T, vc.fptr(ad_ref)), "read")


def stream(interval_s):
    step = int(round(interval_s * 100))
    L.vox_cuda_state_reset_encoder(st)
    L.vox_cuda_state_reset_decoder(st)
    lat, rows, dec_ms = [], [], []
    toks = []
    prompt = np.array([1] + [32] * 38, np.int32)
    prefilled = False
    next_pos, prev = 38, 32
    off = 0
    while off < nf:
        n = min(max(step, 312) if off == 0 else step, nf - off)
        chunk = np.ascontiguousarray(mel[off:off + n])
        t0 = time.perf_counter()
        total = vc.check(L.vox_cuda_encode_chunk(st, vc.fptr(chunk), n), "encode")
        lat.append((time.perf_counter() - t0) * 1e3)
        rows.append(n)
        off += n
        # decoder catches up (voxtral.c:1012-1145): prefill once 39 tokens exist, then one step per new token
        t0 = time.perf_counter()
        if not prefilled and total >= 39:
            vc.check(L.vox_cuda_decoder_prefill_fused(st, vc.iptr(prompt), 0, 38), "prefill")
            prefilled = True
        if prefilled and total > next_pos:
            k = total - next_pos
            out = np.empty(k, np.int32)
            vc.check(L.vox_cuda_decoder_run(st, prev, next_pos, k, vc.iptr(out)), "run")
            toks += out.tolist()
            prev, next_pos = int(out[-1]), total
        dec_ms.append((time.perf_counter() - t0) * 1e3)
    ad = np.empty((T, 3072), np.float32)
    got = L.vox_cuda_adapter_count(st)
    vc.check(L.vox_cuda_adapter_read(st, 0, got, vc.fptr(ad[:got])), "read")
    return np.array(lat), np.array(rows), np.array(dec_ms), toks, ad[:got]


print(f"clip {secs:.0f} s: {nf} mel frames, {T} adapter tokens, {n_ref} decode steps (one-pass)")
for interval in (0.5, 2.0):
    stream(interval)                                     # warm-up
    lat, rows, dec_ms, toks, ad = stream(interval)
    steady = lat[1:-1] if len(lat) > 2 else lat
    d_ad = float(np.abs(ad - ad_ref[:ad.shape[0]]).max())
    same = toks == toks_ref[:n_ref].tolist()
    print(f"-I {interval}: {len(lat)} encoder calls; first chunk ({rows[0]} mel) {lat[0]:.2f} ms; steady chunks ({rows[1]} mel): "
          f"median {np.median(steady):.2f} ms  p90 {np.percentile(steady, 90):.2f}  max {steady.max():.2f}; "
          f"decoder catch-up per chunk median {np.median(dec_ms[1:]):.2f} ms; "
          f"total {lat.sum() + dec_ms.sum():.0f} ms for {secs:.0f} s of audio (RTF {(lat.sum() + dec_ms.sum()) / 1e3 / secs:.4f}); "
          f"adapter max|d| vs one-pass {d_ad:.2e}; tokens identical to one-pass: {same} ({len(toks)}/{n_ref})")

# fixed cost per encoder call: latency vs chunk rows, steady-state cache (750-row window full)
print("latency vs chunk size (window full):")
pts = []
for n in (8, 16, 24, 50, 100, 200, 400, 800):
    L.vox_cuda_state_reset_encoder(st)
    vc.check(L.vox_cuda_encode_chunk(st, vc.fptr(np.ascontiguousarray(mel[:2048])), 2048), "fill")
    ts = []
    off = 2048
    for _ in range(6):
        chunk = np.ascontiguousarray(mel[off:off + n])
        t0 = time.perf_counter()
        vc.check(L.vox_cuda_encode_chunk(st, vc.fptr(chunk), n), "encode")
        ts.append((time.perf_counter() - t0) * 1e3)
        off += n
    pts.append((n, float(np.median(ts[1:]))))
    print(f"  {n:4d} mel ({n // 2:3d} positions): {pts[-1][1]:.3f} ms")
a = np.array(pts)
slope, icpt = np.polyfit(a[:4, 0] / 2, a[:4, 1], 1)
print(f"fixed cost per encoder call (intercept over the 4 smallest sizes): {icpt:.3f} ms; {slope * 1e3:.1f} us per extra position "
      f"(weights are re-read per call: 1.93 GB / 6.5 TB/s = 0.30 ms floor)")
