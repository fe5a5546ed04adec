#!/usr/bin/env python3
"""Encoder probe (GPU): time vox_cuda_encode_chunk on a 60 s clip's mel in the
precise (fp32 CUDA-core) and fast (tcgen05) modes (one process per mode, the mode
is read once) and dump the adapter tokens for comparison.
usage: enc_probe.py <out.npy> [seconds]"""
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

secs = float(sys.argv[2]) if len(sys.argv) > 2 else 60.0
L = vc.lib()
assert L.vox_cuda_init() == 1, vc.last_error()
hm = hostmodel.HostModel(rh.model_dir())
model = vc.check(L.vox_cuda_model_create(C.byref(hm.weights)), "model")
st = vc.check(L.vox_cuda_state_create(model), "state")
x = bench.synth_clip(secs, 1)
buf, nf = bench.frame_buffer(x)
mel = np.empty((nf, 128), np.float32)
vc.check(L.vox_cuda_mel_frames(vc.fptr(buf), nf, vc.fptr(mel)), "mel")
for rep in range(3):
    L.vox_cuda_state_reset_encoder(st)
    t0 = time.perf_counter()
    T = vc.check(L.vox_cuda_encode_chunk(st, vc.fptr(mel), nf), "encode")
    dt = time.perf_counter() - t0
    print(f"mode precise={os.environ.get('VOX_CUDA_ENC_PRECISE', '0')} rep {rep}: {nf} mel -> {T} tokens in {dt * 1e3:.1f} ms "
          f"({nf // 2} positions, {1.929e9 * (nf // 2) / dt / 1e12:.1f} TFLOP/s linear)", flush=True)
ad = np.empty((T, 3072), np.float32)
vc.check(L.vox_cuda_adapter_read(st, 0, T, vc.fptr(ad)), "read")
np.save(sys.argv[1], ad)
