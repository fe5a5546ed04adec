#!/usr/bin/env python3
"""Kernel list of ONE streaming encoder call (BASELINE config 3, -I 0.5: 50 mel frames = 25 positions against a full 750-key
window).  Run under `ncu --metrics gpu__time_duration.sum --csv` with VOX_CUDA_ENC_GRAPH=0 (kernel launches instead of the
captured graph) and summarise with tools/summarize_launches.py; the script prints the launch counter around the last call so
that the tail of the list can be cut out.  usage: stream_launches.py [frames_per_call]"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "voxtral.c_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import refharness as rh  # noqa: E402
import voxcuda as vc  # noqa: E402
import hostmodel  # noqa: E402

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 50
L = vc.lib()
assert L.vox_cuda_init() == 1, vc.last_error()
hm = hostmodel.HostModel(rh.model_dir())
model = vc.check(L.vox_cuda_model_create(C.byref(hm.weights)), "model")
st = vc.check(L.vox_cuda_state_create(model), "state")
rng = np.random.default_rng(0)
mel = (0.3 * rng.standard_normal((4000, 128))).astype(np.float32)
vc.check(L.vox_cuda_encode_chunk(st, vc.fptr(np.ascontiguousarray(mel[:1912])), 1912), "encode")      # 956 positions: window full
off = 1912
for i in range(4):
    n0 = L.vox_cuda_launch_count()
    vc.check(L.vox_cuda_encode_chunk(st, vc.fptr(np.ascontiguousarray(mel[off:off + frames])), frames), "encode")
    vc.check(L.vox_cuda_adapter_discard_before(st, L.vox_cuda_adapter_count(st)), "discard")
    off += frames
    print(f"call {i}: launches {n0} .. {L.vox_cuda_launch_count()}", flush=True)
