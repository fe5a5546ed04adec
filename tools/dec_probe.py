#!/usr/bin/env python3
"""Decoder-step probe (GPU): ms/step from CUDA events plus the megakernel's
cycle breakdown (VOX_CUDA_MEGA_PROF=1).  usage: dec_probe.py [steps] [start_pos] [q8]   (q8: the Q8 container)"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "voxtral.c_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ.setdefault("VOX_CUDA_MEGA_PROF", "1")
import refharness as rh  # noqa: E402
import voxcuda as vc  # noqa: E402
import hostmodel  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
start = int(sys.argv[2]) if len(sys.argv) > 2 else 38
L = vc.lib()
assert L.vox_cuda_init() == 1, vc.last_error()
Q8 = len(sys.argv) > 3 and sys.argv[3] == "q8"
hm = hostmodel.HostModel(rh.model_dir(extra=("--q8",)) if Q8 else rh.model_dir())
model = vc.check(hm.create_model(L), "model")
WEIGHT_BYTES = 6857687040 // 2 + 4 * 2 * (26 * (6144 + 3072 + 18432 + 3072) + 131072) if Q8 else 6857687040   # int8 + scale + row sum
st = vc.check(L.vox_cuda_state_create(model), "state")
rng = np.random.default_rng(0)
ad = (0.5 * rng.standard_normal((steps, 3072))).astype(np.float32)
vc.check(L.vox_cuda_adapter_append(st, vc.fptr(ad), steps), "append")
L.vox_cuda_debug_mega_profile.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
for rep in range(3):
    vc.check(L.vox_cuda_decoder_set_position(st, start, min(start, 8192)), "setpos")
    toks = np.empty(steps, np.int32)
    ms = L.vox_cuda_bench_decoder_steps(st, 32, 0, steps, vc.iptr(toks))
    prof = (C.c_double * 8)()
    line = f"rep {rep}: {ms:.4f} ms/step  ({WEIGHT_BYTES / ms / 1e6:.0f} GB/s)  uniq tokens {len(set(toks.tolist()))}"
    if L.vox_cuda_debug_mega_profile(st, prof) == 0 and prof[0] > 0:
        t = prof[0]
        line += (f" | cycles/step {t / steps:.0f}: wait-tiles {100 * prof[1] / t:.1f}% grid-sync {100 * prof[2] / t:.1f}% "
                 f"(spin {100 * prof[5] / t:.1f}% atom-rtt {100 * prof[6] / t:.1f}%) attention {100 * prof[3] / t:.1f}% prologue {100 * prof[4] / t:.1f}% (Q8 digit prep {100 * prof[7] / t:.1f}%) "
                 f"compute+other {100 * (t - prof[1] - prof[2] - prof[3] - prof[4]) / t:.1f}%")
    print(line, flush=True)

# barrier skew: for barriers 0..11 from layer 10 of the last step (after wo, w13, w2, wo, ...)
L.vox_cuda_debug_mega_barriers.argtypes = [C.c_void_p, C.POINTER(C.c_ulonglong), C.POINTER(C.c_int)]
buf = (C.c_ulonglong * (1024 * 24))()
g = C.c_int()
if L.vox_cuda_debug_mega_barriers(st, buf, C.byref(g)) == 0:
    a = np.array(buf[:g.value * 24], dtype=np.int64).reshape(g.value, 12, 2)
    names = ["wo", "w13", "w2"]     # grid barriers of a layer (qkv -> attention -> wo hand over through counters)
    t0 = a[:, 0, 0].min()
    if os.environ.get("DEC_PROBE_DUMP"):
        np.save(os.environ["DEC_PROBE_DUMP"], a)          # [cta, barrier, (arrive, exit)] ns, for offline analysis
    for k in range(12):
        arr, ex = a[:, k, 0], a[:, k, 1]
        print(f"barrier {k:2d} after {names[k % 3]:4s}: first arrive {arr.min() - t0:7d} ns, last arrive {arr.max() - t0:7d} "
              f"(skew {arr.max() - arr.min():6d}), first exit +{ex.min() - arr.max():5d}, last exit +{ex.max() - arr.max():5d}")
