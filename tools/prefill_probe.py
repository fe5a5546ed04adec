#!/usr/bin/env python3
"""Prompt prefill probe (GPU): wall-clock ms of vox_cuda_decoder_prefill(38 rows) - the call returns after its
scratch is freed, i.e. synchronised.  usage: prefill_probe.py [rows]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "voxtral.c_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import refharness as rh  # noqa: E402
import voxcuda as vc  # noqa: E402
import hostmodel  # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 38
L = vc.lib()
assert L.vox_cuda_init() == 1, vc.last_error()
hm = hostmodel.HostModel(rh.model_dir())
model = vc.check(hm.create_model(L), "model")
st = vc.check(L.vox_cuda_state_create(model), "state")
emb = (0.5 * np.random.default_rng(0).standard_normal((rows, 3072))).astype(np.float32)
ts = []
for rep in range(8):
    L.vox_cuda_state_reset_decoder(st)
    t0 = time.perf_counter()
    vc.check(L.vox_cuda_decoder_prefill(st, vc.fptr(emb), rows), "prefill")
    ts.append(1e3 * (time.perf_counter() - t0))
print(f"prefill of {rows} rows: {['%.2f' % t for t in ts]} ms  (6.05 GB of decoder weights once = {6.05e3 / 6550:.2f} ms at the HBM peak)")
