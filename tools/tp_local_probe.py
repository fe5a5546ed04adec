#!/usr/bin/env python3
"""Lock-step tensor-parallel decode on ONE GPU (vox_cuda_tp_create_local): ms per token for G = 1/2/4/8 with
every rank's kernels serialised on one stream.  Under `ncu --metrics gpu__time_duration.sum` the per-kernel
durations are those of a rank's shard without the NVLink hop (profiles/r01_j_tp_decode.md)."""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "voxtral.c_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=10.0)
    ap.add_argument("--worlds", default="1,2,4,8")
    ap.add_argument("--max-steps", type=int, default=100000)
    args = ap.parse_args()
    import bench
    import hostmodel
    import refharness as rh
    import voxcuda as vc
    L = vc.lib()
    assert L.vox_cuda_init() == 1, vc.last_error()
    hm = hostmodel.HostModel(rh.model_dir())
    model = vc.check(L.vox_cuda_model_create(C.byref(hm.weights)), "model_create")
    buf, nf = bench.frame_buffer(bench.synth_clip(args.seconds, 1))
    toks = np.empty(nf // 8 + 8, np.int32)
    out = {}
    for w in [int(x) for x in args.worlds.split(",")]:
        tp = vc.check(L.vox_cuda_tp_create_local(model, w), "tp_create_local")
        ms = C.c_float()
        for _ in range(2):
            n = vc.check(L.vox_cuda_tp_transcribe(tp, vc.fptr(buf), len(buf), nf, 39, vc.iptr(toks),
                                                  min(len(toks), args.max_steps), C.byref(ms)), "tp_transcribe")
        out[w] = {"steps": int(n), "ms_per_token_all_ranks_serialised": float(ms.value)}
        L.vox_cuda_tp_free(tp)
    print(json.dumps(out))
    L.vox_cuda_model_free(model)


if __name__ == "__main__":
    main()
