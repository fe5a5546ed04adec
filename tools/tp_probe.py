#!/usr/bin/env python3
"""Tensor-parallel decoder probe: run under torchrun, one process per GPU.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/tp_probe.py

Every rank encodes + prefills the same clip, the greedy steps run tensor-parallel (csrc/dec_tp.cu); rank 0
also runs the single-stream path on the clip and all ranks compare ids.  Prints ms per token (max over ranks)."""
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
    ap.add_argument("--seconds", type=float, default=60.0)
    ap.add_argument("--model", default=None)
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    os.environ.setdefault("VOX_CUDA_DEVICE", os.environ.get("LOCAL_RANK", "0"))
    import torch
    import torch.distributed as dist
    import bench
    import hostmodel
    import refharness as rh
    import tp as tpmod
    import voxcuda as vc
    dist.init_process_group("gloo")
    L = vc.lib()
    assert L.vox_cuda_init() == 1, vc.last_error()
    if args.model is None:
        if rank == 0:
            rh.model_dir()
        dist.barrier()
    mdir = args.model or rh.model_dir()
    hm = hostmodel.HostModel(mdir)
    model = vc.check(L.vox_cuda_model_create(C.byref(hm.weights)), "model_create")
    buf, nf = bench.frame_buffer(bench.synth_clip(args.seconds, 1))
    toks = np.empty(nf // 8 + 8, np.int32)
    st = vc.check(L.vox_cuda_state_create(model), "state_create")
    n1 = vc.check(L.vox_cuda_transcribe(st, vc.fptr(buf), len(buf), nf, 39, vc.iptr(toks), len(toks), None), "transcribe")
    single = toks[:n1].copy()
    vc.check(L.vox_cuda_decoder_set_position(st, 38, 38), "set_position")
    ms_single = L.vox_cuda_bench_decoder_steps(st, 32, 38, n1, None)
    L.vox_cuda_state_free(st)
    tp = tpmod.create_attached(L, model, rank, world, dist)
    ms = C.c_float()
    best = 1e9
    equal = True
    for _ in range(args.reps):
        dist.barrier()
        n = vc.check(L.vox_cuda_tp_transcribe(tp, vc.fptr(buf), len(buf), nf, 39, vc.iptr(toks), len(toks), C.byref(ms)), "tp_transcribe")
        equal = equal and n == n1 and bool(np.array_equal(toks[:n], single))
        t = torch.tensor([float(ms.value)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = min(best, float(t.item()))
    eq = torch.tensor([1 if equal else 0])
    dist.all_reduce(eq, op=dist.ReduceOp.MIN)
    dist.barrier()
    L.vox_cuda_tp_free(tp)
    L.vox_cuda_model_free(model)
    if rank == 0:
        print(json.dumps({"tp_gpus": world, "clip_s": args.seconds, "steps": int(n1), "ms_per_token_tp": best,
                          "tok_per_s_tp": 1e3 / best, "ms_per_token_single_gpu": ms_single,
                          "kernel_launches_per_token_per_rank": 1 + 5 * hm.n_dec + 3}))
        print(f"ids_equal_single_stream={bool(eq.item())}", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if eq.item() else 1)


if __name__ == "__main__":
    main()
