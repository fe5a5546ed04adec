"""Host-side plumbing of the tensor-parallel decoder (include/vox_cuda.h, vox_cuda_tp_*): one process per
GPU creates its rank, the 80-byte window handles are all-gathered over torch.distributed (the transport is
free: a file or a socket would do for the C host), then every rank maps its peers' windows."""
import ctypes as C

import numpy as np

import voxcuda as vc


def shard_ranges(rank, world):
    """The rows / columns rank `rank` of `world` owns; mirrors csrc/dec_tp.cu and is what the
    CPU tests check: q / kv head ranges, w1|w3 pair range, wo / w2 column ranges, vocabulary range."""
    assert world in (1, 2, 4, 8) and 0 <= rank < world
    qh, kh = 32 // world, 8 // world
    return {
        "q_heads": (rank * qh, (rank + 1) * qh),
        "kv_heads": (rank * kh, (rank + 1) * kh),
        "ffn_pairs": (rank * 9216 // world, (rank + 1) * 9216 // world),
        "wo_cols": (rank * 4096 // world, (rank + 1) * 4096 // world),
        "w2_cols": (rank * 9216 // world, (rank + 1) * 9216 // world),
        "vocab": (rank * 131072 // world, (rank + 1) * 131072 // world),
    }


def gather_handles(mine, dist, world):
    """all-gather of the per-rank handle bytes; works on the gloo (CPU) and nccl (CUDA) backends"""
    import torch
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.frombuffer(bytearray(mine), dtype=torch.uint8).to(dev)
    outs = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(outs, t)
    return b"".join(bytes(o.cpu().numpy().tobytes()) for o in outs)


def create_attached(L, model, rank, world, dist, barrier=True):
    """Create this process's rank and map every peer's window.  Collective: all ranks call it together (the
    handle all-gather happens even on a rank whose export failed, so nobody is left waiting)."""
    tp = vc.check(L.vox_cuda_tp_create(model, rank, world), "tp_create")
    if world > 1:
        h = (C.c_ubyte * vc.TP_HANDLE_BYTES)()
        rc = L.vox_cuda_tp_export(tp, h)
        allh = gather_handles(bytes(h), dist, world)
        assert len(allh) == world * vc.TP_HANDLE_BYTES
        buf = (C.c_ubyte * len(allh)).from_buffer_copy(allh)
        if rc != 0 or L.vox_cuda_tp_connect(tp, buf) != 0:
            err = vc.last_error()
            L.vox_cuda_tp_free(tp)
            raise RuntimeError("tp attach failed: " + err)
        if barrier:
            dist.barrier()
    return tp
