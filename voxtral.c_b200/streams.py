"""Multi-GPU work split for independent audio streams (BASELINE config 5): one process per
GPU, the full weight set replicated on each, streams dealt to ranks in contiguous balanced
ranges, no collective on the data path.  The only communication is the timing protocol of
bench.py: a barrier on both sides of the timed region, MAX over ranks of the elapsed time and
SUM over ranks of the audio seconds processed.

The reference has no multi-device mode (one vox_ctx_t per process, voxtral.h:188-239); this
mirrors how a server would run N reference processes, one per device."""


def shard_streams(n_streams, rank, world):
    """[begin, end) of the streams rank owns: sizes differ by at most one, earlier ranks get
    the longer shards, empty shards only when n_streams < world."""
    if world <= 0 or not (0 <= rank < world) or n_streams < 0:
        raise ValueError("bad shard arguments")
    base, extra = divmod(n_streams, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def aggregate(dist, device, audio_seconds_local, elapsed_local):
    """(total audio seconds over all ranks, max elapsed over ranks); dist may be None (1 GPU)."""
    if dist is None:
        return float(audio_seconds_local), float(elapsed_local)
    import torch
    t = torch.tensor([audio_seconds_local], device=device, dtype=torch.float64)
    m = torch.tensor([elapsed_local], device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    dist.all_reduce(m, op=dist.ReduceOp.MAX)
    return float(t.item()), float(m.item())
