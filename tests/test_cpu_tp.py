"""Host logic of the tensor-parallel decoder on CPU: the shard table (voxtral.c_b200/tp.py, mirrored by
csrc/dec_tp.cu) partitions every matrix exactly once; the sharded arithmetic (per-rank heads, column-range
partial sums added in rank order, first-maximum merge of vocabulary slices) reproduces the oracle's
unsharded step; the 80-byte handle all-gather works over gloo with world_size 2."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "voxtral.c_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import tp  # noqa: E402


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_shards_partition_every_matrix_once(world):
    for key, total in (("q_heads", 32), ("kv_heads", 8), ("ffn_pairs", 9216), ("wo_cols", 4096),
                       ("w2_cols", 9216), ("vocab", 131072)):
        got = []
        for r in range(world):
            a, b = tp.shard_ranges(r, world)[key]
            got += list(range(a, b))
        assert got == list(range(total)), key
    for r in range(world):
        s = tp.shard_ranges(r, world)
        # a rank's query heads are exactly those of its kv heads (GQA groups of 4 stay together) and its
        # wo columns are those heads' outputs; its w2 columns are its own gate/up pairs
        assert s["q_heads"] == (4 * s["kv_heads"][0], 4 * s["kv_heads"][1])
        assert s["wo_cols"] == (128 * s["q_heads"][0], 128 * s["q_heads"][1])
        assert s["w2_cols"] == s["ffn_pairs"]
    with pytest.raises(AssertionError):
        tp.shard_ranges(0, 3)


@pytest.mark.parametrize("world", [2, 8])
def test_sharded_step_arithmetic_matches_unsharded(world):
    """attention -> wo -> FFN -> w2 -> logits on random tensors of the real head geometry (small
    vocabulary), once whole (oracle functions) and once rank by rank."""
    import vox_oracle as vo
    rng = np.random.default_rng(world)
    F = np.float32
    n_keys, hidden, vocab = 9, 9216, 4096
    q = rng.standard_normal((1, 4096)).astype(F)
    K = rng.standard_normal((n_keys, 1024)).astype(F)
    V = rng.standard_normal((n_keys, 1024)).astype(F)
    wo = (0.02 * rng.standard_normal((3072, 4096))).astype(F)
    w1 = (0.02 * rng.standard_normal((hidden, 3072))).astype(F)
    w3 = (0.02 * rng.standard_normal((hidden, 3072))).astype(F)
    w2 = (0.02 * rng.standard_normal((3072, hidden))).astype(F)
    emb = (0.05 * rng.standard_normal((vocab, 3072))).astype(F)
    x = rng.standard_normal(3072).astype(F)
    nw = (1 + 0.1 * rng.standard_normal(3072)).astype(F)

    att = vo.causal_attention(q, K, V, 32, 8, 128, 8192, n_keys - 1)[0]
    x1 = x + wo @ att
    xn = vo.rms_norm(x1[None], nw)[0]
    h = vo.silu(w1 @ xn) * (w3 @ xn)
    x2 = x1 + w2 @ h
    logits = emb @ vo.rms_norm(x2[None], nw)[0]

    parts_wo, parts_w2, att_s, h_s = [], [], np.zeros(4096, F), np.zeros(hidden, F)
    for r in range(world):
        s = tp.shard_ranges(r, world)
        (q0, q1), (k0, k1) = s["q_heads"], s["kv_heads"]
        a = vo.causal_attention(q[:, 128 * q0:128 * q1], K[:, 128 * k0:128 * k1], V[:, 128 * k0:128 * k1],
                                q1 - q0, k1 - k0, 128, 8192, n_keys - 1)[0]
        att_s[128 * q0:128 * q1] = a
        c0, c1 = s["wo_cols"]
        parts_wo.append(wo[:, c0:c1] @ a)
    assert np.array_equal(att_s, att)                       # heads are independent: bit-identical
    x1_s = x.copy()
    for p in parts_wo:                                      # rank order, as the consumer prologue adds them
        x1_s = x1_s + p
    assert float(np.abs(x1_s - x1).max()) < 1e-4
    xn_s = vo.rms_norm(x1_s[None], nw)[0]
    for r in range(world):
        a0, a1 = tp.shard_ranges(r, world)["ffn_pairs"]
        h_s[a0:a1] = vo.silu(w1[a0:a1] @ xn_s) * (w3[a0:a1] @ xn_s)
        parts_w2.append(w2[:, a0:a1] @ h_s[a0:a1])
    x2_s = x1_s.copy()
    for p in parts_w2:
        x2_s = x2_s + p
    assert float(np.abs(x2_s - x2).max()) < 1e-4
    xf = vo.rms_norm(x2_s[None], nw)[0]
    best = (-np.inf, 2 ** 31 - 1)
    for r in range(world):
        v0, v1 = r * vocab // world, (r + 1) * vocab // world
        sl = emb[v0:v1] @ xf
        i = int(np.argmax(sl))                              # first maximum of the slice
        if sl[i] > best[0] or (sl[i] == best[0] and v0 + i < best[1]):
            best = (sl[i], v0 + i)
    assert best[1] == int(np.argmax(logits))


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = bytes([rank + 1]) * 80
    allh = tp.gather_handles(mine, dist, world)
    dist.barrier()
    q.put((rank, allh))
    dist.destroy_process_group()


def test_gloo_world2_handle_exchange():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for _, allh in res:
        assert allh == bytes([1]) * 80 + bytes([2]) * 80     # rank order, identical on both ranks
