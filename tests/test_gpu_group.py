"""Stream groups (streams batched per decode step on one GPU): the multi-stream
matrix product against a numpy fp32 product, and lock-step group decoding against the
single-stream path (whose ids tests/test_gpu_golden.py pins to the reference BLAS path on the
same 2 s clip).  All through the C ABI."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def _bf16(a):
    """round-to-nearest-even fp32 -> bf16 bit patterns"""
    u = np.ascontiguousarray(a, np.float32).view(np.uint32)
    return ((u + 0x7FFF + ((u >> 16) & 1)) >> 16).astype(np.uint16)


def _widen(b):
    return (b.astype(np.uint32) << 16).view(np.float32)


@pytest.mark.parametrize("B,N,K,ksplit", [(1, 128, 64, 1), (3, 256, 3072, 3), (32, 384, 4096, 6), (33, 256, 1024, 2),
                                          (64, 3072, 9216, 6), (100, 128, 3072, 1)])
def test_streams_product_matches_fp32(full_env, B, N, K, ksplit):
    """C[b][n] = sum_k A[b][k] W[n][k]: fp32 activations x bf16 weights like vox_linear_nobias_bf16
    (voxtral_kernels.c:215-236); the two-term bf16 split of A leaves ~2^-17 relative error per product."""
    e = full_env
    rng = np.random.default_rng(B * 1000 + N)
    A = rng.standard_normal((B, K)).astype(np.float32)
    Wb = _bf16(0.05 * rng.standard_normal((N, K)))
    out = np.empty((B, N), np.float32)
    assert e.L.vox_cuda_gemm_bf16_streams(B, N, K, e.vc.fptr(A), Wb.ctypes.data_as(e.vc.u16p), ksplit, 0, e.vc.fptr(out)) == 0, e.vc.last_error()
    want = (A.astype(np.float64) @ _widen(Wb).astype(np.float64).T)
    scale = float(np.abs(want).max())
    assert float(np.abs(out - want).max()) < 2e-5 * scale


def test_streams_product_swiglu(full_env):
    e = full_env
    rng = np.random.default_rng(7)
    B, N, K = 5, 512, 3072
    A = rng.standard_normal((B, K)).astype(np.float32)
    Wb = _bf16(0.03 * rng.standard_normal((N, K)))
    out = np.empty((B, N // 2), np.float32)
    assert e.L.vox_cuda_gemm_bf16_streams(B, N, K, e.vc.fptr(A), Wb.ctypes.data_as(e.vc.u16p), 1, 1, e.vc.fptr(out)) == 0, e.vc.last_error()
    gu = A.astype(np.float64) @ _widen(Wb).astype(np.float64).T
    g, u = gu[:, 0::2], gu[:, 1::2]
    want = g / (1.0 + np.exp(-g)) * u                      # voxtral_kernels.c:498-503
    assert float(np.abs(out - want).max()) < 2e-4 * max(1.0, float(np.abs(want).max()))


def _clip(seconds, seed):
    import test_gpu_parity as tp
    import vox_oracle as vo
    x = tp.synth_audio(int(16000 * seconds), seed=seed)
    return vo.stream_buffer(x)


def test_group_tokens_equal_single_stream(full_env):
    """4 slots: three clips of different lengths (ragged finish) + one slot left empty.  Every stream's
    greedy ids must equal the single-stream whole-clip call on the same audio; stream 0 is the 2 s
    clip whose single-stream ids are pinned to the reference in test_gpu_golden.py."""
    e = full_env
    clips = [_clip(2.0, 5), _clip(3.1, 11), _clip(2.4, 12)]
    st = e.L.vox_cuda_state_create(e.model)
    assert st, e.vc.last_error()
    want = []
    try:
        for buf, nf in clips:
            toks = np.empty(nf // 8 + 8, np.int32)
            n = e.L.vox_cuda_transcribe(st, e.vc.fptr(buf), len(buf), nf, 39, e.vc.iptr(toks), len(toks), None)
            assert n > 0, e.vc.last_error()
            want.append(toks[:n].tolist())
    finally:
        e.L.vox_cuda_state_free(st)
    g = e.L.vox_cuda_group_create(e.model, 4, 128)
    assert g, e.vc.last_error()
    try:
        assert e.L.vox_cuda_group_streams(g) == 4
        for i, (buf, nf) in enumerate(clips):
            T = e.L.vox_cuda_group_encode(g, i if i < 2 else 3, e.vc.fptr(buf), len(buf), nf)     # slot 2 stays empty
            assert T == len(want[i]) + 38, (T, e.vc.last_error())
        max_tok = 64
        toks = np.full((4, max_tok), -7, np.int32)
        counts = np.zeros(4, np.int32)
        ms = C.c_float()
        steps = e.L.vox_cuda_group_decode(g, 39, e.vc.iptr(toks), max_tok, e.vc.iptr(counts), C.byref(ms))
        assert steps == max(len(w) for w in want), (steps, e.vc.last_error())
        assert counts.tolist() == [len(want[0]), len(want[1]), 0, len(want[2])]
        for i, slot in enumerate((0, 1, 3)):
            assert toks[slot, :counts[slot]].tolist() == want[i], (slot, toks[slot, :counts[slot]].tolist(), want[i])
        assert np.all(toks[2] == -7)                            # nothing written for the empty slot
        # a second decode of the same group starts from a clean slate (idempotence)
        toks2 = np.full((4, max_tok), -7, np.int32)
        assert e.L.vox_cuda_group_decode(g, 39, e.vc.iptr(toks2), max_tok, e.vc.iptr(counts), None) == steps
        assert np.array_equal(toks, toks2)
        # too long a clip for the group's capacity is refused, not truncated
        buf, nf = _clip(12.0, 3)
        assert e.L.vox_cuda_group_encode(g, 0, e.vc.fptr(buf), len(buf), nf) == -1
    finally:
        e.L.vox_cuda_group_free(g)


def test_group_persistent_kernel_matches(full_env, monkeypatch):
    """VOX_CUDA_GROUP_MEGA=1: the whole decode as one persistent cooperative kernel (grid barriers instead of
    launches, csrc/group_mega.cu) must give the ids of the default kernel-per-op path."""
    e = full_env
    clips = [_clip(2.0, 5), _clip(2.7, 21)]

    def run():
        g = e.L.vox_cuda_group_create(e.model, 3, 128)
        assert g, e.vc.last_error()
        try:
            for i, (buf, nf) in enumerate(clips):
                assert e.L.vox_cuda_group_encode(g, 2 * i, e.vc.fptr(buf), len(buf), nf) > 39       # slots 0 and 2
            toks = np.full((3, 32), -7, np.int32)
            counts = np.zeros(3, np.int32)
            assert e.L.vox_cuda_group_decode(g, 39, e.vc.iptr(toks), 32, e.vc.iptr(counts), None) > 0, e.vc.last_error()
            return toks, counts.copy()
        finally:
            e.L.vox_cuda_group_free(g)

    monkeypatch.delenv("VOX_CUDA_GROUP_MEGA", raising=False)
    t0, c0 = run()
    monkeypatch.setenv("VOX_CUDA_GROUP_MEGA", "1")
    t1, c1 = run()
    assert c0.tolist() == c1.tolist() and c0[1] == 0 and c0[0] > 0
    assert np.array_equal(t0, t1)


def test_group_64_streams_60s_properties(full_env):
    """Full bench size: 64 slots x 60 s.  Too large to diff on the CPU, so: (1) all 64 copies of one
    clip give identical ids (no cross-stream leakage, no dependence on the slot), and they equal the
    single-stream ids; (2) a different clip placed in one slot changes only that slot."""
    import bench
    e = full_env
    a, nfa = bench.frame_buffer(bench.synth_clip(60.0, 1))
    b, nfb = bench.frame_buffer(bench.synth_clip(60.0, 2))
    st = e.L.vox_cuda_state_create(e.model)
    assert st, e.vc.last_error()
    try:
        ta = np.empty(nfa // 8 + 8, np.int32)
        na = e.L.vox_cuda_transcribe(st, e.vc.fptr(a), len(a), nfa, 39, e.vc.iptr(ta), len(ta), None)
        tb = np.empty(nfb // 8 + 8, np.int32)
        nb = e.L.vox_cuda_transcribe(st, e.vc.fptr(b), len(b), nfb, 39, e.vc.iptr(tb), len(tb), None)
        assert na == 761 and nb == 761
    finally:
        e.L.vox_cuda_state_free(st)
    g = e.L.vox_cuda_group_create(e.model, 64, 800)
    assert g, e.vc.last_error()
    try:
        for i in range(64):
            buf, nf = (b, nfb) if i == 37 else (a, nfa)
            assert e.L.vox_cuda_group_encode(g, i, e.vc.fptr(buf), len(buf), nf) == 799, e.vc.last_error()
        toks = np.empty((64, 768), np.int32)
        counts = np.zeros(64, np.int32)
        ms = C.c_float()
        assert e.L.vox_cuda_group_decode(g, 39, e.vc.iptr(toks), 768, e.vc.iptr(counts), C.byref(ms)) == 761, e.vc.last_error()
        print(f"64 x 60 s group decode: {ms.value:.1f} ms on the device = {ms.value / 761:.3f} ms per lock step")
        assert np.all(counts == 761)
        for i in range(64):
            ref = tb[:761] if i == 37 else ta[:761]
            assert np.array_equal(toks[i, :761], ref), i
    finally:
        e.L.vox_cuda_group_free(g)


def test_group_fp16_kv_ids_equal_fp32_kv(full_env):
    """K/V rings in fp16 (the reference's GPU default, voxtral.c:189-190).  Not bit-comparable
    with the fp32 BLAS oracle, so the statement is on ids: rows are rounded once (relative 2^-11), scores /
    softmax / output stay fp32, and on the test clips (top-2 logit margins ~1e-1 on this checkpoint) the
    greedy ids equal those of the fp32 rings - which test_group_tokens_equal_single_stream ties to the
    single-stream path and test_gpu_golden.py to the reference.  Ragged lengths + an empty slot as there."""
    e = full_env
    clips = [_clip(2.0, 5), _clip(3.1, 11), _clip(2.4, 12)]
    assert not e.L.vox_cuda_group_create_ex(e.model, 4, 128, 8)           # only 32 and 16 exist

    def run(bits):
        g = e.L.vox_cuda_group_create_ex(e.model, 4, 128, bits)
        assert g, e.vc.last_error()
        try:
            for i, (buf, nf) in enumerate(clips):
                assert e.L.vox_cuda_group_encode(g, i if i < 2 else 3, e.vc.fptr(buf), len(buf), nf) > 39
            toks = np.full((4, 64), -7, np.int32)
            counts = np.zeros(4, np.int32)
            assert e.L.vox_cuda_group_decode(g, 39, e.vc.iptr(toks), 64, e.vc.iptr(counts), None) > 0, e.vc.last_error()
            toks2 = np.full((4, 64), -7, np.int32)                        # second decode: clean slate
            assert e.L.vox_cuda_group_decode(g, 39, e.vc.iptr(toks2), 64, e.vc.iptr(counts), None) > 0
            assert np.array_equal(toks, toks2)
            return toks, counts.copy()
        finally:
            e.L.vox_cuda_group_free(g)

    t32, c32 = run(32)
    t16, c16 = run(16)
    assert c32.tolist() == c16.tolist() and c32[2] == 0 and c32[0] > 0
    assert np.array_equal(t32, t16)
