"""vox_cuda_alternatives_noalloc (csrc/noalloc.cu): the entry point that allocates nothing - its scratch is a
file-scope static __device__ array - against the existing vox_cuda_decoder_alternatives on the same logits rows,
through the C ABI, on several calls in a row (the array is shared between calls and keeps the last call's contents).
Collected last on purpose; passes alone in a fresh process and after the others."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_LOGITS = 262144            # VOX_CUDA_NOALLOC_MAX_LOGITS, include/vox_cuda.h


def test_alternatives_noalloc_matches_decoder_alternatives(full_env):
    e = full_env
    L, vc = e.L, e.vc
    rng = np.random.default_rng(23)
    st = L.vox_cuda_state_create(e.model)
    assert st, vc.last_error()
    # inputs and outputs of the new entry point must be device-accessible: pinned host memory from the library itself
    p_log = L.vox_cuda_host_alloc(131072 * 4)
    p_out = L.vox_cuda_host_alloc(64)
    assert p_log and p_out, vc.last_error()
    logits_pin = np.ctypeslib.as_array(C.cast(p_log, vc.f32p), shape=(131072,))
    out_pin = np.ctypeslib.as_array(C.cast(p_out, vc.i32p), shape=(16,))
    try:
        emb = (0.5 * rng.standard_normal((38, 3072))).astype(np.float32)
        assert L.vox_cuda_decoder_prefill(st, vc.fptr(emb), 38) == 0, vc.last_error()
        logits = np.empty(131072, np.float32)
        ids = np.empty(8, np.int32)
        compared = 0
        for step in range(2):                                   # two decoder steps -> two different logits rows
            x = (0.5 * rng.standard_normal(3072)).astype(np.float32)
            tok = L.vox_cuda_decoder_step(st, vc.fptr(x), vc.fptr(logits))
            assert tok >= 0, vc.last_error()
            logits_pin[:] = logits
            order = np.argsort(-logits[1000:], kind="stable")[:5] + 1000
            cand = [int(i) for i in order if int(i) != tok][:3]
            ratios = 1.0 - np.exp(logits[cand] - logits[tok])
            for n_alt, cutoff in ((4, 1.0), (4, 0.5), (3, 0.2), (2, 0.05), (1, 1.0), (4, 0.0), (4, 0.9)):
                n = L.vox_cuda_decoder_alternatives(st, n_alt, C.c_float(cutoff), vc.iptr(ids))
                assert n >= 0, vc.last_error()
                out_pin[:] = -7
                rc = L.vox_cuda_alternatives_noalloc(p_log, 131072, tok, n_alt, C.c_float(cutoff), p_out, p_out + 32, None, 1)
                assert rc == 0, vc.last_error()
                got_n = int(out_pin[8])
                # ids are integers: the tolerance is exact equality, except where a candidate's ratio sits on the cutoff
                # itself (device expf and libm expf may differ in the last bit there)
                if np.any(np.abs(ratios - cutoff) < 1e-5):
                    continue
                assert got_n == n and out_pin[:got_n].tolist() == ids[:n].tolist(), (step, n_alt, cutoff, out_pin[:8].tolist(), ids[:n].tolist())
                compared += 1
        assert compared >= 10
        # a short row, twice in a row with different contents (stale candidates of the longer calls stay in the array)
        for seed in (1, 2):
            row = np.random.default_rng(seed).standard_normal(5000).astype(np.float32)
            logits_pin[:5000] = row
            best = int(np.argmax(row))
            rc = L.vox_cuda_alternatives_noalloc(p_log, 5000, best, 4, C.c_float(1.0), p_out, p_out + 32, None, 1)
            assert rc == 0, vc.last_error()
            want = [int(i) for i in (np.argsort(-row[1000:], kind="stable") + 1000) if int(i) != best][:3]
            assert int(out_pin[8]) == 3 and out_pin[:3].tolist() == want
        # beyond the static array: rejected with a clear error, nothing launched
        rc = L.vox_cuda_alternatives_noalloc(p_log, MAX_LOGITS + 1, 0, 4, C.c_float(1.0), p_out, p_out + 32, None, 1)
        assert rc == -1 and "at most" in vc.last_error()
    finally:
        L.vox_cuda_host_free(p_log)
        L.vox_cuda_host_free(p_out)
        L.vox_cuda_state_free(st)
