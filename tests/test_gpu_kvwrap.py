"""BASELINE config 4 (30-minute audio) without the 7-hour CPU run: start both engines right
before the 8192-position window wraps, with the SAME injected KV state, and step across the
reference's cache compaction (voxtral_decoder.c:354-385, 661-669).

Reference: kv_cache_max = 8192 + 38 + 1024 = 9254 rows; when kv_cache_len reaches it the cache
is compacted to the last 8192 rows and kv_pos_offset += 1062.  CUDA: a ring indexed by
logical position & 8191, no compaction.  Token ids exact, logits 2e-2 max-abs."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_decoder_window_wrap_matches_reference_across_compaction(full_env):
    e, ref = full_env, full_env.ref
    rng = np.random.default_rng(21)
    ref.set("kv_cache_len", 0); ref.set("kv_pos_offset", 0)
    ref.prefill((0.5 * rng.standard_normal((38, 3072))).astype(np.float32))   # allocates the cache
    kmax = ref.get("kv_cache_max")
    assert kmax == 8192 + 38 + 1024
    K, V = ref.kv_views()
    start = kmax - 2                      # two steps fill the cache, the third one compacts it
    for l in range(26):
        K[l, :start] = 0.3 * rng.standard_normal((start, 1024), dtype=np.float32)
        V[l, :start] = 0.3 * rng.standard_normal((start, 1024), dtype=np.float32)
    ref.set("kv_cache_len", start)
    st = e.L.vox_cuda_state_create(e.model)
    assert st, e.vc.last_error()
    try:
        for l in range(26):
            assert e.L.vox_cuda_decoder_kv_write(st, l, start - 8192, 8192,
                                                 e.vc.fptr(np.ascontiguousarray(K[l, start - 8192:start])),
                                                 e.vc.fptr(np.ascontiguousarray(V[l, start - 8192:start]))) == 0
        # read-back across the ring end (logical addressing)
        kb = np.empty((100, 1024), np.float32); vb = np.empty((100, 1024), np.float32)
        assert e.L.vox_cuda_decoder_kv_read(st, 3, 8150, 100, e.vc.fptr(kb), e.vc.fptr(vb)) == 0
        assert np.array_equal(kb, K[3, 8150:8250]) and np.array_equal(vb, V[3, 8150:8250])
        assert e.L.vox_cuda_decoder_set_position(st, start, 8192) == 0
        logits_c = np.empty(131072, np.float32)
        n_steps = 5
        for i in range(n_steps):
            x = (0.5 * rng.standard_normal(3072)).astype(np.float32)
            tok_r, logits_r = ref.decoder_forward(x)
            tok_c = e.L.vox_cuda_decoder_step(st, e.vc.fptr(x), e.vc.fptr(logits_c))
            d = float(np.abs(logits_c - logits_r).max())
            print(f"pos {start + i}: ref {tok_r} cuda {tok_c} max|dlogit| {d:.2e} "
                  f"(ref cache_len {ref.get('kv_cache_len')} offset {ref.get('kv_pos_offset')})")
            assert tok_c == tok_r and d < 2e-2
        # the reference compacted once; logical positions agree
        assert ref.get("kv_pos_offset") == 1062
        assert ref.get("kv_cache_len") == 8192 + (n_steps - 2)
        assert e.L.vox_cuda_decoder_position(st) == start + n_steps == ref.get("kv_pos_offset") + ref.get("kv_cache_len")
        # the newest rows sit where the reference has them
        kb = np.empty((n_steps, 1024), np.float32)
        assert e.L.vox_cuda_decoder_kv_read(st, 25, start, n_steps, e.vc.fptr(kb), None) == 0
        K2, _ = ref.kv_views()
        phys = ref.get("kv_cache_len") - n_steps
        assert float(np.abs(kb - K2[25, phys:phys + n_steps]).max()) < 2e-2
        # full-window step time (attention over 8192 keys x 26 layers on top of the weight stream)
        ad = (0.5 * rng.standard_normal((24, 3072))).astype(np.float32)
        assert e.L.vox_cuda_adapter_append(st, e.vc.fptr(ad), 24) >= 0, e.vc.last_error()
        toks = np.empty(20, np.int32)
        ms = e.L.vox_cuda_bench_decoder_steps(st, 32, 0, 20, e.vc.iptr(toks))
        print(f"decoder step at full 8192-key window: {ms:.3f} ms")
        assert ms > 0
    finally:
        e.L.vox_cuda_state_free(st)
    ref.set("kv_cache_len", 0); ref.set("kv_pos_offset", 0)
