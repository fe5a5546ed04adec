"""__graft_entry__.smoke(): one small invocation of the hot path on cuda:0,
checked against the oracle (oracle/_ref, the compiled reference BLAS path).

Runs: log-mel of 1 s of audio, then prefill(38) + 2 greedy decoder steps on the
random-init checkpoint, then the same two steps for two streams in lock step
(stream group) and as a 2-rank tensor-parallel decode, all through the C ABI (libvoxcuda.so).  Fails loudly if
the CUDA library is missing - there is no CPU fallback.
"""
import numpy as np


def run():
    import refharness as rh
    import voxcuda as vc

    L = vc.lib()
    if L.vox_cuda_init() != 1:
        raise RuntimeError("smoke: " + vc.last_error())
    if not rh.have_ref():
        raise RuntimeError("smoke: oracle/_ref/libvoxref.so missing (run __graft_entry__.build())")

    # ---- mel (hot path #3) vs the strict-IEEE reference build ----
    rng = np.random.default_rng(0)
    x = (0.1 * rng.standard_normal(16000)).astype(np.float32)
    ref_mel = rh.strict_stream_mel(x)
    n = len(x)
    right = (1280 - n % 1280) % 1280 + 17 * 1280
    buf = np.concatenate([np.zeros(200 + 32 * 1280, np.float32), x, np.zeros(right, np.float32)])
    buf = np.ascontiguousarray(np.concatenate([buf, buf[len(buf) - 2 - np.arange(200)]]).astype(np.float32))
    nf = (len(buf) - 400) // 160
    assert nf == ref_mel.shape[0], (nf, ref_mel.shape)
    mel = np.empty((nf, 128), np.float32)
    vc.check(L.vox_cuda_mel_frames(vc.fptr(buf), nf, vc.fptr(mel)), "mel")
    err = float(np.max(np.abs(mel - ref_mel) / np.maximum(np.abs(ref_mel), 1.0)))
    print(f"smoke: mel {nf} frames, max rel err vs reference {err:.2e}")
    assert err < 1e-5

    # ---- decoder (hot path #1): prefill + 2 steps vs the reference ----
    mdir = rh.model_dir()
    ref = rh.Ref(mdir)
    w = vc.Weights()
    ref.fill_weights(w)
    model = vc.check(L.vox_cuda_model_create(w), "model_create")
    st = vc.check(L.vox_cuda_state_create(model), "state_create")
    try:
        adapter = (0.5 * rng.standard_normal((40, 3072))).astype(np.float32)
        prompt = [1] + [32] * 37
        emb = np.stack([adapter[i] + ref.tok_embedding(t) for i, t in enumerate(prompt)]).astype(np.float32)
        ref.set("kv_cache_len", 0)
        ref.set("kv_pos_offset", 0)
        ref.prefill(emb)
        vc.check(L.vox_cuda_decoder_prefill(st, vc.fptr(emb), 38), "prefill")
        prev = 32
        logits = np.empty(131072, np.float32)
        ref_toks = []
        for i in range(2):
            xin = (adapter[38 + i] + ref.tok_embedding(prev)).astype(np.float32)
            tok_r, logits_r = ref.decoder_forward(xin)
            tok_c = vc.check(L.vox_cuda_decoder_step(st, vc.fptr(xin), vc.fptr(logits)), "step")
            d = float(np.abs(logits - logits_r).max())
            print(f"smoke: step {i} ref token {tok_r} cuda token {tok_c} max|dlogit| {d:.2e}")
            assert tok_c == tok_r and d < 2e-2
            ref_toks.append(tok_r)
            prev = tok_r
        # ---- the same two steps for 2 streams in lock step (stream group, skinny tcgen05 GEMM) ----
        g = vc.check(L.vox_cuda_group_create(model, 2, 64), "group_create")
        try:
            for i in range(2):
                vc.check(L.vox_cuda_group_set_adapter(g, i, vc.fptr(adapter), 40), "group_set_adapter")
            gt = np.empty((2, 8), np.int32)
            counts = np.zeros(2, np.int32)
            steps = vc.check(L.vox_cuda_group_decode(g, 39, vc.iptr(gt), 8, vc.iptr(counts), None), "group_decode")
            print(f"smoke: group of 2 streams, {steps} lock steps, ids {gt[0, :2].tolist()} {gt[1, :2].tolist()}")
            assert steps == 2 and gt[0, :2].tolist() == ref_toks and gt[1, :2].tolist() == ref_toks
        finally:
            L.vox_cuda_group_free(g)
        # ---- and tensor-parallel: 2 ranks on this GPU in lock step (the multi-GPU shard arithmetic + exchange windows) ----
        tp = vc.check(L.vox_cuda_tp_create_local(model, 2), "tp_create_local")
        try:
            ts = L.vox_cuda_tp_state(tp)
            vc.check(L.vox_cuda_adapter_append(ts, vc.fptr(adapter), 40), "adapter_append")
            vc.check(L.vox_cuda_decoder_prefill(ts, vc.fptr(emb), 38), "tp prefill")
            tt = np.empty(2, np.int32)
            vc.check(L.vox_cuda_tp_decode(tp, 32, 38, 2, vc.iptr(tt), None), "tp_decode")
            print(f"smoke: tensor-parallel (2 ranks, lock step) ids {tt.tolist()}")
            assert tt.tolist() == ref_toks
        finally:
            L.vox_cuda_tp_free(tp)
        print("smoke: kernel launches so far:", L.vox_cuda_launch_count())
    finally:
        L.vox_cuda_state_free(st)
        L.vox_cuda_model_free(model)
        ref.close()
    print("smoke: OK")
