"""Q8 weights.  The CUDA backend on a Q8 container (vox_cuda_model_create_q8: int8 rows streamed by the
single-token kernel through the integer tensor cores, encoder / adapter / prefill on the tcgen05 GEMMs with the row scale
in the epilogue) against the reference's OWN Q8 CPU path (/root/reference/voxtral_kernels.c:277-393, loader
voxtral_safetensors.c:393-468) compiled into oracle/_ref/libvoxref.so, on the Q8 file `gen_model --q8` writes (= what the
reference's quantize.py makes of the bf16 checkpoint: tests/test_cpu_q8.py).

Tolerances (BASELINE.json): token ids exact, logits and activations 2e-2 max-abs (measured values are printed)."""
import ctypes as C
import os

import numpy as np
import pytest

import refharness as rh

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
TOL = 2e-2


@pytest.fixture(scope="module")
def q8_env():
    import voxcuda as vc
    if not rh.have_ref():
        pytest.skip("oracle/_ref/libvoxref.so not built")
    mdir = rh.model_dir(extra=("--q8",))
    ref = rh.Ref(mdir)
    L = vc.lib()
    assert L.vox_cuda_init() == 1, vc.last_error()
    w, q = vc.Weights(), vc.WeightsQ8()
    ref.fill_weights(w)
    assert ref.fill_weights_q8(q) == 1                    # ctx->use_q8
    model = L.vox_cuda_model_create_q8(C.byref(w), C.byref(q))
    assert model, vc.last_error()
    assert L.vox_cuda_model_is_q8(model) == 1

    class Env:
        pass

    e = Env()
    e.ref, e.L, e.vc, e.model, e.mdir = ref, L, vc, model, mdir
    yield e
    L.vox_cuda_model_free(model)
    ref.close()


def test_q8_decoder_prefill_and_steps(q8_env):
    """38-row prompt prefill (tcgen05 GEMM + scale epilogue) then free-running single-token steps (int8 tiles on the
    integer tensor cores): token ids and the full logits row of every step against vox_decoder_forward on Q8"""
    e = q8_env
    rng = np.random.default_rng(7)
    n_steps = 8
    adapter = (0.5 * rng.standard_normal((38 + n_steps, 3072))).astype(np.float32)
    prompt = [1] + [32] * 37
    emb = np.stack([adapter[i] + e.ref.tok_embedding(t) for i, t in enumerate(prompt)]).astype(np.float32)
    e.ref.set("kv_cache_len", 0); e.ref.set("kv_pos_offset", 0)
    e.ref.prefill(emb)
    st = e.L.vox_cuda_state_create(e.model)
    assert st, e.vc.last_error()
    try:
        assert e.L.vox_cuda_decoder_prefill(st, e.vc.fptr(emb), 38) == 0, e.vc.last_error()
        # the prompt's K/V rows against the reference's cache
        kref, vref = e.ref.kv_views()
        k = np.empty((38, 1024), np.float32); v = np.empty((38, 1024), np.float32)
        for layer in (0, 25):
            assert e.L.vox_cuda_decoder_kv_read(st, layer, 0, 38, e.vc.fptr(k), e.vc.fptr(v)) == 0
            dk, dv = np.abs(k - kref[layer, :38]).max(), np.abs(v - vref[layer, :38]).max()
            print(f"layer {layer}: prefill max|dK| {dk:.2e} max|dV| {dv:.2e}")
            assert dk < TOL and dv < TOL
        prev_r = prev_c = 32
        logits_c = np.empty(131072, np.float32)
        worst = 0.0
        for i in range(n_steps):
            x_r = (adapter[38 + i] + e.ref.tok_embedding(prev_r)).astype(np.float32)
            tok_r, logits_r = e.ref.decoder_forward(x_r)
            x_c = (adapter[38 + i] + e.ref.tok_embedding(prev_c)).astype(np.float32)
            tok_c = e.L.vox_cuda_decoder_step(st, e.vc.fptr(x_c), e.vc.fptr(logits_c))
            assert tok_c >= 0, e.vc.last_error()
            srt = np.sort(logits_r)
            d = float(np.abs(logits_c - logits_r).max())
            worst = max(worst, d)
            print(f"step {i}: ref {tok_r} cuda {tok_c} max|dlogit| {d:.3e} margin {srt[-1] - srt[-2]:.3e}")
            assert d < TOL
            assert tok_c == tok_r
            prev_r, prev_c = tok_r, tok_c
        print("Q8 decoder: max |dlogit| over", n_steps, "steps =", worst)
    finally:
        e.L.vox_cuda_state_free(st)


def test_q8_encoder_and_adapter(q8_env):
    """first streaming chunk (312 mel -> 156 positions -> 39 adapter rows) against vox_encoder_forward +
    vox_adapter_forward on the Q8 checkpoint"""
    from test_gpu_parity import synth_audio
    e = q8_env
    x = synth_audio(16000 * 2, seed=3)
    mel = e.ref.stream_mel(x)[:312]
    ref_enc = e.ref.encoder_forward(mel)
    ref_ad = e.ref.adapter(ref_enc)
    st = e.L.vox_cuda_state_create(e.model)
    assert st, e.vc.last_error()
    try:
        assert e.L.vox_cuda_encode_chunk(st, e.vc.fptr(mel), mel.shape[0]) == 39, e.vc.last_error()
        ad = np.empty((39, 3072), np.float32)
        assert e.L.vox_cuda_adapter_read(st, 0, 39, e.vc.fptr(ad)) == 39
        d = float(np.abs(ad - ref_ad).max())
        print("Q8 adapter rows: max|d|", d, "ref rms", float(np.sqrt((ref_ad ** 2).mean())))
        assert d < TOL
        # a 56-frame streaming chunk after it: 28 positions -> the small-M path (swap-AB GEMM with the row scale in ITS
        # epilogue) must give the rows of one 368-frame call (184 positions: the large-tile path)
        mel_all = e.ref.stream_mel(x)[:368]
        mel2 = np.ascontiguousarray(mel_all[312:])
        assert e.L.vox_cuda_encode_chunk(st, e.vc.fptr(mel2), mel2.shape[0]) == 46, e.vc.last_error()   # total rows so far
        tail = np.empty((7, 3072), np.float32)
        assert e.L.vox_cuda_adapter_read(st, 39, 7, e.vc.fptr(tail)) == 7
        e.L.vox_cuda_state_reset_encoder(st)
        assert e.L.vox_cuda_encode_chunk(st, e.vc.fptr(np.ascontiguousarray(mel_all)), 368) == 46, e.vc.last_error()
        one = np.empty((46, 3072), np.float32)
        assert e.L.vox_cuda_adapter_read(st, 0, 46, e.vc.fptr(one)) == 46
        d2 = float(np.abs(one[39:] - tail).max())
        print("Q8 streaming chunk (small-M path) vs one call (large tiles): max|d|", d2)
        assert np.abs(one[:39] - ad).max() < 1e-3 and d2 < 1e-3
    finally:
        e.L.vox_cuda_state_free(st)


def _transcribe(e, x, st=None):
    import bench
    buf, nf = bench.frame_buffer(x)
    own = st is None
    if own:
        st = e.L.vox_cuda_state_create(e.model)
        assert st, e.vc.last_error()
    try:
        toks = np.empty(nf // 8 + 8, np.int32)
        n = e.L.vox_cuda_transcribe(st, e.vc.fptr(buf), len(buf), nf, 39, e.vc.iptr(toks), len(toks), None)
        assert n > 0, e.vc.last_error()
        return toks[:n].copy()
    finally:
        if own:
            e.L.vox_cuda_state_free(st)


def _compare_ids(name, got, fx):
    want = fx["ids"]
    assert len(got) == len(want), (len(got), len(want))
    nz = np.nonzero(got != want)[0]
    print(f"{name}: {len(want) - len(nz)} of {len(want)} ids equal the reference's Q8 path; smallest reference top-2 margin "
          f"{float(fx['margins'].min()):.2e}")
    if len(nz):
        i = int(nz[0])
        print(f"first difference at step {i}: cuda {got[i]} ref {want[i]}, reference margin there {float(fx['margins'][i]):.2e}")
    assert len(nz) == 0


def test_q8_jfk_ids(q8_env):
    """config 1 on Q8: samples/jfk.wav (PCM replayed from the fixture) -> ids identical to the reference's stream API"""
    fx = np.load(os.path.join(HERE, "golden", "ids_jfk_q8.npz"))
    x = fx["pcm_s16"].astype(np.float32) / 32768.0
    _compare_ids("jfk.wav Q8", _transcribe(q8_env, x), fx)


def test_q8_clip60_ids(q8_env):
    """config 2 on Q8: the 60 s bench clip, 761 steps"""
    import bench
    path = os.path.join(HERE, "golden", "ids_clip60_q8.npz")
    if not os.path.exists(path):
        pytest.skip("ids_clip60_q8.npz not generated")
    fx = np.load(path)
    _compare_ids("60 s clip Q8", _transcribe(q8_env, bench.synth_clip(60.0, 1)), fx)


def test_q8_fp16_kv_and_alternatives(q8_env):
    """the Q8 kernel's other instantiation (fp16 K/V rings) and the fused alternatives run on a Q8 model: ids of a 4 s
    clip against the fp32-ring run, alternatives of the last step against a numpy scan of the logits row"""
    import bench
    e = q8_env
    x = bench.synth_clip(4.0, 3)
    a = _transcribe(e, x)
    st = e.L.vox_cuda_state_create_ex(e.model, 16)
    assert st, e.vc.last_error()
    try:
        assert e.L.vox_cuda_decoder_set_alt(st, 4, C.c_float(1.0)) == 0, e.vc.last_error()
        b = _transcribe(e, x, st)
        assert len(a) == len(b)
        nz = np.nonzero(a != b)[0]
        print(f"Q8, 4 s clip: {len(a) - len(nz)} of {len(a)} ids equal between fp32 and fp16 K/V rings")
        assert len(nz) == 0 or nz[0] >= 10
        logits = np.empty(131072, np.float32)
        assert e.L.vox_cuda_decoder_read_logits(st, e.vc.fptr(logits)) == 0
        ids = np.empty(8, np.int32)
        n = e.L.vox_cuda_decoder_alternatives(st, 4, C.c_float(1.0), e.vc.iptr(ids))
        assert n >= 0, e.vc.last_error()
        best = int(np.argmax(logits))
        want = [int(i) for i in (np.argsort(-logits[1000:], kind="stable") + 1000) if int(i) != best][:3]
        assert ids[:n].tolist() == want[:n] and n == 3
    finally:
        e.L.vox_cuda_state_free(st)


def test_q8_stream_group_equals_single_stream(q8_env):
    """a Q8 model behind a stream group (lock-step decode on the swap-AB GEMM, row scale in its epilogue, embedding rows
    = int8 row * scale): three clips of different lengths, every stream's ids = the single-stream Q8 run; fp32 and fp16 rings"""
    import bench
    e = q8_env
    clips = [bench.frame_buffer(bench.synth_clip(sec, seed)) for sec, seed in ((2.0, 5), (3.1, 11), (2.4, 12))]
    want = [_transcribe(e, bench.synth_clip(sec, seed)).tolist() for sec, seed in ((2.0, 5), (3.1, 11), (2.4, 12))]
    for bits in (32, 16):
        g = e.L.vox_cuda_group_create_ex(e.model, 3, 128, bits)
        assert g, e.vc.last_error()
        try:
            for i, (buf, nf) in enumerate(clips):
                assert e.L.vox_cuda_group_encode(g, i, e.vc.fptr(buf), len(buf), nf) == len(want[i]) + 38, e.vc.last_error()
            toks = np.full((3, 64), -7, np.int32)
            counts = np.zeros(3, np.int32)
            steps = e.L.vox_cuda_group_decode(g, 39, e.vc.iptr(toks), 64, e.vc.iptr(counts), None)
            assert steps == max(len(w) for w in want), (steps, e.vc.last_error())
            same = [toks[i, :counts[i]].tolist() == want[i] for i in range(3)]
            print(f"Q8 stream group, {bits}-bit K/V: streams equal to single-stream ids: {same}")
            if bits == 32:
                assert all(same)
            else:                                             # fp16 rings: a flip may only come late (free-running)
                for i in range(3):
                    got = toks[i, :counts[i]]
                    nz = np.nonzero(got != np.array(want[i]))[0]
                    assert len(nz) == 0 or nz[0] >= 10
        finally:
            e.L.vox_cuda_group_free(g)


@pytest.mark.parametrize("world,bits", [(2, 32), (4, 16)])
def test_q8_tensor_parallel_lockstep(q8_env, world, bits):
    """a Q8 model behind the tensor-parallel decoder (all ranks on this GPU in lock step, vox_cuda_tp_create_local_ex): row
    scales applied in every shard's GEMV, embedding row = int8 row * scale; ids = the single-stream Q8 ids of the clip"""
    import bench
    e = q8_env
    x = bench.synth_clip(2.0, 5)
    want = _transcribe(e, x).tolist()
    buf, nf = bench.frame_buffer(x)
    tp = e.L.vox_cuda_tp_create_local_ex(e.model, world, bits)
    assert tp, e.vc.last_error()
    try:
        toks = np.empty(nf // 8 + 8, np.int32)
        n = e.L.vox_cuda_tp_transcribe(tp, e.vc.fptr(buf), len(buf), nf, 39, e.vc.iptr(toks), len(toks), None)
        assert n == len(want), e.vc.last_error()
        got = toks[:n].tolist()
        print(f"Q8, TP x{world}, {bits}-bit K/V: {sum(a == b for a, b in zip(got, want))} of {n} ids equal the single-stream Q8 ids")
        if bits == 32:
            assert got == want
        else:
            nz = [i for i, (a, b) in enumerate(zip(got, want)) if a != b]
            assert not nz or nz[0] >= 10
    finally:
        e.L.vox_cuda_tp_free(tp)


def test_q8_dropin_cli(q8_env, tmp_path):
    """the reference's own host (main.c + voxtral*.c, -DUSE_CUDA) on the Q8 directory: its loader takes the Q8 branch
    (ctx->use_q8, voxtral.c:175-181), dropin/vox_cuda_dropin.c binds the *_weight_q8 / *_scale_q8 pairs to
    vox_cuda_model_create_q8, and the text must be the reference's Q8 CPU text"""
    import subprocess
    from test_gpu_ids import PKG, text_of, write_wav
    exe = os.path.join(PKG, "voxtral_cuda")
    if not os.path.exists(exe):
        pytest.skip("voxtral_cuda not built (make -C voxtral.c_b200 dropin needs /root/reference)")
    fx = np.load(os.path.join(HERE, "golden", "ids_jfk_q8.npz"))
    wav = str(tmp_path / "jfk.wav")
    write_wav(wav, fx["pcm_s16"].astype(np.float32) / 32768.0)
    p = subprocess.run([exe, "-d", q8_env.mdir, "-i", wav], capture_output=True, timeout=900)
    assert p.returncode == 0, p.stderr.decode()[-2000:]
    assert p.stdout.decode().strip() == text_of(fx["ids"]).strip()
