#!/usr/bin/env python3
"""Token-id fixtures produced by the reference's OWN stream API (oracle/_ref/libvoxref.so,
the unmodified BLAS path compiled by oracle/build_ref.sh) on the seeded random-init
checkpoint (tools/gen_model.c, seed 1234) - the pin for BASELINE configs 1-4:

  jfk        config 1: samples/jfk.wav through vox_transcribe_audio's call sequence
             (vox_stream_init + feed + finish, /root/reference/voxtral.c:1390-1396), with
             alternatives on (vox_stream_set_alt(4, 0.2) + vox_stream_get_alt) so that
             stream_fill_alts (voxtral.c:955-1010) is pinned through the reference API too
  clip60     config 2: the 60 s bench clip (bench.synth_clip(60, 1)), one feed + finish
  clip60_i05 config 3: the same clip fed in 8000-sample pieces with
             vox_set_processing_interval(0.5) (main.c:200-205 feed rule)
  clip60_i20 config 3: 16000-sample pieces, interval 2.0
  long_head  config 4 (i): the first 62 s of the 30-minute clip (seed 2), fed without
             finish - by causality its ids equal the head of the full run
  jfk_q8, clip60_q8   SURVEY 8(f3): the same two jobs on the Q8 container of the same weights
             (gen_model --q8 = the reference's quantize.py, tests/golden/q8_quantizer.json), i.e. the
             reference's own Q8 CPU path (voxtral_kernels.c:277-393); not part of "all"

Every decoder call of the reference's stream path is recorded by the --wrap recorder in
oracle/ref_shim.c (token id + top-2 logit margin), every encoder call by its row count.
The text the reference returns (" t<id>" pieces of the synthetic tekken.json) is parsed
back and must agree with the recorded ids >= 1000.

Run in the build container (needs /root/reference for jfk.wav and oracle/_ref):
    python tests/golden/make_golden_ids.py all        # forks one process per job
Outputs tests/golden/ids_<job>.npz (a few KB each; jfk also stores the s16 PCM of the
sample so that the GPU box, which has no /root/reference, can replay it).
"""
import ctypes as C
import os
import re
import subprocess
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import refharness as rh  # noqa: E402

JOBS = ("jfk", "clip60", "clip60_i05", "clip60_i20", "long_head")
ALT_N, ALT_CUTOFF = 4, 0.2


def job_audio(job):
    import bench
    if job == "jfk":
        wav = open("/root/reference/samples/jfk.wav", "rb").read()
        x = rh.ref_parse_wav(wav, strict=False)
        s16 = np.round(x * 32768.0).astype(np.int16)
        assert np.array_equal(s16.astype(np.float32) / 32768.0, x)      # mono 16 kHz s16: no resampling
        return x, s16
    if job == "jfk_q8":
        return job_audio("jfk")
    if job.startswith("clip60"):
        return bench.synth_clip(60.0, 1), None
    if job == "long_head":
        return bench.synth_clip(1800.0, 2)[:62 * 16000], None
    raise SystemExit("unknown job " + job)


def run_job(job):
    x, s16 = job_audio(job)
    ref = rh.Ref(rh.model_dir(extra=("--q8",)) if job.endswith("_q8") else rh.model_dir())
    L = ref.L
    L.vox_stream_init.restype = C.c_void_p
    L.vox_stream_init.argtypes = [C.c_void_p]
    L.vox_stream_feed.argtypes = [C.c_void_p, rh.f32p, C.c_int]
    L.vox_stream_finish.argtypes = [C.c_void_p]
    L.vox_stream_free.argtypes = [C.c_void_p]
    L.vox_stream_set_alt.argtypes = [C.c_void_p, C.c_int, C.c_float]
    L.vox_set_processing_interval.argtypes = [C.c_void_p, C.c_float]
    L.vox_stream_get_alt.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.c_int, C.c_int]
    L.refshim_log_tokens.restype = C.POINTER(C.c_int)
    L.refshim_log_margins.restype = C.POINTER(C.c_float)
    L.refshim_enc_call_rows.restype = C.POINTER(C.c_int)
    L.refshim_log_top.restype = C.POINTER(C.c_float)
    L.refshim_log_reset()
    s = L.vox_stream_init(ref.ctx)
    assert s
    if job.startswith("jfk"):
        L.vox_stream_set_alt(s, ALT_N, C.c_float(ALT_CUTOFF))
    piece = len(x)
    if job == "clip60_i05":
        L.vox_set_processing_interval(s, C.c_float(0.5)); piece = 8000
    if job == "clip60_i20":
        L.vox_set_processing_interval(s, C.c_float(2.0)); piece = 16000
    alts = []
    buf = (C.c_char_p * (64 * ALT_N))()

    def drain():
        while True:
            n = L.vox_stream_get_alt(s, buf, 64, ALT_N)
            if n <= 0:
                break
            for i in range(n):
                row = []
                for a in range(ALT_N):
                    t = buf[i * ALT_N + a]
                    row.append(int(re.fullmatch(rb" t(\d+)", t).group(1)) if t else -1)
                alts.append(row)

    t0 = time.time()
    for off in range(0, len(x), piece):
        c = np.ascontiguousarray(x[off:off + piece])
        assert L.vox_stream_feed(s, c.ctypes.data_as(rh.f32p), len(c)) == 0
        drain()
    if job != "long_head":
        assert L.vox_stream_finish(s) == 0
        drain()
    n = L.refshim_log_count()
    ids = np.ctypeslib.as_array(L.refshim_log_tokens(), shape=(n,)).copy()
    margins = np.ctypeslib.as_array(L.refshim_log_margins(), shape=(n,)).copy()
    top = np.ctypeslib.as_array(L.refshim_log_top(), shape=(n, 5)).copy()
    ne = L.refshim_enc_call_count()
    enc_rows = np.ctypeslib.as_array(L.refshim_enc_call_rows(), shape=(ne,)).copy()
    alts = np.array(alts, np.int32).reshape(-1, ALT_N)
    # the text path (tokenizer + queue) and the recorder must tell the same story
    text_ids = ids[ids >= 1000]
    if 2 in ids.tolist():
        text_ids = ids[:ids.tolist().index(2)]
        text_ids = text_ids[text_ids >= 1000]
    assert np.array_equal(alts[:, 0], text_ids), (alts[:8, 0], text_ids[:8])
    out = {"ids": ids.astype(np.int32), "margins": margins.astype(np.float32), "enc_call_rows": enc_rows.astype(np.int32),
           "top_logits": top.astype(np.float32),     # [steps][5]: logit of the id, then the 4 best other text logits
           "alts": alts, "alt_n": np.int32(ALT_N), "alt_cutoff": np.float32(ALT_CUTOFF if job.startswith("jfk") else 0.0),
           "n_samples": np.int64(len(x)), "cpu_seconds": np.float32(time.time() - t0)}
    if s16 is not None:
        out["pcm_s16"] = s16
    L.vox_stream_free(s)
    ref.close()
    path = os.path.join(HERE, f"ids_{job}.npz")
    np.savez_compressed(path, **out)
    print(job, "steps", n, "enc calls", enc_rows.tolist()[:6], "... min margin", float(margins.min()),
          "text tokens", len(alts), "seconds", round(time.time() - t0), "->", path, flush=True)


if __name__ == "__main__":
    which = sys.argv[1:] or ["all"]
    if which == ["all"]:
        env = dict(os.environ, OPENBLAS_NUM_THREADS="2", OMP_NUM_THREADS="2")
        ps = [subprocess.Popen([sys.executable, __file__, j], env=env) for j in JOBS]
        sys.exit(max(p.wait() for p in ps))
    for j in which:
        run_job(j)
