"""SURVEY 8(f3), CPU side: the Q8 container this repo writes (`tools/gen_model.c --q8`) is the one the reference's
quantize.py writes (/root/reference/quantize.py:34-43,86-129; fixture made by running quantize.py itself, see
tests/golden/make_q8_quantizer_fixture.md), the host loader reads it, and the Q8 weight table has the layout the
reference-side hook (compiled C, oracle/ref_shim.c) sees."""
import ctypes as C
import hashlib
import json
import os
import struct
import subprocess

import numpy as np

import refharness as rh

HERE = os.path.dirname(os.path.abspath(__file__))


def _small_q8_dir(tmp_path_factory):
    d = str(tmp_path_factory.mktemp("q8chk"))
    subprocess.check_call([rh.ensure_gen_model(), d, "--enc-layers", "1", "--dec-layers", "1", "--no-tokenizer", "--q8"],
                          stderr=subprocess.DEVNULL)
    return d


def test_gen_model_q8_equals_reference_quantizer(tmp_path_factory):
    fx = json.load(open(os.path.join(HERE, "golden", "q8_quantizer.json")))
    d = _small_q8_dir(tmp_path_factory)
    raw = np.memmap(os.path.join(d, "consolidated.safetensors"), dtype=np.uint8, mode="r")
    (hlen,) = struct.unpack("<Q", bytes(raw[:8]))
    hdr = json.loads(bytes(raw[8:8 + hlen]))
    assert len(hdr) == fx["tensors"]
    assert sum(1 for v in hdr.values() if v["dtype"] == "Q8") == fx["q8_tensors"]
    assert all(v["dtype"] in ("Q8", "F32") for v in hdr.values())
    for name, (dtype, shape, offs) in fx["entries"].items():
        assert [hdr[name]["dtype"], hdr[name]["shape"], hdr[name]["data_offsets"]] == [dtype, shape, offs], name
    data = raw[8 + hlen:]
    assert len(data) == fx["data_bytes"]
    assert hashlib.sha256(data).hexdigest() == fx["data_sha256"]          # scales and int8 rows bit for bit


def test_host_loader_reads_q8(tmp_path_factory):
    """hostmodel.HostModel on a Q8 file: (int8, scale) pairs in the Q8 table, fp32 smalls, ada_scale from q * scale"""
    import hostmodel
    import voxcuda as vc
    d = _small_q8_dir(tmp_path_factory)
    hm = hostmodel.HostModel(d)
    assert hm.q8 and hm.n_enc == 1 and hm.n_dec == 1
    q, sc = hm.q8_parts("layers.0.attention.wq.weight")
    assert q.shape == (4096, 3072) and q.dtype == np.int8 and sc.shape == (4096,)
    # per-row symmetric quantisation: the largest |q| of every row is 127 (quantize.py:34-43)
    assert int(np.abs(q[:64].astype(np.int32)).max(axis=1).min()) == 127
    w8 = hm.weights_q8
    assert C.cast(w8.dec[0].wq.q, C.c_void_p).value == q.ctypes.data
    assert C.cast(w8.dec[0].wq.scale, C.c_void_p).value == sc.ctypes.data
    assert not hm.weights.dec[0].wq                                       # no bf16 pointer in a Q8 model
    assert hm.ada.shape == (1, 3072) and np.isfinite(hm.ada).all()
    down = hm.dense_f32("layers.0.ada_rms_norm_t_cond.0.weight")
    assert down.shape == (32, 3072) and down.dtype == np.float32


def test_q8_weight_table_layout_matches_c():
    import voxcuda as vc
    if not rh.have_ref():
        import pytest
        pytest.skip("oracle/_ref not built")
    L = C.CDLL(rh.REF_SO)
    assert L.refshim_sizeof_weights_q8() == C.sizeof(vc.WeightsQ8)


def test_q8_oracle_restatement_against_reference():
    """oracle/vox_oracle.py quantize_q8 / linear_q8 pinned to the compiled reference: vox_linear_nobias_q8 and
    vox_linear_q8 of libvoxref.so (/root/reference/voxtral_kernels.c:312-364) on random inputs, one row (the fused
    matvec) and several rows (widen + sgemm); and the quantiser against the file gen_model --q8 wrote"""
    import pytest
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(HERE), "oracle"))
    import vox_oracle as vo
    if not rh.have_ref():
        pytest.skip("oracle/_ref not built")
    L = C.CDLL(rh.REF_SO)
    f32p, i8p = C.POINTER(C.c_float), C.POINTER(C.c_int8)
    L.vox_linear_nobias_q8.argtypes = [f32p, f32p, i8p, f32p, C.c_int, C.c_int, C.c_int]
    L.vox_linear_q8.argtypes = [f32p, f32p, i8p, f32p, f32p, C.c_int, C.c_int, C.c_int]
    rng = np.random.default_rng(3)
    for rows, k, n in ((1, 3072, 64), (5, 1280, 96)):
        w = (0.05 * rng.standard_normal((n, k))).astype(np.float32)
        q, sc = vo.quantize_q8(w)
        assert q.dtype == np.int8 and int(np.abs(q.astype(np.int32)).max(axis=1).min()) == 127
        x = rng.standard_normal((rows, k)).astype(np.float32)
        b = rng.standard_normal(n).astype(np.float32)
        y = np.empty((rows, n), np.float32)
        L.vox_linear_nobias_q8(y.ctypes.data_as(f32p), x.ctypes.data_as(f32p), q.ctypes.data_as(i8p), sc.ctypes.data_as(f32p), rows, k, n)
        want = vo.linear_q8(x, q, sc)
        assert np.abs(y - want).max() < 2e-5 * max(1.0, float(np.abs(want).max()))
        L.vox_linear_q8(y.ctypes.data_as(f32p), x.ctypes.data_as(f32p), q.ctypes.data_as(i8p), sc.ctypes.data_as(f32p), b.ctypes.data_as(f32p), rows, k, n)
        assert np.abs(y - vo.linear_q8(x, q, sc, b)).max() < 2e-5 * max(1.0, float(np.abs(want).max()))


def test_q8_oracle_quantiser_equals_the_file(tmp_path_factory):
    """vo.quantize_q8 on the bf16 tensor == the (int8, scale) pair gen_model --q8 wrote for it (itself = quantize.py)"""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(HERE), "oracle"))
    import hostmodel
    import vox_oracle as vo
    dq = _small_q8_dir(tmp_path_factory)
    db = str(tmp_path_factory.mktemp("bfchk"))
    subprocess.check_call([rh.ensure_gen_model(), db, "--enc-layers", "1", "--dec-layers", "1", "--no-tokenizer"], stderr=subprocess.DEVNULL)
    hq, hb = hostmodel.HostModel(dq), hostmodel.HostModel(db)
    for name, shape in (("layers.0.attention.wk.weight", (1024, 3072)), ("layers.0.ada_rms_norm_t_cond.2.weight", (3072, 32))):
        w = hostmodel.bf16_to_f32(hb.u16(name)).reshape(shape)
        q, sc = vo.quantize_q8(w)
        fq, fs = hq.q8_parts(name)
        assert np.array_equal(q, fq) and np.array_equal(sc.view(np.uint32), fs.view(np.uint32)), name
