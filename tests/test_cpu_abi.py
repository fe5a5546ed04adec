"""CPU-side checks of the drop-in boundary: the C-ABI library loads without a
GPU and exports every symbol include/vox_cuda.h declares; the ctypes binding
covers the same set; entry points fail loudly (no CPU fallback) when no CUDA
device is present."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "vox_cuda.h")


def header_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vox_cuda_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_boundary():
    syms = header_symbols()
    for must in ("vox_cuda_init", "vox_cuda_model_create", "vox_cuda_decoder_step",
                 "vox_cuda_decoder_prefill", "vox_cuda_encoder_step", "vox_cuda_mel_frames",
                 "vox_cuda_sgemm_bf16"):
        assert must in syms


def test_library_exports_every_declared_symbol():
    import voxcuda as vc
    L = vc.lib()
    missing = [s for s in header_symbols() if not hasattr(L, s)]
    assert not missing, f"declared in vox_cuda.h but not exported: {missing}"


def test_ctypes_binding_matches_header():
    import voxcuda as vc
    syms = set(header_symbols())
    bound = set(vc.SIGNATURES)
    assert syms == bound, (sorted(syms - bound), sorted(bound - syms))


def test_weight_table_layout_matches_c():
    """sizeof(vox_cuda_weights_t) as seen by the reference-side hook (compiled C)
    must equal the ctypes mirror."""
    import voxcuda as vc
    import refharness as rh
    if not rh.have_ref():
        pytest.skip("oracle/_ref not built")
    L = C.CDLL(rh.REF_SO)
    assert L.refshim_sizeof_weights() == C.sizeof(vc.Weights)


def test_no_cpu_fallback_without_gpu():
    import voxcuda as vc
    L = vc.lib()
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        pytest.skip("GPU present")
    assert L.vox_cuda_init() == 0
    assert b"no CUDA device" in L.vox_cuda_last_error() or b"cannot" in L.vox_cuda_last_error()
    x = np.zeros(400, np.float32)
    out = np.zeros(128, np.float32)
    assert L.vox_cuda_mel_frames(vc.fptr(x), 1, vc.fptr(out)) < 0
    assert not L.vox_cuda_model_create(C.byref(vc.Weights()))


def test_product_does_not_import_oracle():
    """oracle/ is test infrastructure: nothing in the shipped package may reference it, and the
    shared library must not link anything from oracle/_ref."""
    import subprocess
    pkg = os.path.join(ROOT, "voxtral.c_b200")
    bad = []
    for dirpath, _, files in os.walk(pkg):
        if os.sep + "build" in dirpath:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".c", ".cpp")) or f == "Makefile":
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                if re.search(r"vox_oracle|oracle/|libvoxref|refshim_", src):
                    bad.append(os.path.join(dirpath, f))
    assert not bad, bad
    so = os.path.join(pkg, "libvoxcuda.so")
    needed = subprocess.run(["readelf", "-d", so], capture_output=True, text=True).stdout
    assert "voxref" not in needed and "openblas" not in needed
