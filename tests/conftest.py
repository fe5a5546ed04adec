import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "voxtral.c_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def full_env():
    """Reference engine + CUDA backend over the SAME random-init checkpoint.

    The reference (oracle/_ref) loads and mmaps the file; the CUDA weight table
    is filled from the reference's own ctx by the vox_load hook in
    oracle/ref_shim.c - exactly the binding INTEGRATION.md proposes.
    """
    import refharness as rh
    import voxcuda as vc

    if not rh.have_ref():
        pytest.skip("oracle/_ref/libvoxref.so not built")
    mdir = rh.model_dir()
    ref = rh.Ref(mdir)
    L = vc.lib()
    assert L.vox_cuda_init() == 1, vc.last_error()
    w = vc.Weights()
    ref.fill_weights(w)
    model = L.vox_cuda_model_create(w)
    assert model, vc.last_error()

    class Env:
        pass

    e = Env()
    e.ref, e.L, e.vc, e.model, e.mdir, e.weights = ref, L, vc, model, mdir, w
    yield e
    L.vox_cuda_model_free(model)
    ref.close()
