"""GPU parity of the tcgen05 GEMM (gemm_tc.cu) through the C ABI against numpy
on bf16-rounded operands with fp32 accumulation.  Tolerance: 2e-3 relative to
the output scale (fp32 summation order only; the bf16 rounding of A is applied
to the numpy reference too)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

BIAS, GELU, RESIDUAL, SWIGLU = 1, 2, 4, 8


def to_bf16(x):
    u = np.ascontiguousarray(x, np.float32).view(np.uint32)
    r = ((u >> 16) & 1) + 0x7FFF
    return ((u + r) >> 16).astype(np.uint16)


def bf16_to_f32(u):
    return (u.astype(np.uint32) << 16).view(np.float32)


def gelu(v):
    return 0.5 * v * (1.0 + np.tanh(0.7978845608028654 * (v + 0.044715 * v ** 3)))


@pytest.mark.parametrize("M,N,K,flags", [
    (128, 128, 64, 0),
    (1, 128, 128, 0),
    (200, 256, 1280, BIAS),
    (156, 1280, 2048, BIAS | RESIDUAL),
    (333, 1024, 1280, SWIGLU),
    (77, 3072, 5120, GELU),
    (1000, 6144, 1280, BIAS),
    (312, 1280, 384, BIAS | GELU),
])
def test_gemm_tc(M, N, K, flags):
    import voxcuda as vc
    L = vc.lib()
    assert L.vox_cuda_init() == 1, vc.last_error()
    rng = np.random.default_rng(M * 7 + N + K)
    A = rng.standard_normal((M, K)).astype(np.float32)
    Wb = to_bf16(rng.standard_normal((N, K)).astype(np.float32) / np.sqrt(K))
    bias = rng.standard_normal(N).astype(np.float32) if flags & BIAS else None
    nout = N // 2 if flags & SWIGLU else N
    C0 = rng.standard_normal((M, nout)).astype(np.float32)
    out = C0.copy()
    rc = L.vox_cuda_gemm_bf16_tc(M, N, K, vc.fptr(A), Wb.ctypes.data_as(vc.u16p),
                                 vc.fptr(bias) if bias is not None else None, flags, vc.fptr(out))
    assert rc == 0, vc.last_error()
    ref = bf16_to_f32(to_bf16(A)).astype(np.float64) @ bf16_to_f32(Wb).astype(np.float64).T
    if flags & BIAS:
        ref = ref + bias
    if flags & GELU:
        ref = gelu(ref)
    if flags & SWIGLU:
        g, u = ref[:, 0::2], ref[:, 1::2]
        ref = g / (1.0 + np.exp(-g)) * u
    if flags & RESIDUAL:
        ref = ref + C0
    err = np.abs(out - ref).max()
    scale = max(1.0, np.abs(ref).max())
    print(f"M={M} N={N} K={K} flags={flags}: max abs err {err:.3e} (scale {scale:.2f})")
    assert err < 2e-3 * scale
