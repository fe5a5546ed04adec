#!/usr/bin/env python3
"""Time the multi-stream product (gemm_wt.cu) alone on the decoder's five shapes and report the
HBM rate of each (weight bytes / time).  usage: gemm_wt_probe.py [B] [iters]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "voxtral.c_b200"))
import voxcuda as vc  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 200
L = vc.lib()
assert L.vox_cuda_init() == 1, vc.last_error()
shapes = [("wqkv", 6144, 3072, (1, 2, 3, 4, 6), 0), ("wo", 3072, 4096, (1, 2, 4, 6, 8), 0), ("w13", 18432, 3072, (1, 2), 1),
          ("w2", 3072, 9216, (1, 3, 6, 9, 12), 0), ("logits", 131072, 3072, (1,), 0)]
only = os.environ.get("PROBE_ONLY")
if only:
    shapes = [s for s in shapes if s[0] in only.split(",")]
total = {}
for name, N, K, splits, swiglu in shapes:
    for ks in splits:
        if swiglu and ks != 1:
            continue
        us = L.vox_cuda_bench_gemm_streams(B, N, K, ks, swiglu, iters)
        assert us > 0, vc.last_error()
        gb = N * K * 2 / 1e9
        print(f"{name:7s} N={N:6d} K={K:5d} ksplit={ks:2d} items={N // 128 * ks:5d}: {us:8.2f} us  {gb / (us * 1e-6) / 1e3:6.2f} TB/s", flush=True)
        total.setdefault(name, []).append(us)
best = {k: min(v) for k, v in total.items()}
if only:
    sys.exit(0)
step = 26 * (best["wqkv"] + best["wo"] + best["w13"] + best["w2"]) + best["logits"]
print(f"B={B}: best-split GEMM time per decoder step {step / 1e3:.3f} ms = {6.857687 / (step * 1e-6) / 1e3:.2f} TB/s over 6.858 GB")
