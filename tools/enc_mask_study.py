#!/usr/bin/env python3
"""Which encoder products need the hi/lo bf16 operand pair?  (GPU)  Runs tools/enc_probe.py once per
VOX_CUDA_ENC_X1_MASK value (bit = product with single-bf16 activations: 1 wqkv, 2 wo, 4 w1|w3, 8 w2, 16 adapter,
32 attention) on the 60 s bench clip and prints max / mean |adapter rows - fp32 CUDA-core run| and the encoder time.
The fp32 run itself is within ~1e-4 of the reference (tests/test_gpu_parity.py); tolerance 2e-2 max-abs."""
import os
import re
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
masks = [int(m) for m in sys.argv[1:]] or [0, 1, 2, 4, 8, 16, 32, 15, 47, 63]


def run(env, out):
    e = dict(os.environ, **env)
    p = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "enc_probe.py"), out], env=e, capture_output=True, text=True)
    ms = [float(v) for v in re.findall(r"tokens in ([0-9.]+) ms", p.stdout)]
    if p.returncode != 0 or not ms:
        print(p.stdout[-500:], p.stderr[-1500:])
        raise SystemExit(1)
    return min(ms)


ms = run({"VOX_CUDA_ENC_PRECISE": "1"}, "/tmp/ad_precise.npy")
ref = np.load("/tmp/ad_precise.npy")
print(f"fp32 CUDA-core run: {ms:.1f} ms; adapter rows rms {np.sqrt((ref ** 2).mean()):.3f} max {np.abs(ref).max():.2f}", flush=True)
for m in masks:
    ms = run({"VOX_CUDA_ENC_X1_MASK": str(m)}, "/tmp/ad_m.npy")
    d = np.abs(np.load("/tmp/ad_m.npy") - ref)
    print(f"mask {m:2d}: {ms:6.2f} ms  max|d| {d.max():.5f}  mean|d| {d.mean():.6f}  p99.99 {np.quantile(d, 0.9999):.5f}", flush=True)
