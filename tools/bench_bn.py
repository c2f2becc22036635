"""Tile-width sweep of fgdm_conv_gemm on one large linear shape (device time via graph replay; run under gpurun)."""
import os
import sys
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from fgdm_b200 import ops  # noqa: E402
from bench_gemm import timeit  # noqa: E402

bf = torch.bfloat16
m, k = 65536, 2880 // 64 * 64
for n in (1280, 2240):
    x = torch.randn(m, k, device="cuda").to(bf)
    w = (torch.randn(n, k, device="cuda") / k ** 0.5).to(bf)
    for bn in (64, 128, 160, 224, 256, 1160, 1224, 1256):     # 1000 + BN: CTA-pair kernel
        if n % (bn % 1000) and not (bn % 1000 == 224):
            continue
        us = timeit(lambda: ops.conv_gemm(x, w, n, bn_hint=bn), n=5)
        print(f"lin m={m} k={k} n={n} BN={bn}: {us:8.1f} us {2.0 * m * k * n / us / 1e6:7.0f} TFLOP/s")
