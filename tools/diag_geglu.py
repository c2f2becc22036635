"""Tile-width sweep of the GEGLU projections (C -> 8C, gate * gelu epilogue) at the UNet's three widths (run under gpurun)."""
import os
import sys
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fgdm_b200 import ops  # noqa: E402
from tools.bench_gemm import timeit  # noqa: E402

bf = torch.bfloat16
for m, k in [(65536, 320), (16384, 640), (4096, 1280)]:
    n = 8 * k
    x = torch.randn(m, k, device="cuda").to(bf)
    wt = (torch.randn(n, k, device="cuda") / k ** 0.5).to(bf)
    b = torch.randn(n, device="cuda")
    out = torch.empty(m, n // 2, device="cuda", dtype=bf)
    for bn in (0, 128, 160, 224, 256, 1160, 1224, 1256):
        try:
            us = timeit(lambda: ops.conv_gemm(x, wt, n, bias=b, out_mode=ops.OUT_GEGLU, out=out, bn_hint=bn))
        except Exception as e:
            print(f"geglu m={m} k={k} bn={bn}: refused ({str(e)[:80]})")
            continue
        print(f"geglu m={m:6d} k={k:5d} n={n:5d} bn={bn:4d}: {us:7.1f} us {2.0 * m * k * n / us / 1e6:6.0f} TF/s")
