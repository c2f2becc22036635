"""Diagnostic sweep of the short-K transformer linears (run under gpurun): tile width, residual on/off, output rows.
What bounds 65536x320->320 at ~37 us when its MMA time is 12 us and its HBM floor 20 us?"""
import os
import sys
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fgdm_b200 import ops  # noqa: E402
from tools.bench_gemm import timeit  # noqa: E402

bf = torch.bfloat16
for m, k, n in [(65536, 320, 320), (16384, 640, 640), (4096, 1280, 1280)]:
    x = torch.randn(m, k, device="cuda").to(bf)
    wt = (torch.randn(n, k, device="cuda") / k ** 0.5).to(bf)
    b = torch.randn(n, device="cuda")
    r = torch.randn(m, n, device="cuda").to(bf)
    for bn in (0, 64, 128, 160, 224, 256, 1160, 1224, 1256):
        for res in (True, False):
            try:
                fn = lambda: ops.conv_gemm(x, wt, n, bias=b, res0=r if res else None, bn_hint=bn)
                us = timeit(fn)
            except Exception as e:  # a width the launcher refuses for this shape
                print(f"m={m} k={k} n={n} bn={bn} res={int(res)}: refused ({str(e)[:60]})")
                continue
            by = 2.0 * (m * k + n * k + m * n * (2 if res else 1))
            print(f"m={m:6d} k={k:5d} n={n:5d} bn={bn:4d} res={int(res)}: {us:7.1f} us {2.0 * m * k * n / us / 1e6:6.0f} TF/s {by / us / 1e3:6.0f} GB/s")
