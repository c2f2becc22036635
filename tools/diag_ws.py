"""Residual-carrying GEMM timings used for the round-1 A/B experiments on the short-K linears (run under gpurun).
The experiments themselves (residual-tile L2 prefetch, weight-stationary ring) were reverted: see DESIGN.md section 5."""
import os
import sys
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fgdm_b200 import ops  # noqa: E402
from tools.bench_gemm import timeit  # noqa: E402

bf = torch.bfloat16
for m, k, n in [(65536, 320, 320), (65536, 320, 960), (65536, 320, 2560), (65536, 1280, 320), (16384, 640, 640),
                (16384, 2560, 640), (4096, 1280, 1280), (4096, 5120, 1280)]:
    x = torch.randn(m, k, device="cuda").to(bf)
    wt = (torch.randn(n, k, device="cuda") / k ** 0.5).to(bf)
    b = torch.randn(n, device="cuda")
    r = torch.randn(m, n, device="cuda").to(bf)
    us = timeit(lambda: ops.conv_gemm(x, wt, n, bias=b, res0=r))
    print(f"lin  m={m:6d} k={k:5d} n={n:5d}: {us:7.1f} us {2.0 * m * k * n / us / 1e6:6.0f} TF/s")
