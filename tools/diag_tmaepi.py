"""A/B of the TMA epilogue on the short-K transformer linears (run twice: FGDM_TMA_EPI=1 / 0)."""
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fgdm_b200 import ops  # noqa: E402

dev = "cuda"
cases = [(65536, 320, 320, True), (65536, 320, 320, False), (16384, 640, 640, True), (65536, 1280, 320, True),
         (16384, 2560, 640, True), (4096, 1280, 1280, True), (65536, 320, 640, True)]
print("FGDM_TMA_EPI =", os.environ.get("FGDM_TMA_EPI", "1"))
for m, k, n, res in cases:
    x = torch.randn(m, k, device=dev).to(torch.bfloat16)
    w = (torch.randn(n, k, device=dev) / math.sqrt(k)).to(torch.bfloat16)
    b = torch.randn(n, device=dev)
    r = torch.randn(m, n, device=dev).to(torch.bfloat16) if res else None
    y = ops.conv_gemm(x, w, n, bias=b, res0=r)
    ref = x.float() @ w.float().t() + b + (r.float() if res else 0)
    err = ((y.float() - ref).abs().max() / ref.abs().max()).item()
    g = torch.cuda.CUDAGraph()
    out = torch.empty_like(y)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3):
            ops.conv_gemm(x, w, n, bias=b, res0=r, out=out)
    torch.cuda.synchronize()
    with torch.cuda.graph(g):
        for _ in range(20):
            ops.conv_gemm(x, w, n, bias=b, res0=r, out=out)
    g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(5):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 100
    by = 2.0 * (m * k + n * k + m * n * (2 if res else 1))
    print(f"m={m} k={k} n={n} res={int(res)}: {us:7.1f} us  {2.0 * m * n * k / us / 1e6:7.1f} TFLOP/s  {by / us / 1e3:6.0f} GB/s  relerr {err:.2e}")
