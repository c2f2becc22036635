"""Range mode (split-K / stream-K) on the K = 1280 / 2560 linears of the 16x16 and 8x8 levels: FGDM_RANGE_FORCE=<ctas>."""
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fgdm_b200 import ops  # noqa: E402

dev = "cuda"
print("FGDM_RANGE_FORCE =", os.environ.get("FGDM_RANGE_FORCE", "0"), " FGDM_PAIR =", os.environ.get("FGDM_PAIR", "-"))
for m, k, n in [(4096, 1280, 1280), (1024, 1280, 1280), (1024, 2560, 1280), (4096, 5120, 1280), (1024, 5120, 1280), (4096, 2560, 1280)]:
    x = torch.randn(m, k, device=dev).to(torch.bfloat16)
    w = (torch.randn(n, k, device=dev) / math.sqrt(k)).to(torch.bfloat16)
    b = torch.randn(n, device=dev)
    r = torch.randn(m, n, device=dev).to(torch.bfloat16)
    y = ops.conv_gemm(x, w, n, bias=b, res0=r)
    ref = x.float() @ w.float().t() + b + r.float()
    err = ((y.float() - ref).abs().max() / ref.abs().max()).item()
    out = torch.empty_like(y)
    g = torch.cuda.CUDAGraph()
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
    print(f"m={m} k={k} n={n}: {us:7.1f} us  {2.0 * m * n * k / us / 1e6:7.1f} TFLOP/s  relerr {err:.2e}")
