"""Device time of the epilogue-bound GEMM families at the 64x64 / 32x32 levels (QKV, GEGLU, rowvec, lean)."""
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fgdm_b200 import ops  # noqa: E402

dev = "cuda"
bf = torch.bfloat16


def timeit(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            for _ in range(n):
                fn()
    g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(3):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (3 * n)


for nimg, ntok, c, heads in [(16, 4096, 320, 8), (16, 1024, 640, 8), (16, 256, 1280, 8)]:
    d = c // heads
    m = nimg * ntok
    x = torch.randn(m, c, device=dev).to(bf)
    wqkv = (torch.randn(3 * c, c, device=dev) / math.sqrt(c)).to(bf)
    q = torch.empty(nimg, heads, ntok, d, device=dev, dtype=bf)
    k = torch.empty_like(q)
    vt = torch.empty(nimg, heads, d, ntok, device=dev, dtype=bf)
    us = timeit(lambda: ops.conv_gemm(x, wqkv, 3 * c, out_mode=ops.OUT_QKV, rows_per_img=ntok, qkv=(q, k, vt, 0, heads, d, ntok)))
    print(f"QKV   m={m} k={c} n={3 * c}: {us:7.1f} us  {2.0 * m * c * 3 * c / us / 1e6:7.1f} TFLOP/s")
    wg, bg = ops.pack_geglu(torch.randn(8 * c, c, device=dev) / math.sqrt(c), torch.randn(8 * c, device=dev))
    us = timeit(lambda: ops.conv_gemm(x, wg, 8 * c, bias=bg, out_mode=ops.OUT_GEGLU))
    print(f"GEGLU m={m} k={c} n={8 * c}: {us:7.1f} us  {2.0 * m * c * 8 * c / us / 1e6:7.1f} TFLOP/s")
    w1 = (torch.randn(c, c, device=dev) / math.sqrt(c)).to(bf)
    b1 = torch.randn(c, device=dev)
    r = torch.randn(m, c, device=dev).to(bf)
    us = timeit(lambda: ops.conv_gemm(x, w1, c, bias=b1, res0=r))
    print(f"lean  m={m} k={c} n={c} +res: {us:7.1f} us  {2.0 * m * c * c / us / 1e6:7.1f} TFLOP/s")
    us = timeit(lambda: ops.conv_gemm(x, w1, c, bias=b1, res0=r, want_stats=True))
    print(f"lean  m={m} k={c} n={c} +res +row_stats: {us:7.1f} us")
