"""GPU micro-benchmark of fgdm_conv_gemm at the UNet's shapes (device time via CUDA-graph replay; run under gpurun)."""
import os
import sys
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fgdm_b200 import ops  # noqa: E402


def timeit(fn, n=10):
    for _ in range(2):
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
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


bf = torch.bfloat16
# (kind, nb, h, w, cin, cout, residual)
cases = [("conv", 16, 8, 8, 1280, 1280, True), ("conv", 16, 8, 8, 2560, 1280, True), ("conv", 16, 16, 16, 1280, 1280, True),
         ("conv", 16, 32, 32, 640, 640, True), ("conv", 16, 64, 64, 320, 320, True), ("conv", 16, 64, 64, 640, 320, True),
         ("lin", 16, 64, 64, 320, 320, True), ("lin", 16, 32, 32, 640, 640, True), ("lin", 16, 16, 16, 1280, 1280, True),
         ("lin", 16, 8, 8, 1280, 1280, True), ("lin", 16, 64, 64, 1280, 320, True), ("lin", 16, 16, 16, 5120, 1280, True)]
for kind, nb, h, w, cin, cout, res in (cases if __name__ == "__main__" else []):
    m = nb * h * w
    x = torch.randn(nb, h, w, cin, device="cuda").to(bf)
    b = torch.randn(cout, device="cuda")
    r = torch.randn(m, cout, device="cuda").to(bf) if res else None
    if kind == "conv":
        wt = (torch.randn(cout, 9 * cin, device="cuda") / (9 * cin) ** 0.5).to(bf)
        fn = lambda: ops.conv_gemm(x, wt, cout, conv=ops.CONV_3X3, bias=b, res0=r)
        k = 9 * cin
    else:
        wt = (torch.randn(cout, cin, device="cuda") / cin ** 0.5).to(bf)
        xf = x.view(m, cin)
        fn = lambda: ops.conv_gemm(xf, wt, cout, bias=b, res0=r)
        k = cin
    us = timeit(fn)
    fl = 2.0 * m * k * cout
    by = 2.0 * (m * cin + cout * k + m * cout * (2 if res else 1))
    print(f"{kind} m={m:6d} k={k:6d} n={cout:5d}: {us:8.1f} us {fl / us / 1e6:7.0f} TFLOP/s {by / us / 1e3:6.0f} GB/s")
