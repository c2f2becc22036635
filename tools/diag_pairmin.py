"""A/B of the CTA-pair threshold (FGDM_PAIR_MIN, K chunks of 64) on the K = 1280 GEMMs (run under gpurun)."""
import os
import sys
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fgdm_b200 import ops  # noqa: E402
from tools.bench_gemm import timeit  # noqa: E402

bf = torch.bfloat16
print("FGDM_PAIR_MIN =", os.environ.get("FGDM_PAIR_MIN", "(default)"))
for m, k, n, res in [(65536, 1280, 320, True), (4096, 1280, 1280, True), (4096, 1280, 3840, False), (1024, 1280, 1280, True),
                     (16384, 1280, 640, True), (65536, 1280, 640, False)]:
    x = torch.randn(m, k, device="cuda").to(bf)
    wt = (torch.randn(n, k, device="cuda") / k ** 0.5).to(bf)
    b = torch.randn(n, device="cuda")
    r = torch.randn(m, n, device="cuda").to(bf) if res else None
    us = timeit(lambda: ops.conv_gemm(x, wt, n, bias=b, res0=r))
    print(f"lin m={m:6d} k={k:5d} n={n:5d} res={int(res)}: {us:7.1f} us {2.0 * m * k * n / us / 1e6:6.0f} TF/s")
