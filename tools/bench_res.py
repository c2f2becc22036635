"""GPU micro-benchmark: small-K linears with / without the residual operand (epilogue latency study)."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fgdm_b200 import ops  # noqa: E402
from tools.bench_gemm import timeit  # noqa: E402
bf = torch.bfloat16
for m, k, n in [(65536, 320, 320), (16384, 640, 640), (4096, 1280, 1280), (1024, 1280, 1280), (65536, 1280, 320), (16384, 2560, 640)]:
    x = torch.randn(m, k, device="cuda").to(bf)
    w = (torch.randn(n, k, device="cuda") / k ** 0.5).to(bf)
    b = torch.randn(n, device="cuda")
    r = torch.randn(m, n, device="cuda").to(bf)
    t0 = timeit(lambda: ops.conv_gemm(x, w, n, bias=b))
    t1 = timeit(lambda: ops.conv_gemm(x, w, n, bias=b, res0=r))
    t2 = timeit(lambda: ops.conv_gemm(x, w, n, bias=b, res0=r, res1=r))
    print(f"m={m} k={k} n={n} dbg={os.environ.get('FGDM_DBG','0')}: no-res {t0:.1f} us | res0 {t1:.1f} | res0+res1 {t2:.1f}")
