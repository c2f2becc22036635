"""GPU micro-benchmark: cost of the LayerNorm fold on the transformer GEMMs (producer statistics, consumer epilogue).
Run under gpurun."""
import os
import sys
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fgdm_b200 import ops  # noqa: E402
from tools.bench_gemm import timeit  # noqa: E402

bf = torch.bfloat16
for m, c in [(65536, 320), (16384, 640), (4096, 1280)]:
    heads, d, n = 8, c // 8, m // 16
    x = torch.randn(m, c, device="cuda").to(bf)
    r = torch.randn(m, c, device="cuda").to(bf)
    w = (torch.randn(c, c, device="cuda") / c ** 0.5).to(bf)
    b = torch.randn(c, device="cuda")
    wq = (torch.randn(3 * c, c, device="cuda") / c ** 0.5).to(bf)
    bq = torch.randn(3 * c, device="cuda")
    sq = torch.randn(3 * c, device="cuda")
    wg = (torch.randn(8 * c, c, device="cuda") / c ** 0.5).to(bf)
    bg = torch.randn(8 * c, device="cuda")
    sg = torch.randn(8 * c, device="cuda")
    q = torch.empty(16, heads, n, d, device="cuda", dtype=bf)
    k = torch.empty_like(q)
    vt = torch.empty(16, heads, d, n, device="cuda", dtype=bf)
    _, stats = ops.conv_gemm(x, w, c, bias=b, res0=r, want_stats=True)
    ln = (stats, None, c, 1e-5)
    t_prod0 = timeit(lambda: ops.conv_gemm(x, w, c, bias=b, res0=r))
    t_prod1 = timeit(lambda: ops.conv_gemm(x, w, c, bias=b, res0=r, want_stats=True))
    t_q0 = timeit(lambda: ops.conv_gemm(x, wq, 3 * c, bias=bq, rows_per_img=n, out_mode=ops.OUT_QKV,
                                        qkv=(q, k, vt, 0, heads, d, n)))
    t_q1 = timeit(lambda: ops.conv_gemm(x, wq, 3 * c, bias=bq, rows_per_img=n, out_mode=ops.OUT_QKV,
                                        qkv=(q, k, vt, 0, heads, d, n), ln=(stats, sq, c, 1e-5)))
    t_g0 = timeit(lambda: ops.conv_gemm(x, wg, 8 * c, bias=bg, out_mode=ops.OUT_GEGLU))
    t_g1 = timeit(lambda: ops.conv_gemm(x, wg, 8 * c, bias=bg, out_mode=ops.OUT_GEGLU, ln=(stats, sg, c, 1e-5)))
    t_ln = timeit(lambda: ops.layernorm(x, b, b))
    print(f"m={m} c={c} slots={stats.shape[1]}: producer {t_prod0:.1f} -> {t_prod1:.1f} us"
          f" | qkv {t_q0:.1f} -> {t_q1:.1f} | geglu {t_g0:.1f} -> {t_g1:.1f} | layernorm kernel {t_ln:.1f}")
