"""A/B of the q/k/v projection GEMMs (head-split epilogue) at the UNet's shapes: set FGDM_QKV_TMA=0/1 per process."""
import os
import sys
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fgdm_b200 import ops  # noqa: E402

bf = torch.bfloat16
shapes = [(16, 4096, 320, 3), (8, 4096, 320, 3), (16, 1024, 640, 3), (16, 256, 1280, 3), (16, 4096, 320, 1), (16, 1024, 640, 1),
          (16, 256, 1280, 1)]
for images, n, c, sections in shapes:
    heads, d = 8, c // 8
    m = images * n
    x = torch.randn(m, c, device="cuda").to(bf)
    w = (torch.randn(sections * c, c, device="cuda") / c ** 0.5).to(bf)
    q = torch.empty(images, heads, n, d, device="cuda", dtype=bf)
    k = torch.empty_like(q)
    vt = torch.empty(images, heads, d, n, device="cuda", dtype=bf)
    outs = (q, k, vt) if sections == 3 else (q, q, q)
    fn = lambda: ops.conv_gemm(x, w, sections * c, rows_per_img=n, out_mode=ops.OUT_QKV, qkv=outs + (0, heads, d, n))
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(20):
        fn()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 20
    fl = 2.0 * m * c * sections * c
    by = 2.0 * (m * c + m * sections * c + sections * c * c)
    print(f"qkv m={m} k={c} n={sections * c} d={d}: {us:7.1f} us  {fl / us / 1e6:7.1f} TFLOP/s  {by / us / 1e3:7.1f} GB/s")
