"""GPU micro-benchmark of fgdm_attention at the UNet's shapes (run under gpurun; ncu-friendly)."""
import sys
import os
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fgdm_b200 import ops  # noqa: E402

shapes = [(16, 8, 4096, 4096, 40), (16, 8, 1024, 1024, 80), (16, 8, 4096, 77, 40), (16, 8, 1024, 77, 80), (16, 8, 256, 77, 160),
          (16, 8, 256, 256, 160)]
if len(sys.argv) > 1:
    shapes = shapes[:int(sys.argv[1])]
for images, heads, nq, nk, d in shapes:
    q = torch.randn(images, heads, nq, d, device="cuda").to(torch.bfloat16)
    k = torch.randn(images, heads, nk, d, device="cuda").to(torch.bfloat16)
    tp = (nk + 7) // 8 * 8
    vt = torch.randn(images, heads, d, tp, device="cuda").to(torch.bfloat16)
    out = torch.empty(images, nq, heads * d, device="cuda", dtype=torch.bfloat16)
    for _ in range(3):
        ops.attention(q, k, vt, out, images, heads, nq, nk, d, tp)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    n = 10
    for _ in range(n):
        ops.attention(q, k, vt, out, images, heads, nq, nk, d, tp)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / n
    fl = 4.0 * images * heads * nq * nk * d
    tiles = images * heads * ((nq + 127) // 128) * ((nk + 127) // 128)
    print(f"attn images={images} nq={nq} nk={nk} d={d}: {us:8.1f} us  {fl / us / 1e6:7.1f} TFLOP/s  "
          f"{us * 1e-6 * 1.9e9 * 148 / tiles:7.0f} clk/tile/SM@1.9GHz")
