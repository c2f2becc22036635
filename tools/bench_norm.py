"""GPU micro-benchmark of fgdm_groupnorm / fgdm_layernorm at the UNet's shapes (run under gpurun)."""
import os
import sys
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fgdm_b200 import ops  # noqa: E402


def timeit(fn, n=20):
    """device time per call: n calls captured in one CUDA graph (no host launch overhead in the number)"""
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
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


shapes = [(16, 4096, 320), (8, 4096, 320), (8, 4096, 512), (16, 4096, 640), (16, 1024, 640), (16, 1024, 1280), (16, 256, 1280),
          (16, 64, 1280), (16, 64, 2560), (16, 256, 2560), (8, 65536, 256), (8, 262144, 128)]
if len(sys.argv) > 1:
    shapes = shapes[:int(sys.argv[1])]
for nb, hw, c in shapes:
    h = int(hw ** 0.5)
    x = torch.randn(nb, h, h, c, device="cuda").to(torch.bfloat16)
    g = torch.randn(c, device="cuda")
    b = torch.randn(c, device="cuda")
    out = torch.empty_like(x)
    us = timeit(lambda: ops.groupnorm(x, g, b, silu=True, out=out))
    print(f"gn nb={nb} hw={hw} c={c}: {us:8.1f} us  {4.0 * x.numel() / us / 1e3:7.0f} GB/s (algorithmic 4 B/elem)")
for rows, c in ([(65536, 320), (16384, 640), (4096, 1280)] if len(sys.argv) <= 1 else []):
    x = torch.randn(rows, c, device="cuda").to(torch.bfloat16)
    g = torch.randn(c, device="cuda")
    b = torch.randn(c, device="cuda")
    us = timeit(lambda: ops.layernorm(x, g, b))
    print(f"ln rows={rows} c={c}: {us:8.1f} us  {4.0 * x.numel() / us / 1e3:7.0f} GB/s")
