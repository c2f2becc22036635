"""Why is a short kernel slower inside the sampling step than in a loop of itself? Same launches, same buffers, three orders
inside one CUDA graph each: AAAA..BBBB.. (each kernel follows itself), ABAB.. (every launch follows a different kernel),
with ONE weight matrix or 400 of them in rotation (328 MB: every weight comes from HBM, as in the step). Run under gpurun."""
import os
import sys
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fgdm_b200 import ops  # noqa: E402

bf = torch.bfloat16
m, c = 16384, 640
x = torch.randn(m, c, device="cuda").to(bf)
ws = [(torch.randn(c, c, device="cuda") / c ** 0.5).to(bf) for _ in range(400)]      # 400 x 0.8 MB = 328 MB >> L2
b = torch.randn(c, device="cuda")
r = torch.randn(m, c, device="cuda").to(bf)
g, be = torch.randn(c, device="cuda"), torch.randn(c, device="cuda")
y = torch.empty_like(x)
z = torch.empty_like(x)
xs = x.view(16, 32, 32, c)
zs = torch.empty_like(xs)

A = lambda i=0: ops.conv_gemm(x, ws[i % len(ws)], c, bias=b, res0=r, out=y)            # 16384x640->640 + residual (TMA epilogue)
B = lambda i=0: ops.layernorm(y, g, be, out=z)                                          # 16384x640 LayerNorm
C = lambda i=0: ops.groupnorm(xs, g, be, silu=True, out=zs)                             # 16x1024x640 GroupNorm (slab kernel)


def run(seq, n=400):
    for f in set(seq):
        f()
    torch.cuda.synchronize()
    gr, s = torch.cuda.CUDAGraph(), torch.cuda.Stream()
    with torch.cuda.stream(s):
        with torch.cuda.graph(gr, stream=s):
            for i in range(n):
                seq[i % len(seq)](i)
    gr.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(5):
        gr.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (5 * n)


a, bb, cc = run([A]), run([B]), run([C])
print(f"alone: gemm {a:.1f} us, layernorm {bb:.1f} us, groupnorm {cc:.1f} us per launch")
ab = run([A, B])
print(f"gemm, layernorm alternating: {ab:.1f} us per launch (mean of the two alone: {(a + bb) / 2:.1f})")
abc = run([A, B, C])
print(f"gemm, layernorm, groupnorm rotating: {abc:.1f} us per launch (mean alone: {(a + bb + cc) / 3:.1f})")
A2 = lambda i=0: ops.conv_gemm(x, ws[0], c, bias=b, res0=r, out=y)
print(f"gemm with ONE weight matrix in a loop of itself: {run([A2]):.1f} us (400 rotating matrices, 328 MB: {a:.1f})")
