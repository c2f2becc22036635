"""GPU micro-benchmark of the VAE mid-block attention (model.py:178-202; 8 images, N = 4096 tokens, d = 512):
csrc/attention_wide.cu (one launch, no N x N buffer) against the materialised form (per image: S = QK^T fp32 GEMM,
softmax_rows, PV GEMM). Run under gpurun."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fgdm_b200 import ops  # noqa: E402
from fgdm_b200.layers import OUT_F32  # noqa: E402

nb, n, c = 8, 4096, 512
if len(sys.argv) > 1:
    nb = int(sys.argv[1])
q = torch.randn(nb, 1, n, c, device="cuda").to(torch.bfloat16)
k = torch.randn(nb, 1, n, c, device="cuda").to(torch.bfloat16)
vt = torch.randn(nb, 1, c, n, device="cuda").to(torch.bfloat16)
o = torch.empty(nb, n, c, device="cuda", dtype=torch.bfloat16)
o2 = torch.empty_like(o)
s = torch.empty(n, n, device="cuda", dtype=torch.float32)
p = torch.empty(n, n, device="cuda", dtype=torch.bfloat16)


def flash():
    ops.attention(q, k, vt, o, nb, 1, n, n, c, n, scale=float(c) ** -0.5)


def materialised():
    for i in range(nb):
        ops.conv_gemm(q[i, 0], k[i, 0], n, out_mode=OUT_F32, out=s)
        ops.softmax_rows(s, float(c) ** -0.5, out=p)
        ops.conv_gemm(p, vt[i, 0], c, out=o2[i])


def timeit(fn, reps=10):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps


fl = 4.0 * nb * n * n * c
tf, tm = timeit(flash), timeit(materialised)
err = ((o.float() - o2.float()).abs().max() / o2.float().abs().max()).item()
print(f"vae mid attention nb={nb} n={n} d={c}: flash {tf:8.1f} us ({fl / tf / 1e6:6.1f} TFLOP/s algorithmic), "
      f"materialised {tm:8.1f} us; rel-to-max difference {err:.2e}; N x N buffers avoided: {n * n * 6 / 2**20:.0f} MiB per image")
