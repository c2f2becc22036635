"""Small-K linears for an `ncu --set full --import-source on` capture (epilogue study; run under gpurun)."""
import os
import sys
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fgdm_b200 import ops  # noqa: E402

bf = torch.bfloat16
torch.manual_seed(0)
m, k, n = 65536, 320, 320
x = torch.randn(m, k, device="cuda").to(bf)
wt = (torch.randn(n, k, device="cuda") / k ** 0.5).to(bf)
b = torch.randn(n, device="cuda")
r = torch.randn(m, n, device="cuda").to(bf)
for rep in range(3):
    ops.conv_gemm(x, wt, n, bias=b)              # no residual
    ops.conv_gemm(x, wt, n, bias=b, res0=r)      # residual
    torch.cuda.synchronize()
