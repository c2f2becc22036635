"""BASELINE configs[4] (C5): SpatialTransformer CrossAttention + ResBlock microbench at the 64x64 latent (4096 tokens,
77 text tokens), batch 16 rows, against the tensor-core / HBM rooflines. Run under gpurun; prints one line per op and a
JSON summary (committed as profiles/rNN_c5_microbench.json)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fgdm_b200 import ops, synth  # noqa: E402
from fgdm_b200.layers import ResBlock, SpatialTransformer  # noqa: E402

dev = torch.device("cuda")
bf = torch.bfloat16
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {
    "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "hbm_gbs": 6650.0}


def timeit(fn, n=10):
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
    for _ in range(3):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (3 * n)


def profiled(fn):
    fn()
    ops.profile_start()
    fn()
    return ops.profile_stop()


R, HW, C, HEADS, CTX = 16, 64, 320, 8, 768
x = (torch.randn(R, HW, HW, C, device=dev) * 0.5).to(bf)
ctx = torch.randn(R, 77, CTX, device=dev).to(bf)
emb = torch.randn(R, C, device=dev)

st = SpatialTransformer(C, HEADS, C // HEADS, CTX)
st.kv_slot = 0
sd = synth.make_state_dict(dict(st.specs("")), 1)
st.load(sd, "", dev)
kv = [st.project_context(ctx)]
rb = ResBlock(C, 4 * C, C)
sd = synth.make_state_dict(dict(rb.specs("")), 2)
rb.load(sd, "", dev)
rb.emb_off = 0

out = {"config": "C5: R=16 rows, 64x64 latent (4096 tokens), 320 channels, 8 heads x 40, 77 text tokens", "peaks": peaks, "ops": {}}
for name, fn in (("SpatialTransformer", lambda: st.forward(x, kv)), ("ResBlock", lambda: rb.forward(x, None, emb))):
    us = timeit(fn)
    prof = profiled(fn)
    flops = sum(v[2] for sh in prof.values() for v in sh.values())
    nbytes = sum(v[3] for sh in prof.values() for v in sh.values())
    kernels = {k: {"launches": sum(v[0] for v in sh.values()), "ms": round(sum(v[1] for v in sh.values()), 4)} for k, sh in prof.items()}
    tf = flops / us / 1e6
    out["ops"][name] = {"us": round(us, 1), "algorithmic_gflop": round(flops / 1e9, 1), "tflops": round(tf, 1),
                        "frac_of_sustained_bf16": round(tf / peaks["bf16_tflops_sustained"], 3),
                        "frac_of_burst_bf16": round(tf / peaks["bf16_tflops"], 3), "kernels_eager_ms": kernels}
    print(f"{name}: {us:.1f} us  {flops / 1e9:.1f} GFLOP  {tf:.1f} TFLOP/s = {tf / peaks['bf16_tflops_sustained']:.3f} of sustained bf16")
print(json.dumps(out))
