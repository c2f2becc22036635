"""GPU debugging aid: per-block error of the CUDA UNet against the fp32 oracle (run under gpurun)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import fgdm_oracle as orc  # noqa: E402
from fgdm_b200 import golden_cases as gc, synth  # noqa: E402
from fgdm_b200.ldm import instantiate_from_config  # noqa: E402

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
hw = int(sys.argv[1]) if len(sys.argv) > 1 else 16
model = instantiate_from_config(gc.fgdm_config())
sd = synth.make_state_dict(model.param_shapes(), gc.FGDM_WEIGHT_SEED)
model.load_state_dict(sd)
sd_dev = {k: v.cuda() for k, v in sd.items() if k.startswith("model.")}
x = synth.randn((2, 4, hw, hw), 21).cuda()
t = torch.tensor([981, 441], device="cuda")
ctx = synth.randn((2, 77, 768), 22).cuda()
unet = model.model.diffusion_model
unet.trace = []
eps = unet(x, t, context=ctx)
torch.cuda.synchronize()
orc.TRACE = []
eps_ref = orc.unet_forward(sd_dev, x, t, ctx, prefix="model.diffusion_model.")
ref = [(n, v) for n, v in orc.TRACE if not n.endswith("input_blocks.0")]
print("blocks", len(unet.trace), len(ref))
for (n, r), m in zip(ref, unet.trace):
    m = m.float().permute(0, 3, 1, 2)
    err = float((m - r).abs().max() / r.abs().max())
    print(f"{n:45s} shape {tuple(r.shape)} absmax {float(r.abs().max()):9.3f} rel-to-max err {err:.3e}")
print("eps rel-to-max err", float((eps - eps_ref).abs().max() / eps_ref.abs().max()), "absmax", float(eps_ref.abs().max()))
