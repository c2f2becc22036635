"""Time ControlNet.encode_hint (input_hint_block, cldm.py:655-671) for B = 8 hints of 512x512 (run under gpurun)."""
import os
import sys
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fgdm_b200 import synth  # noqa: E402
from fgdm_b200 import golden_cases as gc  # noqa: E402
from fgdm_b200.ldm import instantiate_from_config  # noqa: E402

model = instantiate_from_config(gc.control_config())
model.load_state_dict(synth.make_state_dict(model.param_shapes(), gc.CONTROL_WEIGHT_SEED))
cn = model.control_model
hint = torch.rand(8, 3, 512, 512, device="cuda")
for _ in range(2):
    g = cn.encode_hint(hint)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
e0.record()
for _ in range(5):
    g = cn.encode_hint(hint)
e1.record()
torch.cuda.synchronize()
print(f"encode_hint B=8 512x512: {e0.elapsed_time(e1) / 5:.3f} ms, out {tuple(g.shape)}, checksum {float(g.float().abs().sum()):.1f}")
