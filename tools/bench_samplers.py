"""GPU timing of the widened rows: CLIP text encode, PLMS and DPM-Solver chains vs DDIM on the seg factor
(B=8, 64x64 latents, CFG 7.5; CUDA events, after one warm-up call). Run under gpurun."""
import os
import sys
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fgdm_b200 import golden_cases as gc, synth  # noqa: E402
from fgdm_b200.clip import FrozenCLIPEmbedder  # noqa: E402
from fgdm_b200.ddim import DDIMSampler  # noqa: E402
from fgdm_b200.dpm_solver import DPMSolverSampler  # noqa: E402
from fgdm_b200.ldm import instantiate_from_config  # noqa: E402
from fgdm_b200.plms import PLMSSampler  # noqa: E402

DEV = "cuda"


def timed(fn, reps=2):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


enc = FrozenCLIPEmbedder(device=DEV).load_state_dict(gc.clip_state_dict())
ids = gc.clip_case()["ids"][:1].repeat(16, 1).to(DEV)            # 8 prompts + 8 negative prompts
print(f"CLIP text encode, 16 x 77 tokens: {timed(lambda: enc.encode(ids), 5):.2f} ms")

model = instantiate_from_config(gc.fgdm_config())
model.load_state_dict(synth.make_state_dict(model.param_shapes(), gc.FGDM_WEIGHT_SEED))
B = 8
inp = {k: v.to(DEV) for k, v in synth.prompt_inputs(list(range(B)), (64, 64)).items()}
kw = dict(batch_size=B, shape=[4, 64, 64], conditioning=inp["c"], unconditional_conditioning=inp["uc"][:1].expand(B, -1, -1).contiguous()
          if inp["uc"].shape[0] != B else inp["uc"], unconditional_guidance_scale=7.5, eta=0.0, x_T=inp["xT_seg"], verbose=False)
for name, smp, S in (("DDIM", DDIMSampler(model), 50), ("PLMS", PLMSSampler(model), 50), ("DPM-Solver++(2M)", DPMSolverSampler(model), 20)):
    ms = timed(lambda: smp.sample(S=S, **kw), 1)
    print(f"{name:18s} S={S:3d}: {ms:8.1f} ms per chain ({ms / S:.2f} ms per step, B={B}, 64x64, CFG)")
