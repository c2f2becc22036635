"""GPU parity of the network / sampler level against (a) the golden fixtures produced by the UNMODIFIED
reference (tests/golden/*.npz, written by oracle/make_golden.py in the build container) and (b) the fp32
oracle restatement (oracle/fgdm_oracle.py) at sizes the fixtures do not cover.

Tolerance (BASELINE.json north_star): per-step UNet eps max error relative to the reference's max
magnitude <= 2e-2 with bf16 storage; decoded images PSNR >= 35 dB; integer hand-off bit exact.
"""
import os
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
GOLD = os.path.join(ROOT, "tests", "golden")

from fgdm_b200 import _lib, ops, synth  # noqa: E402
from fgdm_b200 import golden_cases as gc  # noqa: E402
from fgdm_b200.ddim import DDIMSampler  # noqa: E402
from fgdm_b200.ldm import instantiate_from_config  # noqa: E402

DEV = "cuda"
EPS_TOL = 2e-2


def relmax(a, b):
    a, b = torch.as_tensor(a).float().cpu(), torch.as_tensor(b).float().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-12))


def psnr(a, b, peak):
    mse = float(((torch.as_tensor(a).double().cpu() - torch.as_tensor(b).double().cpu()) ** 2).mean())
    return 10.0 * np.log10(peak * peak / max(mse, 1e-20))


def gold(name):
    return np.load(os.path.join(GOLD, name + ".npz"))


@pytest.fixture(scope="module", autouse=True)
def _dev():
    _lib.require_device()
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False


@pytest.fixture(scope="module")
def fgdm():
    model = instantiate_from_config(gc.fgdm_config())
    sd = synth.make_state_dict(model.param_shapes(), gc.FGDM_WEIGHT_SEED)
    model.load_state_dict(sd)
    return model, {k: v.to(DEV) for k, v in sd.items()}


@pytest.fixture(scope="module")
def control():
    model = instantiate_from_config(gc.control_config())
    sd = synth.make_state_dict(model.param_shapes(), gc.CONTROL_WEIGHT_SEED)
    model.load_state_dict(sd)
    return model, {k: v.to(DEV) for k, v in sd.items()}


def test_unet_forward_vs_reference_golden(fgdm):
    model, _ = fgdm
    c = gc.unet_case()
    eps = model.apply_model(c["x"].to(DEV), c["t"].to(DEV), c["ctx"].to(DEV), num=1)
    assert eps.shape == (2, 4, 16, 16) and eps.dtype == torch.float32
    assert relmax(eps, gold("unet_fgdm")["eps"]) < EPS_TOL


def test_unet_pcond_vs_reference_golden(fgdm):
    """Joint FG-DM factor: UNetModel.forward(pcond=...) through apply_model's kwargs (ddpm.py:1130, openaimodel.py:836-844)."""
    import fgdm_oracle as orc
    model, sd = fgdm
    c = gc.pcond_case()
    eps = model.apply_model(c["x"].to(DEV), c["t"].to(DEV), c["ctx"].to(DEV), pcond=c["pcond"].to(DEV))
    assert relmax(eps, gold("unet_pcond")["eps"]) < EPS_TOL
    # sampler: pcond features are hoisted out of the captured step; graph path == per-step apply_model path == oracle
    d = gc.ddim_case()
    pc = c["pcond"].to(DEV)
    kw = dict(S=d["S"], conditioning=d["c"].to(DEV), batch_size=d["B"], shape=list(d["shape"]), verbose=False,
              unconditional_guidance_scale=d["scale"], unconditional_conditioning=d["uc"].to(DEV), eta=0.0,
              x_T=d["x_T"].to(DEV), pcond=pc)
    x0, _ = DDIMSampler(model).sample(**kw)
    x0b, _ = DDIMSampler(model, use_graph=False).sample(**kw)
    assert relmax(x0b, x0) < 1e-2
    with torch.no_grad():                                      # fp32 oracle chain, evaluated on the GPU
        x0o, _ = orc.ddim_sample(
            lambda x, t, cc: orc.unet_forward(sd, x, t, cc, prefix="model.diffusion_model.", pcond=pc),
            d["S"], (d["B"],) + tuple(d["shape"]), kw["conditioning"], kw["unconditional_conditioning"], d["scale"],
            kw["x_T"], model.alphas_cumprod)
    assert psnr(x0, x0o, float(x0o.abs().max())) > 35.0
    x0p, _ = DDIMSampler(model).sample(**{k: v for k, v in kw.items() if k != "pcond"})
    assert relmax(x0p, x0) > 1e-2                              # and differs from the plain (pcond=None) chain


def test_clip_text_encoder_vs_transformers_golden():
    """FrozenCLIPEmbedder (modules.py:137-163) on the B200 kernels vs transformers' CLIPTextModel (golden) and the oracle."""
    import fgdm_oracle as orc
    from fgdm_b200.clip import FrozenCLIPEmbedder
    sd = gc.clip_state_dict()
    enc = FrozenCLIPEmbedder(device=DEV)
    assert enc.param_shapes() == {k: tuple(v) for k, v in gc.clip_param_shapes().items()}
    enc.load_state_dict(sd)
    ids = gc.clip_case()["ids"]
    z = enc.encode(ids.to(DEV))
    g = gold("clip_text")["z"]
    assert z.shape == (2, 77, 768) and z.dtype == torch.bfloat16
    print("clip text: relmax", relmax(z, g), "psnr", psnr(z, g, float(np.abs(g).max())))
    assert relmax(z, g) < 3e-2 and psnr(z, g, float(np.abs(g).max())) > 40.0
    # causal mask: position i ignores later tokens; shorter sequences and a batch of one work
    ids2 = ids.clone()
    ids2[:, 5:] = 1234
    z2 = enc.encode(ids2.to(DEV))
    assert torch.equal(z2[:, :5], z[:, :5]) and relmax(z2[:, 5:], z[:, 5:].float().cpu().numpy()) > 1e-2
    sd_dev = {k: v.to(DEV) for k, v in sd.items()}
    z3 = enc.encode(ids[:1, :20].to(DEV))
    ref3 = orc.clip_text_forward(sd_dev, ids[:1, :20].to(DEV))
    assert relmax(z3, ref3.cpu().numpy()) < 3e-2
    with pytest.raises(NotImplementedError):
        enc.encode(["a photo of a cat"])
    # wired in through the reference's config: cond_stage_config target FrozenCLIPEmbedder (models/config.yaml:74-75)
    cfg = gc.fgdm_config()
    cfg["params"] = dict(cfg["params"], cond_stage_config={"target": "ldm.modules.encoders.modules.FrozenCLIPEmbedder"})
    ldm = instantiate_from_config(cfg)
    ldm.cond_stage_model.load_state_dict(sd)
    assert torch.equal(ldm.get_learned_conditioning(ids.to(DEV)), z)
    ctx = torch.randn(1, 77, 768, device=DEV)
    assert ldm.get_learned_conditioning(ctx) is ctx
    with pytest.raises(ValueError):
        enc.encode(torch.full((1, 77), 60000, dtype=torch.int64))     # id outside the vocabulary


def test_dpm_solver_vs_reference_golden(fgdm):
    """DPMSolverSampler (DPM-Solver++ 2M) on the captured model-call graph, against the reference's solver."""
    from fgdm_b200.dpm_solver import DPMSolverSampler
    model, _ = fgdm
    c, g = gc.dpm_case(), gold("dpm_fgdm")
    kw = dict(S=c["S"], conditioning=c["c"].to(DEV), batch_size=c["B"], shape=list(c["shape"]), verbose=False,
              unconditional_guidance_scale=c["scale"], unconditional_conditioning=c["uc"].to(DEV), eta=0.0,
              x_T=c["x_T"].to(DEV))
    sampler = DPMSolverSampler(model)
    x0, extra = sampler.sample(**kw)
    assert extra is None
    peak = float(np.abs(g["x0"]).max())
    print("dpm fgdm: relmax", relmax(x0, g["x0"]), "psnr", psnr(x0, g["x0"], peak))
    assert relmax(x0, g["x0"]) < 8e-2 and psnr(x0, g["x0"], peak) > 35.0
    x0b, _ = sampler.sample(**kw)
    assert torch.equal(x0b, x0)                                # graph replay is reproducible bit for bit
    x20, _ = sampler.sample(**{**kw, "S": 20})                 # S >= 15: second order to the end
    assert torch.isfinite(x20).all()


def test_plms_loop_vs_reference_golden(fgdm):
    """PLMSSampler (plms.py) on the captured model-call graph + fgdm_plms_step, against the reference's sampler."""
    from fgdm_b200.plms import PLMSSampler
    model, _ = fgdm
    c, g = gc.plms_case(), gold("plms_fgdm")
    kw = dict(S=c["S"], conditioning=c["c"].to(DEV), batch_size=c["B"], shape=list(c["shape"]), verbose=False,
              unconditional_guidance_scale=c["scale"], unconditional_conditioning=c["uc"].to(DEV), eta=0.0,
              x_T=c["x_T"].to(DEV))
    sampler = PLMSSampler(model)
    seen, preds = [], []
    x0, inter = sampler.sample(img_callback=lambda p, i: (seen.append(i), preds.append(p)), **kw)
    assert seen == list(range(g["steps"].shape[0]))
    assert len(inter["x_inter"]) == int(g["n_inter"])
    peak = float(np.abs(g["x0"]).max())
    print("plms fgdm: relmax", relmax(x0, g["x0"]), "psnr", psnr(x0, g["x0"], peak))
    assert relmax(x0, g["x0"]) < 8e-2 and psnr(x0, g["x0"], peak) > 35.0
    assert relmax(inter["pred_x0"][-1], g["pred_x0_last"]) < 8e-2
    x0b, _ = sampler.sample(**kw)                              # graph replay is reproducible bit for bit
    assert torch.equal(x0b, x0)
    with pytest.raises(ValueError):
        sampler.sample(**{**kw, "eta": 0.5})                   # plms.py:25-26


def test_ddim_cfg_loop_vs_reference_golden(fgdm):
    model, _ = fgdm
    c, g = gc.ddim_case(), gold("ddim_fgdm")
    kw = dict(S=c["S"], conditioning=c["c"].to(DEV), batch_size=c["B"], shape=list(c["shape"]), verbose=False,
              unconditional_guidance_scale=c["scale"], unconditional_conditioning=c["uc"].to(DEV), eta=0.0,
              x_T=c["x_T"].to(DEV), num=1)
    sampler = DDIMSampler(model)
    seen = []
    x0, inter = sampler.sample(img_callback=lambda p, i: seen.append(i), **kw)
    assert seen == list(range(c["S"]))
    assert len(inter["x_inter"]) == int(g["n_inter"]) and torch.equal(inter["x_inter"][0], kw["x_T"])
    # after S chained CFG steps (scale 7.5 amplifies eps error) the bar is the north_star's PSNR >= 35 dB;
    # the max-error bound is a looser sanity check
    print("ddim fgdm: relmax", relmax(x0, g["x0"]), "psnr", psnr(x0, g["x0"], float(np.abs(g["x0"]).max())))
    assert relmax(x0, g["x0"]) < 8e-2
    assert relmax(inter["pred_x0"][-1], g["pred_x0_last"]) < 8e-2
    assert psnr(x0, g["x0"], float(np.abs(g["x0"]).max())) > 35.0
    # schedule tables identical to the reference's
    assert (sampler.ddim_timesteps == g["ddim_timesteps"]).all()
    assert np.array_equal(sampler.ddim_alphas.numpy(), g["ddim_alphas"])
    assert np.array_equal(sampler.ddim_alphas_prev, g["ddim_alphas_prev"])
    assert np.array_equal(sampler.ddim_sqrt_one_minus_alphas.numpy(), g["ddim_sqrt_one_minus_alphas"])
    # the generic (apply_model-per-step, no CUDA graph) path computes the same thing
    x0b, _ = DDIMSampler(model, use_graph=False).sample(**kw)
    assert relmax(x0b, x0) < 1e-2
    # graph replay is reproducible bit for bit
    x0c, _ = sampler.sample(**kw)
    assert torch.equal(x0c, x0)


def test_vae_decode_vs_reference_golden(fgdm):
    model, _ = fgdm
    img = model.decode_first_stage(gc.vae_case()["z"].to(DEV))
    ref = gold("vae_decode")["img"]
    assert img.shape == ref.shape
    assert relmax(img, ref) < EPS_TOL
    assert psnr(img, ref, 2.0) > 35.0


def test_vae_decode_flash_vs_materialised_attention(fgdm, monkeypatch):
    """K10: the mid-block AttnBlock (model.py:178-202) through csrc/attention_wide.cu (no N x N buffer) and in the
    reference's form (S = QK^T in fp32, softmax, PV through HBM), both against the fp32 oracle, at the benchmark's 64x64
    latent (N = 4096 tokens, d = 512) for two images at once. Same bars as the 16x16 golden; the two bf16 forms differ
    from each other by their rounding noise only (P unnormalised <= 2^8 vs normalised)."""
    import fgdm_oracle as orc
    from fgdm_b200 import nets
    model, sd = fgdm
    z = (synth.randn((2, 4, 64, 64), 43) * 0.18215 * 3.0).to(DEV)
    assert nets.VAE_FLASH
    img = model.decode_first_stage(z)
    monkeypatch.setattr(nets, "VAE_FLASH", False)
    mat = model.decode_first_stage(z)
    ref = orc.vae_decode(sd, z)
    e_f, e_m = relmax(img, ref), relmax(mat, ref)
    print("vae decode 512^2 vs oracle: flash relmax", e_f, "psnr", psnr(img, ref, 2.0), "| materialised relmax", e_m, "psnr",
          psnr(mat, ref, 2.0), "| flash vs materialised relmax", relmax(img, mat))
    assert torch.isfinite(img).all()
    assert e_f < 3e-2 and psnr(img, ref, 2.0) > 35.0
    assert e_f < 1.5 * e_m + 5e-3            # the flash form is no worse than the form it replaces
    assert relmax(img, mat) < 3e-2


def test_controlldm_apply_model_vs_reference_golden(control):
    model, _ = control
    c, g = gc.control_case(), gold("control_apply")
    cond = {"c_concat": [c["hint"].to(DEV)], "c_crossattn": [c["ctx"].to(DEV)]}
    x, t = c["x"].to(DEV), c["t"].to(DEV)
    assert relmax(model.apply_model(x, t, cond), g["eps"]) < EPS_TOL
    ctrl = model.control_model(x, hint=cond["c_concat"][0], timesteps=t, context=cond["c_crossattn"][0])
    assert len(ctrl) == 13
    assert relmax(ctrl[0].float().permute(0, 3, 1, 2), g["ctrl0"]) < EPS_TOL
    # ctrl[12] is an INTERNAL bf16 tensor (deepest 2x2 map behind 8 bf16 ResBlock/transformer stages), not the
    # north_star's eps: its max-abs error is an extreme-value statistic that moves +-15 % with the summation order of
    # any kernel, so it is held to 3e-2 on the max and EPS_TOL on the RMS; eps itself stays at EPS_TOL above and below.
    c12, g12 = ctrl[12].float().permute(0, 3, 1, 2), torch.as_tensor(g["ctrl12"]).to(DEV)
    assert relmax(c12, g["ctrl12"]) < 3e-2
    assert float((c12 - g12).pow(2).mean().sqrt() / g12.pow(2).mean().sqrt()) < EPS_TOL
    model.control_scales = list(gc.CONTROL_SCALES_VARIANT)
    try:
        assert relmax(model.apply_model(x, t, cond), g["eps_scaled"]) < EPS_TOL
    finally:
        model.control_scales = [1.0] * 13
    model.only_mid_control = True
    try:
        assert relmax(model.apply_model(x, t, cond), g["eps_only_mid"]) < EPS_TOL
    finally:
        model.only_mid_control = False


def test_ddim_hacked_loop_vs_reference_golden(control):
    model, _ = control
    c, g = gc.control_ddim_case(), gold("ddim_control")
    hint = c["hint"].to(DEV)
    cond = {"c_concat": [hint], "c_crossattn": [c["c"].to(DEV)]}
    un = {"c_concat": [hint], "c_crossattn": [c["uc"].to(DEV)]}
    x0, inter = DDIMSampler(model).sample(c["S"], c["B"], list(c["shape"]), cond, verbose=False, eta=0.0,
                                          unconditional_guidance_scale=c["scale"], unconditional_conditioning=un,
                                          x_T=c["x_T"].to(DEV))
    print("ddim control: relmax", relmax(x0, g["x0"]), "psnr", psnr(x0, g["x0"], float(np.abs(g["x0"]).max())))
    assert relmax(x0, g["x0"]) < 8e-2
    assert psnr(x0, g["x0"], float(np.abs(g["x0"]).max())) > 35.0


def test_handoff_bit_exact_vs_reference_golden():
    c, g = gc.handoff_case(), gold("handoff")
    x = c["img"].to(DEV)
    u = ops.image_to_u8(x, 0)
    assert np.array_equal(u.cpu().numpy(), g["seg_u8"])
    assert np.array_equal(ops.resize_u8(u, c["dh"], c["dw"]).cpu().numpy(), g["resized"])
    assert np.array_equal(ops.resize_u8(u, 80, 96).cpu().numpy(), g["resized_odd"])
    assert np.array_equal(ops.image_to_u8(x, 1).cpu().numpy(), g["final_u8"])
    import fgdm_oracle as orc
    # floating point: the reference divides on the GPU (`.float().cuda() / 255.0`, initialize_cn.py:78), where
    # torch multiplies by the reciprocal; the CPU oracle divides. They agree to 1 ulp.
    torch.testing.assert_close(ops.u8_to_hint(u).cpu(), orc.u8_to_hint(g["seg_u8"]), rtol=2.4e-7, atol=0)


@pytest.mark.parametrize("hw,rows", [(32, 2), (64, 2)])
def test_unet_eps_vs_oracle_full_resolution(fgdm, hw, rows):
    """BASELINE configs' latent sizes (256^2 -> 32x32, 512^2 -> 64x64), oracle evaluated in fp32 on the GPU."""
    import fgdm_oracle as orc
    model, sd = fgdm
    x = synth.randn((rows, 4, hw, hw), 81).to(DEV)
    t = torch.tensor([961, 21][:rows], device=DEV)
    ctx = synth.randn((rows, 77, 768), 82).to(DEV)
    eps = model.apply_model(x, t, ctx)
    ref = orc.unet_forward(sd, x, t, ctx, prefix="model.diffusion_model.")
    print("unet", hw, "eps relmax", relmax(eps, ref))
    assert relmax(eps, ref) < EPS_TOL


def test_chain_seg_to_image_vs_oracle(fgdm, control):
    """Whole factor chain at reduced size (128^2 seg map -> resize 256^2 -> image factor 32x32 latent),
    6+6 steps, CFG: final decoded image PSNR vs the oracle chain >= 35 dB; hand-off uint8 maps compared."""
    import fgdm_oracle as orc
    from fgdm_b200.pipeline import FGDMPipeline
    (seg, sd_s), (img, sd_i) = fgdm, control
    S = 6
    pipe = FGDMPipeline(seg, img, seg_steps=S, img_steps=S, seg_scale=7.5, img_scale=7.5, out_hw=(256, 256))
    B = 2
    inp = synth.prompt_inputs(list(range(B)), (16, 16))
    inp["xT_img"] = torch.stack([synth.randn((4, 32, 32), 3000 + i) for i in range(B)])
    d = {k: v.to(DEV) for k, v in inp.items()}
    out = pipe(d["c"], d["uc"], d["xT_seg"], d["c2"], d["uc2"], d["xT_img"], return_all=True)
    # oracle chain (fp32, GPU), restating txt2img_fgdm_inference.py:233-261 + initialize_cn.py:78-99
    ac = seg.alphas_cumprod
    z_seg, _ = orc.ddim_sample(lambda x, t, c: orc.unet_forward(sd_s, x, t, c, prefix="model.diffusion_model."), S,
                               (B, 4, 16, 16), d["c"], d["uc"], 7.5, d["xT_seg"], ac)
    seg_u8 = orc.seg_image_to_u8(orc.vae_decode(sd_s, z_seg))
    seg_rs = np.stack([orc.resize_u8_bilinear(u, 256, 256) for u in seg_u8])
    hint = orc.u8_to_hint(seg_rs).to(DEV)
    z_img, _ = orc.ddim_sample(lambda x, t, c: orc.control_ldm_apply(sd_i, x, t, hint, c), S, (B, 4, 32, 32),
                               d["c2"], d["uc2"], 7.5, d["xT_img"], ac, sequential_cfg=True)
    final = orc.final_image_to_u8(orc.vae_decode(sd_i, z_img))
    p_seg = psnr(out["seg_u8"].float(), torch.from_numpy(seg_rs).float(), 255.0)
    p_img = psnr(out["img_u8"].float(), torch.from_numpy(final).float(), 255.0)
    print("chain: seg map PSNR", p_seg, "final image PSNR", p_img)
    assert p_seg > 35.0 and p_img > 35.0


def test_full_chain_512_50_steps_vs_oracle(fgdm, control):
    """BASELINE config 2's shape at B=1: seg factor 64x64 latent x 50 steps -> decode 512^2 -> uint8 hint ->
    image factor 64x64 x 50 steps -> decode, CFG 7.5 on both; against the fp32 oracle chain (run on the GPU in
    fp32, TF32 off). north_star bar: final decoded-image PSNR >= 35 dB."""
    import fgdm_oracle as orc
    from fgdm_b200.pipeline import FGDMPipeline
    (seg, sd_s), (img, sd_i) = fgdm, control
    S = 50
    pipe = FGDMPipeline(seg, img, seg_steps=S, img_steps=S)
    inp = {k: v.to(DEV) for k, v in synth.prompt_inputs([0], (64, 64)).items()}
    out = pipe(inp["c"], inp["uc"], inp["xT_seg"], inp["c2"], inp["uc2"], inp["xT_img"], return_all=True)
    ac = seg.alphas_cumprod
    z_seg, _ = orc.ddim_sample(lambda x, t, c: orc.unet_forward(sd_s, x, t, c, prefix="model.diffusion_model."), S,
                               (1, 4, 64, 64), inp["c"], inp["uc"], 7.5, inp["xT_seg"], ac)
    dec_seg = orc.vae_decode(sd_s, z_seg)
    seg_u8 = orc.seg_image_to_u8(dec_seg)
    hint = orc.u8_to_hint(seg_u8).to(DEV)
    z_img, _ = orc.ddim_sample(lambda x, t, c: orc.control_ldm_apply(sd_i, x, t, hint, c), S, (1, 4, 64, 64),
                               inp["c2"], inp["uc2"], 7.5, inp["xT_img"], ac, sequential_cfg=True)
    final = orc.final_image_to_u8(orc.vae_decode(sd_i, z_img))
    p_zs = psnr(out["z_seg"], z_seg, float(z_seg.abs().max()))
    p_seg = psnr(out["seg_u8"].float(), torch.from_numpy(seg_u8).float(), 255.0)
    p_img = psnr(out["img_u8"].float(), torch.from_numpy(final).float(), 255.0)
    print("full chain: seg latent PSNR", p_zs, "seg map PSNR", p_seg, "final image PSNR", p_img)
    assert p_seg > 35.0 and p_img > 35.0
