"""TEST INFRASTRUCTURE ONLY — runs in the BUILD CONTAINER (needs /root/reference; not on the GPU box).

1. builds the UNMODIFIED reference models (oracle/ref_harness.py) and loads the seeded synthetic weights
   of fgdm_b200.synth.make_state_dict (names/shapes checked equal to the reference's state_dict);
2. checks the restatement oracle/fgdm_oracle.py against the reference's own forward code;
3. writes the reference's outputs as golden fixtures to tests/golden/*.npz (inputs are re-derived from
   the seeds stored alongside, weights from the weight seed).

Usage: python oracle/make_golden.py [--only fgdm|control|misc|pcond|plms|dpm|clip]
"""
from __future__ import annotations

import argparse
import os
import sys
import time

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import fgdm_oracle as orc  # noqa: E402
import ref_harness as rh  # noqa: E402
from fgdm_b200 import golden_cases as gc  # noqa: E402
from fgdm_b200 import synth  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def rel(a, b):
    a, b = torch.as_tensor(a).double(), torch.as_tensor(b).double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-12))


def check(name, got, ref, tol=2e-4):
    r = rel(got, ref)
    print(f"  oracle vs reference  {name:34s} rel-to-max err {r:.3e}")
    assert r < tol, f"{name}: oracle restatement disagrees with the reference ({r})"
    return r


def save(name, **arrays):
    os.makedirs(GOLD, exist_ok=True)
    path = os.path.join(GOLD, name + ".npz")
    np.savez_compressed(path, **{k: (v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v))
                                 for k, v in arrays.items()})
    print(f"  wrote {os.path.relpath(path, ROOT)} ({os.path.getsize(path) / 1024:.0f} KiB)")


def load_weights(ref_model, shapes, seed):
    ref_shapes = {k: tuple(v.shape) for k, v in ref_model.state_dict().items()}
    for k, s in shapes.items():
        assert k in ref_shapes, f"{k} not a reference parameter"
        assert ref_shapes[k] == tuple(s), (k, ref_shapes[k], s)
    for pre in ("model.diffusion_model.", "control_model.", "first_stage_model.decoder.",
                "first_stage_model.post_quant_conv."):
        ours = {k for k in shapes if k.startswith(pre)}
        theirs = {k for k in ref_shapes if k.startswith(pre)}
        assert ours == theirs or not ours, (pre, sorted(theirs - ours)[:5], sorted(ours - theirs)[:5])
    sd = synth.make_state_dict(shapes, seed)
    missing, unexpected = ref_model.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected[:5]
    return sd


def golden_fgdm():
    print("== FG-DM factor (LatentDiffusion + UNetModel w/ Adapter + AutoencoderKL) ==")
    t0 = time.time()
    model = rh.build_fgdm()
    shapes = gc.fgdm_param_shapes()
    sd = load_weights(model, shapes, gc.FGDM_WEIGHT_SEED)
    print(f"  reference built + weights loaded in {time.time() - t0:.0f}s")
    unet = model.model.diffusion_model
    with torch.no_grad():
        # --- one UNet forward (per-row different timesteps)
        case = gc.unet_case()
        eps_ref = unet(case["x"], case["t"], context=case["ctx"])
        eps_orc = orc.unet_forward(sd, case["x"], case["t"], case["ctx"], prefix="model.diffusion_model.")
        check("UNetModel.forward", eps_orc, eps_ref)
        am = model.apply_model(case["x"], case["t"], case["ctx"], num=1)
        assert torch.equal(am, eps_ref)
        save("unet_fgdm", eps=eps_ref, weight_seed=gc.FGDM_WEIGHT_SEED)
        golden_pcond(model, sd)
        golden_plms(model, sd)
        # --- DDIM loop with CFG through the reference sampler
        from ldm.models.diffusion.ddim import DDIMSampler
        case = gc.ddim_case()
        sampler = rh.cpu_sampler(DDIMSampler(model))
        steps = []
        orig = sampler.p_sample_ddim

        def spy(*a, **k):
            out = orig(*a, **k)
            steps.append(out[0].clone())
            return out
        sampler.p_sample_ddim = spy
        x0, inter = sampler.sample(S=case["S"], conditioning=case["c"], batch_size=case["B"], shape=case["shape"],
                                   verbose=False, unconditional_guidance_scale=case["scale"],
                                   unconditional_conditioning=case["uc"], eta=0.0, x_T=case["x_T"], num=1)
        rec = []
        x0_orc, inter_orc = orc.ddim_sample(
            lambda x, t, c: orc.unet_forward(sd, x, t, c, prefix="model.diffusion_model."), case["S"],
            (case["B"],) + tuple(case["shape"]), case["c"], case["uc"], case["scale"], case["x_T"],
            model.alphas_cumprod, record=rec)
        check("DDIMSampler.sample final latent", x0_orc, x0, tol=1e-3)
        for i, (s_ref, (_, s_orc)) in enumerate(zip(steps, rec)):
            check(f"  step {i} x_prev", s_orc, s_ref, tol=1e-3)
        assert len(inter["x_inter"]) == len(inter_orc["x_inter"])
        check("intermediates pred_x0[-1]", inter_orc["pred_x0"][-1], inter["pred_x0"][-1], tol=1e-3)
        save("ddim_fgdm", x0=x0, steps=torch.stack(steps), pred_x0_last=inter["pred_x0"][-1],
             n_inter=len(inter["x_inter"]), ddim_timesteps=sampler.ddim_timesteps,
             ddim_alphas=np.asarray(sampler.ddim_alphas), ddim_alphas_prev=np.asarray(sampler.ddim_alphas_prev),
             ddim_sqrt_one_minus_alphas=np.asarray(sampler.ddim_sqrt_one_minus_alphas),
             ddim_sigmas=np.asarray(sampler.ddim_sigmas))
        # schedule tables for the step counts the configs use
        tabs = {}
        for S in (10, 20, 50):
            sampler.make_schedule(S, ddim_eta=0.0, verbose=False)
            tabs[f"t{S}"] = sampler.ddim_timesteps
            tabs[f"a{S}"] = np.asarray(sampler.ddim_alphas)
            tabs[f"ap{S}"] = np.asarray(sampler.ddim_alphas_prev)
            tabs[f"s1m{S}"] = np.asarray(sampler.ddim_sqrt_one_minus_alphas)
            o = orc.make_ddim_schedule(model.alphas_cumprod, S)
            assert (o["timesteps"] == tabs[f"t{S}"]).all()
            assert np.array_equal(np.asarray(o["alphas"]), tabs[f"a{S}"])
            assert np.array_equal(o["alphas_prev"], tabs[f"ap{S}"])
            assert np.array_equal(np.asarray(o["sqrt_one_minus_alphas"]), tabs[f"s1m{S}"])
        save("schedules", betas=model.betas, alphas_cumprod=model.alphas_cumprod,
             alphas_cumprod_prev=model.alphas_cumprod_prev, **tabs)
        ob = orc.make_schedule_buffers()
        assert torch.equal(ob["alphas_cumprod"], model.alphas_cumprod) and torch.equal(ob["betas"], model.betas)
        # --- VAE decode
        case = gc.vae_case()
        img_ref = model.decode_first_stage(case["z"])
        img_orc = orc.vae_decode(sd, case["z"])
        check("decode_first_stage", img_orc, img_ref)
        save("vae_decode", img=img_ref)


def golden_pcond(model=None, sd=None):
    """UNetModel.forward(pcond=...) (openaimodel.py:836-844): the joint FG-DM factor's adapter input."""
    if model is None:
        model = rh.build_fgdm()
        sd = load_weights(model, gc.fgdm_param_shapes(), gc.FGDM_WEIGHT_SEED)
    unet = model.model.diffusion_model
    with torch.no_grad():
        case = gc.pcond_case()
        eps_ref = unet(case["x"], case["t"], context=case["ctx"], pcond=case["pcond"])
        eps_orc = orc.unet_forward(sd, case["x"], case["t"], case["ctx"], prefix="model.diffusion_model.",
                                   pcond=case["pcond"])
        check("UNetModel.forward(pcond)", eps_orc, eps_ref)
        am = model.apply_model(case["x"], case["t"], case["ctx"], pcond=case["pcond"])
        assert torch.equal(am, eps_ref)
        save("unet_pcond", eps=eps_ref, weight_seed=gc.FGDM_WEIGHT_SEED)


def golden_plms(model=None, sd=None):
    """PLMSSampler.sample (plms.py) through the reference's own sampler class."""
    if model is None:
        model = rh.build_fgdm()
        sd = load_weights(model, gc.fgdm_param_shapes(), gc.FGDM_WEIGHT_SEED)
    from ldm.models.diffusion.plms import PLMSSampler
    with torch.no_grad():
        case = gc.plms_case()
        sampler = rh.cpu_sampler(PLMSSampler(model))
        steps = []
        orig = sampler.p_sample_plms

        def spy(*a, **k):
            out = orig(*a, **k)
            steps.append(out[0].clone())
            return out
        sampler.p_sample_plms = spy
        x0, inter = sampler.sample(S=case["S"], conditioning=case["c"], batch_size=case["B"], shape=case["shape"],
                                   verbose=False, unconditional_guidance_scale=case["scale"],
                                   unconditional_conditioning=case["uc"], eta=0.0, x_T=case["x_T"])
        rec = []
        x0_orc, inter_orc = orc.plms_sample(
            lambda x, t, c: orc.unet_forward(sd, x, t, c, prefix="model.diffusion_model."), case["S"],
            (case["B"],) + tuple(case["shape"]), case["c"], case["uc"], case["scale"], case["x_T"],
            model.alphas_cumprod, record=rec)
        check("PLMSSampler.sample final latent", x0_orc, x0, tol=1e-3)
        for i, (s_ref, (_, s_orc)) in enumerate(zip(steps, rec)):
            check(f"  plms step {i} x_prev", s_orc, s_ref, tol=1e-3)
        assert len(inter["x_inter"]) == len(inter_orc["x_inter"])
        save("plms_fgdm", x0=x0, steps=torch.stack(steps), pred_x0_last=inter["pred_x0"][-1],
             n_inter=len(inter["x_inter"]))


def golden_dpm(model=None, sd=None):
    """DPMSolverSampler.sample (dpm_solver/sampler.py) through the reference's own solver."""
    if model is None:
        model = rh.build_fgdm()
        sd = load_weights(model, gc.fgdm_param_shapes(), gc.FGDM_WEIGHT_SEED)
    from ldm.models.diffusion.dpm_solver import sampler as ds
    ds.DPMSolverSampler.register_buffer = lambda self, name, attr: setattr(self, name, attr)   # reference forces cuda
    with torch.no_grad():
        case = gc.dpm_case()
        smp = ds.DPMSolverSampler(model)
        x0, _ = smp.sample(S=case["S"], conditioning=case["c"], batch_size=case["B"], shape=case["shape"],
                           verbose=False, unconditional_guidance_scale=case["scale"],
                           unconditional_conditioning=case["uc"], eta=0.0, x_T=case["x_T"])
        rec = []
        x0_orc = orc.dpm_solver_sample(
            lambda x, t, c: orc.unet_forward(sd, x, t, c, prefix="model.diffusion_model."), case["S"],
            (case["B"],) + tuple(case["shape"]), case["c"], case["uc"], case["scale"], case["x_T"],
            model.alphas_cumprod, record=rec)
        check("DPMSolverSampler.sample final latent", x0_orc, x0, tol=1e-3)
        save("dpm_fgdm", x0=x0, steps=torch.stack(rec))


def golden_clip():
    """FrozenCLIPEmbedder.forward (modules.py:137-163) = transformers CLIPTextModel(...).last_hidden_state: the
    dependency itself, built from its config with the seeded synthetic weights (no checkpoint download)."""
    from transformers import CLIPTextConfig, CLIPTextModel
    cfg = CLIPTextConfig(vocab_size=49408, hidden_size=768, intermediate_size=3072, num_hidden_layers=12,
                         num_attention_heads=12, max_position_embeddings=77, hidden_act="quick_gelu",
                         layer_norm_eps=1e-5, projection_dim=768)
    model = CLIPTextModel(cfg).eval()
    sd = gc.clip_state_dict()
    ref_sd = model.state_dict()
    stripped = {k[len("transformer."):]: v for k, v in sd.items()}
    for k, v in stripped.items():
        assert k in ref_sd and tuple(ref_sd[k].shape) == tuple(v.shape), k
    missing = [k for k in ref_sd if k not in stripped and not k.endswith("position_ids")]
    assert not missing, missing
    model.load_state_dict(stripped, strict=False)
    case = gc.clip_case()
    with torch.no_grad():
        z_ref = model(input_ids=case["ids"]).last_hidden_state
        z_orc = orc.clip_text_forward(sd, case["ids"])
    check("CLIPTextModel.last_hidden_state", z_orc, z_ref)
    save("clip_text", z=z_ref, weight_seed=gc.CLIP_WEIGHT_SEED)


def golden_control():
    print("== image factor (ControlLDM = ControlNet + ControlledUnetModel) ==")
    t0 = time.time()
    model = rh.build_controlldm()
    shapes = gc.control_param_shapes()
    sd = load_weights(model, shapes, gc.CONTROL_WEIGHT_SEED)
    print(f"  reference built + weights loaded in {time.time() - t0:.0f}s")
    with torch.no_grad():
        case = gc.control_case()
        cond = {"c_concat": [case["hint"]], "c_crossattn": [case["ctx"]]}
        eps_ref = model.apply_model(case["x"], case["t"], cond)
        eps_orc = orc.control_ldm_apply(sd, case["x"], case["t"], case["hint"], case["ctx"])
        check("ControlLDM.apply_model", eps_orc, eps_ref)
        ctrl_ref = model.control_model(x=case["x"], hint=case["hint"], timesteps=case["t"], context=case["ctx"])
        ctrl_orc = orc.controlnet_forward(sd, case["x"], case["hint"], case["t"], case["ctx"], prefix="control_model.")
        for i, (a, b) in enumerate(zip(ctrl_orc, ctrl_ref)):
            check(f"  ControlNet out[{i}]", a, b)
        # control_scales / only_mid_control variants
        model.control_scales = list(gc.CONTROL_SCALES_VARIANT)
        eps_sc = model.apply_model(case["x"], case["t"], cond)
        check("apply_model (control_scales)", orc.control_ldm_apply(sd, case["x"], case["t"], case["hint"], case["ctx"],
                                                                    control_scales=list(gc.CONTROL_SCALES_VARIANT)), eps_sc)
        model.control_scales = [1.0] * 13
        model.only_mid_control = True
        eps_mid = model.apply_model(case["x"], case["t"], cond)
        check("apply_model (only_mid_control)", orc.control_ldm_apply(sd, case["x"], case["t"], case["hint"], case["ctx"],
                                                                      only_mid_control=True), eps_mid)
        model.only_mid_control = False
        save("control_apply", eps=eps_ref, eps_scaled=eps_sc, eps_only_mid=eps_mid,
             ctrl_absmax=np.asarray([float(c.abs().max()) for c in ctrl_ref]), ctrl0=ctrl_ref[0], ctrl12=ctrl_ref[12])
        # --- ddim_hacked sampler
        from controlnet.cldm.ddim_hacked import DDIMSampler
        case = gc.control_ddim_case()
        sampler = rh.cpu_sampler(DDIMSampler(model))
        cond = {"c_concat": [case["hint"]], "c_crossattn": [case["c"]]}
        un = {"c_concat": [case["hint"]], "c_crossattn": [case["uc"]]}
        x0, inter = sampler.sample(case["S"], case["B"], case["shape"], cond, verbose=False, eta=0.0,
                                   unconditional_guidance_scale=case["scale"], unconditional_conditioning=un,
                                   x_T=case["x_T"])
        x0_orc, _ = orc.ddim_sample(
            lambda x, t, c: orc.control_ldm_apply(sd, x, t, case["hint"], c), case["S"],
            (case["B"],) + tuple(case["shape"]), case["c"], case["uc"], case["scale"], case["x_T"],
            model.alphas_cumprod, sequential_cfg=True)
        check("ddim_hacked sample final latent", x0_orc, x0, tol=1e-3)
        save("ddim_control", x0=x0, pred_x0_last=inter["pred_x0"][-1])
        # --- ControlNet guess mode (initialize_cn.py:87,92 with guess_mode=True): the unconditional half has
        # c_concat=None (ddim_hacked.py:191 -> cldm.py:841-842, control=None) and the scales decay 0.825**(12-i)
        gscales = list(gc.GUESS_MODE_SCALES)
        model.control_scales = gscales
        un_g = {"c_concat": None, "c_crossattn": [case["uc"]]}
        xg, _ = sampler.sample(case["S"], case["B"], case["shape"], cond, verbose=False, eta=0.0,
                               unconditional_guidance_scale=case["scale"], unconditional_conditioning=un_g,
                               x_T=case["x_T"])
        model.control_scales = [1.0] * 13
        xg_orc, _ = orc.ddim_sample(
            lambda x, t, c: orc.control_ldm_apply(sd, x, t, case["hint"] if c is case["c"] else None, c,
                                                  control_scales=gscales), case["S"],
            (case["B"],) + tuple(case["shape"]), case["c"], case["uc"], case["scale"], case["x_T"],
            model.alphas_cumprod, sequential_cfg=True)
        check("ddim_hacked guess-mode final latent", xg_orc, xg, tol=1e-3)
        assert float((xg - x0).abs().max()) > 1e-3, "guess mode must differ from the guided-uncond run"
        save("ddim_control_guess", x0=xg)


def golden_misc():
    print("== factor hand-off (uint8 quantisation, cv2 bilinear resize) ==")
    import cv2
    case = gc.handoff_case()
    x = case["img"]
    # scripts/txt2img_fgdm_inference.py:245,249-259
    xs = torch.clamp((x + 1.0) / 2.0, min=0.0, max=1.0)
    u_ref = []
    for s in xs:
        from einops import rearrange
        u_ref.append((255. * rearrange(s.cpu().numpy(), 'c h w -> h w c')).astype(np.uint8))
    u_ref = np.stack(u_ref)
    assert np.array_equal(orc.seg_image_to_u8(x), u_ref)
    rs_ref = np.stack([cv2.resize(u, (case["dw"], case["dh"]), interpolation=cv2.INTER_LINEAR) for u in u_ref])
    rs_orc = np.stack([orc.resize_u8_bilinear(u, case["dh"], case["dw"]) for u in u_ref])
    print("  cv2.resize vs oracle: mismatching bytes =", int((rs_ref != rs_orc).sum()))
    assert np.array_equal(rs_ref, rs_orc)
    rs2_ref = np.stack([cv2.resize(u, (96, 80), interpolation=cv2.INTER_LINEAR) for u in u_ref])
    rs2_orc = np.stack([orc.resize_u8_bilinear(u, 80, 96) for u in u_ref])
    assert np.array_equal(rs2_ref, rs2_orc)
    # controlnet/initialize_cn.py:78-80 and :98-99
    import einops
    control = torch.from_numpy(rs_ref.copy()).float() / 255.0
    control = einops.rearrange(control, 'b h w c -> b c h w').clone()
    assert torch.equal(orc.u8_to_hint(rs_ref), control)
    fin_ref = (einops.rearrange(x, 'b c h w -> b h w c') * 127.5 + 127.5).cpu().numpy().clip(0, 255).astype(np.uint8)
    assert np.array_equal(orc.final_image_to_u8(x), fin_ref)
    save("handoff", seg_u8=u_ref, resized=rs_ref, resized_odd=rs2_ref, final_u8=fin_ref)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="all")
    a = ap.parse_args()
    torch.set_num_threads(os.cpu_count() or 8)
    if a.only in ("all", "misc"):
        golden_misc()
    if a.only in ("all", "fgdm"):
        golden_fgdm()
    if a.only == "pcond":
        golden_pcond()
    if a.only == "plms":
        golden_plms()
    if a.only in ("all", "dpm"):
        golden_dpm()
    if a.only in ("all", "clip"):
        golden_clip()
    if a.only in ("all", "control"):
        golden_control()
    print("golden fixtures written; oracle restatement agrees with the reference")
