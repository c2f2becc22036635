"""Drop-in boundary: the reference's OWN driver statements run against the fgdm_b200 classes.

The function bodies below are the statement sequences of
  scripts/txt2img_fgdm_inference.py:23-38 (load_model_from_config), :185-190, :216-259 (sampling loop)
  controlnet/cldm/model.py:24-28 (create_model), controlnet/initialize_cn.py:25-41 (initialize_controlnet),
  controlnet/initialize_cn.py:74-102 (process)
copied as they stand in the reference (only file paths / argparse values are test fixtures), with the
`target:` strings of the YAMLs resolved through fgdm_b200.ldm.TARGET_ALIASES. No GPU here: the C-ABI library is
replaced by a stub whose kernels do nothing (tensors keep whatever torch.empty gave them), so this checks the
host-side contract — return values that are unpacked, nn.Module look-alikes, kwargs that are passed — not numbers.
"""
import os
import random
import sys
import types

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

from fgdm_b200 import _lib, ops, synth  # noqa: E402
from fgdm_b200 import golden_cases as gc  # noqa: E402
from fgdm_b200.ldm import instantiate_from_config  # noqa: E402


class _StubLib:
    """Every fgdm_* entry point 'succeeds' without touching memory."""

    def __getattr__(self, name):
        if name == "fgdm_gemm_stats_slots":
            return lambda *a: 4
        if name == "fgdm_launch_count":
            return lambda *a: 0
        if name == "fgdm_last_error":
            return lambda *a: b""
        if name.startswith("fgdm_"):
            return lambda *a: 0
        raise AttributeError(name)


class StubTextEncoder(torch.nn.Module):
    """Stands in for the reference's FrozenCLIPEmbedder object (tokenizer + CLIP weights are downloads): strings ->
    [B, 77, 768], like modules.py:137-163."""

    def __init__(self, **kw):
        super().__init__()
        self.proj = torch.nn.Linear(4, 4)           # something for load_state_dict(strict=False) to see

    def encode(self, text):
        assert isinstance(text, (list, tuple)) and all(isinstance(s, str) for s in text)
        return torch.stack([torch.full((77, 768), float(len(s) % 7)) for s in text])

    forward = encode


SMALL_UNET = dict(gc.UNET_PARAMS, model_channels=64)
SMALL_CN = dict(gc.CONTROLNET_PARAMS, model_channels=64)
SMALL_VAE = dict(gc.VAE_PARAMS, ddconfig=dict(gc.VAE_PARAMS["ddconfig"], ch=64))
COND_STAGE = dict(target="test_boundary_cpu.StubTextEncoder")


def _fgdm_yaml():
    """models/config.yaml with the reference's own target strings (resolved through TARGET_ALIASES)."""
    cfg = gc.fgdm_config()
    cfg["target"] = "ldm.models.diffusion.ddpm.LatentDiffusion"
    p = cfg["params"]
    p["unet_config"] = dict(target="ldm.modules.diffusionmodules.openaimodel.UNetModel", params=dict(SMALL_UNET))
    p["first_stage_config"] = dict(target="ldm.models.autoencoder.AutoencoderKL", params=dict(SMALL_VAE))
    p["cond_stage_config"] = dict(COND_STAGE)
    p["device"] = "cpu"
    return types.SimpleNamespace(model=cfg)


def _control_yaml():
    cfg = gc.control_config()
    cfg["target"] = "cldm.cldm.ControlLDM"
    p = cfg["params"]
    p["control_stage_config"] = dict(target="cldm.cldm.ControlNet", params=dict(SMALL_CN))
    p["unet_config"] = dict(target="cldm.cldm.ControlledUnetModel", params=dict(SMALL_UNET))
    p["first_stage_config"] = dict(target="ldm.models.autoencoder.AutoencoderKL", params=dict(SMALL_VAE))
    p["cond_stage_config"] = dict(COND_STAGE)
    p["device"] = "cpu"
    return types.SimpleNamespace(model=cfg)


@pytest.fixture()
def stubbed(monkeypatch):
    sys.modules.setdefault("test_boundary_cpu", sys.modules[__name__])
    monkeypatch.setattr(_lib, "load", lambda: _StubLib())
    monkeypatch.setattr(_lib, "require_device", lambda: None)
    monkeypatch.setattr(ops, "_stream", lambda: 0)
    monkeypatch.setenv("FGDM_USE_GRAPH", "0")          # CUDA graphs need a device; the per-step path is the reference's flow
    monkeypatch.setattr(torch.Tensor, "cuda", lambda self, *a, **k: self)
    torch.manual_seed(0)


def _reference_style_ckpt(model, seed, path, extra=True, wrap=True):
    """A checkpoint with the reference's key set: ours + the modules a reference ckpt also holds."""
    sd = synth.make_state_dict(model.param_shapes(), seed)
    if extra:
        sd["first_stage_model.encoder.conv_in.weight"] = torch.zeros(8, 3, 3, 3)
        sd["first_stage_model.quant_conv.weight"] = torch.zeros(8, 8, 1, 1)
        sd["first_stage_model.loss.logvar"] = torch.zeros(())
        sd["cond_stage_model.proj.weight"] = torch.zeros(4, 4)
        sd["cond_stage_model.proj.bias"] = torch.zeros(4)
        sd["betas"] = torch.zeros(1000)
        sd["alphas_cumprod"] = torch.zeros(1000)
        sd["model_ema.decay"] = torch.tensor(0.9999)
        sd["model_ema.num_updates"] = torch.tensor(0)
    torch.save({"state_dict": sd} if wrap else sd, path)
    return sd


def test_txt2img_fgdm_inference_call_sites(stubbed, tmp_path, capsys):
    import cv2
    from contextlib import nullcontext
    from einops import rearrange
    from fgdm_b200.ddim import DDIMSampler
    from fgdm_b200.plms import PLMSSampler  # noqa: F401  (imported by the script)

    config = _fgdm_yaml()
    ckpt = str(tmp_path / "fgdm.ckpt")
    _reference_style_ckpt(instantiate_from_config(config.model), 3, ckpt)

    # ---- scripts/txt2img_fgdm_inference.py:23-38 ------------------------------------------------
    def load_model_from_config(config, ckpt, verbose=False):
        print(f"Loading model from {ckpt}")
        pl_sd = torch.load(ckpt, map_location="cpu")
        sd = pl_sd["state_dict"]
        model = instantiate_from_config(config.model)
        m, u = model.load_state_dict(sd, strict=False)
        if len(m) > 0 and verbose:
            print("missing keys:")
            print(m)
        if len(u) > 0 and verbose:
            print("unexpected keys:")
            print(u)

        model.cuda()
        model.eval()
        return model, m, u

    model, m, u = load_model_from_config(config, ckpt, verbose=True)
    assert m == [] and u == []                                  # encoder / loss / EMA / schedule keys are not "unexpected"

    opt = types.SimpleNamespace(plms=False, n_samples=2, scale=7.5, C=4, H=128, W=128, ddim_steps=4, ddim_eta=0.0,
                                n_iter=1, fixed_code=False, precision="full", prompt="a photo of a cat")
    # ---- :185-190 -------------------------------------------------------------------------------
    device = torch.device("cuda") if torch.cuda.is_available() else torch.device("cpu")
    model = model.to(device)

    if opt.plms:
        sampler = PLMSSampler(model)
    else:
        sampler = DDIMSampler(model)
    data = [opt.prompt]
    start_noise = None
    precision_scope = nullcontext
    # ---- :216-259 -------------------------------------------------------------------------------
    with torch.no_grad():
        with precision_scope("cuda"):
            with model.ema_scope():
                uc = None
                if opt.scale != 1.0:
                    uc = model.get_learned_conditioning(opt.n_samples * [""])
                for n, pr in enumerate(data):
                    prompt = pr
                    for niter in range(opt.n_iter):
                        c = model.get_learned_conditioning(opt.n_samples * [prompt])
                        shape = [opt.C, opt.H // 8, opt.W // 8]
                        samples_ddim, _ = sampler.sample(S=opt.ddim_steps,
                                                         conditioning=c,
                                                         batch_size=opt.n_samples,
                                                         shape=shape,
                                                         num=1,
                                                         x_T=start_noise,
                                                         verbose=False,
                                                         unconditional_guidance_scale=opt.scale,
                                                         unconditional_conditioning=uc,
                                                         eta=opt.ddim_eta)

                        x_samples_ddim = model.decode_first_stage(samples_ddim)
                        x_samples_ddim = torch.clamp((x_samples_ddim + 1.0) / 2.0, min=0.0, max=1.0)
                        gen_samples = []
                        for x_sample in x_samples_ddim:
                            x_sample = 255. * rearrange(x_sample.cpu().numpy(), 'c h w -> h w c')
                            num = x_sample.shape[-1] // 3
                            for idx in range(num):
                                x_sample_idx = np.nan_to_num(x_sample[:, :, idx * 3:(idx + 1) * 3]).astype(np.uint8)
                                resized_img = cv2.resize(x_sample_idx, (512, 512), interpolation=cv2.INTER_LINEAR)
                                gen_samples.append(resized_img)
                        gen_samples = np.stack(gen_samples)[:5]
    assert c.shape == (2, 77, 768) and uc.shape == (2, 77, 768)
    assert samples_ddim.shape == (2, 4, 16, 16) and x_samples_ddim.shape == (2, 3, 128, 128)
    assert gen_samples.shape == (2, 512, 512, 3)


def test_initialize_cn_call_sites(stubbed, tmp_path):
    import einops
    from fgdm_b200.ddim import DDIMSampler          # controlnet.cldm.ddim_hacked.DDIMSampler's drop-in

    yaml = _control_yaml()
    pth = str(tmp_path / "fgdm_control_sd15_seg.pth")
    _reference_style_ckpt(instantiate_from_config(yaml.model), 4, pth, wrap=False)

    # ---- controlnet/cldm/model.py:9-28 ----------------------------------------------------------
    def get_state_dict(d):
        return d.get('state_dict', d)

    def load_state_dict(ckpt_path, location='cpu'):
        state_dict = get_state_dict(torch.load(ckpt_path, map_location=torch.device(location)))
        state_dict = get_state_dict(state_dict)
        return state_dict

    def create_model(config_path):
        config = yaml                                # OmegaConf.load(config_path)
        model = instantiate_from_config(config.model).cpu()
        return model

    # ---- controlnet/initialize_cn.py:25-41 ------------------------------------------------------
    def initialize_controlnet(cond='seg'):
        config_path = './controlnet/models/cldm_v15_canny.yaml'
        model = create_model(config_path).cpu()
        if cond == 'seg':
            m, u = model.load_state_dict(load_state_dict(pth, location='cpu'), strict=False)
        else:
            raise NotImplementedError
        print("Missing keys: ", m)
        model = model.cuda()
        ddim_sampler = DDIMSampler(model)
        return model, ddim_sampler, m, u

    cn_model, cn_ddim_sampler, m, u = initialize_controlnet(cond='seg')
    assert list(m) == [] and list(u) == []
    cn_model.low_vram_shift(is_diffusing=False)      # config.save_memory call sites (cldm.py:933-943)
    cn_model.low_vram_shift(is_diffusing=True)

    # ---- controlnet/initialize_cn.py:74-102 -----------------------------------------------------
    def process(model, ddim_sampler, input_image, prompt, a_prompt, n_prompt, num_samples, num_repeats, image_resolution,
                detect_resolution, ddim_steps, guess_mode, strength, scale, seed, eta, class_map=None, spath='', index=0):
        with torch.no_grad():
            mask = None
            B, H, W, C = input_image.shape
            control = torch.from_numpy(input_image.copy()).float().cuda() / 255.0
            control = torch.cat([control for _ in range(num_repeats)], dim=0)
            control = einops.rearrange(control, 'b h w c -> b c h w').clone()

            if seed == -1:
                seed = random.randint(0, 65535)
            torch.manual_seed(seed)                   # pytorch_lightning.seed_everything

            cond = {"c_concat": [control], "c_crossattn": [model.get_learned_conditioning([prompt + ', ' + a_prompt] * num_samples)]}
            un_cond = {"c_concat": None if guess_mode else [control], "c_crossattn": [model.get_learned_conditioning([n_prompt] * num_samples)]}
            shape = (4, H // 8, W // 8)

            model.control_scales = [strength * (0.825 ** float(12 - i)) for i in range(13)] if guess_mode else ([strength] * 13)
            samples, intermediates = ddim_sampler.sample(ddim_steps, num_samples,
                                                         shape, cond, verbose=False, eta=eta,
                                                         unconditional_guidance_scale=scale,
                                                         unconditional_conditioning=un_cond,
                                                         cond_mask=mask)

            x_samples = model.decode_first_stage(samples)
            x_samples = (einops.rearrange(x_samples, 'b c h w -> b h w c') * 127.5 + 127.5).cpu().numpy().clip(0, 255).astype(np.uint8)

            results = [x_samples[i] for i in range(num_samples)]
        return results

    gen_samples = np.random.randint(0, 255, (2, 128, 128, 3), dtype=np.uint8)
    for guess_mode in (False, True):
        results = process(cn_model, cn_ddim_sampler, gen_samples, "a cat", 'best quality, extremely detailed', 'lowres', 2, 1,
                          512, eta=0.0, detect_resolution=512, ddim_steps=2, guess_mode=guess_mode, strength=1.0, scale=9.0,
                          seed=-1, spath="x", index=0)
        assert len(results) == 2 and results[0].shape == (128, 128, 3) and results[0].dtype == np.uint8


def test_load_state_dict_torch_semantics(stubbed):
    """strict=True raises listing the keys; strict=False returns (missing, unexpected) and leaves the net unusable
    until the missing weights arrive; shape mismatches raise under both."""
    model = instantiate_from_config(_fgdm_yaml().model)
    sd = synth.make_state_dict(model.param_shapes(), 9)
    r = model.load_state_dict(sd)                                   # strict=True default, complete dict
    assert r.missing_keys == [] and r.unexpected_keys == [] and "All keys matched" in repr(r)
    k0 = "model.diffusion_model.out.2.weight"
    part = {k: v for k, v in sd.items() if k != k0}
    part["model.diffusion_model.not_a_layer.weight"] = torch.zeros(1)
    with pytest.raises(RuntimeError, match="not_a_layer|out.2.weight"):
        model.load_state_dict(part, strict=True)
    model2 = instantiate_from_config(_fgdm_yaml().model)
    m, u = model2.load_state_dict(part, strict=False)
    assert m == [k0] and u == ["model.diffusion_model.not_a_layer.weight"]
    with pytest.raises(RuntimeError, match="missing"):
        model2.apply_model(torch.zeros(1, 4, 16, 16), torch.zeros(1, dtype=torch.int64), torch.zeros(1, 77, 768))
    bad = dict(sd)
    bad[k0] = torch.zeros(4, 64, 5, 5)
    with pytest.raises(RuntimeError, match="size mismatch"):
        model2.load_state_dict(bad, strict=False)
    unet = model.model.diffusion_model
    assert unet.cpu() is unet and unet.cuda() is unet and unet.eval() is unet and model.cpu() is model


def test_use_ema_loads_the_ema_shadows(stubbed):
    """LitEma shadows (`model_ema.<name without dots>`, ema.py:16-21) are what sampling inside ema_scope() uses."""
    cfg = _fgdm_yaml().model
    cfg["params"]["use_ema"] = True
    model = instantiate_from_config(cfg)
    sd = synth.make_state_dict(model.param_shapes(), 9)
    pre = "model.diffusion_model."
    marker = pre + "out.2.bias"
    for k in list(sd):
        if k.startswith(pre):
            sd["model_ema." + k[len("model."):].replace(".", "")] = sd[k] + 1.0
    m, u = model.load_state_dict(sd, strict=False)
    assert m == [] and u == [] and model._ema_loaded
    got = model.model.diffusion_model.out[2].b32[:4]
    assert torch.allclose(got, sd[marker] + 1.0)                     # the EMA value, not the raw one
    with model.ema_scope("sampling"):
        pass


def test_get_learned_conditioning_variants(stubbed):
    from fgdm_b200.clip import FrozenCLIPEmbedder
    model = instantiate_from_config(_fgdm_yaml().model)
    assert model.get_learned_conditioning(["a", "bb"]).shape == (2, 77, 768)
    ctx = torch.zeros(2, 77, 768)
    assert model.get_learned_conditioning(ctx) is ctx
    # a ready-made encoder object is kept as is (the reference's own FrozenCLIPEmbedder instance)
    cfg = _fgdm_yaml().model
    enc = StubTextEncoder()
    cfg["params"]["cond_stage_config"] = enc
    assert instantiate_from_config(cfg).cond_stage_model is enc
    # no encoder configured: strings are an error, not silently ignored
    cfg = _fgdm_yaml().model
    cfg["params"]["cond_stage_config"] = dict(target="torch.nn.Identity")
    with pytest.raises(NotImplementedError):
        instantiate_from_config(cfg).get_learned_conditioning(["a"])
    # our CLIP tower tokenises strings through an injected tokenizer exactly as modules.py:154-156 calls it
    calls = {}

    def tok(text, **kw):
        calls.update(kw, text=text)
        return {"input_ids": torch.full((len(text), 77), 49407, dtype=torch.int64)}

    clip = FrozenCLIPEmbedder(device="cpu", tokenizer=tok, num_hidden_layers=1)
    ids = clip.tokenize(["x", "y"])
    assert ids.shape == (2, 77) and calls["padding"] == "max_length" and calls["max_length"] == 77 and calls["truncation"]
    with pytest.raises(NotImplementedError, match="offline|tokenizer"):
        FrozenCLIPEmbedder(device="cpu", num_hidden_layers=1, version="no/such-clip").tokenize(["x"])
