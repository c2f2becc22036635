"""Diffusion-model wrappers with the reference's method signatures:

  LatentDiffusion   ldm/models/diffusion/ddpm.py:512-573 (ctor), apply_model :1035-1136,
                    decode_first_stage :831-889, get_learned_conditioning :663-674, register_schedule :175-227
  ControlLDM        controlnet/cldm/cldm.py:816-849
  DiffusionWrapper  ldm/models/diffusion/ddpm.py:1821-1848

plus `instantiate_from_config` (ldm/util.py:78-93), the reference's plugin mechanism: a maintainer swaps
the YAML `target:` strings to `fgdm_b200.ldm.LatentDiffusion`, `fgdm_b200.nets.UNetModel`, ... and
`load_state_dict`s the reference checkpoint unchanged.
"""
from __future__ import annotations

import importlib

import numpy as np
import torch

from . import _lib, ops
from .nets import AutoencoderKL, ControlledUnetModel, ControlNet, UNetModel  # noqa: F401  (targets)

# reference target string -> our class, so the reference's own YAMLs can be used verbatim
TARGET_ALIASES = {
    "ldm.models.diffusion.ddpm.LatentDiffusion": "fgdm_b200.ldm.LatentDiffusion",
    "ldm.modules.diffusionmodules.openaimodel.UNetModel": "fgdm_b200.nets.UNetModel",
    "ldm.models.autoencoder.AutoencoderKL": "fgdm_b200.nets.AutoencoderKL",
    "controlnet.ldm.models.autoencoder.AutoencoderKL": "fgdm_b200.nets.AutoencoderKL",
    "controlnet.cldm.cldm.ControlLDM": "fgdm_b200.ldm.ControlLDM",
    "controlnet.cldm.cldm.ControlNet": "fgdm_b200.nets.ControlNet",
    "controlnet.cldm.cldm.ControlledUnetModel": "fgdm_b200.nets.ControlledUnetModel",
    "cldm.cldm.ControlLDM": "fgdm_b200.ldm.ControlLDM",
    "cldm.cldm.ControlNet": "fgdm_b200.nets.ControlNet",
    "cldm.cldm.ControlledUnetModel": "fgdm_b200.nets.ControlledUnetModel",
    "ldm.modules.encoders.modules.FrozenCLIPEmbedder": "fgdm_b200.clip.FrozenCLIPEmbedder",
    "controlnet.ldm.modules.encoders.modules.FrozenCLIPEmbedder": "fgdm_b200.clip.FrozenCLIPEmbedder",
}


def get_obj_from_str(string, reload=False):
    """ldm/util.py:88-93."""
    string = TARGET_ALIASES.get(string, string)
    module, cls = string.rsplit(".", 1)
    return getattr(importlib.import_module(module, package=None), cls)


def instantiate_from_config(config):
    """ldm/util.py:78-85."""
    if "target" not in config:
        if config in ("__is_first_stage__", "__is_unconditional__"):
            return None
        raise KeyError("Expected key `target` to instantiate.")
    return get_obj_from_str(config["target"])(**config.get("params", dict()))


class IdentityCondStage:
    """Stand-in used when the cond stage is `torch.nn.Identity` or omitted: synthetic [B,77,768] contexts
    pass through (the CLIP text encoder runs once per prompt, outside this path)."""

    def encode(self, c):
        return c

    def __call__(self, c):
        return c


class DiffusionWrapper:
    """ddpm.py:1821-1848: routes `c_crossattn` to the UNet's `context`."""

    def __init__(self, diff_model_config, conditioning_key):
        self.diffusion_model = instantiate_from_config(diff_model_config)
        self.conditioning_key = conditioning_key
        if conditioning_key not in (None, "crossattn"):
            raise NotImplementedError(f"conditioning_key={conditioning_key!r}: FG-DM sampling uses 'crossattn'")

    def __call__(self, x, t, c_concat=None, c_crossattn=None, **kw):
        if self.conditioning_key is None:
            raise NotImplementedError("unconditional UNet call is outside the FG-DM sampling path")
        cc = c_crossattn[0] if len(c_crossattn) == 1 else torch.cat(c_crossattn, 1)
        return self.diffusion_model(x, t, context=cc, **kw)


class LatentDiffusion:
    def __init__(self, first_stage_config=None, cond_stage_config=None, unet_config=None, timesteps=1000,
                 beta_schedule="linear", linear_start=1e-4, linear_end=2e-2, cosine_s=8e-3, given_betas=None,
                 scale_factor=1.0, conditioning_key=None, parameterization="eps", channels=3, image_size=256,
                 num_timesteps_cond=None, cond_stage_key="image", cond_stage_trainable=False, concat_mode=True,
                 cond_stage_forward=None, scale_by_std=False, first_stage_key="image", use_ema=True, device="cuda",
                 **ignored):
        if parameterization != "eps":
            raise NotImplementedError("only eps-parameterisation is on the FG-DM sampling path")
        if beta_schedule != "linear":
            raise NotImplementedError("only the linear beta schedule (models/config.yaml) is provided")
        self.parameterization = parameterization
        self.scale_factor = float(scale_factor)
        self.channels, self.image_size = channels, image_size
        self.device = torch.device(device)
        self.model = DiffusionWrapper(unet_config, conditioning_key)
        self.first_stage_model = instantiate_from_config(first_stage_config) if first_stage_config else None
        self.cond_stage_model = IdentityCondStage()
        if isinstance(cond_stage_config, dict) and str(cond_stage_config.get("target", "")).endswith("FrozenCLIPEmbedder"):
            # models/config.yaml:74-75 / ddpm.py:626-640: the text encoder, on the same kernels (fgdm_b200/clip.py)
            params = dict(cond_stage_config.get("params", {}), device=str(self.device))
            self.cond_stage_model = instantiate_from_config(dict(cond_stage_config, params=params))
        self.register_schedule(given_betas, timesteps, linear_start, linear_end)

    # ddpm.py:175-203 with util.py:21-25 (float64 numpy -> float32 buffers)
    def register_schedule(self, given_betas=None, timesteps=1000, linear_start=1e-4, linear_end=2e-2):
        if given_betas is not None:
            betas = np.asarray(given_betas, dtype=np.float64)
        else:
            betas = (torch.linspace(linear_start ** 0.5, linear_end ** 0.5, timesteps, dtype=torch.float64) ** 2).numpy()
        alphas_cumprod = np.cumprod(1.0 - betas, axis=0)
        alphas_cumprod_prev = np.append(1.0, alphas_cumprod[:-1])
        self.num_timesteps = int(betas.shape[0])
        self.linear_start, self.linear_end = linear_start, linear_end
        # host copies (the sampler builds its tables on the host) — `.device` reports the model device
        self.betas = torch.tensor(betas, dtype=torch.float32)
        self.alphas_cumprod = torch.tensor(alphas_cumprod, dtype=torch.float32)
        self.alphas_cumprod_prev = torch.tensor(alphas_cumprod_prev, dtype=torch.float32)

    # -- weights ----------------------------------------------------------------------------------
    def _parts(self):
        parts = {"model.diffusion_model.": self.model.diffusion_model}
        if self.first_stage_model is not None:
            parts["first_stage_model."] = self.first_stage_model
        return parts

    def param_shapes(self):
        out = {}
        for pre, net in self._parts().items():
            out.update({pre + k: v for k, v in net.param_shapes().items()})
        return out

    def load_state_dict(self, sd, strict=False):
        """Accepts the reference's checkpoint keys (`model.diffusion_model.*`, `first_stage_model.*`,
        `control_model.*`); encoder / cond-stage / loss keys are ignored."""
        _lib.require_device()
        for pre, net in self._parts().items():
            net.load_state_dict(sd, strict=False, prefix=pre, device=self.device)
        pre = "cond_stage_model."
        if hasattr(self.cond_stage_model, "text_model") and any(k.startswith(pre) for k in sd):
            self.cond_stage_model.load_state_dict({k[len(pre):]: v for k, v in sd.items() if k.startswith(pre)})
        return self

    def eval(self):
        return self

    def cuda(self, *a):
        return self

    def to(self, *a, **k):
        return self

    def ema_scope(self, *a, **k):
        import contextlib
        return contextlib.nullcontext()

    # -- reference methods ------------------------------------------------------------------------
    def get_learned_conditioning(self, c):
        """ddpm.py:663-674. Tensors (pre-encoded / synthetic contexts) pass through."""
        if torch.is_tensor(c) and c.is_floating_point():
            return c
        if hasattr(self.cond_stage_model, "text_model"):
            return self.cond_stage_model.encode(c)            # token ids -> [B,77,768] (FrozenCLIPEmbedder)
        raise NotImplementedError("no text encoder configured (cond_stage_config: FrozenCLIPEmbedder); pass [B,77,768] "
                                  "context tensors")

    def apply_model(self, x_noisy, t, cond, return_ids=False, **kw):
        """ddpm.py:1035-1044,1130-1136: tensor/list cond -> {'c_crossattn': [c]} -> DiffusionWrapper."""
        if return_ids:
            raise NotImplementedError("return_ids (joint caption head) is outside the FG-DM sampling path")
        if not isinstance(cond, dict):
            if not isinstance(cond, list):
                cond = [cond]
            cond = {"c_crossattn": cond}
        kw.pop("num", None)
        with torch.no_grad():
            out = self.model(x_noisy, t, **cond, **kw)
        return out.to(x_noisy.dtype)

    def decode_first_stage(self, z, predict_cids=False, force_not_quantize=False, n=None):
        """ddpm.py:831-889: z / scale_factor -> AutoencoderKL.decode. Returns fp32 NCHW in ~[-1, 1]."""
        if predict_cids:
            raise NotImplementedError("VQ codebook decoding is outside the FG-DM path (KL-VAE only)")
        return self.first_stage_model.decode(z, in_scale=1.0 / self.scale_factor)


class ControlLDM(LatentDiffusion):
    """controlnet/cldm/cldm.py:816-849."""

    def __init__(self, control_stage_config, control_key="hint", only_mid_control=False, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.control_model = instantiate_from_config(control_stage_config)
        self.control_key = control_key
        self.only_mid_control = only_mid_control
        self.control_scales = [1.0] * 13

    def _parts(self):
        parts = super()._parts()
        parts["control_model."] = self.control_model
        return parts

    def apply_model(self, x_noisy, t, cond, *args, **kwargs):
        """cldm.py:836-849. The control residual adds of ControlledUnetModel.forward (cldm.py:39-47) and
        the `c * scale` of :846 ride the zero-conv epilogues."""
        assert isinstance(cond, dict)
        unet = self.model.diffusion_model
        cc = cond["c_crossattn"]
        ctx = cc[0] if len(cc) == 1 else torch.cat(cc, 1)
        with torch.no_grad():
            if cond.get("c_concat") is None:
                eps = unet(x_noisy, timesteps=t, context=ctx, control=None, only_mid_control=self.only_mid_control)
            else:
                hint = cond["c_concat"]
                hint = hint[0] if len(hint) == 1 else torch.cat(hint, 1)
                eps = self.control_step(x_noisy, t, unet.project_context(ctx), self.control_model.project_context(ctx),
                                        self.control_model.encode_hint(hint), x_noisy.shape[0])
        return eps.to(x_noisy.dtype)

    def control_step(self, x, t, kv_unet, kv_ctrl, guided_hint, rows):
        """One ControlNet + controlled-UNet evaluation on `rows` rows (x may hold rows/2 images)."""
        unet, cn = self.model.diffusion_model, self.control_model
        h, hs, emb_all, kvs = unet.encode(x, t, kv=kv_unet, rows=rows)
        n = len(hs)
        if self.only_mid_control:
            ctrl = cn.forward(x, timesteps=t, kv=kv_ctrl, rows=rows, guided_hint=guided_hint,
                              scales=self.control_scales, add_to=[None] * n + [h], only_mid=True)
            return unet.decode(ctrl[-1], hs, emb_all, kvs)
        ctrl = cn.forward(x, timesteps=t, kv=kv_ctrl, rows=rows, guided_hint=guided_hint,
                          scales=self.control_scales, add_to=hs + [h])
        return unet.decode(ctrl[-1], ctrl[:-1], emb_all, kvs)
