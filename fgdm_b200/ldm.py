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
import os

import numpy as np
import torch

from . import _lib, ops
from .nets import (AutoencoderKL, ControlledUnetModel, ControlNet, IncompatibleKeys,  # noqa: F401  (targets)
                   UNetModel)

# reference target string -> our class, so the reference's own YAMLs can be used verbatim
# FGDM_OVERLAP_CONTROL=0: ControlNet after the UNet encoder on one stream, like the reference (A/B switch)
OVERLAP_CONTROL = os.environ.get("FGDM_OVERLAP_CONTROL", "1") != "0"

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


def get_obj_from_str(string, reload=False, alias=True):
    """ldm/util.py:88-93 (+ the alias table; alias=False resolves the string as written)."""
    if alias:
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
        self.use_ema = bool(use_ema)
        self._ema_loaded = False
        self.cond_stage_forward = cond_stage_forward
        self.cond_stage_model = self._make_cond_stage(cond_stage_config)
        self.register_schedule(given_betas, timesteps, linear_start, linear_end)

    def _make_cond_stage(self, config):
        """ddpm.py:618-640 `instantiate_cond_stage`. Three cases:
        * the FrozenCLIPEmbedder target (models/config.yaml:74-75) -> fgdm_b200.clip.FrozenCLIPEmbedder, the text
          transformer on the B200 kernels (token ids, or strings when a CLIPTokenizer is available / injected);
        * an already constructed object with `encode` (e.g. the reference's own FrozenCLIPEmbedder instance, built
          by the maintainer where `ldm` + its tokenizer files are importable) is kept as is;
        * any other `target:` is instantiated through the plugin loader exactly as the reference would (ldm/util.py)
          — `torch.nn.Identity` / None / "__is_unconditional__" give the pass-through for pre-encoded contexts."""
        if config is None or config in ("__is_unconditional__", "__is_first_stage__"):
            return IdentityCondStage()
        if not isinstance(config, (dict,)) and not hasattr(config, "get"):
            if hasattr(config, "encode") or callable(config):
                return config                                         # a ready-made encoder object
            raise TypeError(f"cond_stage_config: expected a config mapping or an encoder object, got {type(config)}")
        target = str(config.get("target", ""))
        if target in ("torch.nn.Identity", ""):
            return IdentityCondStage()
        is_clip = TARGET_ALIASES.get(target, target) == "fgdm_b200.clip.FrozenCLIPEmbedder"
        if is_clip and os.environ.get("FGDM_KEEP_REFERENCE_CLIP", "0") != "1":
            params = dict(config.get("params", {}) or {}, device=str(self.device))
            return instantiate_from_config(dict(target=target, params=params))
        # the reference's own class, resolved as written (needs `ldm` / its dependencies importable)
        return get_obj_from_str(target, alias=False)(**(config.get("params", None) or {}))

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

    # keys of a reference checkpoint that belong to modules this inference-only path does not hold (VAE encoder,
    # training losses, schedule buffers of ddpm.py:175-227): the reference's own module tree has them, so they are
    # neither "missing" nor "unexpected" there and are not reported here either.
    _UNUSED_PREFIXES = ("first_stage_model.encoder.", "first_stage_model.quant_conv.", "first_stage_model.loss.",
                        "model_ema.", "cond_stage_model.")
    _SCHEDULE_BUFFERS = ("betas", "alphas_cumprod", "alphas_cumprod_prev", "sqrt_alphas_cumprod",
                         "sqrt_one_minus_alphas_cumprod", "log_one_minus_alphas_cumprod", "sqrt_recip_alphas_cumprod",
                         "sqrt_recipm1_alphas_cumprod", "posterior_variance", "posterior_log_variance_clipped",
                         "posterior_mean_coef1", "posterior_mean_coef2", "lvlb_weights", "logvar", "scale_factor")

    def _ema_state_dict(self, sd):
        """LitEma (ldm/modules/ema.py:16-21) keeps a shadow of every DiffusionWrapper parameter under
        `model_ema.<name without dots>`; map them back onto `model.diffusion_model.*` names."""
        out = {}
        for k in self.model.diffusion_model.param_shapes():
            s_name = "model_ema." + ("diffusion_model." + k).replace(".", "")
            if s_name not in sd:
                return None
            out["model.diffusion_model." + k] = sd[s_name]
        return out

    def load_state_dict(self, sd, strict=True):
        """torch semantics: returns IncompatibleKeys(missing_keys, unexpected_keys) so that the reference's
        `m, u = model.load_state_dict(sd, strict=False)` (scripts/txt2img_fgdm_inference.py:28,
        controlnet/initialize_cn.py:28-35) runs unchanged. Accepts the reference's checkpoint keys
        (`model.diffusion_model.*`, `first_stage_model.*`, `control_model.*`, `cond_stage_model.*`).
        use_ema=True with `model_ema.*` shadows in the checkpoint: the EMA weights are what the reference samples
        with (every sampling call site is inside `with model.ema_scope()`, ddpm.py:230-241), so they are the ones
        packed for the UNet; without shadows in the checkpoint the raw weights are used and a warning is printed."""
        _lib.require_device()
        missing, consumed = [], set()
        unet_sd = sd
        if self.use_ema:
            ema = self._ema_state_dict(sd)
            if ema is not None:
                unet_sd, self._ema_loaded = ema, True
            else:
                print("fgdm_b200: use_ema=True but the checkpoint holds no (complete) model_ema.* shadows; "
                      "sampling with the raw weights")
        for pre, net in self._parts().items():
            src = unet_sd if pre == "model.diffusion_model." else sd
            r = net.load_state_dict(src, strict=False, prefix=pre, device=self.device)
            missing += r.missing_keys
            consumed.update(pre + k for k in net.param_shapes())
        cs = self.cond_stage_model
        pre = "cond_stage_model."
        if hasattr(cs, "text_model") and hasattr(cs, "param_shapes"):          # our CLIP text tower
            want = [pre + k for k in cs.param_shapes()]
            if any(k in sd for k in want):
                miss = [k for k in want if k not in sd]
                missing += miss
                if not miss:
                    cs.load_state_dict({k[len(pre):]: v for k, v in sd.items() if k.startswith(pre)})
            consumed.update(want)
        elif isinstance(cs, torch.nn.Module):                                   # the reference's own encoder, kept
            sub = {k[len(pre):]: v for k, v in sd.items() if k.startswith(pre)}
            if sub:
                cs.load_state_dict(sub, strict=False)
        unexpected = [k for k in sd if k not in consumed and not k.startswith(self._UNUSED_PREFIXES)
                      and k not in self._SCHEDULE_BUFFERS]
        if strict and (missing or unexpected):
            raise RuntimeError(f"Error(s) in loading state_dict for {type(self).__name__}: "
                               f"missing {missing[:8]}{' ...' if len(missing) > 8 else ''}, "
                               f"unexpected {unexpected[:8]}{' ...' if len(unexpected) > 8 else ''}")
        return IncompatibleKeys(missing, unexpected)

    # nn.Module look-alikes the reference's drivers call (txt2img_fgdm_inference.py:36-37,185; cldm/model.py:26;
    # initialize_cn.py:27,39). The packed weights live on the B200 only, so device moves are no-ops.
    def eval(self):
        return self

    def cuda(self, *a):
        return self

    def cpu(self):
        return self

    def to(self, *a, **k):
        return self

    def low_vram_shift(self, is_diffusing):
        """cldm.py:933-943 pages sub-models between host and device under config.save_memory; a 180 GB B200 keeps
        all of them resident (4.4 GB of bf16 weights), so there is nothing to shift."""
        return None

    def ema_scope(self, context=None):
        """ddpm.py:230-241. Inference-only: the EMA weights (when present) were packed at load_state_dict."""
        import contextlib
        return contextlib.nullcontext()

    # -- reference methods ------------------------------------------------------------------------
    def get_learned_conditioning(self, c):
        """ddpm.py:663-674: `cond_stage_model.encode(c)` (or `cond_stage_forward`), result on the model device.
        Pre-encoded float tensors ([B,77,768] contexts) pass through; strings / token ids go to the text encoder."""
        if torch.is_tensor(c) and c.is_floating_point():
            return c
        cs = self.cond_stage_model
        if isinstance(cs, IdentityCondStage):
            raise NotImplementedError("no text encoder configured (cond_stage_config: FrozenCLIPEmbedder or an encoder "
                                      "object); pass [B,77,768] context tensors")
        if self.cond_stage_forward is None:
            out = cs.encode(c) if callable(getattr(cs, "encode", None)) else cs(c)
        else:
            out = getattr(cs, self.cond_stage_forward)(c)
        return out.to(self.device) if torch.is_tensor(out) else out

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

    def control_step(self, x, t, kv_unet, kv_ctrl, guided_hint, rows, shared_halves=False, emb_unet=None,
                     emb_ctrl=None):
        """One ControlNet + controlled-UNet evaluation on `rows` rows (x may hold rows/2 images). shared_halves: rows r
        and r + rows/2 share latent, timestep and hint (the sampler's CFG batch), see _UNetBase._encode.
        The ControlNet's encoder does not depend on the UNet's (cldm.py:836-849 runs them one after the other): it is
        issued on a forked side stream under the library's second scratch set, so the ~150 kernels of each overlap the
        other's launch gaps and tails; the zero convolutions, which add the UNet's skips, join the two."""
        unet, cn = self.model.diffusion_model, self.control_model
        cn_args = dict(timesteps=t, kv=kv_ctrl, rows=rows, guided_hint=guided_hint, shared_halves=shared_halves,
                       emb_all=emb_ctrl)
        if OVERLAP_CONTROL and x.is_cuda:
            main = torch.cuda.current_stream()
            if getattr(self, "_side", None) is None:
                self._side = torch.cuda.Stream()
            self._side.wait_stream(main)
            with torch.cuda.stream(self._side):
                ops.set_scratch_slot(1)
                try:
                    feats = cn.features(x, **cn_args)
                finally:
                    ops.set_scratch_slot(0)
            h, hs, emb_all, kvs = unet.encode(x, t, kv=kv_unet, rows=rows, shared_halves=shared_halves, emb_all=emb_unet)
            main.wait_stream(self._side)
        else:
            h, hs, emb_all, kvs = unet.encode(x, t, kv=kv_unet, rows=rows, shared_halves=shared_halves, emb_all=emb_unet)
            feats = cn.features(x, **cn_args)
        n = len(hs)
        if self.only_mid_control:
            ctrl = cn.control_outputs(feats, self.control_scales, [None] * n + [h], only_mid=True)
            return unet.decode(ctrl[-1], hs, emb_all, kvs)
        ctrl = cn.control_outputs(feats, self.control_scales, hs + [h])
        return unet.decode(ctrl[-1], ctrl[:-1], emb_all, kvs)
