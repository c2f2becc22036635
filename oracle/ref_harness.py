"""TEST INFRASTRUCTURE ONLY — imports the UNMODIFIED reference (/root/reference) in this container.

Used by oracle/make_golden.py to (1) validate oracle/fgdm_oracle.py against the reference's own code
and (2) generate the committed fixtures under tests/golden/. /root/reference does not exist on the
GPU box, so nothing in tests/, bench.py or smoke() imports this module at run time.

The reference needs five non-arithmetic packages that are absent here (no network): omegaconf,
pytorch_lightning, taming, clip/kornia/open_clip, matplotlib. They are stubbed in sys.modules before
import; none of them is on the arithmetic path.
"""
from __future__ import annotations

import os
import sys
import types

import torch
import torch.nn as nn
import yaml

REF = os.environ.get("FGDM_REFERENCE", "/root/reference")


def _mod(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def install_stubs():
    if "pytorch_lightning" in sys.modules and getattr(sys.modules["pytorch_lightning"], "_fgdm_stub", False):
        return

    class ListConfig(list):
        pass

    class DictConfig(dict):
        pass

    oc = _mod("omegaconf", ListConfig=ListConfig, DictConfig=DictConfig)
    oc.OmegaConf = type("OmegaConf", (), {"load": staticmethod(lambda p: yaml.safe_load(open(p))),
                                          "create": staticmethod(lambda d: d)})
    _mod("omegaconf.listconfig", ListConfig=ListConfig)

    class LightningModule(nn.Module):
        @property
        def device(self):
            try:
                return next(self.parameters()).device
            except StopIteration:
                return torch.device("cpu")

        def log(self, *a, **k):
            pass

        def log_dict(self, *a, **k):
            pass

    def rank_zero_only(fn):
        return fn

    def seed_everything(seed):
        import random
        import numpy as np
        random.seed(seed)
        np.random.seed(seed)
        torch.manual_seed(seed)
        return seed

    pl = _mod("pytorch_lightning", LightningModule=LightningModule, seed_everything=seed_everything,
              _fgdm_stub=True, Callback=object, Trainer=object)
    _mod("pytorch_lightning.utilities")
    _mod("pytorch_lightning.utilities.distributed", rank_zero_only=rank_zero_only)
    _mod("pytorch_lightning.utilities.rank_zero", rank_zero_only=rank_zero_only)
    _mod("pytorch_lightning.callbacks", ModelCheckpoint=object, Callback=object, LearningRateMonitor=object)
    pl.utilities = sys.modules["pytorch_lightning.utilities"]

    _mod("taming")
    _mod("taming.modules")
    _mod("taming.modules.vqvae")
    _mod("taming.modules.vqvae.quantize", VectorQuantizer2=object)
    _mod("clip")
    _mod("kornia")
    _mod("open_clip")
    mpl = _mod("matplotlib")
    mpl.use = lambda *a, **k: None
    _mod("matplotlib.pyplot")
    _mod("matplotlib.cm")
    _mod("matplotlib.colors")
    mpl.pyplot = sys.modules["matplotlib.pyplot"]
    try:
        import xformers  # noqa: F401
    except Exception:
        pass
    if REF not in sys.path:
        sys.path.insert(0, REF)


def _instantiate(cfg):
    from ldm.util import instantiate_from_config
    return instantiate_from_config(cfg)


def reinit_zero_modules(model: nn.Module, seed: int = 1, std: float = 0.02):
    """The reference zero-initialises its output convs (openaimodel.py:249-251,727; attention.py:269-273;
    cldm.py:670,789-790): a fresh model returns exactly 0. Re-fill every all-zero weight with N(0, std^2)."""
    g = torch.Generator().manual_seed(seed)
    n = 0
    for name, p in sorted(model.named_parameters()):
        if p.dim() >= 2 and float(p.detach().abs().max()) == 0.0:
            p.data.copy_(torch.randn(p.shape, generator=g) * std)
            n += 1
    return n


def build_fgdm(seed: int = 0, unet_overrides: dict | None = None, vae_overrides: dict | None = None):
    """Reference LatentDiffusion from models/config.yaml with an Identity cond stage."""
    install_stubs()
    cfg = yaml.safe_load(open(os.path.join(REF, "models/config.yaml")))["model"]
    cfg["params"]["cond_stage_config"] = {"target": "torch.nn.Identity"}
    cfg["params"]["unet_config"]["params"]["use_checkpoint"] = False
    cfg["params"].pop("scheduler_config", None)
    if unet_overrides:
        cfg["params"]["unet_config"]["params"].update(unet_overrides)
    if vae_overrides:
        cfg["params"]["first_stage_config"]["params"]["ddconfig"].update(vae_overrides)
    torch.manual_seed(seed)
    model = _instantiate(cfg)
    model.eval()
    return model


def build_controlldm(seed: int = 0, overrides: dict | None = None, vae_overrides: dict | None = None):
    """Reference ControlLDM from controlnet/models/cldm_v15_canny.yaml with an Identity cond stage."""
    install_stubs()
    cfg = yaml.safe_load(open(os.path.join(REF, "controlnet/models/cldm_v15_canny.yaml")))["model"]
    cfg["params"]["cond_stage_config"] = {"target": "torch.nn.Identity"}
    for k in ("control_stage_config", "unet_config"):
        cfg["params"][k]["params"]["use_checkpoint"] = False
        if overrides:
            cfg["params"][k]["params"].update(overrides)
    if vae_overrides:
        cfg["params"]["first_stage_config"]["params"]["ddconfig"].update(vae_overrides)
    torch.manual_seed(seed)
    model = _instantiate(cfg)
    model.eval()
    return model


def cpu_sampler(sampler):
    """The reference's DDIMSampler.register_buffer forces .to('cuda') (ddim.py:20-24); patch for CPU."""
    sampler.register_buffer = lambda name, attr: setattr(sampler, name, attr)
    return sampler
