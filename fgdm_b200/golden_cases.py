"""Seeded definitions of the golden parity cases (inputs + weight seeds + model configs).

Shared by oracle/make_golden.py (build container: runs the real reference, writes tests/golden/*.npz)
and tests/ (GPU box: re-derives the same inputs/weights and compares the CUDA path with the fixtures).
Pure data/plumbing: no arithmetic of the hot path lives here.
"""
from __future__ import annotations

import torch

from . import synth

FGDM_WEIGHT_SEED = 11
CONTROL_WEIGHT_SEED = 12
CONTROL_SCALES_VARIANT = tuple(0.5 + 0.1 * i for i in range(13))
GUESS_MODE_SCALES = tuple(1.0 * (0.825 ** float(12 - i)) for i in range(13))   # initialize_cn.py:92, strength 1

# models/config.yaml:33-48 and controlnet/models/cldm_v15_canny.yaml:21-53
UNET_PARAMS = dict(image_size=32, in_channels=4, out_channels=4, model_channels=320, attention_resolutions=[4, 2, 1],
                   num_res_blocks=2, channel_mult=[1, 2, 4, 4], num_heads=8, use_spatial_transformer=True,
                   transformer_depth=1, context_dim=768, use_checkpoint=True, legacy=False)
CONTROLNET_PARAMS = dict(image_size=32, in_channels=4, hint_channels=3, model_channels=320,
                         attention_resolutions=[4, 2, 1], num_res_blocks=2, channel_mult=[1, 2, 4, 4], num_heads=8,
                         use_spatial_transformer=True, transformer_depth=1, context_dim=768, use_checkpoint=True,
                         legacy=False)
# models/config.yaml:50-71
VAE_PARAMS = dict(embed_dim=4, monitor="val/rec_loss",
                  ddconfig=dict(double_z=True, z_channels=4, resolution=256, in_channels=3, out_ch=3, ch=128,
                                ch_mult=[1, 2, 4, 4], num_res_blocks=2, attn_resolutions=[], dropout=0.0),
                  lossconfig=dict(target="torch.nn.Identity"))


def fgdm_config():
    """models/config.yaml with our target strings (the reference's own strings also resolve through
    fgdm_b200.ldm.TARGET_ALIASES)."""
    return dict(target="fgdm_b200.ldm.LatentDiffusion", params=dict(
        linear_start=0.00085, linear_end=0.0120, num_timesteps_cond=1, log_every_t=200, timesteps=1000,
        first_stage_key="image", cond_stage_key="caption", image_size=32, channels=4, cond_stage_trainable=False,
        conditioning_key="crossattn", scale_factor=0.18215, use_ema=False,
        unet_config=dict(target="fgdm_b200.nets.UNetModel", params=dict(UNET_PARAMS)),
        first_stage_config=dict(target="fgdm_b200.nets.AutoencoderKL", params=dict(VAE_PARAMS)),
        cond_stage_config=dict(target="torch.nn.Identity")))


def control_config():
    """controlnet/models/cldm_v15_canny.yaml with our target strings."""
    return dict(target="fgdm_b200.ldm.ControlLDM", params=dict(
        linear_start=0.00085, linear_end=0.0120, num_timesteps_cond=1, log_every_t=200, timesteps=1000,
        first_stage_key="jpg", cond_stage_key="txt", control_key="hint", image_size=64, channels=4,
        cond_stage_trainable=False, conditioning_key="crossattn", scale_factor=0.18215, use_ema=False,
        only_mid_control=False,
        control_stage_config=dict(target="fgdm_b200.nets.ControlNet", params=dict(CONTROLNET_PARAMS)),
        unet_config=dict(target="fgdm_b200.nets.ControlledUnetModel", params=dict(UNET_PARAMS)),
        first_stage_config=dict(target="fgdm_b200.nets.AutoencoderKL", params=dict(VAE_PARAMS)),
        cond_stage_config=dict(target="torch.nn.Identity")))


def fgdm_param_shapes():
    from .ldm import instantiate_from_config
    return instantiate_from_config(fgdm_config()).param_shapes()


def control_param_shapes():
    from .ldm import instantiate_from_config
    return instantiate_from_config(control_config()).param_shapes()


# ---- cases (16x16 latents: UNet levels 16/8/4/2, attention at 256/64/16/4 tokens) ----------------
def unet_case():
    return dict(x=synth.randn((2, 4, 16, 16), 21), t=torch.tensor([981, 441], dtype=torch.int64),
                ctx=synth.randn((2, 77, 768), 22))


def pcond_case():
    """FG-DM joint ("pcond") factor: the adapter is fed by another factor's clean latent instead of the noisy latent
    (openaimodel.py:836-844). One prompt, its two CFG rows: the reference's `h + fa[k]` broadcasts a 1-row feature."""
    return dict(x=synth.randn((2, 4, 16, 16), 23), t=torch.tensor([621, 621], dtype=torch.int64),
                ctx=synth.randn((2, 77, 768), 24), pcond=synth.randn((1, 4, 16, 16), 25))


def ddim_case():
    return dict(S=5, B=1, shape=(4, 16, 16), scale=7.5, x_T=synth.randn((1, 4, 16, 16), 31),
                c=synth.randn((1, 77, 768), 32), uc=synth.randn((1, 77, 768), 33))


def plms_case():
    """PLMS (plms.py): 6 steps exercise the Euler start and the Adams-Bashforth orders 2, 3, 4."""
    return dict(S=6, B=1, shape=(4, 16, 16), scale=7.5, x_T=synth.randn((1, 4, 16, 16), 34),
                c=synth.randn((1, 77, 768), 35), uc=synth.randn((1, 77, 768), 36))


def dpm_case():
    """DPM-Solver++(2M) via DPMSolverSampler: 5 steps (< 15: the final step drops to first order)."""
    return dict(S=5, B=1, shape=(4, 16, 16), scale=7.5, x_T=synth.randn((1, 4, 16, 16), 37),
                c=synth.randn((1, 77, 768), 38), uc=synth.randn((1, 77, 768), 39))


CLIP_WEIGHT_SEED = 7


def clip_param_shapes(cfg=None):
    """State-dict names/shapes of transformers' CLIPTextModel for openai/clip-vit-large-patch14 (text tower) under
    FrozenCLIPEmbedder's attribute name `transformer` (modules.py:142)."""
    c = cfg or dict(vocab_size=49408, hidden_size=768, intermediate_size=3072, num_hidden_layers=12,
                    max_position_embeddings=77)
    h, f, p = c["hidden_size"], c["intermediate_size"], "transformer.text_model."
    sh = {p + "embeddings.token_embedding.weight": (c["vocab_size"], h),
          p + "embeddings.position_embedding.weight": (c["max_position_embeddings"], h)}
    for i in range(c["num_hidden_layers"]):
        q = f"{p}encoder.layers.{i}."
        for n in ("q_proj", "k_proj", "v_proj", "out_proj"):
            sh[q + f"self_attn.{n}.weight"] = (h, h)
            sh[q + f"self_attn.{n}.bias"] = (h,)
        for n in ("layer_norm1", "layer_norm2"):
            sh[q + n + ".weight"] = (h,)
            sh[q + n + ".bias"] = (h,)
        sh[q + "mlp.fc1.weight"], sh[q + "mlp.fc1.bias"] = (f, h), (f,)
        sh[q + "mlp.fc2.weight"], sh[q + "mlp.fc2.bias"] = (h, f), (h,)
    sh[p + "final_layer_norm.weight"] = (h,)
    sh[p + "final_layer_norm.bias"] = (h,)
    return sh


def clip_state_dict():
    """Seeded synthetic CLIP text weights; q/k projections x3 so that the attention is not near-uniform."""
    sd = synth.make_state_dict(clip_param_shapes(), CLIP_WEIGHT_SEED)
    for k in sd:
        if k.endswith("q_proj.weight") or k.endswith("k_proj.weight"):
            sd[k] = sd[k] * 3.0
    return sd


def clip_case():
    """Two prompts of 77 token ids: BOS, random words, EOS then EOS padding (the tokenizer's max_length padding)."""
    g = torch.Generator(device="cpu").manual_seed(91)
    ids = torch.randint(1000, 40000, (2, 77), generator=g, dtype=torch.int64)
    ids[:, 0] = 49406
    ids[0, 9:] = 49407
    ids[1, 40:] = 49407
    return dict(ids=ids)


def vae_case():
    return dict(z=synth.randn((1, 4, 16, 16), 41) * 0.18215 * 3.0)


def _hint(b, hw, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    u8 = torch.randint(0, 256, (b, hw, hw, 3), generator=g, dtype=torch.uint8)
    return (u8.float() / 255.0).permute(0, 3, 1, 2).contiguous()


def control_case():
    return dict(x=synth.randn((2, 4, 16, 16), 51), t=torch.tensor([801, 201], dtype=torch.int64),
                ctx=synth.randn((2, 77, 768), 52), hint=_hint(2, 128, 53))


def control_ddim_case():
    return dict(S=4, B=1, shape=(4, 16, 16), scale=9.0, x_T=synth.randn((1, 4, 16, 16), 61),
                c=synth.randn((1, 77, 768), 62), uc=synth.randn((1, 77, 768), 63), hint=_hint(1, 128, 64))


def handoff_case():
    return dict(img=synth.randn((2, 3, 64, 64), 71) * 0.8, dh=128, dw=128)
