"""Synthetic weights and inputs (there is no network: no checkpoints, no datasets).

`make_state_dict(shapes, seed)` fills a state_dict that uses the reference's parameter names/shapes
with a seeded, CPU-generated random init, so the build container (where the real reference produces the
golden fixtures) and the GPU box (where the CUDA path is checked against them) hold bit-identical
weights without shipping gigabytes. The init is PyTorch's default for conv/linear (U(+-1/sqrt(fan_in)))
with two deliberate changes that make parity checks sensitive:
  * the reference zero-initialises every block's last conv (openaimodel.py:249-251,727; attention.py:
    269-273; cldm.py:670,789-790) so a fresh model outputs exactly 0 — these get N(0, 0.02^2) instead;
  * to_q / to_k get a 3x gain so softmax rows are peaked rather than uniform, and norm affines are
    perturbed around (1, 0) so gamma/beta handling is exercised.
"""
from __future__ import annotations

import math
import re

import torch

_ZERO_MODULES = re.compile(r"(out_layers\.3|(?<!attn_1\.)proj_out|^out\.2|\.out\.2|zero_convs\.\d+\.0|"
                           r"middle_block_out\.0|input_hint_block\.14)\.(weight|bias)$")


def make_state_dict(shapes: dict, seed: int = 0, dtype=torch.float32) -> dict:
    """shapes: name -> shape (e.g. `model.param_shapes()`). Deterministic for a given (names, seed)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    sd = {}
    names = list(shapes.keys())
    for name in names:
        shape = tuple(shapes[name])
        leaf = name.rsplit(".", 1)[-1]
        base = name[: -len(leaf) - 1]
        wshape = tuple(shapes.get(base + ".weight", shape))
        is_norm = len(wshape) == 1
        t = torch.empty(shape, dtype=torch.float32)
        if is_norm:
            t.normal_(0.0, 0.1, generator=g)
            if leaf == "weight":
                t += 1.0
        elif _ZERO_MODULES.search(name):
            t.normal_(0.0, 0.02, generator=g)
        else:
            fan_in = 1
            for s in wshape[1:]:
                fan_in *= s
            bound = 1.0 / math.sqrt(max(fan_in, 1))
            if leaf == "weight" and (base.endswith("to_q") or base.endswith("to_k")):
                bound *= 3.0
            t.uniform_(-bound, bound, generator=g)
        sd[name] = t.to(dtype)
    return sd


def randn(shape, seed: int) -> torch.Tensor:
    """Host-side N(0,1) tensor with its own seed (results independent of how prompts are sharded)."""
    return torch.randn(shape, generator=torch.Generator(device="cpu").manual_seed(seed), dtype=torch.float32)


def prompt_inputs(indices, latent_hw, ctx_len=77, ctx_dim=768, base=0):
    """SURVEY.md §8(d) synthetic per-prompt inputs: context c_i (seed 1000+i), shared unconditional
    context (seed 999), x_T^seg_i (seed 2000+i), x_T^img_i (seed 3000+i), image-factor contexts c'_i
    (seed 4000+i) and negative context (seed 998)."""
    h, w = latent_hw
    n = len(indices)
    c = torch.stack([randn((ctx_len, ctx_dim), 1000 + base + i) for i in indices])
    uc = randn((ctx_len, ctx_dim), 999).expand(n, -1, -1).contiguous()
    xT_seg = torch.stack([randn((4, h, w), 2000 + base + i) for i in indices])
    xT_img = torch.stack([randn((4, 64, 64), 3000 + base + i) for i in indices])
    c2 = torch.stack([randn((ctx_len, ctx_dim), 4000 + base + i) for i in indices])
    uc2 = randn((ctx_len, ctx_dim), 998).expand(n, -1, -1).contiguous()
    return dict(c=c, uc=uc, xT_seg=xT_seg, xT_img=xT_img, c2=c2, uc2=uc2)
