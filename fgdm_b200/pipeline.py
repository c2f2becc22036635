"""The factor-graph chain of scripts/txt2img_fgdm_inference.py (:216-292) + controlnet/initialize_cn.py
(:74-102), kept on the device end to end:

  seg factor  : DDIMSampler.sample on the FG-DM LatentDiffusion (50 steps, CFG 7.5)     :233-242
  decode      : decode_first_stage -> clamp((x+1)/2,0,1)*255 -> uint8 (truncation)      :244-251
  resize      : cv2.resize(.., (512,512), INTER_LINEAR) semantics, on the GPU           :257-259
  hint        : uint8 / 255 -> [B,3,512,512]                                  initialize_cn.py:78-80
  image factor: ddim_hacked sampler on ControlLDM (hint + text), CFG                     :86-92
  decode      : (x*127.5+127.5).clip(0,255).astype(uint8)                                :98-99

The reference round-trips the seg map through host numpy/OpenCV between the two factors; here the
hand-off is three small kernels (fgdm_image_to_u8 / fgdm_resize_u8 / fgdm_u8_to_hint) with identical
integer semantics. Prompts are independent, so multi-GPU = every rank runs the chain on its slice and
the uint8 results are all-gathered (NCCL over NVLink) — the only collective.
"""
from __future__ import annotations

import torch

from . import ops
from .ddim import DDIMSampler


class FGDMPipeline:
    def __init__(self, seg_model, img_model, seg_steps=50, img_steps=50, seg_scale=7.5, img_scale=7.5,
                 out_hw=(512, 512)):
        self.seg_model, self.img_model = seg_model, img_model
        self.seg_sampler = DDIMSampler(seg_model)
        self.img_sampler = DDIMSampler(img_model)
        self.seg_steps, self.img_steps = seg_steps, img_steps
        self.seg_scale, self.img_scale = seg_scale, img_scale
        self.out_hw = out_hw
        self.device = seg_model.device

    @torch.no_grad()
    def seg_factor(self, c, uc, x_T):
        """-> (latent, uint8 seg map [B,H,W,3] at the seg factor's own resolution)."""
        b = x_T.shape[0]
        z, _ = self.seg_sampler.sample(S=self.seg_steps, conditioning=c, batch_size=b, shape=list(x_T.shape[1:]),
                                       verbose=False, unconditional_guidance_scale=self.seg_scale,
                                       unconditional_conditioning=uc, eta=0.0, x_T=x_T, num=1)
        img = self.seg_model.decode_first_stage(z)
        return z, ops.image_to_u8(img, 0)

    @torch.no_grad()
    def handoff(self, seg_u8):
        """uint8 seg map -> ControlNet hint [B,3,512,512] fp32 in [0,1] (resize only if sizes differ)."""
        dh, dw = self.out_hw
        if seg_u8.shape[1] != dh or seg_u8.shape[2] != dw:
            seg_u8 = ops.resize_u8(seg_u8, dh, dw)
        return seg_u8, ops.u8_to_hint(seg_u8)

    @torch.no_grad()
    def image_factor(self, hint, c, uc, x_T):
        b = x_T.shape[0]
        cond = {"c_concat": [hint], "c_crossattn": [c]}
        un = {"c_concat": [hint], "c_crossattn": [uc]}
        z, _ = self.img_sampler.sample(self.img_steps, b, list(x_T.shape[1:]), cond, verbose=False, eta=0.0,
                                       unconditional_guidance_scale=self.img_scale, unconditional_conditioning=un,
                                       x_T=x_T)
        img = self.img_model.decode_first_stage(z)
        return z, ops.image_to_u8(img, 1)

    @torch.no_grad()
    def __call__(self, c, uc, xT_seg, c2, uc2, xT_img, return_all=False):
        """All inputs are device tensors (fp32). Returns uint8 images [B,512,512,3] (device)."""
        z_seg, seg_u8 = self.seg_factor(c, uc, xT_seg)
        seg_u8, hint = self.handoff(seg_u8)
        z_img, img_u8 = self.image_factor(hint, c2, uc2, xT_img)
        if return_all:
            return dict(z_seg=z_seg, seg_u8=seg_u8, hint=hint, z_img=z_img, img_u8=img_u8)
        return img_u8

    @torch.no_grad()
    def run_host(self, host_in: dict, host_out: torch.Tensor, stream=None):
        """End-to-end call with HOST buffers: pinned inputs -> device, chain, uint8 images -> pinned host_out.
        Returns (h2d_bytes, d2h_bytes)."""
        dev = self.device
        names = ("c", "uc", "xT_seg", "c2", "uc2", "xT_img")
        d = {k: host_in[k].to(dev, non_blocking=True) for k in names}
        h2d = sum(host_in[k].numel() * host_in[k].element_size() for k in names)
        img = self(d["c"], d["uc"], d["xT_seg"], d["c2"], d["uc2"], d["xT_img"])
        host_out.copy_(img, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return h2d, img.numel()


def all_gather_images(img_u8: torch.Tensor) -> torch.Tensor:
    """[B_local,H,W,3] uint8 per rank -> [world*B_local,H,W,3] on every rank (rank-major order). One
    in-place all-gather (NCCL over NVLink/NVSwitch on GPUs; gloo in the CPU tests)."""
    import torch.distributed as dist
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size() == 1:
        return img_u8
    world = dist.get_world_size()
    out = torch.empty((world,) + tuple(img_u8.shape), device=img_u8.device, dtype=img_u8.dtype)
    dist.all_gather_into_tensor(out.view(-1), img_u8.contiguous().view(-1))
    return out.view((world * img_u8.shape[0],) + tuple(img_u8.shape[1:]))


def shard_indices(n_prompts: int, rank: int, world: int):
    """Contiguous block partition of prompt indices; remainder spread over the first ranks."""
    base, rem = divmod(n_prompts, world)
    start = rank * base + min(rank, rem)
    return list(range(start, start + base + (1 if rank < rem else 0)))
