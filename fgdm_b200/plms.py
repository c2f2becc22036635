"""PLMSSampler with the reference's signature and semantics (ldm/models/diffusion/plms.py), on the B200 path.

`make_schedule` (:24-56) shares DDIM's tables and insists on eta = 0; `plms_sampling` (:115-170) keeps the last three
CFG-combined eps; `p_sample_plms` (:172-236) is the pseudo improved Euler on the first step (two model calls) and
Adams-Bashforth of order 2, 3, 4 afterwards. Here the model call is ONE captured CUDA graph replayed per evaluation
(step-invariant work hoisted as in ddim.py), and CFG combine + multistep blend + x_{t-1} update is one kernel
(`fgdm_plms_step`). Options outside the reference's default call raise NotImplementedError.
"""
from __future__ import annotations

import numpy as np
import torch

from . import ops
from .ddim import DDIMSampler, _StepGraph, uncond_hint_matches
from .ldm import ControlLDM, LatentDiffusion

# e' = w0 e_t + w1 old[-1] + w2 old[-2] + w3 old[-3]   (plms.py:221-229)
AB_WEIGHTS = {1: (3 / 2, -1 / 2, 0.0, 0.0), 2: (23 / 12, -16 / 12, 5 / 12, 0.0), 3: (55 / 24, -59 / 24, 37 / 24, -9 / 24)}


class PLMSSampler(DDIMSampler):
    def make_schedule(self, ddim_num_steps, ddim_discretize="uniform", ddim_eta=0., verbose=True):
        if ddim_eta != 0:
            raise ValueError('ddim_eta must be 0 for PLMS')           # plms.py:25-26
        return super().make_schedule(ddim_num_steps, ddim_discretize, ddim_eta, verbose)

    @torch.no_grad()
    def sample(self, S, batch_size, shape, conditioning=None, callback=None, normals_sequence=None,
               img_callback=None, quantize_x0=False, eta=0., mask=None, x0=None, temperature=1., noise_dropout=0.,
               score_corrector=None, corrector_kwargs=None, verbose=True, x_T=None, log_every_t=100,
               unconditional_guidance_scale=1., unconditional_conditioning=None, **kwargs):
        """plms.py:58-113."""
        bad = dict(mask=mask is not None, x0=x0 is not None, quantize_x0=quantize_x0, noise_dropout=noise_dropout > 0,
                   score_corrector=score_corrector is not None, temperature=temperature != 1.)
        bad = [k for k, v in bad.items() if v]
        if bad:
            raise NotImplementedError(f"PLMSSampler.sample: options outside the FG-DM sampling hot path: {bad}")
        self.make_schedule(ddim_num_steps=S, ddim_eta=eta, verbose=False)
        C, H, W = shape
        return self.plms_sampling(conditioning, (batch_size, C, H, W), callback=callback, img_callback=img_callback,
                                  x_T=x_T, log_every_t=log_every_t,
                                  unconditional_guidance_scale=unconditional_guidance_scale,
                                  unconditional_conditioning=unconditional_conditioning, pcond=kwargs.get("pcond"))

    @torch.no_grad()
    def plms_sampling(self, cond, shape, x_T=None, callback=None, img_callback=None, log_every_t=100,
                      unconditional_guidance_scale=1., unconditional_conditioning=None, pcond=None, **kwargs):
        """plms.py:115-170."""
        model = self.model
        if not isinstance(model, LatentDiffusion):
            raise NotImplementedError("PLMSSampler drives fgdm_b200.ldm models")
        device = model.device
        b = shape[0]
        img = torch.randn(shape, device=device) if x_T is None else x_T.to(device=device, dtype=torch.float32)
        time_range = np.flip(self.ddim_timesteps)
        total_steps = time_range.shape[0]
        intermediates = {'x_inter': [img], 'pred_x0': [img]}
        cfg = not (unconditional_conditioning is None or unconditional_guidance_scale == 1.)
        scale = float(unconditional_guidance_scale) if cfg else 1.0
        ctx_c, hint_c = self._split_cond(cond)
        ctx_u, hint_u = self._split_cond(unconditional_conditioning)
        if not uncond_hint_matches(cond, unconditional_conditioning, hint_c, hint_u):
            raise NotImplementedError("PLMS: an unconditional c_concat that differs from the conditional one (ControlNet "
                                      "guess mode) is only provided by DDIMSampler")
        if isinstance(model, ControlLDM) and (hint_c is None or pcond is not None):
            raise NotImplementedError("ControlLDM needs a c_concat hint and takes no pcond")
        key = ("plms", b, tuple(shape[1:]), ctx_c.shape[1], cfg,
               (tuple(model.control_scales), model.only_mid_control) if isinstance(model, ControlLDM) else None,
               0 if pcond is None else pcond.shape[0])
        g = self._graphs.get(key)
        if g is None:
            g = self._graphs[key] = _StepGraph(model, b, shape[1:], ctx_c.shape[1], cfg,
                                               pcond_rows=0 if pcond is None else pcond.shape[0])
        g.set_conditioning(ctx_c, ctx_u, hint_c, pcond)
        coef = self.coef_table.to(device)
        g.x.copy_(img)
        ring = [torch.empty_like(g.x) for _ in range(4)]              # old_eps (3) + the slot being written
        old_eps = []
        x_keep, tmp = torch.empty_like(g.x), torch.empty_like(g.x)
        for i, step in enumerate(time_range):
            index = total_steps - i - 1
            t_next = time_range[min(i + 1, total_steps - 1)]
            e_u, e_c = g.eval_eps(step)
            e_t = ring[i % 4]
            if len(old_eps) == 0:
                # pseudo improved Euler (2nd order), plms.py:216-220: x' from e_t, second model call at (x', t_next)
                x_keep.copy_(g.x)
                ops.plms_step(e_u, e_c, [], (1.0, 0.0, 0.0, 0.0), x_keep, coef, index, scale, e_out=e_t, x_prev=g.x,
                              pred_x0=tmp)
                e_u, e_c = g.eval_eps(t_next)
                ops.plms_step(e_u, e_c, [e_t], (0.5, 0.5, 0.0, 0.0), x_keep, coef, index, scale, e_out=None, x_prev=g.x,
                              pred_x0=g.pred_x0)
            else:
                hist = old_eps[::-1]
                ops.plms_step(e_u, e_c, hist, AB_WEIGHTS[len(hist)], g.x, coef, index, scale, e_out=e_t, x_prev=g.x,
                              pred_x0=g.pred_x0)
            old_eps.append(e_t)
            if len(old_eps) >= 4:
                old_eps.pop(0)
            if callback:
                callback(i)
            if img_callback:
                img_callback(g.pred_x0.clone(), i)
            if index % log_every_t == 0 or index == total_steps - 1:
                intermediates['x_inter'].append(g.x.clone())
                intermediates['pred_x0'].append(g.pred_x0.clone())
        return g.x.clone(), intermediates
