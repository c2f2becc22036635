"""DPMSolverSampler with the reference's signature (ldm/models/diffusion/dpm_solver/sampler.py:8-82), on the B200 path.

The reference class fixes the solver configuration: NoiseScheduleVP('discrete', alphas_cumprod) (dpm_solver.py:99-108),
a classifier-free `model_wrapper` around `apply_model` with FLOAT model timesteps (t - 1/N) * 1000 (:278-287,321-344),
DPM-Solver++ (`predict_x0=True`, no thresholding) in multistep mode, order 2, uniform time steps and a first-order final
step when S < 15 (:1060-1096). Its x_t updates (:504-535, :755-796) are scalar combinations of x and of the last two
x0-predictions, so a step here is: ONE captured-graph replay of the model call, `fgdm_plms_step` (CFG combine + the
x0-prediction (x - sigma e) / alpha), `fgdm_lincomb3` (the update). Schedule scalars are computed on the host in fp32 like
the reference's tensors. Returns `(x, None)` as the reference does.
"""
from __future__ import annotations

import numpy as np
import torch

from . import ops
from .ddim import DDIMSampler, _StepGraph, uncond_hint_matches
from .ldm import ControlLDM, LatentDiffusion


class DPMSolverSampler(object):
    def __init__(self, model, **kwargs):
        self.model = model
        self.alphas_cumprod = model.alphas_cumprod.detach().cpu().to(torch.float32)
        self._graphs = {}

    def register_buffer(self, name, attr):
        setattr(self, name, attr)

    def _marginals(self, ts):
        """marginal_log_mean_coeff / alpha / std / lambda (dpm_solver.py:125-156) at the continuous times `ts`."""
        la_knots = 0.5 * torch.log(self.alphas_cumprod)
        n = la_knots.shape[0]
        t_knots = torch.linspace(0., 1., n + 1)[1:]
        la = torch.tensor(np.interp(ts.double().numpy(), t_knots.double().numpy(), la_knots.double().numpy()),
                          dtype=torch.float32)
        return dict(alpha=torch.exp(la), sigma=torch.sqrt(1. - torch.exp(2. * la)),
                    lam=la - 0.5 * torch.log(1. - torch.exp(2. * la)))

    @torch.no_grad()
    def sample(self, S, batch_size, shape, conditioning=None, callback=None, normals_sequence=None, img_callback=None,
               quantize_x0=False, eta=0., mask=None, x0=None, temperature=1., noise_dropout=0., score_corrector=None,
               corrector_kwargs=None, verbose=True, x_T=None, log_every_t=100, unconditional_guidance_scale=1.,
               unconditional_conditioning=None, **kwargs):
        model = self.model
        if not isinstance(model, LatentDiffusion):
            raise NotImplementedError("DPMSolverSampler drives fgdm_b200.ldm models")
        if mask is not None or x0 is not None or score_corrector is not None:
            raise NotImplementedError("DPMSolverSampler.sample: the reference ignores mask/x0/score_corrector here too")
        if S < 2:
            raise AssertionError("steps >= order (dpm_solver.py:1064)")
        C, H, W = shape
        b = batch_size
        device = model.device
        n = self.alphas_cumprod.shape[0]
        ts = torch.linspace(1.0, 1.0 / n, S + 1)                       # get_time_steps('time_uniform'), :429-430
        mg = self._marginals(ts)
        t_model = ((ts - 1. / n) * 1000.).tolist()                      # get_model_input_time, :278-287
        cfg = not (unconditional_conditioning is None or unconditional_guidance_scale == 1.)
        scale = float(unconditional_guidance_scale) if cfg else 1.0
        ctx_c, hint_c = DDIMSampler._split_cond(conditioning)
        ctx_u, hint_u = DDIMSampler._split_cond(unconditional_conditioning)
        if not uncond_hint_matches(conditioning, unconditional_conditioning, hint_c, hint_u):
            raise NotImplementedError("DPM-Solver: an unconditional c_concat that differs from the conditional one "
                                      "(ControlNet guess mode) is only provided by DDIMSampler")
        pcond = kwargs.get("pcond")
        if isinstance(model, ControlLDM) and (hint_c is None or pcond is not None):
            raise NotImplementedError("ControlLDM needs a c_concat hint and takes no pcond")
        key = (b, (C, H, W), ctx_c.shape[1], cfg,
               (tuple(model.control_scales), model.only_mid_control) if isinstance(model, ControlLDM) else None,
               0 if pcond is None else pcond.shape[0])
        g = self._graphs.get(key)
        if g is None:
            g = self._graphs[key] = _StepGraph(model, b, (C, H, W), ctx_c.shape[1], cfg, float_t=True,
                                               pcond_rows=0 if pcond is None else pcond.shape[0])
        g.set_conditioning(ctx_c, ctx_u, hint_c, pcond)
        img = torch.randn((b, C, H, W), device=device) if x_T is None else x_T.to(device=device, dtype=torch.float32)
        g.x.copy_(img)
        # rows {sigma_t, alpha_t, 0, 0, 0}: fgdm_plms_step's pred_x0 = (x - sigma e) / alpha is data_prediction_fn (:386-399)
        coef = torch.zeros((S + 1, 5), dtype=torch.float32)
        coef[:, 0], coef[:, 1] = mg["sigma"], mg["alpha"]
        coef = coef.to(device)
        m = [torch.empty_like(g.x), torch.empty_like(g.x)]              # model_prev_list (x0-predictions)
        junk = torch.empty_like(g.x)
        lam, sig, alp = mg["lam"], mg["sigma"], mg["alpha"]

        def predict(i, dst):
            e_u, e_c = g.eval_eps(t_model[i])
            ops.plms_step(e_u, e_c, [], (1.0, 0.0, 0.0, 0.0), g.x, coef, i, scale, x_prev=junk, pred_x0=dst)

        def first(s, t, m_s):                                           # dpm_solver_first_update, predict_x0 branch
            h = lam[t] - lam[s]
            ops.lincomb3(float(sig[t] / sig[s]), g.x, float(-(alp[t] * torch.expm1(-h))), m_s, out=g.x)

        def second(p1, p0, t, m1, m0):                                  # multistep_dpm_solver_second_update
            h_0, h = lam[p0] - lam[p1], lam[t] - lam[p0]
            r0 = h_0 / h
            a = alp[t] * (torch.exp(-h) - 1.)
            # x_t = (sigma_t / sigma_p0) x - a m0 - 0.5 a (m0 - m1) / r0
            ops.lincomb3(float(sig[t] / sig[p0]), g.x, float(-a - 0.5 * a / r0), m0, float(0.5 * a / r0), m1, out=g.x)

        predict(0, m[0])
        first(0, 1, m[0])
        predict(1, m[1])
        tprev = [0, 1]
        for step in range(2, S + 1):
            order = min(2, S + 1 - step) if S < 15 else 2               # lower_order_final, :1080-1083
            if order == 1:
                first(tprev[1], step, m[1])
            else:
                second(tprev[0], tprev[1], step, m[0], m[1])
            m[0], m[1] = m[1], m[0]                                     # shift: old model_prev_0 becomes model_prev_1
            tprev = [tprev[1], step]
            if step < S:
                predict(step, m[1])
            else:
                m[1] = m[0]
            if callback:
                callback(step)
        return g.x.clone(), None
