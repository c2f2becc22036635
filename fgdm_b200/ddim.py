"""DDIMSampler with the reference's signature and semantics, driving the B200 path.

Mirrors ldm/models/diffusion/ddim.py (`DDIMSampler.make_schedule` :26-55, `sample` :57-114,
`ddim_sampling` :116-177, `p_sample_ddim` :179-286) and controlnet/cldm/ddim_hacked.py (:54-231), whose
only difference on this path is that CFG is evaluated as two sequential `apply_model` calls (:190-191)
— identical arithmetic to the batched form used here.

Fast path (models from fgdm_b200.ldm): one DDIM step = [device-side loop bookkeeping, UNet (+ControlNet)
forward on the 2B CFG rows, fused CFG-combine + x_{t-1} update] captured ONCE in a CUDA graph and replayed
S times with no host work in between. Step-invariant work is hoisted out of the loop: text K/V projections
of all cross-attentions and the ControlNet hint stem. Options outside the reference's default sampling
call (mask/x0 inpainting, score_corrector, quantize_x0, noise_dropout, inference_loss) raise
NotImplementedError. eta > 0, ControlNet guess mode (uncond c_concat=None) and foreign models take the per-step
generic path (`p_sample_ddim` through `apply_model`, same kernels, same arithmetic as the reference's loop).
"""
from __future__ import annotations

import os

import numpy as np
import torch

from . import ops
from .ldm import ControlLDM, LatentDiffusion


def make_ddim_timesteps(ddim_discr_method, num_ddim_timesteps, num_ddpm_timesteps, verbose=True):
    """ldm/modules/diffusionmodules/util.py:46-60."""
    if ddim_discr_method == "uniform":
        c = num_ddpm_timesteps // num_ddim_timesteps
        ddim_timesteps = np.asarray(list(range(0, num_ddpm_timesteps, c)))
    elif ddim_discr_method == "quad":
        ddim_timesteps = ((np.linspace(0, np.sqrt(num_ddpm_timesteps * .8), num_ddim_timesteps)) ** 2).astype(int)
    else:
        raise NotImplementedError(f'There is no ddim discretization method called "{ddim_discr_method}"')
    return ddim_timesteps + 1


def make_ddim_sampling_parameters(alphacums, ddim_timesteps, eta, verbose=True):
    """util.py:63-74. alphacums: float32 torch tensor on the host. alphas stays float32; alphas_prev and
    sigmas pass through float64 numpy exactly as in the reference."""
    alphas = alphacums[ddim_timesteps]
    alphas_prev = np.asarray([alphacums[0]] + alphacums[ddim_timesteps[:-1]].tolist())
    a = alphas.numpy()
    sigmas = eta * np.sqrt((1 - alphas_prev) / (1 - a) * (1 - a / alphas_prev))
    return sigmas, alphas, alphas_prev


def uncond_hint_matches(cond, uncond, hint_c, hint_u):
    """True when the unconditional half may reuse the conditional half's ControlNet hint: no unconditional
    conditioning at all, a non-dict one (FG-DM factor: no hint anywhere), or a dict whose c_concat equals cond's.
    False for ControlNet guess mode (c_concat=None next to a cond hint) and for differing hints."""
    if uncond is None or not isinstance(uncond, dict) or not isinstance(cond, dict):
        return True
    if hint_u is None or hint_c is None:
        return hint_u is None and hint_c is None
    return hint_u is hint_c or (hint_u.shape == hint_c.shape and torch.equal(hint_u, hint_c))


# FGDM_EMB_TABLE=0: time_embed + emb_layers run inside every step (default: once per schedule, before the loop)
EMB_TABLES = os.environ.get("FGDM_EMB_TABLE", "1") != "0"
EMB_TABLE_STEPS = 256

# FGDM_SHARE_CFG_HEAD=0: evaluate every CFG row through the whole encoder, like the reference (A/B switch)
SHARE_CFG_HEAD = os.environ.get("FGDM_SHARE_CFG_HEAD", "1") != "0"


class _StepGraph:
    """One captured DDIM step for fixed (model, B, latent shape, context length, CFG on/off)."""

    def __init__(self, model, B, shape, ctx_len, cfg, S_max=1000, pcond_rows=0, float_t=False):
        dev = model.device
        self.model, self.B, self.cfg = model, B, cfg
        self.R = 2 * B if cfg else B
        C, H, W = shape
        self.x = torch.zeros((B, C, H, W), device=dev, dtype=torch.float32)
        self.pred_x0 = torch.zeros_like(self.x)
        self.t = torch.zeros((self.R,), device=dev, dtype=torch.float32 if float_t else torch.int64)
        self.index = torch.zeros((1,), device=dev, dtype=torch.int32)
        self.counter = torch.zeros((1,), device=dev, dtype=torch.int32)
        self.t_table = torch.zeros((S_max,), device=dev, dtype=torch.int64)
        self.idx_table = torch.zeros((S_max,), device=dev, dtype=torch.int32)
        self.coef = torch.zeros((S_max, 5), device=dev, dtype=torch.float32)
        self.scale = 1.0
        self.ctx = torch.zeros((self.R, ctx_len, model.model.diffusion_model.context_dim), device=dev,
                               dtype=torch.bfloat16)
        self.is_control = isinstance(model, ControlLDM)
        unet = model.model.diffusion_model
        self.kv_unet = unet.project_context(self.ctx)
        if self.is_control:
            self.kv_ctrl = model.control_model.project_context(self.ctx)
            self.guided = torch.zeros((B, H, W, model.control_model.model_channels), device=dev,
                                      dtype=torch.bfloat16)
            self.control_sig = (tuple(model.control_scales), model.only_mid_control)
        # joint FG-DM factor: adapter features of the pcond latent, step-invariant (openaimodel.py:836-844)
        self.fa = None
        if pcond_rows:
            self.fa = unet.encode_prompt(torch.zeros((pcond_rows, C, H, W), device=dev, dtype=torch.float32))
        # time-embedding vectors of a whole schedule, one row per DDIM index (filled by run(); the captured step picks
        # row `index` on the device), for schedules of up to EMB_TABLE_STEPS steps
        self.emb_unet = torch.zeros((EMB_TABLE_STEPS, unet.w_emb.shape[0]), device=dev, dtype=torch.float32)
        self.emb_ctrl = (torch.zeros((EMB_TABLE_STEPS, model.control_model.w_emb.shape[0]), device=dev,
                                     dtype=torch.float32) if self.is_control else None)
        self.use_emb_table = False
        self.graph = None
        self.launches_per_step = 0

    def _refill(self, kvs, new):
        for (k, vt, _), (k2, vt2, _) in zip(kvs, new):
            k.copy_(k2)
            vt.copy_(vt2)

    def set_conditioning(self, cond_ctx, uncond_ctx, hint=None, pcond=None):
        """Hoisted, step-invariant work (once per sample call): attention.py:162-163 K/V projections of the
        text context for every cross-attention; cldm.py:796 hint stem; adapter features of a pcond latent."""
        if self.fa is not None:
            for dst, src in zip(self.fa, self.model.model.diffusion_model.encode_prompt(pcond.to(self.x.device))):
                dst.copy_(src)
        if self.cfg:
            self.ctx[:self.B].copy_(uncond_ctx)          # rows [0,B) uncond, [B,2B) cond: ddim.py:226
            self.ctx[self.B:].copy_(cond_ctx)
        else:
            self.ctx.copy_(cond_ctx)
        unet = self.model.model.diffusion_model
        self._refill(self.kv_unet, unet.project_context(self.ctx))
        if self.is_control:
            self._refill(self.kv_ctrl, self.model.control_model.project_context(self.ctx))
            self.guided.copy_(self.model.control_model.encode_hint(hint))

    def _eps(self, emb_table=False):
        """The model call of one step on the static buffers: eps for the R CFG rows of (self.x, self.t)."""
        m = self.model
        # CFG rows r and r + B read the same latent self.x[r], the same timestep (self.t is filled with ONE value per
        # step) and the same hint: the networks evaluate the context-free head of the encoder once (nets._encode)
        share = self.cfg and SHARE_CFG_HEAD
        eu = ec = None
        if emb_table:        # row `index` of the precomputed tables, broadcast to the R rows (stride-0 view)
            eu = torch.index_select(self.emb_unet, 0, self.index).expand(self.R, -1)
            ec = torch.index_select(self.emb_ctrl, 0, self.index).expand(self.R, -1) if self.is_control else None
        if self.is_control:
            return m.control_step(self.x, self.t, self.kv_unet, self.kv_ctrl, self.guided, self.R, shared_halves=share,
                                  emb_unet=eu, emb_ctrl=ec)
        return m.model.diffusion_model.forward(self.x, self.t, kv=self.kv_unet, rows=self.R, adapter_features=self.fa,
                                               shared_halves=share, emb_all=eu)

    def _step(self):
        ops.ddim_advance(self.counter, self.t_table, self.idx_table, self.t, self.index)
        eps = self._eps(self.use_emb_table)
        e_u = eps[:self.B]
        e_c = eps[self.B:] if self.cfg else None
        ops.cfg_ddim_step(e_u, e_c, self.x, self.coef, 0, self.scale, index_ptr=self.index, x_prev=self.x,
                          pred_x0=self.pred_x0)

    def eval_eps(self, t):
        """Multi-step samplers (PLMS): replay a captured graph of the model call alone at timestep t on self.x.
        Returns (e_uncond, e_cond) views of the graph's static output (valid until the next replay)."""
        if getattr(self, "eps_graph", None) is None:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                n0 = ops.launch_count()
                self._eps()                                   # eager warm-up (function attributes, allocator)
                self.eps_launches = ops.launch_count() - n0
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            self.eps_graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.eps_graph):
                self.eps_out = self._eps()
        self.t.fill_(float(t) if self.t.is_floating_point() else int(t))
        self.eps_graph.replay()
        ops.count_launches(self.eps_launches)
        return self.eps_out[:self.B], (self.eps_out[self.B:] if self.cfg else None)

    def capture(self, scale, use_emb_table=False):
        self.scale = float(scale)
        self.use_emb_table = use_emb_table
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            self.counter.zero_()
            n0 = ops.launch_count()
            self._step()                                  # eager warm-up (function attributes, allocator)
            self.launches_per_step = ops.launch_count() - n0
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._step()

    def run(self, x_T, ts, coef, scale):
        """ts: int64 numpy [S] in visiting order; coef: fp32 [S,5] rows indexed by `index`."""
        S = len(ts)
        table = EMB_TABLES and S <= EMB_TABLE_STEPS
        if self.graph is None or scale != self.scale or table != self.use_emb_table:
            self.capture(scale, table)
        dev = self.x.device
        if table:            # index S-1-i belongs to the i-th visited timestep: rows in ascending DDIM index
            t_by_index = torch.as_tensor(np.ascontiguousarray(ts[::-1]), dtype=torch.int64).to(dev)
            self.emb_unet[:S].copy_(self.model.model.diffusion_model.emb_table(t_by_index))
            if self.is_control:
                self.emb_ctrl[:S].copy_(self.model.control_model.emb_table(t_by_index))
        self.t_table[:S].copy_(torch.as_tensor(np.ascontiguousarray(ts), dtype=torch.int64), non_blocking=True)
        self.idx_table[:S].copy_(torch.arange(S - 1, -1, -1, dtype=torch.int32), non_blocking=True)
        self.coef[:S].copy_(coef, non_blocking=True)
        self.counter.zero_()
        self.x.copy_(x_T)
        for i in range(S):
            self.graph.replay()
            ops.count_launches(self.launches_per_step)
            yield i


class DDIMSampler(object):
    def __init__(self, model, schedule="linear", **kwargs):
        super().__init__()
        self.model = model
        self.ddpm_num_timesteps = model.num_timesteps
        self.schedule = schedule
        self._graphs = {}
        # FGDM_USE_GRAPH=0: every step goes through p_sample_ddim / apply_model (the reference's own control flow)
        self.use_graph = kwargs.get("use_graph", os.environ.get("FGDM_USE_GRAPH", "1") != "0")

    def register_buffer(self, name, attr):
        setattr(self, name, attr)

    def make_schedule(self, ddim_num_steps, ddim_discretize="uniform", ddim_eta=0., verbose=True):
        """ddim.py:26-55. Tables are kept on the host (fp32 / numpy, same dtypes as the reference) and
        packed into the [S,5] coefficient table the fused update kernel reads."""
        self.ddim_timesteps = make_ddim_timesteps(ddim_discretize, ddim_num_steps, self.ddpm_num_timesteps, verbose)
        ac = self.model.alphas_cumprod.detach().cpu().to(torch.float32)
        assert ac.shape[0] == self.ddpm_num_timesteps, 'alphas have to be defined for each timestep'
        self.betas = self.model.betas.detach().cpu().to(torch.float32)
        self.alphas_cumprod = ac
        self.alphas_cumprod_prev = self.model.alphas_cumprod_prev.detach().cpu().to(torch.float32)
        sigmas, alphas, alphas_prev = make_ddim_sampling_parameters(ac, self.ddim_timesteps, ddim_eta, verbose)
        self.ddim_sigmas = sigmas
        self.ddim_alphas = alphas
        self.ddim_alphas_prev = alphas_prev
        self.ddim_sqrt_one_minus_alphas = torch.sqrt(1. - alphas)
        # p_sample_ddim (ddim.py:254-262) turns each scalar into an fp32 tensor before the arithmetic
        a_t = alphas.to(torch.float32)
        a_prev = torch.tensor(alphas_prev, dtype=torch.float32)
        sig = torch.tensor(np.asarray(sigmas, dtype=np.float64), dtype=torch.float32)
        self.coef_table = torch.stack([self.ddim_sqrt_one_minus_alphas.to(torch.float32), a_t.sqrt(), a_prev.sqrt(),
                                       (1. - a_prev - sig ** 2).sqrt(), sig], dim=1).contiguous()

    # ------------------------------------------------------------------------------------------
    @torch.no_grad()
    def sample(self, S, batch_size, shape, conditioning=None, callback=None, normals_sequence=None,
               img_callback=None, quantize_x0=False, eta=0., mask=None, x0=None, temperature=1., noise_dropout=0.,
               score_corrector=None, corrector_kwargs=None, verbose=True, x_T=None, log_every_t=100,
               unconditional_guidance_scale=1., unconditional_conditioning=None, inference_loss=False,
               dynamic_threshold=None, ucg_schedule=None, **kwargs):
        if conditioning is not None:
            if isinstance(conditioning, dict):
                ctmp = conditioning[list(conditioning.keys())[0]]
                while isinstance(ctmp, list):
                    ctmp = ctmp[0]
                cbs = ctmp.shape[0]
                if cbs != batch_size:
                    print(f"Warning: Got {cbs} conditionings but batch-size is {batch_size}")
            elif conditioning.shape[0] != batch_size:
                print(f"Warning: Got {conditioning.shape[0]} conditionings but batch-size is {batch_size}")
        bad = dict(mask=mask is not None, x0=x0 is not None, quantize_x0=quantize_x0, noise_dropout=noise_dropout > 0,
                   score_corrector=score_corrector is not None, inference_loss=inference_loss,
                   dynamic_threshold=dynamic_threshold is not None, ucg_schedule=ucg_schedule is not None,
                   return_conds=bool(kwargs.get("return_conds")),
                   composable_diffusion="composable_diffusion" in kwargs,
                   augmented_conditoning="augmented_conditoning" in kwargs)
        bad = [k for k, v in bad.items() if v]
        if bad:
            raise NotImplementedError(f"DDIMSampler.sample: options outside the FG-DM sampling hot path: {bad}")
        self.make_schedule(ddim_num_steps=S, ddim_eta=eta, verbose=False)
        C, H, W = shape
        size = (batch_size, C, H, W)
        return self.ddim_sampling(conditioning, size, callback=callback, img_callback=img_callback,
                                  temperature=temperature, x_T=x_T, log_every_t=log_every_t,
                                  unconditional_guidance_scale=unconditional_guidance_scale,
                                  unconditional_conditioning=unconditional_conditioning, eta=eta,
                                  pcond=kwargs.get("pcond"))

    @staticmethod
    def _split_cond(c):
        """Tensor (FG-DM factor) or {'c_concat': [hint], 'c_crossattn': [ctx]} (ControlNet factor)."""
        if c is None:
            return None, None
        if isinstance(c, dict):
            cc = c["c_crossattn"]
            ctx = cc[0] if len(cc) == 1 else torch.cat(cc, 1)
            hint = c.get("c_concat")
            if hint is not None:
                hint = hint[0] if len(hint) == 1 else torch.cat(hint, 1)
            return ctx, hint
        if isinstance(c, list):
            return (c[0] if len(c) == 1 else torch.cat(c, 1)), None
        return c, None

    @torch.no_grad()
    def ddim_sampling(self, cond, shape, x_T=None, callback=None, img_callback=None, log_every_t=100,
                      temperature=1., unconditional_guidance_scale=1., unconditional_conditioning=None, eta=0.,
                      pcond=None, **kwargs):
        """pcond: latent [P,4,h,w] of another factor for the adapter (ddpm.py:1536-1542 passes it down to
        UNetModel.forward through apply_model's kwargs)."""
        model = self.model
        device = model.device
        b = shape[0]
        img = torch.randn(shape, device=device) if x_T is None else x_T.to(device=device, dtype=torch.float32)
        timesteps = self.ddim_timesteps
        time_range = np.flip(timesteps)
        total_steps = timesteps.shape[0]
        intermediates = {'x_inter': [img], 'pred_x0': [img]}
        cfg = not (unconditional_conditioning is None or unconditional_guidance_scale == 1.)
        ctx_c, hint_c = self._split_cond(cond)
        ctx_u, hint_u = self._split_cond(unconditional_conditioning)
        # The captured step feeds ONE hint to both CFG halves. ControlNet guess mode (initialize_cn.py:87: un_cond has
        # c_concat=None, so ddim_hacked.py:191 evaluates the uncond half with control=None) and any other uncond hint
        # that differs from the cond hint take the sequential generic path below, which calls apply_model per half.
        same_hint = uncond_hint_matches(cond, unconditional_conditioning, hint_c, hint_u)
        fast = (self.use_graph and isinstance(model, LatentDiffusion) and eta == 0. and ctx_c is not None
                and same_hint and (not isinstance(model, ControlLDM) or hint_c is not None))
        if fast:
            if pcond is not None and isinstance(model, ControlLDM):
                raise NotImplementedError("pcond belongs to the FG-DM UNet (adapter), not to the ControlNet factor")
            key = (b, tuple(shape[1:]), ctx_c.shape[1], cfg,
                   (tuple(model.control_scales), model.only_mid_control) if isinstance(model, ControlLDM) else None,
                   0 if pcond is None else pcond.shape[0])
            g = self._graphs.get(key)
            if g is None:
                g = self._graphs[key] = _StepGraph(model, b, shape[1:], ctx_c.shape[1], cfg,
                                                   pcond_rows=0 if pcond is None else pcond.shape[0])
            g.set_conditioning(ctx_c, ctx_u, hint_c, pcond)
            for i in g.run(img, time_range, self.coef_table, unconditional_guidance_scale if cfg else 1.0):
                index = total_steps - i - 1
                if callback:
                    callback(i)
                if img_callback:
                    img_callback(g.pred_x0.clone(), i)
                if index % log_every_t == 0 or index == total_steps - 1:
                    intermediates['x_inter'].append(g.x.clone())
                    intermediates['pred_x0'].append(g.pred_x0.clone())
            return g.x.clone(), intermediates

        # generic path: any object with the reference's apply_model (still our kernels for CFG + update)
        coef = self.coef_table.to(device)
        for i, step in enumerate(time_range):
            index = total_steps - i - 1
            ts = torch.full((b,), int(step), device=device, dtype=torch.long)
            img, pred_x0 = self.p_sample_ddim(img, cond, ts, index=index, temperature=temperature,
                                              unconditional_guidance_scale=unconditional_guidance_scale,
                                              unconditional_conditioning=unconditional_conditioning, coef=coef,
                                              **({} if pcond is None else {"pcond": pcond}))
            if callback:
                callback(i)
            if img_callback:
                img_callback(pred_x0, i)
            if index % log_every_t == 0 or index == total_steps - 1:
                intermediates['x_inter'].append(img)
                intermediates['pred_x0'].append(pred_x0)
        return img, intermediates

    @torch.no_grad()
    def p_sample_ddim(self, x, c, t, index, repeat_noise=False, use_original_steps=False, quantize_denoised=False,
                      temperature=1., noise_dropout=0., score_corrector=None, corrector_kwargs=None,
                      unconditional_guidance_scale=1., unconditional_conditioning=None, coef=None, **kwargs):
        """ddim.py:179-286 / ddim_hacked.py:180-231 for one step through `model.apply_model`."""
        if use_original_steps or quantize_denoised or noise_dropout > 0. or score_corrector is not None:
            raise NotImplementedError("p_sample_ddim: option outside the FG-DM sampling hot path")
        b = x.shape[0]
        if coef is None:
            coef = self.coef_table.to(x.device)
        mk = {"pcond": kwargs["pcond"]} if kwargs.get("pcond") is not None else {}   # forwarded like ddim.py:192,241
        if unconditional_conditioning is None or unconditional_guidance_scale == 1.:
            e_u, e_c = self.model.apply_model(x, t, c, **mk), None
        elif isinstance(c, dict):                         # ddim_hacked.py:190-191: two sequential calls
            e_c = self.model.apply_model(x, t, c, **mk)
            e_u = self.model.apply_model(x, t, unconditional_conditioning, **mk)
        else:                                             # ddim.py:222-242: one batched call
            x_in = torch.cat([x] * 2)
            t_in = torch.cat([t] * 2)
            c_in = torch.cat([unconditional_conditioning, c])
            e_u, e_c = self.model.apply_model(x_in, t_in, c_in, **mk).chunk(2)
        sigma = float(coef[index, 4])
        noise = torch.randn_like(x) if sigma != 0. else None     # ddim.py:265 (sigma = 0 when eta = 0)
        x_prev, pred_x0 = ops.cfg_ddim_step(e_u.float().contiguous(), None if e_c is None else e_c.float().contiguous(),
                                            x.float().contiguous(), coef, index, float(unconditional_guidance_scale),
                                            noise=noise, temperature=float(temperature))
        return x_prev, pred_x0
