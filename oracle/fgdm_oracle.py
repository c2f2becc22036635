"""TEST INFRASTRUCTURE ONLY — fp32 restatement of FG-DM's sampling hot path (the parity oracle).

This file is the checker, never the product: only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / `--impl reference` leg import it. Nothing under fgdm_b200/ imports it.

It restates, as plain functional torch fp32 over a state_dict that uses the REFERENCE's parameter
names and shapes, exactly the arithmetic the reference (DeepakSridhar/fgdm, /root/reference) performs on
this path. The reference is pure PyTorch; its own code cannot travel to the GPU box, so this
restatement does. Pinning: the reference ships no tests or golden vectors ("parity unpinned" by the
reference's own tests); instead oracle/make_golden.py runs the UNMODIFIED reference
modules in the build container on shared seeded weights, checks this restatement against them
(max |diff| <= 1e-4 relative) and commits the reference's outputs as fixtures under tests/golden/.

Every function cites the reference file:line it follows (paths relative to /root/reference).
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F

# ----------------------------------------------------------------------------------------------
# small building blocks
# ----------------------------------------------------------------------------------------------


def _conv(sd, p, x, stride=1, padding=None):
    w = sd[p + ".weight"]
    if padding is None:
        padding = w.shape[-1] // 2
    return F.conv2d(x, w, sd.get(p + ".bias"), stride=stride, padding=padding)


def _linear(sd, p, x):
    return F.linear(x, sd[p + ".weight"], sd.get(p + ".bias"))


def _gn(sd, p, x, eps):
    """GroupNorm32: ldm/modules/diffusionmodules/util.py:208-225 (32 groups, computed in fp32)."""
    return F.group_norm(x.float(), 32, sd[p + ".weight"], sd[p + ".bias"], eps)


def _ln(sd, p, x):
    return F.layer_norm(x, (x.shape[-1],), sd[p + ".weight"], sd[p + ".bias"], 1e-5)


def timestep_embedding(t, dim, max_period=10000):
    """ldm/modules/diffusionmodules/util.py:160-180."""
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(0, half, dtype=torch.float32) / half).to(t.device)
    args = t[:, None].float() * freqs[None]
    return torch.cat([torch.cos(args), torch.sin(args)], dim=-1)


def time_embed(sd, p, t, model_channels):
    """openaimodel.py:538-542,827-828: Linear, SiLU, Linear on the sinusoidal embedding."""
    e = timestep_embedding(t, model_channels)
    return _linear(sd, p + "time_embed.2", F.silu(_linear(sd, p + "time_embed.0", e)))


def resblock(sd, p, x, emb):
    """ResBlock._forward, openaimodel.py:275-301 (no up/down, no scale-shift norm); GN eps 1e-5."""
    h = _conv(sd, p + ".in_layers.2", F.silu(_gn(sd, p + ".in_layers.0", x, 1e-5)))
    h = h + _linear(sd, p + ".emb_layers.1", F.silu(emb))[:, :, None, None]
    h = _conv(sd, p + ".out_layers.3", F.silu(_gn(sd, p + ".out_layers.0", h, 1e-5)))
    if (p + ".skip_connection.weight") in sd:
        x = _conv(sd, p + ".skip_connection", x)
    return x + h


def cross_attention(sd, p, x, context, heads):
    """CrossAttention.forward, ldm/modules/attention.py:177-216 (first return value only; the
    head-averaged logits are dead at inference, attention.py:235-238)."""
    context = x if context is None else context
    q = F.linear(x, sd[p + ".to_q.weight"])
    k = F.linear(context, sd[p + ".to_k.weight"])
    v = F.linear(context, sd[p + ".to_v.weight"])
    b, n, c = q.shape
    d = c // heads

    def split(t):
        return t.reshape(b, -1, heads, d).permute(0, 2, 1, 3)

    q, k, v = split(q), split(k), split(v)
    sim = torch.einsum("bhid,bhjd->bhij", q, k) * (d ** -0.5)
    attn = sim.softmax(dim=-1)
    out = torch.einsum("bhij,bhjd->bhid", attn, v)
    out = out.permute(0, 2, 1, 3).reshape(b, n, c)
    return _linear(sd, p + ".to_out.0", out)


def feed_forward(sd, p, x):
    """FeedForward with GEGLU, attention.py:37-64 (exact erf GELU)."""
    h = _linear(sd, p + ".net.0.proj", x)
    a, gate = h.chunk(2, dim=-1)
    return _linear(sd, p + ".net.2", a * F.gelu(gate))


def spatial_transformer(sd, p, x, context, heads):
    """SpatialTransformer.forward attention.py:275-292 + BasicTransformerBlock._forward :234-240.
    Normalize = GroupNorm eps 1e-6 (attention.py:76-77)."""
    b, c, h, w = x.shape
    x_in = x
    t = _conv(sd, p + ".proj_in", _gn(sd, p + ".norm", x, 1e-6))
    t = t.permute(0, 2, 3, 1).reshape(b, h * w, c)
    bp = p + ".transformer_blocks.0"
    t = cross_attention(sd, bp + ".attn1", _ln(sd, bp + ".norm1", t), None, heads) + t
    t = cross_attention(sd, bp + ".attn2", _ln(sd, bp + ".norm2", t), context, heads) + t
    t = feed_forward(sd, bp + ".ff", _ln(sd, bp + ".norm3", t)) + t
    t = t.reshape(b, h, w, c).permute(0, 3, 1, 2)
    return _conv(sd, p + ".proj_out", t) + x_in


def adapter_forward(sd, p, x, channels=(320, 640, 1280, 1280), nums_rb=2):
    """Adapter.forward ldm/modules/encoders/adapter.py:334-346 with ResnetBlock.forward :301-313,
    as instantiated at openaimodel.py:556 (ksize=1, sk=True, use_conv=False -> AvgPool2d down)."""
    feats = []
    x = _conv(sd, p + ".conv_in", x)
    for i in range(len(channels)):
        for j in range(nums_rb):
            bp = f"{p}.body.{i * nums_rb + j}"
            if i != 0 and j == 0:
                x = F.avg_pool2d(x, 2, 2)
            if (bp + ".in_conv.weight") in sd:
                x = _conv(sd, bp + ".in_conv", x)
            h = _conv(sd, bp + ".block2", F.relu(_conv(sd, bp + ".block1", x)))
            x = h + x
        feats.append(x)
    return feats


# ----------------------------------------------------------------------------------------------
# UNet structure (openaimodel.py:469-734): which sub-modules each TimestepEmbedSequential holds
# ----------------------------------------------------------------------------------------------


def unet_layout(cfg):
    """Returns (input_blocks, middle, output_blocks); each block is a list of ('res'|'st'|'down'|'up'|
    'conv_in', ...) in the order the reference constructs them (openaimodel.py:559-721)."""
    mc, mult, nrb = cfg["model_channels"], cfg["channel_mult"], cfg["num_res_blocks"]
    attn_res = cfg["attention_resolutions"]
    inputs = [["conv_in"]]
    ds = 1
    for level in range(len(mult)):
        for _ in range(nrb):
            blk = ["res"]
            if ds in attn_res:
                blk.append("st")
            inputs.append(blk)
        if level != len(mult) - 1:
            inputs.append(["down"])
            ds *= 2
    middle = ["res", "st", "res"]
    outputs = []
    for level in reversed(range(len(mult))):
        for i in range(nrb + 1):
            blk = ["res"]
            if ds in attn_res:
                blk.append("st")
            if level and i == nrb:
                blk.append("up")
                ds //= 2
            outputs.append(blk)
    return inputs, middle, outputs


TRACE = None   # debugging aid: set to a list to collect every block output (NCHW fp32) in execution order


def _run_block(sd, p, kinds, h, emb, context, heads):
    """TimestepEmbedSequential.forward, openaimodel.py:85-93."""
    h = _run_block_inner(sd, p, kinds, h, emb, context, heads)
    if TRACE is not None:
        TRACE.append((p, h))
    return h


def _run_block_inner(sd, p, kinds, h, emb, context, heads):
    for j, kind in enumerate(kinds):
        q = f"{p}.{j}"
        if kind == "conv_in":
            h = _conv(sd, q, h)
        elif kind == "res":
            h = resblock(sd, q, h, emb)
        elif kind == "st":
            h = spatial_transformer(sd, q, h, context, heads)
        elif kind == "down":   # Downsample with conv, openaimodel.py:146-180
            h = _conv(sd, q + ".op", h, stride=2, padding=1)
        elif kind == "up":     # Upsample, openaimodel.py:95-130: nearest 2x then conv3x3
            h = _conv(sd, q + ".conv", F.interpolate(h, scale_factor=2, mode="nearest"))
    return h


DEFAULT_UNET_CFG = dict(in_channels=4, out_channels=4, model_channels=320, attention_resolutions=[4, 2, 1],
                        num_res_blocks=2, channel_mult=[1, 2, 4, 4], num_heads=8, context_dim=768)


def unet_forward(sd, x, t, context, cfg=None, prefix="", adapter=True, pcond=None):
    """FG-DM UNetModel.forward, openaimodel.py:808-884 (pcond=None => the adapter sees the noisy
    latent itself, :836-844; pcond given => the adapter sees that tensor (another factor's latent) and its features
    broadcast over the rows of h exactly as `h + fa[k]` does; features added after input blocks 2,5,8,11, :855-857).
    adapter=False restates forward_original (:753-806), the plain SD UNet."""
    cfg = cfg or DEFAULT_UNET_CFG
    p = prefix
    heads = cfg["num_heads"]
    inputs, middle, outputs = unet_layout(cfg)
    emb = time_embed(sd, p, t, cfg["model_channels"])
    h = x.float()
    fa = adapter_forward(sd, p + "adapter", h if pcond is None else pcond.float()) if adapter else None
    hs = []
    k = 0
    for idx, kinds in enumerate(inputs):
        h = _run_block(sd, f"{p}input_blocks.{idx}", kinds, h, emb, context, heads)
        if adapter and (idx + 1) % 3 == 0:
            h = h + fa[k]
            k += 1
            if TRACE is not None:
                TRACE[-1] = (TRACE[-1][0], h)
        hs.append(h)
    h = _run_block(sd, p + "middle_block", middle, h, emb, context, heads)
    for idx, kinds in enumerate(outputs):
        h = torch.cat([h, hs.pop()], dim=1)
        h = _run_block(sd, f"{p}output_blocks.{idx}", kinds, h, emb, context, heads)
    return _conv(sd, p + "out.2", F.silu(_gn(sd, p + "out.0", h, 1e-5)))


def hint_stem(sd, p, hint):
    """ControlNet.input_hint_block, controlnet/cldm/cldm.py:655-671 (8 convs, SiLU between)."""
    strides = [1, 1, 2, 1, 2, 1, 2, 1]
    h = hint
    for i, s in enumerate(strides):
        h = _conv(sd, f"{p}.{2 * i}", h, stride=s, padding=1)
        if i != len(strides) - 1:
            h = F.silu(h)
    return h


def controlnet_forward(sd, x, hint, t, context, cfg=None, prefix=""):
    """ControlNet.forward, controlnet/cldm/cldm.py:792-813 -> 13 control tensors."""
    cfg = cfg or DEFAULT_UNET_CFG
    p = prefix
    heads = cfg["num_heads"]
    inputs, middle, _ = unet_layout(cfg)
    emb = time_embed(sd, p, t, cfg["model_channels"])
    guided = hint_stem(sd, p + "input_hint_block", hint.float())
    outs = []
    h = x.float()
    for idx, kinds in enumerate(inputs):
        h = _run_block(sd, f"{p}input_blocks.{idx}", kinds, h, emb, context, heads)
        if idx == 0:
            h = h + guided
        outs.append(_conv(sd, f"{p}zero_convs.{idx}.0", h))
    h = _run_block(sd, p + "middle_block", middle, h, emb, context, heads)
    outs.append(_conv(sd, p + "middle_block_out.0", h))
    return outs


def controlled_unet_forward(sd, x, t, context, control, cfg=None, prefix="", only_mid_control=False):
    """ControlledUnetModel.forward, controlnet/cldm/cldm.py:27-50."""
    cfg = cfg or DEFAULT_UNET_CFG
    p = prefix
    heads = cfg["num_heads"]
    inputs, middle, outputs = unet_layout(cfg)
    emb = time_embed(sd, p, t, cfg["model_channels"])
    control = None if control is None else list(control)
    hs = []
    h = x.float()
    for idx, kinds in enumerate(inputs):
        h = _run_block(sd, f"{p}input_blocks.{idx}", kinds, h, emb, context, heads)
        hs.append(h)
    h = _run_block(sd, p + "middle_block", middle, h, emb, context, heads)
    if control is not None:
        h = h + control.pop()
    for idx, kinds in enumerate(outputs):
        if only_mid_control or control is None:
            h = torch.cat([h, hs.pop()], dim=1)
        else:
            h = torch.cat([h, hs.pop() + control.pop()], dim=1)
        h = _run_block(sd, f"{p}output_blocks.{idx}", kinds, h, emb, context, heads)
    return _conv(sd, p + "out.2", F.silu(_gn(sd, p + "out.0", h, 1e-5)))


def control_ldm_apply(sd, x, t, hint, context, control_scales=None, only_mid_control=False, cfg=None,
                      unet_prefix="model.diffusion_model.", control_prefix="control_model."):
    """ControlLDM.apply_model, controlnet/cldm/cldm.py:836-849. hint=None is the `c_concat is None` branch (:841-842,
    ControlNet guess mode's unconditional half): the UNet runs with control=None."""
    scales = control_scales or [1.0] * 13
    if hint is None:
        return controlled_unet_forward(sd, x, t, context, None, cfg, unet_prefix, only_mid_control)
    control = controlnet_forward(sd, x, hint, t, context, cfg, control_prefix)
    control = [c * s for c, s in zip(control, scales)]
    return controlled_unet_forward(sd, x, t, context, control, cfg, unet_prefix, only_mid_control)


# ----------------------------------------------------------------------------------------------
# VAE decoder (ldm/modules/diffusionmodules/model.py:462-569; GroupNorm eps 1e-6, swish)
# ----------------------------------------------------------------------------------------------

DEFAULT_VAE_CFG = dict(ch=128, out_ch=3, ch_mult=[1, 2, 4, 4], num_res_blocks=2, z_channels=4, embed_dim=4)


def _vae_res(sd, p, x):
    """ResnetBlock.forward model.py:121-141 (temb is None in the decoder)."""
    h = _conv(sd, p + ".conv1", F.silu(_gn(sd, p + ".norm1", x, 1e-6)))
    h = _conv(sd, p + ".conv2", F.silu(_gn(sd, p + ".norm2", h, 1e-6)))
    if (p + ".nin_shortcut.weight") in sd:
        x = _conv(sd, p + ".nin_shortcut", x)
    return x + h


def _vae_attn(sd, p, x):
    """AttnBlock.forward model.py:172-202: single head, d = C."""
    h = _gn(sd, p + ".norm", x, 1e-6)
    q, k, v = _conv(sd, p + ".q", h), _conv(sd, p + ".k", h), _conv(sd, p + ".v", h)
    b, c, hh, ww = q.shape
    q = q.reshape(b, c, hh * ww).permute(0, 2, 1)
    k = k.reshape(b, c, hh * ww)
    w_ = torch.bmm(q, k) * (int(c) ** (-0.5))
    w_ = F.softmax(w_, dim=2)
    v = v.reshape(b, c, hh * ww)
    h = torch.bmm(v, w_.permute(0, 2, 1)).reshape(b, c, hh, ww)
    return x + _conv(sd, p + ".proj_out", h)


def vae_decode(sd, z, cfg=None, prefix="first_stage_model.", scale_factor=0.18215):
    """LatentDiffusion.decode_first_stage (ldm/models/diffusion/ddpm.py:831-889: z / scale_factor) ->
    AutoencoderKL.decode (ldm/models/autoencoder.py:330-333) -> Decoder.forward (model.py:535-569)."""
    cfg = cfg or DEFAULT_VAE_CFG
    p = prefix
    nres = len(cfg["ch_mult"])
    z = z.float() * (1.0 / scale_factor)
    z = _conv(sd, p + "post_quant_conv", z)
    h = _conv(sd, p + "decoder.conv_in", z)
    h = _vae_res(sd, p + "decoder.mid.block_1", h)
    h = _vae_attn(sd, p + "decoder.mid.attn_1", h)
    h = _vae_res(sd, p + "decoder.mid.block_2", h)
    for lvl in reversed(range(nres)):
        for i in range(cfg["num_res_blocks"] + 1):
            h = _vae_res(sd, f"{p}decoder.up.{lvl}.block.{i}", h)
        if lvl != 0:
            h = _conv(sd, f"{p}decoder.up.{lvl}.upsample.conv", F.interpolate(h, scale_factor=2.0, mode="nearest"))
    h = F.silu(_gn(sd, p + "decoder.norm_out", h, 1e-6))
    return _conv(sd, p + "decoder.conv_out", h)


# ----------------------------------------------------------------------------------------------
# schedules and the DDIM loop
# ----------------------------------------------------------------------------------------------


def make_schedule_buffers(timesteps=1000, linear_start=0.00085, linear_end=0.0120):
    """DDPM.register_schedule ldm/models/diffusion/ddpm.py:175-203 with make_beta_schedule('linear')
    util.py:21-25 (float64 numpy, stored as float32)."""
    betas = (torch.linspace(linear_start ** 0.5, linear_end ** 0.5, timesteps, dtype=torch.float64) ** 2).numpy()
    alphas_cumprod = np.cumprod(1.0 - betas, axis=0)
    alphas_cumprod_prev = np.append(1.0, alphas_cumprod[:-1])
    return dict(betas=torch.tensor(betas, dtype=torch.float32),
                alphas_cumprod=torch.tensor(alphas_cumprod, dtype=torch.float32),
                alphas_cumprod_prev=torch.tensor(alphas_cumprod_prev, dtype=torch.float32))


def make_ddim_schedule(alphas_cumprod, S, eta=0.0, num_ddpm=1000):
    """DDIMSampler.make_schedule ldm/models/diffusion/ddim.py:26-55 with util.py:46-74 ('uniform').
    alphas_cumprod: float32 torch tensor (the model buffer). Returns numpy float arrays exactly as the
    reference holds them (alphas: torch f32 indexed -> here numpy f32; alphas_prev passes through a
    float64 numpy array built from float32 values; sigmas likewise)."""
    c = num_ddpm // S
    ddim_timesteps = np.asarray(list(range(0, num_ddpm, c))) + 1
    ac = alphas_cumprod.cpu()
    alphas = ac[ddim_timesteps]                                   # torch f32
    alphas_prev = np.asarray([ac[0]] + ac[ddim_timesteps[:-1]].tolist())   # numpy f64 of f32 values
    sigmas = eta * np.sqrt((1 - alphas_prev) / (1 - alphas) * (1 - alphas / alphas_prev))
    sqrt_one_minus_alphas = np.sqrt(1.0 - alphas)
    return dict(timesteps=ddim_timesteps, alphas=alphas, alphas_prev=alphas_prev,
                sigmas=torch.as_tensor(sigmas), sqrt_one_minus_alphas=torch.as_tensor(sqrt_one_minus_alphas))


def ddim_update(x, e_t, sched, index, noise=None, temperature=1.0):
    """p_sample_ddim's update, ddim.py:249-268 (= ddim_hacked.py:207-230): scalars become fp32
    [b,1,1,1] tensors via torch.full (:254-257)."""
    b, dev = x.shape[0], x.device
    a_t = torch.full((b, 1, 1, 1), float(sched["alphas"][index]), device=dev)
    a_prev = torch.full((b, 1, 1, 1), float(sched["alphas_prev"][index]), device=dev)
    sigma_t = torch.full((b, 1, 1, 1), float(sched["sigmas"][index]), device=dev)
    s1m = torch.full((b, 1, 1, 1), float(sched["sqrt_one_minus_alphas"][index]), device=dev)
    pred_x0 = (x - s1m * e_t) / a_t.sqrt()
    dir_xt = (1.0 - a_prev - sigma_t ** 2).sqrt() * e_t
    if noise is None:
        noise = torch.zeros_like(x)
    x_prev = a_prev.sqrt() * pred_x0 + dir_xt + sigma_t * noise * temperature
    return x_prev, pred_x0


def ddim_sample(apply_model, S, shape, cond, uncond, scale, x_T, alphas_cumprod, eta=0.0, log_every_t=100,
                sequential_cfg=False, record=None, max_steps=None):
    """DDIMSampler.sample/ddim_sampling/p_sample_ddim with CFG.
    sequential_cfg=False: ldm/models/diffusion/ddim.py:116-177,221-243 (one batched call on cat([uc, c])).
    sequential_cfg=True:  controlnet/cldm/ddim_hacked.py:150-192 (cond call then uncond call).
    apply_model(x, t, c) -> eps. eta must be 0 here (the noise term vanishes; the reference still draws
    and discards randn, ddim.py:265). record: optional list receiving (e_t, x_prev) per step."""
    assert eta == 0.0
    sched = make_ddim_schedule(alphas_cumprod, S, eta)
    ts = sched["timesteps"]
    img = x_T
    b = shape[0]
    inter = {"x_inter": [img], "pred_x0": [img]}
    total = ts.shape[0]
    for i, step in enumerate(np.flip(ts)):
        if max_steps is not None and i >= max_steps:   # timing aid: run only the first steps of the loop
            break
        index = total - i - 1
        t = torch.full((b,), int(step), device=img.device, dtype=torch.long)
        if uncond is None or scale == 1.0:
            e_t = apply_model(img, t, cond)
        elif sequential_cfg:
            e_c = apply_model(img, t, cond)
            e_u = apply_model(img, t, uncond)
            e_t = e_u + scale * (e_c - e_u)
        else:
            x_in = torch.cat([img] * 2)
            t_in = torch.cat([t] * 2)
            c_in = torch.cat([uncond, cond])
            e_u, e_c = apply_model(x_in, t_in, c_in).chunk(2)
            e_t = e_u + scale * (e_c - e_u)
        img, pred_x0 = ddim_update(img, e_t, sched, index)
        if record is not None:
            record.append((e_t, img))
        if index % log_every_t == 0 or index == total - 1:
            inter["x_inter"].append(img)
            inter["pred_x0"].append(pred_x0)
    return img, inter


# ----------------------------------------------------------------------------------------------
# DPM-Solver++ (2M): DPMSolverSampler.sample, ldm/models/diffusion/dpm_solver/sampler.py:23-82, which fixes the
# configuration to NoiseScheduleVP('discrete'), model_type "noise", classifier-free guidance, predict_x0=True,
# multistep order 2, skip_type "time_uniform", lower_order_final=True (dpm_solver.py line numbers below).
# ----------------------------------------------------------------------------------------------
def dpm_schedule(alphas_cumprod):
    """NoiseScheduleVP('discrete', alphas_cumprod=...), dpm_solver.py:99-108."""
    log_alphas = 0.5 * torch.log(alphas_cumprod.float())
    n = log_alphas.shape[0]
    return dict(N=n, t_array=torch.linspace(0., 1., n + 1)[1:], log_alpha=log_alphas)


def dpm_marginals(sched, t):
    """marginal_log_mean_coeff / alpha / std / lambda at continuous t in [1/N, 1], dpm_solver.py:125-156 with the
    piecewise-linear interpolate_fn (:1132-1171; inside the knot range it is plain linear interpolation)."""
    xp, yp = sched["t_array"].double().numpy(), sched["log_alpha"].double().numpy()
    la = torch.tensor(np.interp(np.asarray(t, dtype=np.float64), xp, yp), dtype=torch.float32)
    return dict(log_alpha=la, alpha=torch.exp(la), sigma=torch.sqrt(1. - torch.exp(2. * la)),
                lam=la - 0.5 * torch.log(1. - torch.exp(2. * la)))


def dpm_solver_sample(apply_model, S, shape, cond, uncond, scale, x_T, alphas_cumprod, record=None):
    """DPM_Solver.sample(method='multistep', order=2), dpm_solver.py:1060-1096; model_wrapper's classifier-free
    model_fn (:321-344) with t_input = (t - 1/N) * 1000 (:278-287, a FLOAT timestep); data_prediction_fn (:386-399);
    dpm_solver_first_update (:504-535) and multistep_dpm_solver_second_update (:755-796), predict_x0 branches."""
    sched = dpm_schedule(alphas_cumprod)
    n = sched["N"]
    ts = torch.linspace(1.0, 1.0 / n, S + 1)                    # get_time_steps('time_uniform'), :429-430
    mg = dpm_marginals(sched, ts.numpy())
    b = shape[0]
    x = x_T

    def model_fn(x, i):                                         # -> x0 prediction at timesteps[i]
        t_in = ((ts[i] - 1. / n) * 1000.).to(x.device).expand(b)
        if uncond is None or scale == 1.0:
            e = apply_model(x, t_in, cond)
        else:
            e_u, e_c = apply_model(torch.cat([x] * 2), torch.cat([t_in] * 2), torch.cat([uncond, cond])).chunk(2)
            e = e_u + scale * (e_c - e_u)
        return (x - mg["sigma"][i] * e) / mg["alpha"][i]

    def first(x, s, t, m_s):
        h = mg["lam"][t] - mg["lam"][s]
        return (mg["sigma"][t] / mg["sigma"][s]) * x - (mg["alpha"][t] * torch.expm1(-h)) * m_s

    def second(x, p1, p0, t, m1, m0):
        h_0 = mg["lam"][p0] - mg["lam"][p1]
        h = mg["lam"][t] - mg["lam"][p0]
        r0 = h_0 / h
        d1 = (1. / r0) * (m0 - m1)
        a = mg["alpha"][t] * (torch.exp(-h) - 1.)
        return (mg["sigma"][t] / mg["sigma"][p0]) * x - a * m0 - 0.5 * a * d1

    assert S >= 2
    models, tprev = [model_fn(x, 0)], [0]
    x = first(x, 0, 1, models[-1])                              # init_order = 1
    if record is not None:
        record.append(x)
    models.append(model_fn(x, 1))
    tprev.append(1)
    for step in range(2, S + 1):
        order = min(2, S + 1 - step) if S < 15 else 2           # lower_order_final
        if order == 1:
            x = first(x, tprev[-1], step, models[-1])
        else:
            x = second(x, tprev[0], tprev[1], step, models[0], models[1])
        if record is not None:
            record.append(x)
        tprev[0], models[0] = tprev[1], models[1]
        tprev[1] = step
        if step < S:
            models[1] = model_fn(x, step)
    return x


# ----------------------------------------------------------------------------------------------
# text conditioning: FrozenCLIPEmbedder.forward, ldm/modules/encoders/modules.py:137-163
# ----------------------------------------------------------------------------------------------
# The arithmetic lives in a third-party dependency that is not vendored in the reference: Hugging Face
# `transformers` (pinned transformers==4.19.2, fgdm.yml:175), class CLIPTextModel / CLIPTextTransformer
# (models/clip/modeling_clip.py). Restated from its published algorithm: token + learned position embeddings; 12
# pre-LayerNorm blocks of [causal multi-head self-attention (q scaled by d^-1/2 before q k^T), + residual,
# MLP fc1 -> quick_gelu (x * sigmoid(1.702 x)) -> fc2, + residual]; final LayerNorm; the embedder returns
# `last_hidden_state` (no pooling, modules.py:160). Pinned against transformers' own CLIPTextModel run in the build
# container (oracle/make_golden.py --only clip; tests/golden/clip_text.npz).
CLIP_TEXT_CFG = dict(vocab_size=49408, hidden_size=768, intermediate_size=3072, num_hidden_layers=12,
                     num_attention_heads=12, max_position_embeddings=77, layer_norm_eps=1e-5)


def clip_text_forward(sd, ids, cfg=None, prefix="transformer.text_model."):
    """ids: int64 [B, L<=77] -> last_hidden_state fp32 [B, L, hidden]."""
    cfg = cfg or CLIP_TEXT_CFG
    p = prefix
    b, L = ids.shape
    heads = cfg["num_attention_heads"]
    eps = cfg["layer_norm_eps"]
    h = sd[p + "embeddings.token_embedding.weight"].float()[ids] + \
        sd[p + "embeddings.position_embedding.weight"].float()[:L][None]
    c = h.shape[-1]
    d = c // heads
    mask = torch.full((L, L), float("-inf"), device=h.device).triu(1)          # causal: token i sees j <= i

    def lin(name, x):
        return F.linear(x, sd[name + ".weight"].float(), sd[name + ".bias"].float())

    for i in range(cfg["num_hidden_layers"]):
        q_ = f"{p}encoder.layers.{i}."
        r = h
        x = F.layer_norm(h, (c,), sd[q_ + "layer_norm1.weight"].float(), sd[q_ + "layer_norm1.bias"].float(), eps)
        q = (lin(q_ + "self_attn.q_proj", x) * d ** -0.5).view(b, L, heads, d).transpose(1, 2)
        k = lin(q_ + "self_attn.k_proj", x).view(b, L, heads, d).transpose(1, 2)
        v = lin(q_ + "self_attn.v_proj", x).view(b, L, heads, d).transpose(1, 2)
        a = torch.softmax(q @ k.transpose(-1, -2) + mask, dim=-1) @ v
        h = r + lin(q_ + "self_attn.out_proj", a.transpose(1, 2).reshape(b, L, c))
        r = h
        x = F.layer_norm(h, (c,), sd[q_ + "layer_norm2.weight"].float(), sd[q_ + "layer_norm2.bias"].float(), eps)
        x = lin(q_ + "mlp.fc1", x)
        x = x * torch.sigmoid(1.702 * x)
        h = r + lin(q_ + "mlp.fc2", x)
    return F.layer_norm(h, (c,), sd[p + "final_layer_norm.weight"].float(), sd[p + "final_layer_norm.bias"].float(), eps)


def plms_sample(apply_model, S, shape, cond, uncond, scale, x_T, alphas_cumprod, log_every_t=100, record=None):
    """PLMSSampler.sample/plms_sampling/p_sample_plms, ldm/models/diffusion/plms.py:58-236 (eta must be 0, :25-26).
    Same schedule tables as DDIM (:24-56). The first step is the pseudo improved Euler (two model calls, :216-220), then
    Adams-Bashforth of order 2, 3, 4 on the CFG-combined eps history (:221-229). record: receives (e_t, x_prev)."""
    sched = make_ddim_schedule(alphas_cumprod, S, 0.0)
    ts = sched["timesteps"]
    time_range = np.flip(ts)
    img = x_T
    b = shape[0]
    inter = {"x_inter": [img], "pred_x0": [img]}
    total = ts.shape[0]
    old_eps = []

    def model_output(x, t):                                   # plms.py:178-192
        if uncond is None or scale == 1.0:
            return apply_model(x, t, cond)
        e_u, e_c = apply_model(torch.cat([x] * 2), torch.cat([t] * 2), torch.cat([uncond, cond])).chunk(2)
        return e_u + scale * (e_c - e_u)

    for i, step in enumerate(time_range):
        index = total - i - 1
        t = torch.full((b,), int(step), device=img.device, dtype=torch.long)
        t_next = torch.full((b,), int(time_range[min(i + 1, len(time_range) - 1)]), device=img.device, dtype=torch.long)
        e_t = model_output(img, t)
        if len(old_eps) == 0:
            x_prev, _ = ddim_update(img, e_t, sched, index)
            e_t_next = model_output(x_prev, t_next)
            e_t_prime = (e_t + e_t_next) / 2
        elif len(old_eps) == 1:
            e_t_prime = (3 * e_t - old_eps[-1]) / 2
        elif len(old_eps) == 2:
            e_t_prime = (23 * e_t - 16 * old_eps[-1] + 5 * old_eps[-2]) / 12
        else:
            e_t_prime = (55 * e_t - 59 * old_eps[-1] + 37 * old_eps[-2] - 9 * old_eps[-3]) / 24
        img, pred_x0 = ddim_update(img, e_t_prime, sched, index)
        old_eps.append(e_t)
        if len(old_eps) >= 4:
            old_eps.pop(0)
        if record is not None:
            record.append((e_t, img))
        if index % log_every_t == 0 or index == total - 1:
            inter["x_inter"].append(img)
            inter["pred_x0"].append(pred_x0)
    return img, inter


# ----------------------------------------------------------------------------------------------
# factor hand-off (scripts/txt2img_fgdm_inference.py:244-261; controlnet/initialize_cn.py:77-80,98-99)
# ----------------------------------------------------------------------------------------------


def seg_image_to_u8(x):
    """txt2img_fgdm_inference.py:245,249-251: clamp((x+1)/2, 0, 1); 255 * (c h w -> h w c); astype(uint8)
    (truncation). x: [B,3,H,W] float -> numpy uint8 [B,H,W,3]."""
    x = torch.clamp((x.float() + 1.0) / 2.0, min=0.0, max=1.0)
    arr = 255.0 * x.permute(0, 2, 3, 1).cpu().numpy()
    return arr.astype(np.uint8)


def final_image_to_u8(x):
    """initialize_cn.py:98-99: (b c h w -> b h w c) * 127.5 + 127.5, clip(0, 255), astype(uint8)."""
    arr = (x.float().permute(0, 2, 3, 1) * 127.5 + 127.5).cpu().numpy().clip(0, 255)
    return arr.astype(np.uint8)


def resize_u8_bilinear(img, dh, dw):
    """cv2.resize(img, (dw, dh), interpolation=cv2.INTER_LINEAR) for 8-bit images
    (txt2img_fgdm_inference.py:257-259). OpenCV (third-party, opencv-python 4.x; algorithm in
    modules/imgproc/src/resize.cpp) uses half-pixel centres, coefficients rounded to 11-bit fixed point
    (INTER_RESIZE_COEF_SCALE = 2048), a horizontal pass in int32 and the vertical pass
    (((b0 * (r0 >> 4)) >> 16) + ((b1 * (r1 >> 4)) >> 16) + 2) >> 2. img: uint8 [H,W,C]."""
    sh, sw, c = img.shape

    def coeffs(dst, src, clamp_frac):
        scale = np.float32(src) / np.float32(dst)
        f = (np.arange(dst, dtype=np.float32) + np.float32(0.5)) * scale - np.float32(0.5)
        s = np.floor(f).astype(np.int32)
        f = (f - s).astype(np.float32)
        if clamp_frac:
            # horizontal pass only (resize.cpp: `if (sx < 0) fx = 0, sx = 0; if (sx >= ssize.width-1) fx = 0, ...`)
            lo = s < 0
            f[lo] = 0
            s[lo] = 0
            hi = s >= src - 1
            f[hi] = 0
            s[hi] = src - 1
        # vertical pass: the fraction is kept and BOTH source rows are clipped into the image
        c0 = np.rint((np.float32(1.0) - f) * np.float32(2048)).astype(np.int32)
        c1 = np.rint(f * np.float32(2048)).astype(np.int32)
        return np.clip(s, 0, src - 1), np.clip(s + 1, 0, src - 1), c0, c1

    sx0, sx1, a0, a1 = coeffs(dw, sw, True)
    sy0, sy1, b0, b1 = coeffs(dh, sh, False)
    src = img.astype(np.int32)
    rows = src[:, sx0, :] * a0[None, :, None] + src[:, sx1, :] * a1[None, :, None]     # [SH, DW, C]
    r0, r1 = rows[sy0], rows[sy1]
    out = (((b0[:, None, None] * (r0 >> 4)) >> 16) + ((b1[:, None, None] * (r1 >> 4)) >> 16) + 2) >> 2
    return np.clip(out, 0, 255).astype(np.uint8)


def u8_to_hint(u8):
    """initialize_cn.py:78-80: uint8 [B,H,W,3] -> float / 255 -> [B,3,H,W]."""
    t = torch.from_numpy(np.ascontiguousarray(u8)).float() / 255.0
    return t.permute(0, 3, 1, 2).contiguous()
