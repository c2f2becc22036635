"""Networks of the FG-DM sampling hot path, constructor- and state_dict-compatible with the reference:

  UNetModel            ldm/modules/diffusionmodules/openaimodel.py:469-884 (FG-DM: Adapter self-prompting)
  ControlNet           controlnet/cldm/cldm.py:545-813
  ControlledUnetModel  controlnet/cldm/cldm.py:27-50
  AutoencoderKL        ldm/models/autoencoder.py (decode only) + Decoder, ldm/modules/diffusionmodules/model.py:462-569

They are the `target:` classes a maintainer points `instantiate_from_config` (ldm/util.py:78-93) at.
Inputs/outputs at the model boundary are the reference's NCHW fp32 tensors; inside, everything is NHWC
bf16 and every arithmetic op is a kernel of libfgdm_b200.so. Only inference (`no_grad`) is provided.
"""
from __future__ import annotations

import os
from collections import namedtuple

import torch

from . import _lib, ops
from .layers import (ACT_NONE, ACT_RELU, ACT_SILU, CONV_1X1, CONV_3X3, OUT_F32, OUT_QKV, Conv2d, Downsample, Layer,
                     Linear, Norm, ResBlock, Seq, SpatialTransformer, TimeEmbed, Upsample, batch_emb_layers, bf16,
                     repeat_rows)


# FGDM_VAE_FLASH=0: the VAE mid-block attention materialises S and P per image as the reference does (A/B switch)
VAE_FLASH = os.environ.get("FGDM_VAE_FLASH", "1") != "0"


def _unsupported(**flags):
    bad = {k: v for k, v in flags.items() if v}
    if bad:
        raise NotImplementedError(f"fgdm_b200 implements the SD-v1.4/FG-DM inference configuration only; "
                                  f"unsupported constructor options: {bad}")


class IncompatibleKeys(namedtuple("IncompatibleKeys", ["missing_keys", "unexpected_keys"])):
    """What torch.nn.Module.load_state_dict returns (torch/nn/modules/module.py `_IncompatibleKeys`)."""

    def __repr__(self):
        if not self.missing_keys and not self.unexpected_keys:
            return "<All keys matched successfully>"
        return super().__repr__()


class _Net(Layer):
    """Common module-ish surface (enough of nn.Module for the reference's call sites)."""

    def __init__(self):
        super().__init__()
        self.device = None
        self._loaded = False

    def param_shapes(self) -> dict:
        return dict(self.specs(""))

    def load_state_dict(self, sd, strict=True, prefix="", device=None):
        """torch.nn.Module.load_state_dict semantics (the reference's call sites unpack `m, u = ...`,
        scripts/txt2img_fgdm_inference.py:28, controlnet/initialize_cn.py:28-35): returns
        `IncompatibleKeys(missing_keys, unexpected_keys)`; strict=True raises RuntimeError when either list is
        non-empty. A shape mismatch raises RuntimeError under both settings (as torch does). With strict=False and
        missing keys nothing is packed: the net stays unloaded and the first forward raises, naming the keys —
        there are no default-initialised weights to fall back to on this path."""
        dev = torch.device(device or self.device or "cuda")
        _lib.require_device()
        want = self.param_shapes()
        missing = [prefix + k for k in want if prefix + k not in sd]
        unexpected = [k for k in sd if k.startswith(prefix) and k[len(prefix):] not in want]
        bad = [f"size mismatch for {prefix + k}: checkpoint {tuple(sd[prefix + k].shape)} vs model {tuple(shp)}"
               for k, shp in want.items() if prefix + k in sd and tuple(sd[prefix + k].shape) != tuple(shp)]
        errs = list(bad)
        if strict and unexpected:
            errs.append("Unexpected key(s) in state_dict: " + ", ".join(f'"{k}"' for k in unexpected[:8])
                        + (" ..." if len(unexpected) > 8 else ""))
        if strict and missing:
            errs.append("Missing key(s) in state_dict: " + ", ".join(f'"{k}"' for k in missing[:8])
                        + (" ..." if len(missing) > 8 else ""))
        if errs:
            raise RuntimeError(f"Error(s) in loading state_dict for {type(self).__name__}:\n\t" + "\n\t".join(errs))
        self._missing = missing
        if not missing:
            with torch.no_grad():
                self.load(sd, prefix, dev)
                self._finalize(dev)
            self.device = dev
            self._loaded = True
        return IncompatibleKeys(missing, unexpected)

    def _finalize(self, dev):
        pass

    # nn.Module look-alikes used by reference-side glue
    def eval(self):
        return self

    def cuda(self, *a):
        return self

    def cpu(self):
        """The packed weights live on the B200 only; `.cpu()` (controlnet/cldm/model.py:26) is a no-op."""
        return self

    def to(self, *a, **k):
        return self

    def parameters(self):
        return iter(())

    def __call__(self, *a, **k):
        if not self._loaded:
            miss = getattr(self, "_missing", None)
            raise RuntimeError("weights not loaded: " + (f"{len(miss)} keys were missing from the state_dict, e.g. {miss[:3]}"
                                                         if miss else "call load_state_dict(reference_state_dict) first"))
        with torch.no_grad():
            return self.forward(*a, **k)


# ---------------------------------------------------------------------------------------------
# UNet family
# ---------------------------------------------------------------------------------------------
class _AdapterBlock(Layer):
    """ldm/modules/encoders/adapter.py:280-313 with ksize=1, sk=True, use_conv=False."""

    def __init__(self, cin, cout, down):
        super().__init__()
        self.down = down
        self.in_conv = self.add("in_conv", Conv2d(cin, cout, 1)) if cin != cout else None
        self.block1 = self.add("block1", Conv2d(cout, cout, 3))
        self.block2 = self.add("block2", Conv2d(cout, cout, 1))

    def forward(self, x):
        if self.down:
            x = ops.resample(x, 1)                       # AvgPool2d(2), adapter.py:270-277
        if self.in_conv is not None:
            x = self.in_conv.gemm(x)
        h = self.block1.gemm(x, act=ACT_RELU)
        return self.block2.gemm(h, res0=x)


class Adapter(Layer):
    """Adapter(cin, channels=[320,640,1280,1280], nums_rb=2, ksize=1, sk=True, use_conv=False):
    adapter.py:316-346 as instantiated at openaimodel.py:556."""

    def __init__(self, cin, channels=(320, 640, 1280, 1280), nums_rb=2):
        super().__init__()
        self.channels, self.nums_rb = list(channels), nums_rb
        body = []
        for i in range(len(channels)):
            for j in range(nums_rb):
                if i != 0 and j == 0:
                    body.append(_AdapterBlock(channels[i - 1], channels[i], True))
                else:
                    body.append(_AdapterBlock(channels[i], channels[i], False))
        self.body = self.add("body", Seq(body))
        self.conv_in = self.add("conv_in", Conv2d(cin, channels[0], 3))

    def forward(self, x_nchw, rows):
        """x_nchw: fp32 [nb_in, cin, h, w]; `rows` output images (input image index = row % nb_in)."""
        x = self.conv_in.from_nchw(x_nchw, rows)
        feats = []
        for i in range(len(self.channels)):
            for j in range(self.nums_rb):
                x = self.body[i * self.nums_rb + j].forward(x)
            feats.append(x)
        return feats


class _UNetBase(_Net):
    def _build(self, in_channels, model_channels, out_channels, num_res_blocks, attention_resolutions, channel_mult,
               num_heads, context_dim, decoder=True):
        mc = model_channels
        self.in_channels, self.model_channels, self.out_channels = in_channels, mc, out_channels
        self.num_heads, self.context_dim = num_heads, context_dim
        ted = 4 * mc
        self.time_embed = self.add("time_embed", TimeEmbed(mc))
        self._res, self._sts = [], []
        self.trace = None

        def res(cin, cout):
            r = ResBlock(cin, ted, cout)
            self._res.append(r)
            return r

        def st(ch):
            s = SpatialTransformer(ch, num_heads, ch // num_heads, context_dim)
            s.kv_slot = len(self._sts)
            self._sts.append(s)
            return s

        blocks = [Seq([Conv2d(in_channels, mc, 3)])]
        chans = [mc]
        ch, ds = mc, 1
        for level, mult in enumerate(channel_mult):
            for _ in range(num_res_blocks):
                layers = [res(ch, mult * mc)]
                ch = mult * mc
                if ds in attention_resolutions:
                    layers.append(st(ch))
                blocks.append(Seq(layers))
                chans.append(ch)
            if level != len(channel_mult) - 1:
                blocks.append(Seq([Downsample(ch)]))
                chans.append(ch)
                ds *= 2
        self.input_blocks = self.add("input_blocks", Seq(blocks))
        self.input_chans = list(chans)
        self.middle_block = self.add("middle_block", Seq([res(ch, ch), st(ch), res(ch, ch)]))
        self.mid_ch = ch
        if not decoder:
            return
        outs = []
        for level, mult in list(enumerate(channel_mult))[::-1]:
            for i in range(num_res_blocks + 1):
                ich = chans.pop()
                layers = [res(ch + ich, mc * mult)]
                ch = mc * mult
                if ds in attention_resolutions:
                    layers.append(st(ch))
                if level and i == num_res_blocks:
                    layers.append(Upsample(ch))
                    ds //= 2
                outs.append(Seq(layers))
        self.output_blocks = self.add("output_blocks", Seq(outs))
        self.out = self.add("out", Seq({0: Norm(ch), 2: Conv2d(mc, out_channels, 3)}))

    def _finalize(self, dev):
        self.w_emb, self.b_emb = batch_emb_layers(self._res, dev)

    # -- step-invariant work: text K/V projections of every cross-attention (attention.py:162-163) --
    def project_context(self, context):
        """context: [R, L, context_dim] (any float dtype) -> per-SpatialTransformer (k, vt, L)."""
        ctx = context.to(bf16).contiguous()
        return [s.project_context(ctx) for s in self._sts]

    def emb_table(self, timesteps):
        """[S] timesteps -> [S, total] fp32: every ResBlock's emb_layers(time_embed(timestep_embedding(t))) for a whole
        sampling schedule at once (openaimodel.py:538-542,238-244 depend on t only, so the sampler evaluates them
        before the loop instead of once per step). Row-wise identical to _emb on the same timestep."""
        return self._emb(timesteps)

    def _emb(self, timesteps):
        # int64 steps (DDIM/PLMS) or DPM-Solver's fractional float32 timesteps (util.py:160-180 takes either)
        ts = timesteps.to(torch.float32 if timesteps.is_floating_point() else torch.int64).contiguous()
        emb = self.time_embed.forward(ts)
        return ops.skinny_linear(emb, self.w_emb, self.b_emb, in_act=True)   # emb_layers = SiLU, Linear

    def _run(self, blk, h, h1, emb_all, kvs, res1=None):
        """TimestepEmbedSequential.forward (openaimodel.py:85-93). h1: concat partner of the first ResBlock.
        res1 is added by the block's LAST layer's epilogue."""
        n = len(blk.children)
        for j, layer in enumerate(blk.children.values()):
            r1 = res1 if j == n - 1 else None
            if isinstance(layer, ResBlock):
                h = layer.forward(h, h1, emb_all, res1=r1)
                h1 = None
            elif isinstance(layer, SpatialTransformer):
                h = layer.forward(h, kvs, res1=r1)
            else:
                assert r1 is None
                h = layer.forward(h)
        if self.trace is not None:          # debugging aid: NHWC bf16 block outputs in execution order
            self.trace.append(h)
        return h

    def _shareable(self, x, rows):
        """Can the CFG halves share the context-free head of the encoder? rows == 2 * unique latents, input block 1 is
        [ResBlock, SpatialTransformer], and no block trace is being recorded."""
        if rows != 2 * x.shape[0] or self.trace is not None or len(self.input_blocks.children) < 2:
            return False
        kinds = [type(c) for c in self.input_blocks[1].children.values()]
        return kinds == [ResBlock, SpatialTransformer]

    def _encode(self, x, rows, emb_all, kvs, fa=None, after_conv_in=None, shared_halves=False):
        """shared_halves: the caller guarantees that rows r and r + rows/2 have the same latent x[r % nb_in] AND the
        same timestep (the sampler's CFG batch, ddim.py:222-226: x_in = cat([x] * 2), t_in = cat([t] * 2)). Until the
        first cross-attention nothing depends on the text context, so conv_in, the first ResBlock and the first
        transformer's GroupNorm / proj_in / self-attention are evaluated once per unique latent and repeated."""
        start = 1
        if shared_halves and self._shareable(x, rows):
            nb = x.shape[0]
            h0 = self.input_blocks[0][0].from_nchw(x, nb)
            if after_conv_in is not None:
                h0 = after_conv_in(h0)
            res, st = self.input_blocks[1].children.values()
            h = st.forward_shared(res.forward(h0, None, emb_all[:nb]), kvs, copies=2)
            hs = [repeat_rows(h0, 2), h]
            start = 2
        else:
            h = self.input_blocks[0][0].from_nchw(x, rows)
            if after_conv_in is not None:
                h = after_conv_in(h)
            hs = [h]
        k = 0
        for idx in range(start, len(self.input_blocks.children)):
            r1 = None
            if fa is not None and (idx + 1) % 3 == 0:
                r1 = fa[k]
                k += 1
            h = self._run(self.input_blocks[idx], h, None, emb_all, kvs, res1=r1)
            hs.append(h)
        if fa is not None:
            assert k == len(fa), "Wrong features_adapter"      # openaimodel.py:859
        return h, hs

    def _decode(self, h, hs, emb_all, kvs):
        for idx in range(len(self.output_blocks.children)):
            h = self._run(self.output_blocks[idx], h, hs.pop(), emb_all, kvs)
        n, c = self.out[0], self.out[2]
        h = ops.groupnorm(h, n.g, n.b, eps=1e-5, silu=True)
        return c.to_nchw(h)


class UNetModel(_UNetBase):
    """FG-DM UNetModel (openaimodel.py:469-498 kwargs). forward(x, timesteps, context) -> eps, :808-884."""

    def __init__(self, image_size=32, in_channels=4, model_channels=320, out_channels=4, num_res_blocks=2,
                 attention_resolutions=(4, 2, 1), dropout=0, channel_mult=(1, 2, 4, 4), conv_resample=True, dims=2,
                 num_classes=None, use_checkpoint=False, use_fp16=False, num_heads=-1, num_head_channels=-1,
                 num_heads_upsample=-1, use_scale_shift_norm=False, resblock_updown=False,
                 use_new_attention_order=False, use_spatial_transformer=False, transformer_depth=1, context_dim=None,
                 n_embed=None, legacy=True, no_prompting=False, adapter_channels=None, use_time_adapter=False):
        super().__init__()
        _unsupported(num_classes=num_classes is not None, use_scale_shift_norm=use_scale_shift_norm,
                     resblock_updown=resblock_updown, no_spatial_transformer=not use_spatial_transformer,
                     transformer_depth=transformer_depth != 1, n_embed=n_embed is not None, dims=dims != 2,
                     num_head_channels=num_head_channels != -1, use_time_adapter=use_time_adapter,
                     no_conv_resample=not conv_resample, dropout=dropout != 0)
        self.no_prompting = no_prompting
        self._build(in_channels, model_channels, out_channels, num_res_blocks, list(attention_resolutions),
                    list(channel_mult), num_heads, context_dim)
        if not no_prompting:
            acin = in_channels if adapter_channels is None else adapter_channels
            self.adapter = self.add("adapter", Adapter(acin))     # openaimodel.py:549-556
        else:
            self.adapter = None

    def encode_prompt(self, pcond):
        """Adapter features of a `pcond` latent (openaimodel.py:836-844: `prompt = pcond`): step-invariant, so the
        sampler computes them once per sample call and hands them back through `adapter_features`."""
        if self.adapter is None:
            raise NotImplementedError("pcond needs the FG-DM adapter (no_prompting=False)")
        pc = pcond.float().contiguous()
        return self.adapter.forward(pc, pc.shape[0])

    def forward(self, x, timesteps=None, context=None, y=None, pcond=None, kv=None, rows=None, adapter_features=None,
                shared_halves=False, emb_all=None, **kwargs):
        """x: fp32 [nb_in, 4, h, w]; rows >= nb_in output rows (row r reads x[r % nb_in] — lets the sampler
        feed ONE copy of x for the cond and uncond halves instead of the reference's torch.cat([x]*2),
        ddim.py:222). timesteps: int64 [rows]; context [rows, L, 768] or kv = project_context(context).
        pcond: fp32 [P, 4, h, w] latent of another factor fed to the adapter instead of x (the joint FG-DM factor);
        row r takes the features of pcond[r % P] (P = 1 is the reference's broadcast of `h + fa[k]`)."""
        if y is not None:
            raise NotImplementedError("class labels are outside the FG-DM sampling path")
        x = x.float().contiguous()
        rows = rows or x.shape[0]
        kvs = kv if kv is not None else self.project_context(context)
        if emb_all is None:             # else: the per-ResBlock time-embedding vectors of these timesteps, precomputed
            emb_all = self._emb(timesteps)
        # prompt = h (:836-844): the adapter sees only the latent, so its features are identical for the cond and
        # uncond CFG rows -> computed once per unique latent, added modulo by the GEMM epilogue (res1_rows)
        if adapter_features is not None:
            fa = adapter_features
        elif self.adapter is not None:
            fa = self.encode_prompt(pcond) if pcond is not None else self.adapter.forward(x, x.shape[0])
        else:
            fa = None
        h, hs = self._encode(x, rows, emb_all, kvs, fa, shared_halves=shared_halves)
        h = self._run(self.middle_block, h, None, emb_all, kvs)
        return self._decode(h, hs, emb_all, kvs)


class _PadConv(Conv2d):
    """A 3x3 conv whose channel counts are zero-padded to multiples of 64 so that it runs on the tensor-core implicit
    GEMM (the kernel's K chunks are 64 channels): padded input channels meet zero weights, padded output channels have
    zero weights and bias, and SiLU(0) = 0 keeps them zero for the next layer."""

    def __init__(self, cin, cout, stride, cin_p, cout_p):
        super().__init__(cin, cout, 3, stride)
        self.cin_p, self.cout_p = cin_p, cout_p

    def _load(self, sd, prefix, dev):
        w = self._get(sd, prefix + "weight", self.params["weight"]).to(dev, torch.float32)
        b = self._get(sd, prefix + "bias", self.params["bias"]).to(dev, torch.float32)
        self.b = torch.zeros((self.cout_p,), device=dev, dtype=torch.float32)
        self.b[:self.cout] = b
        if self.cin <= 4:        # first layer: one 64-wide K chunk from fgdm_im2col_small (k = tap*cin + c)
            self.w = torch.zeros((self.cout_p, 64), device=dev, dtype=bf16)
            self.w[:self.cout, :9 * self.cin] = ops.pack_conv_weight(w)
        else:
            wp = torch.zeros((self.cout_p, self.cin_p, 3, 3), device=dev, dtype=torch.float32)
            wp[:self.cout, :self.cin] = w
            self.w = ops.pack_conv_weight(wp)

    def forward(self, x):
        """x: NCHW fp32 (first layer) or NHWC bf16 with cin_p channels -> NHWC bf16 with cout_p channels, SiLU applied."""
        if self.cin <= 4:
            nb, _, h, w = x.shape
            return ops.conv_gemm(ops.im2col_small(x), self.w, self.cout_p, bias=self.b, act=ACT_SILU).view(nb, h, w, -1)
        nb, h, w, _ = x.shape
        if self.stride == 2:
            y = ops.conv_gemm(ops.resample(x, 2), self.w, self.cout_p, conv=ops.CONV_3X3_S2, nb=nb, h=h // 2, w=w // 2,
                              bias=self.b, act=ACT_SILU)
            return y.view(nb, h // 2, w // 2, -1)
        return ops.conv_gemm(x, self.w, self.cout_p, conv=CONV_3X3, bias=self.b, act=ACT_SILU).view(nb, h, w, -1)


class _HintStem(Layer):
    """ControlNet.input_hint_block, controlnet/cldm/cldm.py:655-671: eight 3x3 convs (3 -> 16 -> 16 -> 32 -> 32 -> 96 -> 96
    -> 256 -> model_channels, three of them stride 2) with SiLU between them. Step-invariant (once per sample call), but at
    512x512 the CUDA-core direct conv took 12 ms per call for B = 8; with channels padded to 64 / 128 all eight run on
    the tensor-core GEMM."""
    CH = [(None, 16, 1), (16, 16, 1), (16, 32, 2), (32, 32, 1), (32, 96, 2), (96, 96, 1), (96, 256, 2)]

    def __init__(self, hint_channels, mc):
        super().__init__()
        pad = lambda c: (c + 63) // 64 * 64
        self.convs = []
        for i, (ci, co, s) in enumerate(self.CH):
            ci = ci or hint_channels
            self.convs.append(self.add(str(2 * i), _PadConv(ci, co, s, pad(ci), pad(co))))
        self.last = self.add("14", Conv2d(256, mc, 3))

    def forward(self, hint):
        """hint: fp32 NCHW [B,3,H,W] in [0,1] -> NHWC bf16 [B,H/8,W/8,mc]."""
        h = hint.float().contiguous()
        for c in self.convs:
            h = c.forward(h)
        return self.last.gemm(h)


class ControlNet(_UNetBase):
    """controlnet/cldm/cldm.py:545-813. forward(x, hint, timesteps, context) -> 13 control tensors."""

    def __init__(self, image_size=32, in_channels=4, model_channels=320, hint_channels=3, num_res_blocks=2,
                 attention_resolutions=(4, 2, 1), dropout=0, channel_mult=(1, 2, 4, 4), conv_resample=True, dims=2,
                 use_checkpoint=False, use_fp16=False, num_heads=-1, num_head_channels=-1, num_heads_upsample=-1,
                 use_scale_shift_norm=False, resblock_updown=False, use_new_attention_order=False,
                 use_spatial_transformer=False, transformer_depth=1, context_dim=None, n_embed=None, legacy=True,
                 disable_self_attentions=None, num_attention_blocks=None, disable_middle_self_attn=False,
                 use_linear_in_transformer=False):
        super().__init__()
        _unsupported(use_scale_shift_norm=use_scale_shift_norm, resblock_updown=resblock_updown,
                     no_spatial_transformer=not use_spatial_transformer, transformer_depth=transformer_depth != 1,
                     dims=dims != 2, num_head_channels=num_head_channels != -1,
                     use_linear_in_transformer=use_linear_in_transformer, dropout=dropout != 0,
                     disable_self_attentions=disable_self_attentions is not None,
                     num_attention_blocks=num_attention_blocks is not None)
        self._build(in_channels, model_channels, None, num_res_blocks, list(attention_resolutions),
                    list(channel_mult), num_heads, context_dim, decoder=False)
        self.zero_convs = self.add("zero_convs", Seq([Seq([Conv2d(c, c, 1)]) for c in self.input_chans]))
        self.input_hint_block = self.add("input_hint_block", _HintStem(hint_channels, model_channels))
        self.middle_block_out = self.add("middle_block_out", Seq([Conv2d(self.mid_ch, self.mid_ch, 1)]))

    def encode_hint(self, hint):
        return self.input_hint_block.forward(hint)

    def features(self, x, hint=None, timesteps=None, context=None, kv=None, rows=None, guided_hint=None,
                 shared_halves=False, emb_all=None):
        """The encoder + middle block (cldm.py:796-811 without the zero convolutions): the 12 skip tensors and the middle
        tensor, NHWC bf16. Independent of the UNet it controls, so a caller may run it next to the UNet's encoder."""
        x = x.float().contiguous()
        rows = rows or x.shape[0]
        kvs = kv if kv is not None else self.project_context(context)
        if emb_all is None:
            emb_all = self._emb(timesteps)
        if guided_hint is None:
            guided_hint = self.encode_hint(hint)
        gb = guided_hint.shape[0]

        def add_hint(h):                                          # h += guided_hint, cldm.py:803-805
            for r0 in range(0, h.shape[0], gb):
                ops.add_bf16(h[r0:r0 + gb], guided_hint, out=h[r0:r0 + gb])
            return h

        # the halves may only share the encoder head if they also share the hint (guided_hint rows repeat modulo gb)
        share = shared_halves and (rows // 2) % gb == 0
        h, hs = self._encode(x, rows, emb_all, kvs, None, after_conv_in=add_hint, shared_halves=share)
        h = self._run(self.middle_block, h, None, emb_all, kvs)
        return hs + [h]

    def control_outputs(self, feats, scales=None, add_to=None, only_mid=False):
        """The 13 tensors `zero_conv(h) * scale` (cldm.py:801-811, scaling of :846 folded into the epilogue). With
        add_to = [hs..., h_mid] of the UNet encoder, `add_to[k] + scale*control[k]` instead (the adds of cldm.py:39-47
        ride the same epilogue)."""
        n = len(feats) - 1
        scales = scales or [1.0] * (n + 1)
        outs = []
        for k, hk in enumerate(feats):
            if only_mid and k < n:
                outs.append(None)
                continue
            zc = self.zero_convs[k][0] if k < n else self.middle_block_out[0]
            outs.append(zc.gemm(hk, alpha=float(scales[k]), res0=None if add_to is None else add_to[k]))
        return outs

    def forward(self, x, hint=None, timesteps=None, context=None, kv=None, rows=None, guided_hint=None,
                scales=None, add_to=None, only_mid=False, shared_halves=False, emb_all=None, **kwargs):
        """Returns the 13 control tensors (see control_outputs). Outputs are NHWC bf16."""
        feats = self.features(x, hint, timesteps, context, kv, rows, guided_hint, shared_halves, emb_all)
        return self.control_outputs(feats, scales, add_to, only_mid)


class ControlledUnetModel(_UNetBase):
    """controlnet/cldm/cldm.py:27-50: plain SD UNet whose skips / mid receive control residuals."""

    def __init__(self, image_size=32, in_channels=4, model_channels=320, out_channels=4, num_res_blocks=2,
                 attention_resolutions=(4, 2, 1), dropout=0, channel_mult=(1, 2, 4, 4), conv_resample=True, dims=2,
                 num_classes=None, use_checkpoint=False, use_fp16=False, num_heads=-1, num_head_channels=-1,
                 num_heads_upsample=-1, use_scale_shift_norm=False, resblock_updown=False,
                 use_new_attention_order=False, use_spatial_transformer=False, transformer_depth=1, context_dim=None,
                 n_embed=None, legacy=True, **ignored):
        super().__init__()
        _unsupported(num_classes=num_classes is not None, use_scale_shift_norm=use_scale_shift_norm,
                     resblock_updown=resblock_updown, no_spatial_transformer=not use_spatial_transformer,
                     transformer_depth=transformer_depth != 1, n_embed=n_embed is not None, dims=dims != 2,
                     num_head_channels=num_head_channels != -1, dropout=dropout != 0)
        self._build(in_channels, model_channels, out_channels, num_res_blocks, list(attention_resolutions),
                    list(channel_mult), num_heads, context_dim)

    def encode(self, x, timesteps, context=None, kv=None, rows=None, shared_halves=False, emb_all=None):
        x = x.float().contiguous()
        rows = rows or x.shape[0]
        kvs = kv if kv is not None else self.project_context(context)
        if emb_all is None:
            emb_all = self._emb(timesteps)
        h, hs = self._encode(x, rows, emb_all, kvs, shared_halves=shared_halves)
        h = self._run(self.middle_block, h, None, emb_all, kvs)
        return h, hs, emb_all, kvs

    def decode(self, h, hs, emb_all, kvs):
        return self._decode(h, list(hs), emb_all, kvs)

    def forward(self, x, timesteps=None, context=None, control=None, only_mid_control=False, kv=None, rows=None,
                **kwargs):
        """control: 13 NHWC bf16 tensors from ControlNet.forward (already scaled), or None."""
        h, hs, emb_all, kvs = self.encode(x, timesteps, context, kv, rows)
        if control is not None:
            control = list(control)
            h = ops.add_bf16(h, control.pop())
            if not only_mid_control:
                hs = [ops.add_bf16(a, c) for a, c in zip(hs, control)]
        return self._decode(h, hs, emb_all, kvs)


# ---------------------------------------------------------------------------------------------
# KL-VAE decoder
# ---------------------------------------------------------------------------------------------
class _VaeRes(Layer):
    """ResnetBlock, model.py:82-141 (temb_channels = 0 in the decoder); GroupNorm eps 1e-6."""

    def __init__(self, cin, cout):
        super().__init__()
        self.norm1 = self.add("norm1", Norm(cin))
        self.conv1 = self.add("conv1", Conv2d(cin, cout, 3))
        self.norm2 = self.add("norm2", Norm(cout))
        self.conv2 = self.add("conv2", Conv2d(cout, cout, 3))
        self.nin = self.add("nin_shortcut", Conv2d(cin, cout, 1)) if cin != cout else None

    def forward(self, x):
        h = self.conv1.gemm(ops.groupnorm(x, self.norm1.g, self.norm1.b, eps=1e-6, silu="exact"))
        h = ops.groupnorm(h, self.norm2.g, self.norm2.b, eps=1e-6, silu="exact")
        s = self.nin.gemm(x) if self.nin is not None else x
        return self.conv2.gemm(h, res0=s)


class _VaeAttn(Layer):
    """AttnBlock, model.py:150-202: single head with d = C = 512 over h*w tokens."""

    def __init__(self, c):
        super().__init__()
        self.c = c
        self.norm = self.add("norm", Norm(c))
        self.q = self.add("q", Conv2d(c, c, 1))
        self.k = self.add("k", Conv2d(c, c, 1))
        self.v = self.add("v", Conv2d(c, c, 1))
        self.proj_out = self.add("proj_out", Conv2d(c, c, 1))

    def load(self, sd, prefix, dev):
        super().load(sd, prefix, dev)
        self.w_qkv = torch.cat([self.q.w, self.k.w, self.v.w], 0).contiguous()
        self.b_qkv = torch.cat([self.q.b, self.k.b, self.v.b], 0).contiguous()
        self.q.w = self.k.w = self.v.w = None

    def forward(self, x):
        nb, h, w, c = x.shape
        n = h * w
        dev = x.device
        xn = ops.groupnorm(x, self.norm.g, self.norm.b, eps=1e-6, silu=False)
        q = torch.empty((nb, 1, n, c), device=dev, dtype=bf16)
        k = torch.empty((nb, 1, n, c), device=dev, dtype=bf16)
        vt = torch.empty((nb, 1, c, n), device=dev, dtype=bf16)
        ops.conv_gemm(xn.view(nb * n, c), self.w_qkv, 3 * c, bias=self.b_qkv, out_mode=OUT_QKV, rows_per_img=n,
                      qkv=(q, k, vt, 0, 1, c, n))
        o = torch.empty((nb, n, c), device=dev, dtype=bf16)
        if VAE_FLASH and c % 128 == 0 and 256 <= c <= 1024 and n % 8 == 0:
            # one launch for every image, S and P never leave the SM (csrc/attention_wide.cu)
            ops.attention(q, k, vt, o, nb, 1, n, n, c, n, scale=float(c) ** -0.5)
            return self.proj_out.gemm(o.view(nb, h, w, c), res0=x)
        s = torch.empty((n, n), device=dev, dtype=torch.float32)                         # the reference's form: N x N in HBM
        p = torch.empty((n, n), device=dev, dtype=bf16)
        for i in range(nb):
            ops.conv_gemm(q[i, 0], k[i, 0], n, out_mode=OUT_F32, out=s)                 # S = Q K^T
            ops.softmax_rows(s, float(c) ** -0.5, out=p)                                # softmax(S * c^-1/2)
            ops.conv_gemm(p, vt[i, 0], c, out=o[i])                                     # O = P V
        return self.proj_out.gemm(o.view(nb, h, w, c), res0=x)


class _Mid(Layer):
    def __init__(self, c):
        super().__init__()
        self.block_1 = self.add("block_1", _VaeRes(c, c))
        self.attn_1 = self.add("attn_1", _VaeAttn(c))
        self.block_2 = self.add("block_2", _VaeRes(c, c))


class _UpLevel(Layer):
    def __init__(self, cin, cout, nblocks, upsample):
        super().__init__()
        blocks = []
        for _ in range(nblocks):
            blocks.append(_VaeRes(cin, cout))
            cin = cout
        self.block = self.add("block", Seq(blocks))
        self.upsample = self.add("upsample", Upsample(cout)) if upsample else None


class Decoder(Layer):
    """model.py:462-569."""

    def __init__(self, ch=128, out_ch=3, ch_mult=(1, 2, 4, 4), num_res_blocks=2, z_channels=4, **ignored):
        super().__init__()
        nres = len(ch_mult)
        block_in = ch * ch_mult[nres - 1]
        self.out_ch = out_ch
        self.conv_in = self.add("conv_in", Conv2d(z_channels, block_in, 3))
        self.mid = self.add("mid", _Mid(block_in))
        ups = {}
        for lvl in reversed(range(nres)):
            block_out = ch * ch_mult[lvl]
            ups[lvl] = _UpLevel(block_in, block_out, num_res_blocks + 1, lvl != 0)
            block_in = block_out
        self.up = self.add("up", Seq({lvl: ups[lvl] for lvl in range(nres)}))
        self.norm_out = self.add("norm_out", Norm(block_in))
        self.conv_out = self.add("conv_out", Conv2d(block_in, out_ch, 3))
        self.nres = nres

    def forward(self, z):
        """z: fp32 NCHW [B, z_channels, h, w] -> fp32 NCHW [B, out_ch, 8h, 8w]."""
        h = self.conv_in.from_nchw(z)
        h = self.mid.block_1.forward(h)
        h = self.mid.attn_1.forward(h)
        h = self.mid.block_2.forward(h)
        for lvl in reversed(range(self.nres)):
            up = self.up[lvl]
            for blk in up.block.children.values():
                h = blk.forward(h)
            if up.upsample is not None:
                h = up.upsample.forward(h)
        h = ops.groupnorm(h, self.norm_out.g, self.norm_out.b, eps=1e-6, silu="exact")
        return self.conv_out.to_nchw(h)


class _Conv1x1F32(Layer):
    """post_quant_conv (autoencoder.py:305,331): 1x1 conv on the 4-channel fp32 latent."""

    def __init__(self, c):
        super().__init__()
        self.c = c
        self.params = {"weight": (c, c, 1, 1), "bias": (c,)}

    def _load(self, sd, prefix, dev):
        self.w = self._get(sd, prefix + "weight", self.params["weight"]).to(dev, torch.float32).reshape(self.c, self.c).contiguous()
        self.b = self._get(sd, prefix + "bias", (self.c,)).to(dev, torch.float32).contiguous()


class AutoencoderKL(_Net):
    """ldm/models/autoencoder.py AutoencoderKL — decode only (sampling never encodes). Encoder, quant_conv
    and loss keys of a reference state_dict are ignored (load with strict=False)."""

    def __init__(self, ddconfig=None, lossconfig=None, embed_dim=4, ckpt_path=None, ignore_keys=(), image_key="image",
                 colorize_nlabels=None, monitor=None, **ignored):
        super().__init__()
        dd = dict(ddconfig or {})
        self.embed_dim = embed_dim
        self.decoder = self.add("decoder", Decoder(ch=dd.get("ch", 128), out_ch=dd.get("out_ch", 3),
                                                   ch_mult=tuple(dd.get("ch_mult", (1, 2, 4, 4))),
                                                   num_res_blocks=dd.get("num_res_blocks", 2),
                                                   z_channels=dd.get("z_channels", 4)))
        if dd.get("attn_resolutions"):
            raise NotImplementedError("VAE attn_resolutions other than [] are outside the FG-DM configuration")
        self.post_quant_conv = self.add("post_quant_conv", _Conv1x1F32(embed_dim))

    def decode(self, z, n=None, in_scale=1.0):
        """autoencoder.py:330-333; in_scale folds decode_first_stage's z / scale_factor (ddpm.py:839)."""
        if not self._loaded:
            raise RuntimeError("weights not loaded")
        with torch.no_grad():
            pq = self.post_quant_conv
            z = ops.latent_affine(z.float().contiguous(), pq.w, pq.b, in_scale)
            return self.decoder.forward(z)

    def forward(self, z):
        return self.decode(z)
