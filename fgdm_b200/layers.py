"""Host-side layer objects of the B200 path: parameter specs (reference names/shapes), weight packing
and forward functions that only launch C-ABI kernels (fgdm_b200.ops).

Activations are NHWC bf16 `[nb, h, w, c]`; tokens are the same memory viewed `[nb*h*w, c]`, so the
reference's `rearrange(b c h w -> b (h w) c)` (ldm/modules/attention.py:283) costs nothing here.
Parameter NAMES and SHAPES are the reference's (so its state_dicts load unchanged); the packed device
layouts are ours. No layer falls back to torch arithmetic.
"""
from __future__ import annotations

import os

import torch

from . import ops
from .ops import (ACT_NONE, ACT_RELU, ACT_SILU, CONV_1X1, CONV_3X3, CONV_3X3_S2, CONV_UP2_P0, OUT_BF16, OUT_F32, OUT_GEGLU,
                  OUT_NCHW_F32, OUT_QKV)

bf16 = torch.bfloat16


class Layer:
    """A node of the parameter tree. Children are registered in `self.children` (name -> Layer); leaves
    declare `self.params` (name -> shape)."""

    def __init__(self):
        self.children: dict[str, "Layer"] = {}
        self.params: dict[str, tuple] = {}

    def add(self, name: str, layer: "Layer") -> "Layer":
        self.children[str(name)] = layer
        return layer

    def specs(self, prefix: str = ""):
        for n, s in self.params.items():
            yield prefix + n, tuple(s)
        for n, c in self.children.items():
            yield from c.specs(prefix + n + ".")

    def load(self, sd, prefix: str, dev):
        self._load(sd, prefix, dev)
        for n, c in self.children.items():
            c.load(sd, prefix + n + ".", dev)

    def _load(self, sd, prefix, dev):
        pass

    @staticmethod
    def _get(sd, key, shape):
        if key not in sd:
            raise KeyError(f"missing parameter {key}")
        t = sd[key]
        if tuple(t.shape) != tuple(shape):
            raise ValueError(f"parameter {key}: shape {tuple(t.shape)} != expected {tuple(shape)}")
        return t


class Conv2d(Layer):
    """nn.Conv2d / conv_nd parameters; packed [Cout][kh*kw][Cin] bf16 for the implicit-GEMM kernels."""

    def __init__(self, cin, cout, k=3, stride=1):
        super().__init__()
        self.cin, self.cout, self.k, self.stride = cin, cout, k, stride
        self.params = {"weight": (cout, cin, k, k), "bias": (cout,)}

    def _load(self, sd, prefix, dev):
        w = self._get(sd, prefix + "weight", self.params["weight"])
        b = self._get(sd, prefix + "bias", self.params["bias"])
        self.w = ops.pack_conv_weight(w.to(dev, torch.float32))
        self.b = b.to(dev, torch.float32).contiguous()
        if self.k == 3 and self.cin <= 4 and self.cout % 32 == 0:
            # tiny Cin: K = 9*cin zero-padded to one 64-wide K chunk (A comes from fgdm_im2col_small)
            self.w64 = torch.zeros((self.cout, 64), device=dev, dtype=bf16)
            self.w64[:, :9 * self.cin] = self.w
        if self.k == 3 and self.cout <= 4 and self.cin % 64 == 0:
            # tiny Cout: rows zero-padded to one 32-wide N tile, stored NCHW fp32 by the epilogue
            self.w32 = torch.zeros((32, 9 * self.cin), device=dev, dtype=bf16)
            self.w32[:self.cout] = self.w
            self.b32 = torch.zeros((32,), device=dev, dtype=torch.float32)
            self.b32[:self.cout] = self.b

    def from_nchw(self, x, rows=None, in_scale=1.0, **kw):
        """conv3x3 on an NCHW fp32 tensor with cin <= 4 -> NHWC bf16 [rows, h, w, cout] (im2col + tensor-core GEMM)."""
        nb_in, _, h, w = x.shape
        rows = rows or nb_in
        a = ops.im2col_small(x, rows, in_scale)
        return ops.conv_gemm(a, self.w64, self.cout, bias=self.b, **kw).view(rows, h, w, self.cout)

    def to_nchw(self, x):
        """conv3x3 NHWC bf16 -> NCHW fp32 [nb, cout, h, w] with cout <= 4 (tensor-core GEMM, N padded to 32)."""
        return ops.conv_gemm(x, self.w32, 32, conv=CONV_3X3, bias=self.b32, out_mode=OUT_NCHW_F32, n_store=self.cout)

    # tensor-core implicit GEMM; x may be a virtual concat (x, x1)
    def gemm(self, x, x1=None, **kw):
        nb, h, w, _ = x.shape
        if self.k == 1:
            y = ops.conv_gemm(x, self.w, self.cout, a1=x1, bias=self.b, nb=1, h=1, w=nb * h * w, **kw)
            if kw.get("want_stats"):
                return y[0].view(nb, h, w, -1), y[1]
            return y.view(nb, h, w, -1) if y is not None else None
        if self.stride == 2:
            ph = ops.resample(x, 2)
            y = ops.conv_gemm(ph, self.w, self.cout, conv=CONV_3X3_S2, nb=nb, h=h // 2, w=w // 2, bias=self.b, **kw)
            return y.view(nb, h // 2, w // 2, -1)
        y = ops.conv_gemm(x, self.w, self.cout, conv=CONV_3X3, a1=x1, bias=self.b, **kw)
        return y.view(nb, h, w, -1)


class Linear(Layer):
    def __init__(self, cin, cout, bias=True):
        super().__init__()
        self.cin, self.cout, self.has_bias = cin, cout, bias
        self.params = {"weight": (cout, cin)}
        if bias:
            self.params["bias"] = (cout,)

    def _load(self, sd, prefix, dev):
        self.w32 = self._get(sd, prefix + "weight", self.params["weight"]).to(dev, torch.float32).contiguous()
        self.b = (self._get(sd, prefix + "bias", self.params["bias"]).to(dev, torch.float32).contiguous()
                  if self.has_bias else None)
        self.w = None

    def pack(self):
        """bf16 copy for the tensor-core path (fp32 copy is dropped)."""
        self.w = self.w32.to(bf16).contiguous()
        self.w32 = None
        return self


class Norm(Layer):
    """GroupNorm / LayerNorm affine parameters (fp32 on device)."""

    def __init__(self, c):
        super().__init__()
        self.c = c
        self.params = {"weight": (c,), "bias": (c,)}

    def _load(self, sd, prefix, dev):
        self.g = self._get(sd, prefix + "weight", (self.c,)).to(dev, torch.float32).contiguous()
        self.b = self._get(sd, prefix + "bias", (self.c,)).to(dev, torch.float32).contiguous()


LN_EPS = 1e-5   # nn.LayerNorm default (attention.py:226-228)


def fold_layernorm(w32, bias, norm):
    """LayerNorm folded into the Linear that consumes it: y = LN(x) W^T + b
         = rstd * (x (W*gamma)^T - mean * s) + t,  s[n] = sum_k (W*gamma)[n,k],  t[n] = sum_k beta[k] W[n,k] + b[n].
    Returns (W*gamma as bf16, s from the bf16-rounded weights (what the tensor core multiplies), t), fp32 vectors."""
    wf = (w32 * norm.g[None, :]).to(bf16).contiguous()
    s = wf.float().sum(1).contiguous()
    t = w32 @ norm.b
    if bias is not None:
        t = t + bias
    return wf, s, t.contiguous()


class Seq(Layer):
    """nn.Sequential / nn.ModuleList naming: children keyed by integer position."""

    def __init__(self, items: dict | list):
        super().__init__()
        if isinstance(items, list):
            items = dict(enumerate(items))
        self.items = {}
        for i, l in items.items():
            if l is not None:
                self.add(i, l)
            self.items[i] = l

    def __getitem__(self, i):
        return self.children[str(i)]


# ---------------------------------------------------------------------------------------------
# ResBlock (ldm/modules/diffusionmodules/openaimodel.py:183-301; controlnet fork :162-275)
# ---------------------------------------------------------------------------------------------
class ResBlock(Layer):
    def __init__(self, cin, emb_ch, cout):
        super().__init__()
        self.cin, self.cout = cin, cout
        self.in_layers = self.add("in_layers", Seq({0: Norm(cin), 2: Conv2d(cin, cout, 3)}))
        self.emb_layers = self.add("emb_layers", Seq({1: Linear(emb_ch, cout)}))
        self.out_layers = self.add("out_layers", Seq({0: Norm(cout), 3: Conv2d(cout, cout, 3)}))
        self.skip = self.add("skip_connection", Conv2d(cin, cout, 1)) if cin != cout else None
        self.emb_off = -1   # column offset into the batched emb_layers output (set by the owning net)

    def forward(self, x, x1, emb_all, res1=None):
        """x (+ optional concat partner x1): NHWC bf16. emb_all: fp32 [R, total] batched emb_layers output.
        Returns skip(x) + h (+ res1), openaimodel.py:275-301."""
        n1, c1 = self.in_layers[0], self.in_layers[2]
        n2, c2 = self.out_layers[0], self.out_layers[3]
        h = ops.groupnorm(x, n1.g, n1.b, x1=x1, eps=1e-5, silu=True)
        rv = emb_all[:, self.emb_off:self.emb_off + self.cout]
        h = c1.gemm(h, rowvec=rv)                                   # conv + bias + emb[:, :, None, None]
        h = ops.groupnorm(h, n2.g, n2.b, eps=1e-5, silu=True)
        s = self.skip.gemm(x, x1) if self.skip is not None else x
        return c2.gemm(h, res0=s, res1=res1)


# ---------------------------------------------------------------------------------------------
# CrossAttention / BasicTransformerBlock / SpatialTransformer (ldm/modules/attention.py:152-292)
# ---------------------------------------------------------------------------------------------
class CrossAttention(Layer):
    def __init__(self, query_dim, context_dim, heads, dim_head):
        super().__init__()
        inner = heads * dim_head
        self.heads, self.d, self.inner = heads, dim_head, inner
        self.is_self = context_dim is None
        cdim = query_dim if context_dim is None else context_dim
        self.to_q = self.add("to_q", Linear(query_dim, inner, bias=False))
        self.to_k = self.add("to_k", Linear(cdim, inner, bias=False))
        self.to_v = self.add("to_v", Linear(cdim, inner, bias=False))
        self.to_out = self.add("to_out", Seq({0: Linear(inner, query_dim)}))

    ln = None         # the nn.LayerNorm in front of this attention (set by BasicTransformerBlock before load)
    ln_fold = False   # True: the LayerNorm is folded into the q / qkv projection (fold_layernorm)

    def load(self, sd, prefix, dev):
        super().load(sd, prefix, dev)
        q, k, v = self.to_q.w32, self.to_k.w32, self.to_v.w32
        wq = torch.cat([q, k, v], 0) if self.is_self else q           # K/V of cross-attention project the text context
        if self.ln_fold:
            wq, self.ln_s, self.ln_t = fold_layernorm(wq, None, self.ln)
        else:
            wq, self.ln_s, self.ln_t = wq.to(bf16).contiguous(), None, None
        if self.is_self:
            self.w_qkv = wq
        else:
            self.w_q = wq
            self.w_kv = torch.cat([k, v], 0).to(bf16).contiguous()
        self.to_q.w32 = self.to_k.w32 = self.to_v.w32 = None
        self.to_out[0].pack()

    def project_context(self, ctx):
        """ctx: bf16 [R, L, cdim] -> (k [R,heads,L,d], vt [R,heads,d,Lpad], L). Step-invariant."""
        r, L, cd = ctx.shape
        lp = (L + 7) // 8 * 8
        k = torch.empty((r, self.heads, L, self.d), device=ctx.device, dtype=bf16)
        vt = torch.empty((r, self.heads, self.d, lp), device=ctx.device, dtype=bf16)
        ops.conv_gemm(ctx.view(r * L, cd), self.w_kv, 2 * self.inner, out_mode=OUT_QKV, rows_per_img=L,
                      qkv=(None, k, vt, 1, self.heads, self.d, lp))
        return k, vt, L

    def forward(self, x, stats, nimg, ntok, kv=None, want_stats=False):
        """x + to_out(attention(LayerNorm(x))). x: RAW tokens [nimg*ntok, C]; with ln_fold the LayerNorm happens inside
        the q/k/v projection from `stats` (per-row partial sums written by the GEMM that produced x), otherwise in the
        LayerNorm kernel. want_stats: also return the statistics of the result for the next folded LayerNorm."""
        dev = x.device
        if self.ln_fold:
            xn, kw = x, dict(bias=self.ln_t, ln=(stats, self.ln_s, x.shape[-1], LN_EPS))
        else:
            xn, kw = ops.layernorm(x, self.ln.g, self.ln.b), {}
        q = torch.empty((nimg, self.heads, ntok, self.d), device=dev, dtype=bf16)
        if self.is_self:
            tp = (ntok + 7) // 8 * 8
            k = torch.empty((nimg, self.heads, ntok, self.d), device=dev, dtype=bf16)
            vt = torch.empty((nimg, self.heads, self.d, tp), device=dev, dtype=bf16)
            ops.conv_gemm(xn, self.w_qkv, 3 * self.inner, out_mode=OUT_QKV, rows_per_img=ntok,
                          qkv=(q, k, vt, 0, self.heads, self.d, tp), **kw)
            nk = ntok
        else:
            k, vt, nk = kv
            tp = vt.shape[-1]
            ops.conv_gemm(xn, self.w_q, self.inner, out_mode=OUT_QKV, rows_per_img=ntok,
                          qkv=(q, q, q, 0, self.heads, self.d, tp), **kw)
        a = torch.empty((nimg * ntok, self.inner), device=dev, dtype=bf16)
        ops.attention(q, k, vt, a, nimg, self.heads, ntok, nk, self.d, tp)
        o = self.to_out[0]
        y = ops.conv_gemm(a, o.w, o.cout, bias=o.b, res0=x, want_stats=want_stats)
        return y if want_stats else (y, None)


class GEGLU(Layer):
    def __init__(self, dim, inner):
        super().__init__()
        self.inner = inner
        self.proj = self.add("proj", Linear(dim, 2 * inner))

    ln = None         # norm3 of the owning BasicTransformerBlock
    ln_fold = False   # True: norm3 is folded into the projection

    def load(self, sd, prefix, dev):
        super().load(sd, prefix, dev)
        w, b = self.proj.w32, self.proj.b
        self.ln_s = None
        if self.ln_fold:
            wg = w * self.ln.g[None, :]
            _, self.ln_s = ops.pack_geglu(wg, wg.to(bf16).float().sum(1))    # row sums of the bf16 weights, interleaved alike
            w, b = wg, w @ self.ln.b + b                                      # gamma-folded weights, bias = beta.W^T + b
        self.w, self.b = ops.pack_geglu(w, b)
        self.proj.w32 = None


class FeedForward(Layer):
    def __init__(self, dim, mult=4):
        super().__init__()
        inner = dim * mult
        self.net = self.add("net", Seq({0: GEGLU(dim, inner), 2: Linear(inner, dim)}))

    def load(self, sd, prefix, dev):
        super().load(sd, prefix, dev)
        self.net[2].pack()

    def forward(self, x, stats):
        """x + net(LayerNorm(x)); the LayerNorm (norm3) is folded into the GEGLU projection or a kernel of its own."""
        g = self.net[0]
        if g.ln_fold:
            h = ops.conv_gemm(x, g.w, 2 * g.inner, bias=g.b, out_mode=OUT_GEGLU, ln=(stats, g.ln_s, x.shape[-1], LN_EPS))
        else:
            h = ops.conv_gemm(ops.layernorm(x, g.ln.g, g.ln.b), g.w, 2 * g.inner, bias=g.b, out_mode=OUT_GEGLU)
        o = self.net[2]
        return ops.conv_gemm(h, o.w, o.cout, bias=o.b, res0=x)


class BasicTransformerBlock(Layer):
    def __init__(self, dim, heads, d_head, context_dim):
        super().__init__()
        self.attn1 = self.add("attn1", CrossAttention(dim, None, heads, d_head))
        self.ff = self.add("ff", FeedForward(dim))
        self.attn2 = self.add("attn2", CrossAttention(dim, context_dim, heads, d_head))
        self.norm1 = self.add("norm1", Norm(dim))
        self.norm2 = self.add("norm2", Norm(dim))
        self.norm3 = self.add("norm3", Norm(dim))

    # Which LayerNorms are folded into the GEMM that consumes them (fold_layernorm). The fold moves work into GEMM
    # epilogues; the K = 320 projections of the 64x64 level are epilogue-bound already, and there the fold into the wide
    # q/k/v and GEGLU projections costs more than the LayerNorm kernel it removes (tools/bench_lnfold.py).
    FOLD_MIN_DIM = int(os.environ.get("FGDM_LN_FOLD_MIN", "640"))
    # norm1 -> q/k/v on its own switch: the head-aligned QKV epilogue (gemm.cu tma_epi == 2) is lean enough to carry it
    FOLD_MIN_DIM_ATTN1 = int(os.environ.get("FGDM_LN_FOLD_MIN_ATTN1", str(FOLD_MIN_DIM)))

    def load(self, sd, prefix, dev):
        for n in ("norm1", "norm2", "norm3"):
            self.children[n].load(sd, prefix + n + ".", dev)
        on = os.environ.get("FGDM_LN_FOLD", "1") != "0"           # 0: every LayerNorm runs as its own kernel
        wide = on and self.norm1.c >= self.FOLD_MIN_DIM
        wide1 = on and self.norm1.c >= self.FOLD_MIN_DIM_ATTN1
        for consumer, norm, fold in ((self.attn1, self.norm1, wide1), (self.attn2, self.norm2, on),
                                     (self.ff.net[0], self.norm3, wide)):
            consumer.ln, consumer.ln_fold = norm, fold
        for n in ("attn1", "ff", "attn2"):
            self.children[n].load(sd, prefix + n + ".", dev)

    @property
    def wants_stats(self):
        """Does the GEMM that produces this block's input have to write LayerNorm statistics?"""
        return self.attn1.ln_fold

    def forward(self, t, stats, nimg, ntok, kv):
        """BasicTransformerBlock._forward, attention.py:234-240. `stats`: per-row {sum, sumsq} partials of t written by
        the GEMM that produced it; each residual GEMM hands the next folded LayerNorm its statistics the same way."""
        t, stats = self.attn1.forward(t, stats, nimg, ntok, want_stats=self.attn2.ln_fold)
        t, stats = self.attn2.forward(t, stats, nimg, ntok, kv, want_stats=self.ff.net[0].ln_fold)
        return self.ff.forward(t, stats)


def repeat_rows(a, copies):
    """[n, ...] -> [copies * n, ...] with row r = a[r % n] (one copy kernel; torch.cat([x] * 2) of ddim.py:222)."""
    return torch.cat([a] * copies, 0)


class SpatialTransformer(Layer):
    def __init__(self, ch, heads, d_head, context_dim):
        super().__init__()
        self.ch = ch
        self.norm = self.add("norm", Norm(ch))
        self.proj_in = self.add("proj_in", Conv2d(ch, heads * d_head, 1))
        self.blocks = self.add("transformer_blocks", Seq([BasicTransformerBlock(heads * d_head, heads, d_head,
                                                                                   context_dim)]))
        self.proj_out = self.add("proj_out", Conv2d(heads * d_head, ch, 1))
        self.kv_slot = -1   # index into the per-call list of projected contexts

    def project_context(self, ctx):
        return self.blocks[0].attn2.project_context(ctx)

    def forward(self, x, kvs, res1=None):
        """SpatialTransformer.forward attention.py:275-292; GroupNorm eps 1e-6 (attention.py:76-77)."""
        nb, h, w, c = x.shape
        xn = ops.groupnorm(x, self.norm.g, self.norm.b, eps=1e-6, silu=False)
        blk = self.blocks[0]
        t, stats = self.proj_in.gemm(xn, want_stats=True) if blk.wants_stats else (self.proj_in.gemm(xn), None)
        t = blk.forward(t.view(nb * h * w, -1), stats, nb, h * w, kvs[self.kv_slot])
        return self.proj_out.gemm(t.view(nb, h, w, -1), res0=x, res1=res1)

    def forward_shared(self, x, kvs, copies=2):
        """forward() for `copies` x nb output rows whose inputs x are IDENTICAL across the copies (the uncond / cond CFG
        halves of one latent, ddim.py:222-226) and differ only in their text context: GroupNorm, proj_in, LayerNorm, the
        q/k/v projection, the self-attention and its output projection see no context (attention.py:234-236), so they
        run once on the nb unique rows; the rows are repeated in front of the cross-attention (attention.py:237)."""
        nb, h, w, c = x.shape
        xn = ops.groupnorm(x, self.norm.g, self.norm.b, eps=1e-6, silu=False)
        blk = self.blocks[0]
        t, stats = self.proj_in.gemm(xn, want_stats=True) if blk.wants_stats else (self.proj_in.gemm(xn), None)
        t, stats = blk.attn1.forward(t.view(nb * h * w, -1), stats, nb, h * w, want_stats=blk.attn2.ln_fold)
        t = repeat_rows(t, copies)
        stats = None if stats is None else repeat_rows(stats, copies)
        t, stats = blk.attn2.forward(t, stats, copies * nb, h * w, kvs[self.kv_slot], want_stats=blk.ff.net[0].ln_fold)
        t = blk.ff.forward(t, stats)
        return self.proj_out.gemm(t.view(copies * nb, h, w, -1), res0=repeat_rows(x, copies))


class Downsample(Layer):
    """openaimodel.py:146-180 (conv3x3 stride 2 pad 1)."""

    def __init__(self, ch):
        super().__init__()
        self.op = self.add("op", Conv2d(ch, ch, 3, stride=2))

    def forward(self, x):
        return self.op.gemm(x)


# FGDM_UP_PHASE=0: Upsample as nearest-2x kernel + 3x3 conv on the upsampled tensor (A/B switch). The sub-pixel form needs
# enough rows per phase GEMM to fill the SMs (FGDM_UP_PHASE_MIN_ROWS input pixels; the UNet's 8x8 -> 16x16 level has 1024).
UP_PHASE = os.environ.get("FGDM_UP_PHASE", "1") != "0"
UP_PHASE_MIN_ROWS = int(os.environ.get("FGDM_UP_PHASE_MIN_ROWS", "4096"))


def _tile_geometry_ok(nb, h, w, bm=128):
    """Mirror of gemm.cu tile_box: can [nb, h, w] images be cut into 128-row A tiles?"""
    if nb == 1 and h == 1:
        return True
    if w >= bm:
        return w % bm == 0
    if bm % w != 0:
        return False
    rows = bm // w
    return h % rows == 0 if rows <= h else rows % h == 0


class Upsample(Layer):
    """openaimodel.py:95-130; VAE model.py:42-57: nearest 2x then conv3x3."""

    def __init__(self, ch):
        super().__init__()
        self.ch = ch
        self.conv = self.add("conv", Conv2d(ch, ch, 3))

    def load(self, sd, prefix, dev):
        super().load(sd, prefix, dev)
        # Sub-pixel form: output pixel (2i+a, 2j+b) of conv3x3(nearest2x(x)) only sees input rows {i-1+a, i+a} and
        # columns {j-1+b, j+b}; the 3x3 taps that land on the same input pixel are summed (in fp32, rounded once):
        # four 4-tap convolutions on the LOW-resolution tensor = 4/9 of the flops, and no upsampled tensor in HBM.
        w = self._get(sd, prefix + "conv.weight", (self.ch, self.ch, 3, 3)).to(dev, torch.float32)
        sets = (((0,), (1, 2)), ((0, 1), (2,)))                  # a (or b) -> 3x3 taps behind 2x2 tap r (or s)
        self.w_phase = []
        for a in range(2):
            for b in range(2):
                taps = [sum(w[:, :, ky, kx] for ky in sets[a][r] for kx in sets[b][s]) for r in range(2) for s in range(2)]
                self.w_phase.append(torch.stack(taps, 1).reshape(self.ch, 4 * self.ch).to(bf16).contiguous())   # [O][(r,s)][I]

    def forward(self, x):
        nb, h, w, c = x.shape
        if UP_PHASE and nb * h * w >= UP_PHASE_MIN_ROWS and _tile_geometry_ok(nb, h, w):
            packed = torch.empty((nb, h, w, 4 * c), device=x.device, dtype=bf16)      # [.., phase, c]
            for ph in range(4):
                ops.conv_gemm(x, self.w_phase[ph], c, conv=CONV_UP2_P0 + ph, bias=self.conv.b, out=packed.view(-1, 4 * c),
                              ldo=4 * c, col_off=ph * c)
            return ops.resample(packed, 3)                        # -> [nb, 2h, 2w, c]
        return self.conv.gemm(ops.resample(x, 0))


class TimeEmbed(Layer):
    """time_embed = Linear, SiLU, Linear (openaimodel.py:538-542) on timestep_embedding (util.py:160-180)."""

    def __init__(self, mc):
        super().__init__()
        self.mc = mc
        self.add("0", Linear(mc, 4 * mc))
        self.add("2", Linear(4 * mc, 4 * mc))

    def _load(self, sd, prefix, dev):
        self.freqs = ops.timestep_freqs(self.mc).to(dev)

    def load(self, sd, prefix, dev):
        super().load(sd, prefix, dev)
        self.children["0"].pack()
        self.children["2"].pack()

    def forward(self, t):
        l0, l2 = self.children["0"], self.children["2"]
        e = ops.timestep_embedding(t, self.freqs, self.mc)
        e = ops.skinny_linear(e, l0.w, l0.b, out_act=True)
        return ops.skinny_linear(e, l2.w, l2.b)


def batch_emb_layers(resblocks, dev):
    """Concatenate every ResBlock's emb_layers Linear (openaimodel.py:238-244) into one fp32 matrix so a
    single skinny GEMM per step produces all per-block time-embedding vectors."""
    ws, bs, off = [], [], 0
    for rb in resblocks:
        lin = rb.emb_layers[1]
        ws.append(lin.w32)
        bs.append(lin.b)
        rb.emb_off = off
        off += rb.cout
        lin.w32 = None
    return torch.cat(ws, 0).to(bf16).contiguous(), torch.cat(bs, 0).contiguous()
