"""GPU parity tests of every C-ABI kernel against a plain torch fp32 statement of the same op.

Tolerances: bf16 storage (8 mantissa bits) with fp32 accumulation -> max error relative to the
output's max magnitude <= 1.5e-2 for GEMM-shaped ops, 1e-2 for norms; exact for integer work.
"""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from fgdm_b200 import _lib, ops

DEV = "cuda"


def relerr(a, b):
    a, b = a.float(), b.float()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-6)).item()


def nhwc(x):  # NCHW fp32 -> NHWC bf16
    return x.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16)


def nchw(x):  # NHWC -> NCHW fp32
    return x.float().permute(0, 3, 1, 2).contiguous()


@pytest.fixture(scope="module", autouse=True)
def _dev():
    _lib.require_device()
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(0)


@pytest.mark.parametrize("m,k,n,bn", [(128, 64, 64, 0), (256, 128, 128, 0), (1232, 768, 640, 0), (4096, 320, 320, 0),
                                      (8192, 1280, 1280, 0), (512, 640, 2560, 128), (300, 192, 96, 0), (2048, 256, 640, 224),
                                      (16384, 640, 640, 0), (1000, 128, 320, 224)])
def test_linear(m, k, n, bn):
    x = torch.randn(m, k, device=DEV)
    w = torch.randn(n, k, device=DEV) / math.sqrt(k)
    b = torch.randn(n, device=DEV)
    xb, wb = x.to(torch.bfloat16), w.to(torch.bfloat16)
    y = ops.conv_gemm(xb, wb, n, bias=b, bn_hint=bn)
    ref = xb.float() @ wb.float().t() + b
    assert y.shape == (m, n)
    assert relerr(y, ref) < 1e-2


@pytest.mark.parametrize("m,k,n,bn", [(4096, 1280, 1280, 1256), (1536, 640, 640, 1160), (65536, 320, 640, 1224)])
def test_linear_cta_pair(m, k, n, bn):
    """bn_hint = 1000 + BN selects the cta_group::2 kernel: a 2-CTA cluster issues one 256-row tcgen05.mma, each CTA
    stages its 128 rows of A and half of the B tile, the leader's barriers cover both (include/fgdm_b200.h)."""
    x = torch.randn(m, k, device=DEV).to(torch.bfloat16)
    w = (torch.randn(n, k, device=DEV) / math.sqrt(k)).to(torch.bfloat16)
    b = torch.randn(n, device=DEV)
    r0 = torch.randn(m, n, device=DEV).to(torch.bfloat16)
    y = ops.conv_gemm(x, w, n, bias=b, res0=r0, bn_hint=bn)
    ref = x.float() @ w.float().t() + b + r0.float()
    assert relerr(y, ref) < 1e-2


def test_linear_epilogues():
    m, k, n = 512, 256, 320
    x = torch.randn(m, k, device=DEV).to(torch.bfloat16)
    w = (torch.randn(n, k, device=DEV) / math.sqrt(k)).to(torch.bfloat16)
    b = torch.randn(n, device=DEV)
    r0 = torch.randn(m, n, device=DEV).to(torch.bfloat16)
    r1 = torch.randn(m, n, device=DEV).to(torch.bfloat16)
    rv = torch.randn(4, n, device=DEV)
    y = ops.conv_gemm(x, w, n, bias=b, rowvec=rv, rows_per_img=128, res0=r0, res1=r1, alpha=0.5, act=ops.ACT_RELU)
    ref = (x.float() @ w.float().t() + b + rv.repeat_interleave(128, 0)) * 0.5
    ref = ref.relu() + r0.float() + r1.float()
    assert relerr(y, ref) < 1e-2
    y32 = ops.conv_gemm(x, w, n, out_mode=ops.OUT_F32)
    assert y32.dtype == torch.float32
    assert relerr(y32, x.float() @ w.float().t()) < 1e-4


@pytest.mark.parametrize("nb,h,w,cin,cout", [(2, 8, 8, 64, 64), (2, 16, 16, 128, 160), (1, 32, 32, 64, 320),
                                             (2, 64, 64, 320, 320), (1, 128, 128, 128, 128), (1, 4, 4, 64, 64),
                                             (3, 8, 8, 64, 96)])
def test_conv3x3(nb, h, w, cin, cout):
    x = torch.randn(nb, cin, h, w, device=DEV)
    wt = torch.randn(cout, cin, 3, 3, device=DEV) / math.sqrt(9 * cin)
    b = torch.randn(cout, device=DEV)
    xb = nhwc(x)
    y = ops.conv_gemm(xb, ops.pack_conv_weight(wt), cout, conv=ops.CONV_3X3, bias=b).view(nb, h, w, cout)
    ref = F.conv2d(nchw(xb), wt.to(torch.bfloat16).float(), b, padding=1)
    assert relerr(nchw(y), ref) < 1e-2


@pytest.mark.parametrize("nb,h,w,cin,cout", [(16, 8, 8, 1280, 1280), (16, 16, 16, 640, 1280), (4, 8, 8, 2560, 1280),
                                             (16, 16, 16, 1280, 1280), (20, 16, 16, 320, 640)])
def test_conv3x3_split_k(nb, h, w, cin, cout):
    """Small-M / deep-K layers (the UNet's 8x8 and 16x16 levels) take the split-K path: partial tiles are summed in
    fixed split order by the last unit to finish, so results are correct AND bit-identical run to run."""
    x = torch.randn(nb, cin, h, w, device=DEV)
    wt = torch.randn(cout, cin, 3, 3, device=DEV) / math.sqrt(9 * cin)
    b = torch.randn(cout, device=DEV)
    res = torch.randn(nb, h, w, cout, device=DEV).to(torch.bfloat16)
    xb, wp = nhwc(x), ops.pack_conv_weight(wt)
    y = ops.conv_gemm(xb, wp, cout, conv=ops.CONV_3X3, bias=b, res0=res.view(-1, cout)).view(nb, h, w, cout)
    ref = F.conv2d(nchw(xb), wt.to(torch.bfloat16).float(), b, padding=1) + nchw(res)
    assert relerr(nchw(y), ref) < 1e-2
    for _ in range(3):
        y2 = ops.conv_gemm(xb, wp, cout, conv=ops.CONV_3X3, bias=b, res0=res.view(-1, cout)).view(nb, h, w, cout)
        assert torch.equal(y, y2)


def test_conv3x3_concat_and_temb():
    nb, h, w, c0, c1, cout = 2, 16, 16, 128, 64, 128
    x0 = torch.randn(nb, c0, h, w, device=DEV)
    x1 = torch.randn(nb, c1, h, w, device=DEV)
    wt = torch.randn(cout, c0 + c1, 3, 3, device=DEV) / math.sqrt(9 * (c0 + c1))
    b = torch.randn(cout, device=DEV)
    temb = torch.randn(nb, cout, device=DEV)
    a0, a1 = nhwc(x0), nhwc(x1)
    y = ops.conv_gemm(a0, ops.pack_conv_weight(wt), cout, conv=ops.CONV_3X3, a1=a1, bias=b, rowvec=temb)
    ref = F.conv2d(torch.cat([nchw(a0), nchw(a1)], 1), wt.to(torch.bfloat16).float(), b, padding=1)
    ref = ref + temb[:, :, None, None]
    assert relerr(nchw(y.view(nb, h, w, cout)), ref) < 1e-2


@pytest.mark.parametrize("nb,h,w,c", [(2, 16, 16, 64), (2, 64, 64, 320), (1, 8, 8, 128)])
def test_conv3x3_stride2(nb, h, w, c):
    x = torch.randn(nb, c, h, w, device=DEV)
    wt = torch.randn(c, c, 3, 3, device=DEV) / math.sqrt(9 * c)
    b = torch.randn(c, device=DEV)
    xb = nhwc(x)
    ph = ops.resample(xb, 2)
    y = ops.conv_gemm(ph, ops.pack_conv_weight(wt), c, conv=ops.CONV_3X3_S2, nb=nb, h=h // 2, w=w // 2, bias=b)
    ref = F.conv2d(nchw(xb), wt.to(torch.bfloat16).float(), b, stride=2, padding=1)
    assert relerr(nchw(y.view(nb, h // 2, w // 2, c)), ref) < 1e-2


def test_geglu():
    m, k, inner = 512, 320, 1280
    x = torch.randn(m, k, device=DEV).to(torch.bfloat16)
    w = torch.randn(2 * inner, k, device=DEV) / math.sqrt(k)
    b = torch.randn(2 * inner, device=DEV)
    wi, bi = ops.pack_geglu(w, b)
    y = ops.conv_gemm(x, wi, 2 * inner, bias=bi, out_mode=ops.OUT_GEGLU)
    proj = x.float() @ w.to(torch.bfloat16).float().t() + b
    a, g = proj.chunk(2, dim=-1)
    assert y.shape == (m, inner)
    assert relerr(y, a * F.gelu(g)) < 1e-2


@pytest.mark.parametrize("m,k0,c", [(4096, 320, 320), (1024, 640, 640), (256, 1280, 1280), (200, 3072, 1280), (1000, 64, 320)])
def test_layernorm_folded_into_gemm(m, k0, c):
    """LayerNorm -> Linear as one GEMM: the producing GEMM writes per-row {sum, sumsq} partials, the consumer applies
    rstd * (acc - mean * s[n]) + t[n] with gamma folded into its weights (BasicTransformerBlock, attention.py:234-240)."""
    from types import SimpleNamespace
    from fgdm_b200.layers import fold_layernorm, LN_EPS
    heads, d = 8, c // 8
    x0 = torch.randn(m, k0, device=DEV).to(torch.bfloat16)
    w0 = (torch.randn(c, k0, device=DEV) / math.sqrt(k0)).to(torch.bfloat16)
    b0 = torch.randn(c, device=DEV) + 1.5                      # DC offset: mean comparable to the spread
    res = torch.randn(m, c, device=DEV).to(torch.bfloat16)
    t, stats = ops.conv_gemm(x0, w0, c, bias=b0, res0=res, want_stats=True)
    tref = x0.float() @ w0.float().t() + b0 + res.float()
    assert relerr(t, tref) < 1e-2
    tot = stats.sum(1)
    assert torch.allclose(tot[:, 0], t.float().sum(1), rtol=1e-4, atol=1e-2)
    assert torch.allclose(tot[:, 1], (t.float() ** 2).sum(1), rtol=1e-4, atol=1e-2)

    norm = SimpleNamespace(g=torch.randn(c, device=DEV) * 0.3 + 1.0, b=torch.randn(c, device=DEV) * 0.2)
    xn = F.layer_norm(t.float(), (c,), norm.g, norm.b, LN_EPS)
    ln = lambda s: (stats, s, c, LN_EPS)

    # plain Linear with bias
    w = torch.randn(c, c, device=DEV) / math.sqrt(c)
    b = torch.randn(c, device=DEV)
    wf, s, tb = fold_layernorm(w, b, norm)
    y = ops.conv_gemm(t, wf, c, bias=tb, ln=ln(s))
    assert relerr(y, xn @ w.t() + b) < 1.5e-2

    # q/k/v projection (head-major layouts), no bias in the reference layer
    if m % 16 == 0:
        images, n = 2, m // 2
        w = torch.randn(3 * c, c, device=DEV) / math.sqrt(c)
        wf, s, tb = fold_layernorm(w, None, norm)
        q = torch.empty(images, heads, n, d, device=DEV, dtype=torch.bfloat16)
        k = torch.empty_like(q)
        vt = torch.empty(images, heads, d, n, device=DEV, dtype=torch.bfloat16)
        ops.conv_gemm(t, wf, 3 * c, bias=tb, rows_per_img=n, out_mode=ops.OUT_QKV, qkv=(q, k, vt, 0, heads, d, n), ln=ln(s))
        ref = (xn @ w.t()).view(images, n, 3, heads, d)
        assert relerr(q, ref[:, :, 0].permute(0, 2, 1, 3)) < 1.5e-2
        assert relerr(k, ref[:, :, 1].permute(0, 2, 1, 3)) < 1.5e-2
        assert relerr(vt, ref[:, :, 2].permute(0, 2, 3, 1)) < 1.5e-2

    # GEGLU
    inner = 4 * c
    w = torch.randn(2 * inner, c, device=DEV) / math.sqrt(c)
    b = torch.randn(2 * inner, device=DEV)
    wg = w * norm.g[None, :]
    wi, bi = ops.pack_geglu(wg, w @ norm.b + b)
    _, si = ops.pack_geglu(wg, wg.to(torch.bfloat16).float().sum(1))
    y = ops.conv_gemm(t, wi, 2 * inner, bias=bi, out_mode=ops.OUT_GEGLU, ln=ln(si))
    a, g = (xn @ w.t() + b).chunk(2, dim=-1)
    assert relerr(y, a * F.gelu(g)) < 1.5e-2


def _attn_ref(q, k, v, scale):
    s = torch.einsum("bhid,bhjd->bhij", q.float(), k.float()) * scale
    return torch.einsum("bhij,bhjd->bhid", s.softmax(-1), v.float())


@pytest.mark.parametrize("images,heads,nq,nk,d", [(2, 8, 4096, 4096, 40), (2, 8, 1024, 1024, 80), (2, 8, 256, 256, 160),
                                                  (2, 8, 64, 64, 160), (2, 8, 4096, 77, 40), (2, 8, 256, 77, 160),
                                                  (1, 2, 128, 128, 8), (1, 4, 64, 77, 32), (1, 1, 256, 384, 64)])
def test_attention(images, heads, nq, nk, d):
    q = torch.randn(images, heads, nq, d, device=DEV).to(torch.bfloat16)
    k = torch.randn(images, heads, nk, d, device=DEV).to(torch.bfloat16)
    v = torch.randn(images, heads, nk, d, device=DEV).to(torch.bfloat16)
    tok_pad = (nk + 7) // 8 * 8
    vt = torch.zeros(images, heads, d, tok_pad, device=DEV, dtype=torch.bfloat16)
    vt[..., :nk] = v.transpose(-1, -2)
    out = torch.empty(images, nq, heads * d, device=DEV, dtype=torch.bfloat16)
    ops.attention(q, k, vt, out, images, heads, nq, nk, d, tok_pad)
    ref = _attn_ref(q, k, v, d ** -0.5).permute(0, 2, 1, 3).reshape(images, nq, heads * d)
    assert relerr(out, ref) < 1.5e-2


@pytest.mark.parametrize("images,heads,nq,nk,d", [(2, 1, 4096, 4096, 512), (1, 1, 1024, 1024, 512), (1, 2, 256, 256, 256),
                                                  (1, 1, 200, 300, 384), (2, 1, 128, 77, 512), (1, 1, 64, 64, 1024)])
def test_attention_wide_heads(images, heads, nq, nk, d):
    """d > 192 (the VAE mid-block AttnBlock, model.py:178-202: one head of 512 channels): csrc/attention_wide.cu streams
    Q / K chunks, no N x N buffer. Ragged key / query counts exercise the TMA zero-fill + mask path."""
    torch.manual_seed(11)
    q = torch.randn(images, heads, nq, d, device=DEV).to(torch.bfloat16)
    k = torch.randn(images, heads, nk, d, device=DEV).to(torch.bfloat16)
    v = torch.randn(images, heads, nk, d, device=DEV).to(torch.bfloat16)
    tok_pad = (nk + 7) // 8 * 8
    vt = torch.zeros(images, heads, d, tok_pad, device=DEV, dtype=torch.bfloat16)
    vt[..., :nk] = v.transpose(-1, -2)
    out = torch.full((images, nq, heads * d), float("nan"), device=DEV, dtype=torch.bfloat16)
    ops.attention(q, k, vt, out, images, heads, nq, nk, d, tok_pad)
    ref = _attn_ref(q, k, v, d ** -0.5).permute(0, 2, 1, 3).reshape(images, nq, heads * d)
    assert torch.isfinite(out.float()).all()
    assert relerr(out, ref) < 1.5e-2


@pytest.mark.parametrize("scale_q,step", [(3.0, 0.0), (0.3, 12.0), (0.3, 3.0)])
def test_attention_wide_heads_adversarial(scale_q, step):
    """Wide logits (softmax nearly one-hot) and a row maximum that rises on every key tile: the lazy rescale of O in
    TMEM and the running sum of attention_wide.cu."""
    torch.manual_seed(12)
    n, d = 1024, 512
    u = F.normalize(torch.randn(d, device=DEV), dim=0)
    ramp = (torch.arange(n, device=DEV) // 128).float() * (step * 0.6931)
    q = torch.randn(1, 1, n, d, device=DEV) * scale_q + (u * (d ** 0.5) if step else 0.0)
    k = torch.randn(1, 1, n, d, device=DEV) * scale_q + u[None, None, None, :] * ramp[None, None, :, None]
    v = torch.randn(1, 1, n, d, device=DEV)
    q, k, v = q.to(torch.bfloat16), k.to(torch.bfloat16), v.to(torch.bfloat16)
    vt = v.transpose(-1, -2).contiguous()
    out = torch.empty(1, n, d, device=DEV, dtype=torch.bfloat16)
    ops.attention(q, k, vt, out, 1, 1, n, n, d, n)
    ref = _attn_ref(q, k, v, d ** -0.5).permute(0, 2, 1, 3).reshape(1, n, d)
    assert torch.isfinite(out.float()).all()
    assert relerr(out, ref) < 1.5e-2


def test_attention_causal_and_text_kernels():
    """Causal mask (CLIP text transformer), token+position gather, quick-GELU epilogue."""
    images, heads, n, d = 3, 12, 77, 64
    q = torch.randn(images, heads, n, d, device=DEV).to(torch.bfloat16)
    k = torch.randn(images, heads, n, d, device=DEV).to(torch.bfloat16)
    v = torch.randn(images, heads, n, d, device=DEV).to(torch.bfloat16)
    tp = 80
    vt = torch.zeros(images, heads, d, tp, device=DEV, dtype=torch.bfloat16)
    vt[..., :n] = v.transpose(-1, -2)
    out = torch.empty(images, n, heads * d, device=DEV, dtype=torch.bfloat16)
    ops.attention_causal(q, k, vt, out, images, heads, n, d, tp)
    s = torch.einsum("bhid,bhjd->bhij", q.float(), k.float()) * d ** -0.5
    s = s + torch.full((n, n), float("-inf"), device=DEV).triu(1)
    ref = torch.einsum("bhij,bhjd->bhid", s.softmax(-1), v.float()).permute(0, 2, 1, 3).reshape(images, n, heads * d)
    assert relerr(out, ref) < 1e-2
    for n2 in (200, 300):                                  # more than one KV tile / query tile
        q2, k2, v2 = (torch.randn(1, 2, n2, d, device=DEV).to(torch.bfloat16) for _ in range(3))
        tp2 = (n2 + 7) // 8 * 8
        vt2 = torch.zeros(1, 2, d, tp2, device=DEV, dtype=torch.bfloat16)
        vt2[..., :n2] = v2.transpose(-1, -2)
        o2 = torch.empty(1, n2, 2 * d, device=DEV, dtype=torch.bfloat16)
        ops.attention_causal(q2, k2, vt2, o2, 1, 2, n2, d, tp2)
        s2 = torch.einsum("bhid,bhjd->bhij", q2.float(), k2.float()) * d ** -0.5
        s2 = s2 + torch.full((n2, n2), float("-inf"), device=DEV).triu(1)
        r2 = torch.einsum("bhij,bhjd->bhid", s2.softmax(-1), v2.float()).permute(0, 2, 1, 3).reshape(1, n2, 2 * d)
        assert relerr(o2, r2) < 1e-2
    # embeddings
    tok, pos = torch.randn(1000, 768, device=DEV), torch.randn(77, 768, device=DEV)
    ids = torch.randint(0, 1000, (4, 77), device=DEV)
    e = ops.embed_tokens(ids, tok, pos)
    assert relerr(e.view(4, 77, 768), tok[ids] + pos[None]) < 5e-3
    # quick-GELU epilogue
    x = torch.randn(154, 768, device=DEV).to(torch.bfloat16)
    w = (torch.randn(3072, 768, device=DEV) / math.sqrt(768)).to(torch.bfloat16)
    b = torch.randn(3072, device=DEV)
    y = ops.conv_gemm(x, w, 3072, bias=b, act=ops.ACT_QUICK_GELU)
    h = x.float() @ w.float().t() + b
    assert relerr(y, h * torch.sigmoid(1.702 * h)) < 1e-2


def test_qkv_projection_layouts():
    images, n, c, heads = 2, 256, 320, 8
    d = c // heads
    x = torch.randn(images * n, c, device=DEV).to(torch.bfloat16)
    w = (torch.randn(3 * c, c, device=DEV) / math.sqrt(c)).to(torch.bfloat16)
    q = torch.empty(images, heads, n, d, device=DEV, dtype=torch.bfloat16)
    k = torch.empty_like(q)
    vt = torch.empty(images, heads, d, n, device=DEV, dtype=torch.bfloat16)
    ops.conv_gemm(x, w, 3 * c, rows_per_img=n, out_mode=ops.OUT_QKV, qkv=(q, k, vt, 0, heads, d, n))
    ref = (x.float() @ w.float().t()).view(images, n, 3, heads, d)
    assert relerr(q, ref[:, :, 0].permute(0, 2, 1, 3)) < 1e-2
    assert relerr(k, ref[:, :, 1].permute(0, 2, 1, 3)) < 1e-2
    assert relerr(vt, ref[:, :, 2].permute(0, 2, 3, 1)) < 1e-2


@pytest.mark.parametrize("images,n,c,mode", [(2, 4096, 320, "qkv"), (3, 1024, 640, "qkv_ln"), (2, 256, 1280, "qkv_ln"),
                                             (2, 1024, 320, "q_ln"), (2, 128, 640, "kv_pad"), (16, 4096, 320, "qkv")])
def test_qkv_head_aligned_tma_epilogue(images, n, c, mode):
    """The head-aligned TMA epilogue of the q/k/v projections (160-column tiles, dhead 40 / 80 / 160; gemm.cu tma_epi == 2)
    against the fp32 reference and BIT-equal to the generic per-lane epilogue (forced with bn_hint = 128)."""
    from types import SimpleNamespace
    from fgdm_b200.layers import LN_EPS, fold_layernorm
    heads, d = 8, c // 8
    m = images * n
    torch.manual_seed(images * 7 + n)
    x = (torch.randn(m, c, device=DEV) * 1.5 + 0.3).to(torch.bfloat16)
    sections = {"qkv": 3, "qkv_ln": 3, "q_ln": 1, "kv_pad": 2}[mode]
    w = torch.randn(sections * c, c, device=DEV) / math.sqrt(c)
    tp = n + 8 if mode == "kv_pad" else n
    kw, xref = {}, x.float()
    if mode.endswith("_ln"):
        # statistics as a producing GEMM writes them: per-row {sum, sumsq} partials in 2 slots
        xf = x.float()
        half = c // 2
        stats = torch.stack([torch.stack([xf[:, :half].sum(1), (xf[:, :half] ** 2).sum(1)], -1),
                             torch.stack([xf[:, half:].sum(1), (xf[:, half:] ** 2).sum(1)], -1)], 1).contiguous()
        norm = SimpleNamespace(g=torch.randn(c, device=DEV) * 0.3 + 1.0, b=torch.randn(c, device=DEV) * 0.2)
        wf, s, tb = fold_layernorm(w, None, norm)
        kw = dict(bias=tb, ln=(stats, s, c, LN_EPS))
        xref = F.layer_norm(xf, (c,), norm.g, norm.b, LN_EPS)
    else:
        wf = w.to(torch.bfloat16)
        w = wf.float()
    ref = (xref @ w.t()).view(images, n, sections, heads, d)

    def run(bn_hint):
        q = torch.full((images, heads, n, d), float("nan"), device=DEV, dtype=torch.bfloat16)
        k = torch.full_like(q, float("nan"))
        vt = torch.zeros(images, heads, d, tp, device=DEV, dtype=torch.bfloat16)
        first = 1 if mode == "kv_pad" else 0
        outs = (q, q, q) if sections == 1 else (None, k, vt) if first else (q, k, vt)
        # the cross-attention q projection passes the CONTEXT's padded length (80 for 77 tokens) as tok_pad
        ops.conv_gemm(x, wf, sections * c, rows_per_img=n, out_mode=ops.OUT_QKV,
                      qkv=outs + (first, heads, d, 80 if sections == 1 else tp),
                      bn_hint=bn_hint, **kw)
        return q, k, vt

    q, k, vt = run(0)
    q2, k2, vt2 = run(128)
    tol = 1.5e-2 if kw else 1e-2
    if sections == 3:
        assert relerr(q, ref[:, :, 0].permute(0, 2, 1, 3)) < tol
        assert relerr(k, ref[:, :, 1].permute(0, 2, 1, 3)) < tol
        assert relerr(vt, ref[:, :, 2].permute(0, 2, 3, 1)) < tol
        assert torch.equal(q, q2) and torch.equal(k, k2) and torch.equal(vt, vt2)
    elif sections == 1:
        assert relerr(q, ref[:, :, 0].permute(0, 2, 1, 3)) < tol
        assert torch.equal(q, q2)
    else:
        assert relerr(k, ref[:, :, 0].permute(0, 2, 1, 3)) < tol
        assert relerr(vt[..., :n], ref[:, :, 1].permute(0, 2, 3, 1)) < tol
        assert float(vt[..., n:].abs().max()) == 0.0                   # the padding columns stay untouched
        assert torch.equal(k, k2) and torch.equal(vt, vt2)


@pytest.mark.parametrize("nb,h,w,c0,c1,silu,eps", [(2, 64, 64, 320, 0, True, 1e-5), (2, 16, 16, 1280, 640, True, 1e-5),
                                                   (1, 8, 8, 64, 0, False, 1e-6), (1, 128, 128, 128, 0, True, 1e-6),
                                                   (3, 32, 32, 640, 320, False, 1e-5),
                                                   # the benchmark's 64x64 level (single-launch re-read kernel, 8 / 16-byte slabs)
                                                   (16, 64, 64, 320, 0, True, 1e-5), (16, 64, 64, 640, 320, True, 1e-5),
                                                   (8, 64, 64, 320, 0, False, 1e-5), (8, 64, 64, 512, 0, "exact", 1e-6),
                                                   (16, 64, 64, 320, 320, True, 1e-5), (12, 48, 48, 320, 0, True, 1e-5)])
def test_groupnorm(nb, h, w, c0, c1, silu, eps):
    x0 = nhwc(torch.randn(nb, c0, h, w, device=DEV) * 2 + 0.5)
    x1 = nhwc(torch.randn(nb, c1, h, w, device=DEV)) if c1 else None
    g = torch.randn(c0 + c1, device=DEV)
    b = torch.randn(c0 + c1, device=DEV)
    y = ops.groupnorm(x0, g, b, x1=x1, eps=eps, silu=silu)
    xin = nchw(x0) if x1 is None else torch.cat([nchw(x0), nchw(x1)], 1)
    ref = F.group_norm(xin, 32, g, b, eps)
    if silu:
        ref = F.silu(ref)
    assert relerr(nchw(y), ref) < 1e-2


@pytest.mark.parametrize("nb,h,w,c", [(16, 16, 16, 1280), (16, 32, 32, 640), (2, 64, 64, 128), (3, 16, 32, 64), (1, 8, 8, 64)])
def test_upsample_subpixel_phases(nb, h, w, c, monkeypatch):
    """Upsample (openaimodel.py:95-130, model.py:42-57): nearest 2x + conv3x3 as four 4-tap sub-pixel convolutions on the
    low-resolution tensor (FGDM_CONV_UP2_P*, pixel shuffle = fgdm_resample mode 3) against F.conv2d(F.interpolate(x)) and
    against the kernel path that materialises the upsampled tensor."""
    from fgdm_b200 import layers
    torch.manual_seed(21)
    up = layers.Upsample(c)
    sd = {"conv.weight": torch.randn(c, c, 3, 3) * (9 * c) ** -0.5, "conv.bias": torch.randn(c) * 0.1}
    up.load(sd, "", DEV)
    x = torch.randn(nb, c, h, w, device=DEV)
    xb = nhwc(x)
    monkeypatch.setattr(layers, "UP_PHASE", True)
    monkeypatch.setattr(layers, "UP_PHASE_MIN_ROWS", 1)
    y = up.forward(xb)
    monkeypatch.setattr(layers, "UP_PHASE", False)
    y_old = up.forward(xb)
    ref = F.conv2d(F.interpolate(nchw(xb), scale_factor=2, mode="nearest"), sd["conv.weight"].to(DEV), sd["conv.bias"].to(DEV),
                   padding=1)
    assert y.shape == y_old.shape == (nb, 2 * h, 2 * w, c)
    e_new, e_old = relerr(nchw(y), ref), relerr(nchw(y_old), ref)
    print(f"upsample {nb}x{h}x{w}x{c}: sub-pixel {e_new:.2e}, materialised {e_old:.2e}")
    assert e_new < 1.5e-2 and e_old < 1.5e-2


@pytest.mark.parametrize("rows,c", [(4096, 320), (1001, 320), (1024, 640), (256, 1280), (100, 64)])
def test_layernorm(rows, c):
    x = (torch.randn(rows, c, device=DEV) * 3 + 1).to(torch.bfloat16)
    g = torch.randn(c, device=DEV)
    b = torch.randn(c, device=DEV)
    y = ops.layernorm(x, g, b)
    assert relerr(y, F.layer_norm(x.float(), (c,), g, b, 1e-5)) < 1e-2


def test_softmax_rows():
    s = torch.randn(512, 4096, device=DEV) * 20
    p = ops.softmax_rows(s, 0.1)
    assert relerr(p, (s * 0.1).softmax(-1)) < 1e-2


def test_cfg_ddim():
    b = 8
    eu, ec, x = (torch.randn(b, 4, 64, 64, device=DEV) for _ in range(3))
    table = torch.rand(50, 5, device=DEV) + 0.1
    xp, p0 = ops.cfg_ddim_step(eu, ec, x, table, 7, 7.5)
    s1m, sa, sap, dirc, sig = table[7]
    e = eu + 7.5 * (ec - eu)
    x0 = (x - s1m * e) / sa
    torch.testing.assert_close(p0, x0, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(xp, sap * x0 + dirc * e, rtol=1e-5, atol=1e-5)
    idx = torch.tensor([7], device=DEV, dtype=torch.int32)
    xp2, _ = ops.cfg_ddim_step(eu, ec, x, table, 0, 7.5, index_ptr=idx)
    assert torch.equal(xp, xp2)


def test_lincomb3_and_float_timestep_embedding():
    x, y, z = (torch.randn(2, 4, 64, 64, device=DEV) for _ in range(3))
    torch.testing.assert_close(ops.lincomb3(0.7, x, -1.3, y, 0.25, z), 0.7 * x - 1.3 * y + 0.25 * z, rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(ops.lincomb3(0.7, x, -1.3, y), 0.7 * x - 1.3 * y, rtol=1e-6, atol=1e-6)
    xc = x.clone()
    ops.lincomb3(2.0, xc, 1.0, y, out=xc)
    torch.testing.assert_close(xc, 2.0 * x + y, rtol=1e-6, atol=1e-6)
    freqs = ops.timestep_freqs(320).to(DEV)
    t = torch.tensor([999.0, 749.25, 0.0, 12.5], device=DEV)
    e = ops.timestep_embedding(t, freqs, 320)
    arg = t[:, None] * freqs[None]
    torch.testing.assert_close(e, torch.cat([arg.cos(), arg.sin()], -1), rtol=1e-4, atol=2e-4)
    ti = torch.tensor([999, 749, 0, 12], device=DEV)
    torch.testing.assert_close(ops.timestep_embedding(ti, freqs, 320), ops.timestep_embedding(ti.float(), freqs, 320))


def test_plms_step():
    """fgdm_plms_step vs plms.py:178-234 in torch (CFG combine, Adams-Bashforth blend, x_prev / pred_x0)."""
    torch.manual_seed(3)
    n = (2, 4, 64, 64)
    e_u, e_c, x = (torch.randn(n, device=DEV) for _ in range(3))
    old = [torch.randn(n, device=DEV) for _ in range(3)]           # old_eps[-3], [-2], [-1]
    coef = torch.rand(7, 5, device=DEV) + 0.5
    s, idx = 7.5, 4
    s1m, sa, sap, dr, _ = coef[idx].tolist()
    e_t = e_u + s * (e_c - e_u)
    cases = {0: e_t, 1: (3 * e_t - old[-1]) / 2, 2: (23 * e_t - 16 * old[-1] + 5 * old[-2]) / 12,
             3: (55 * e_t - 59 * old[-1] + 37 * old[-2] - 9 * old[-3]) / 24}
    weights = {0: (1.0, 0.0, 0.0, 0.0), 1: (3 / 2, -1 / 2, 0.0, 0.0), 2: (23 / 12, -16 / 12, 5 / 12, 0.0),
               3: (55 / 24, -59 / 24, 37 / 24, -9 / 24)}
    for order, ep in cases.items():
        e_out = torch.empty_like(x)
        xp, p0 = ops.plms_step(e_u, e_c, old[::-1][:order], weights[order], x, coef, idx, s, e_out=e_out)
        ref0 = (x - s1m * ep) / sa
        torch.testing.assert_close(e_out, e_t, rtol=1e-6, atol=1e-6)
        torch.testing.assert_close(p0, ref0, rtol=1e-5, atol=2e-5)
        torch.testing.assert_close(xp, sap * ref0 + dr * ep, rtol=1e-5, atol=2e-5)
    with pytest.raises(RuntimeError):
        ops.plms_step(e_u, e_c, [], weights[2], x, coef, idx, s)   # weight without its history tensor


def test_timestep_embedding_and_skinny_linear():
    t = torch.tensor([1, 21, 981, 500], device=DEV, dtype=torch.int64)
    freqs = ops.timestep_freqs(320).to(DEV)
    emb = ops.timestep_embedding(t, freqs, 320)
    args = t[:, None].float() * freqs[None]
    ref = torch.cat([torch.cos(args), torch.sin(args)], -1)
    torch.testing.assert_close(emb, ref, rtol=1e-4, atol=2e-4)
    w = torch.randn(1280, 320, device=DEV) / 18
    b = torch.randn(1280, device=DEV)
    wb = w.to(torch.bfloat16)
    y = ops.skinny_linear(emb, wb, b, in_act=True, out_act=True)
    torch.testing.assert_close(y, F.silu(F.silu(emb) @ wb.float().t() + b), rtol=1e-3, atol=1e-3)
    x19 = torch.randn(19, 320, device=DEV)
    torch.testing.assert_close(ops.skinny_linear(x19, wb, None), x19 @ wb.float().t(), rtol=1e-3, atol=1e-3)
    w2 = (torch.randn(2000, 1280, device=DEV) / 36).to(torch.bfloat16)
    x16 = torch.randn(16, 1280, device=DEV)
    torch.testing.assert_close(ops.skinny_linear(x16, w2, None, in_act=True), F.silu(x16) @ w2.float().t(),
                               rtol=1e-3, atol=1e-3)


def test_resample_modes():
    x = nhwc(torch.randn(2, 64, 8, 8, device=DEV))
    up = ops.resample(x, 0)
    assert torch.equal(nchw(up), F.interpolate(nchw(x), scale_factor=2, mode="nearest"))
    pool = ops.resample(x, 1)
    assert relerr(nchw(pool), F.avg_pool2d(nchw(x), 2)) < 1e-2
    ph = ops.resample(x, 2)
    xr = x.view(2, 4, 2, 4, 2, 64).permute(2, 4, 0, 1, 3, 5).reshape(4, 2, 4, 4, 64)
    assert torch.equal(ph, xr)
    assert relerr(ops.add_bf16(x, x), x.float() * 2) < 1e-2


def test_small_convs():
    x = torch.randn(2, 4, 32, 32, device=DEV)
    w = torch.randn(320, 4, 3, 3, device=DEV) / 6
    b = torch.randn(320, device=DEV)
    y = ops.conv_small_cin(x, ops.pack_conv_weight(w), b, 320)
    assert relerr(nchw(y), F.conv2d(x, w.to(torch.bfloat16).float(), b, padding=1)) < 1e-2
    h3 = torch.rand(1, 3, 64, 64, device=DEV)
    w3 = torch.randn(16, 3, 3, 3, device=DEV) / 5
    y3 = ops.conv_small_cin(h3, ops.pack_conv_weight(w3), None, 16, act=ops.ACT_SILU)
    assert relerr(nchw(y3), F.silu(F.conv2d(h3, w3.to(torch.bfloat16).float(), None, padding=1))) < 1e-2
    # direct, stride 2
    xd = nhwc(torch.randn(2, 16, 32, 32, device=DEV))
    wd = torch.randn(32, 16, 3, 3, device=DEV) / 12
    bd = torch.randn(32, device=DEV)
    yd = ops.conv_direct(xd, ops.pack_conv_weight(wd), bd, 32, stride=2, act=ops.ACT_SILU)
    assert relerr(nchw(yd), F.silu(F.conv2d(nchw(xd), wd.to(torch.bfloat16).float(), bd, stride=2, padding=1))) < 1e-2
    # small cout
    for cin, cout in [(320, 4), (128, 3), (64, 4)]:
        xc = nhwc(torch.randn(2, cin, 16, 16, device=DEV))
        wc = torch.randn(cout, cin, 3, 3, device=DEV) / math.sqrt(9 * cin)
        bc = torch.randn(cout, device=DEV)
        yc = ops.conv_small_cout(xc, ops.pack_conv_weight(wc), bc, cout)
        assert relerr(yc, F.conv2d(nchw(xc), wc.to(torch.bfloat16).float(), bc, padding=1)) < 1e-2
    z = torch.randn(2, 4, 8, 8, device=DEV)
    wq = torch.randn(4, 4, device=DEV)
    bq = torch.randn(4, device=DEV)
    # tensor-core routes for the tiny-channel convs: im2col + GEMM (cin 4), NCHW-fp32 epilogue (cout 4 / 3)
    from fgdm_b200.layers import Conv2d
    for (ci, co, hw2, rows) in [(4, 320, 32, 4), (320, 4, 16, 2), (128, 3, 32, 1)]:
        cv = Conv2d(ci, co, 3)
        wt = torch.randn(co, ci, 3, 3, device=DEV) / math.sqrt(9 * ci)
        bt = torch.randn(co, device=DEV)
        cv.load({"weight": wt, "bias": bt}, "", DEV)
        if ci == 4:
            xin = torch.randn(2, 4, hw2, hw2, device=DEV)
            out = cv.from_nchw(xin, rows, in_scale=0.5)
            ref = F.conv2d((xin * 0.5).repeat(rows // 2, 1, 1, 1), wt.to(torch.bfloat16).float(), bt, padding=1)
            assert relerr(nchw(out), ref) < 1e-2
        else:
            xin = nhwc(torch.randn(rows, ci, hw2, hw2, device=DEV))
            out = cv.to_nchw(xin)
            assert out.shape == (rows, co, hw2, hw2) and out.dtype == torch.float32
            assert relerr(out, F.conv2d(nchw(xin), wt.to(torch.bfloat16).float(), bt, padding=1)) < 1e-2
    ya = ops.latent_affine(z, wq, bq, 1 / 0.18215)
    torch.testing.assert_close(ya, F.conv2d(z / 0.18215, wq[:, :, None, None], bq), rtol=1e-4, atol=1e-4)


def test_u8_handoff_bit_exact():
    x = torch.randn(2, 3, 64, 64, device=DEV) * 0.8
    u0 = ops.image_to_u8(x, 0)
    ref0 = (255.0 * torch.clamp((x + 1.0) / 2.0, 0.0, 1.0)).permute(0, 2, 3, 1).cpu().numpy().astype("uint8")
    assert (u0.cpu().numpy() == ref0).all()
    u1 = ops.image_to_u8(x, 1)
    ref1 = (x.permute(0, 2, 3, 1) * 127.5 + 127.5).cpu().numpy().clip(0, 255).astype("uint8")
    assert (u1.cpu().numpy() == ref1).all()
    hint = ops.u8_to_hint(u0)
    assert torch.equal(hint, (u0.float() / 255.0).permute(0, 3, 1, 2))
    cv2 = pytest.importorskip("cv2")
    up = ops.resize_u8(u0, 128, 128).cpu().numpy()
    for i in range(2):
        ref = cv2.resize(u0[i].cpu().numpy(), (128, 128), interpolation=cv2.INTER_LINEAR)
        assert (up[i] == ref).all()
    same = ops.resize_u8(u0, 64, 64)
    assert torch.equal(same, u0)
