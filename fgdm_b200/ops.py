"""Tensor-level wrappers over the C ABI. torch is used only for device memory and the current stream.

Activations are NHWC bf16 tensors `[nb, h, w, c]` (tokens: `[nb, n, c]`); see include/fgdm_b200.h.
Every function launches on torch's current CUDA stream and never synchronises.
"""
from __future__ import annotations

import ctypes as C
import math

import torch

from . import _lib
from ._lib import (ACT_NONE, ACT_QUICK_GELU, ACT_RELU, ACT_SILU, CONV_1X1, CONV_3X3, CONV_3X3_S2, CONV_UP2_P0, OUT_BF16, OUT_F32, OUT_GEGLU,
                   OUT_NCHW_F32, OUT_QKV, GemmArgs)

bf16 = torch.bfloat16


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t):
    return None if t is None else t.data_ptr()


_REPLAYED = 0


_PROF = None   # when a list: every launch is bracketed by CUDA events (bench.py's per-kernel roofline pass)


def profile_start():
    global _PROF
    _PROF = []


def profile_stop():
    """-> {kernel: {shape_key: [launches, ms, algorithmic flops, algorithmic bytes]}} (synchronises)."""
    global _PROF
    rec, _PROF = _PROF, None
    torch.cuda.synchronize()
    out = {}
    for name, (key, flops, nbytes), e0, e1 in rec:
        slot = out.setdefault(name, {}).setdefault(key, [0, 0.0, 0.0, 0.0])
        slot[0] += 1
        slot[1] += e0.elapsed_time(e1)
        slot[2] += flops
        slot[3] += nbytes
    return out


def _call(name, work, *args):
    fn = getattr(_lib.load(), name)
    if _PROF is None:
        _lib.check(fn(*args), name)
        return
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    rc = fn(*args)
    e1.record()
    _lib.check(rc, name)
    _PROF.append((name, work, e0, e1))


def launch_count() -> int:
    """Kernels of libfgdm_b200.so launched so far: direct launches + those replayed through CUDA graphs."""
    return int(_lib.load().fgdm_launch_count()) + _REPLAYED


def set_scratch_slot(slot: int) -> None:
    """Scratch set (0 / 1) of the launches that follow; see fgdm_set_scratch_slot in include/fgdm_b200.h."""
    rc = _lib.load().fgdm_set_scratch_slot(int(slot))
    if rc:
        raise RuntimeError(_lib.load().fgdm_last_error().decode())


def count_launches(n: int) -> None:
    """Account for `n` of our kernels executed by a CUDA-graph replay."""
    global _REPLAYED
    _REPLAYED += n


def conv_gemm(a0, weight, n, *, conv=CONV_1X1, a1=None, nb=None, h=None, w=None, bias=None, rowvec=None,
              rows_per_img=0, res0=None, res1=None, alpha=1.0, act=ACT_NONE, out_mode=OUT_BF16, out=None,
              ldo=None, col_off=0, qkv=None, bn_hint=0, n_store=0, res1_rows=0, want_stats=False, ln=None):
    """Implicit-GEMM conv / Linear. a0 (and a1) are NHWC bf16 `[nb, h, w, c]` or token `[rows, c]` tensors.

    For CONV_3X3_S2, a0 is the phase-split tensor `[4, nb, h, w, c]` (see resample(mode=2)) and
    (nb, h, w) describe the OUTPUT. Returns the output tensor (or None in QKV mode).

    Folded LayerNorm (see include/fgdm_b200.h): `want_stats=True` makes a bf16-output GEMM also return the per-row
    partial statistics of its output, `(out, stats[m, slots, 2] fp32)`; a consumer passes
    `ln=(stats, ln_s[n] fp32, channels, eps)` together with gamma-folded weights and `bias = beta.W^T + b`.
    """
    c0 = a0.shape[-1]
    c1 = a1.shape[-1] if a1 is not None else 0
    if nb is None:
        if a0.dim() == 4:
            nb, h, w = a0.shape[0], a0.shape[1], a0.shape[2]
        else:  # [rows, c] or [b, n, c] tokens: one flat "image row"
            nb, h, w = 1, 1, a0.numel() // c0
    m = nb * h * w
    args = GemmArgs()
    args.a0, args.a1, args.c0, args.c1 = _p(a0), _p(a1), c0, c1
    args.nb, args.h, args.w, args.conv = nb, h, w, conv
    args.weight, args.n = _p(weight), n
    args.bias, args.rowvec = _p(bias), _p(rowvec)
    args.ld_rowvec = rowvec.stride(0) if rowvec is not None else 0
    args.rows_per_img = rows_per_img
    args.res0, args.ld_res0 = _p(res0), (res0.stride(-2) if res0 is not None else 0)
    args.res1, args.ld_res1 = _p(res1), (res1.stride(-2) if res1 is not None else 0)
    args.alpha, args.act, args.out_mode = alpha, act, out_mode
    args.bn_hint = bn_hint
    if res1 is not None and res1_rows == 0:
        r1 = res1.numel() // res1.shape[-1]
        if r1 < m:                       # shared by several images (adapter features of the CFG halves): modulo rows
            assert m % r1 == 0, (m, r1)
            res1_rows = r1
    args.res1_rows = res1_rows
    if out_mode == OUT_QKV:
        q, k, vt, first, heads, dhead, tok_pad = qkv
        args.q_out, args.k_out, args.vt_out = _p(q), _p(k), _p(vt)
        args.qkv_first, args.heads, args.dhead, args.tok_pad = first, heads, dhead, tok_pad
    elif out_mode == OUT_NCHW_F32:
        if out is None:
            out = torch.empty((nb, n_store, h, w), device=a0.device, dtype=torch.float32)
        args.out, args.ldo, args.n_store = _p(out), 0, n_store
        if rows_per_img == 0:
            args.rows_per_img = h * w
    else:
        n_out = n // 2 if out_mode == OUT_GEGLU else n
        if out is None:
            out = torch.empty((m, n_out), device=a0.device, dtype=torch.float32 if out_mode == OUT_F32 else bf16)
            ldo = n_out
        elif ldo is None:
            ldo = out.shape[-1]
        args.out, args.ldo, args.col_off = _p(out), ldo, col_off
    if ln is not None:
        st, ln_s, ln_c, ln_eps = ln
        args.ln_stats, args.ln_slots, args.ln_s = _p(st), st.shape[1], _p(ln_s)
        args.ln_channels, args.ln_eps = ln_c, ln_eps
    stats = None
    if want_stats:
        slots = _lib.load().fgdm_gemm_stats_slots(C.byref(args))
        if slots <= 0:
            raise RuntimeError("fgdm_gemm_stats_slots: " + _lib.last_error())
        stats = torch.empty((m, slots, 2), device=a0.device, dtype=torch.float32)
        args.row_stats = _p(stats)
    taps = 1 if conv == CONV_1X1 else (4 if conv >= CONV_UP2_P0 else 9)
    kdim = taps * (c0 + c1)
    nres = (res0 is not None) + (res1 is not None)
    # profile key: shape + the epilogue variant (out mode, residuals, rowvec, LayerNorm fold consumer / producer)
    var = "".join(("" if out_mode == OUT_BF16 else " o%d" % out_mode, " r%d" % nres if nres else "", " rv" if rowvec is not None else "",
                   " ln" if ln is not None else "", " st" if want_stats else ""))
    _call("fgdm_conv_gemm", ("gemm taps=%d m=%d k=%d n=%d%s" % (taps, m, kdim, n, var), 2.0 * m * n * kdim,
                             2.0 * (m * (c0 + c1) + n * kdim + m * n * (1 + nres))), C.byref(args), _stream())
    return (out, stats) if want_stats else out


def attention(q, k, vt, out, images, heads, nq, nk, d, tok_pad, scale=None):
    scale = d ** -0.5 if scale is None else scale
    _call("fgdm_attention", ("attn d=%d nq=%d nk=%d" % (d, nq, nk), 4.0 * images * heads * nq * nk * d, 2.0 * images * heads * d * (2 * nq + 2 * nk)), _p(q), _p(k), _p(vt), _p(out), images, heads, nq, nk, d, tok_pad,
                                         scale, _stream())
    return out


def attention_causal(q, k, vt, out, images, heads, n, d, tok_pad, scale=None):
    """Causal self-attention (CLIP text transformer): query i attends to keys j <= i."""
    scale = d ** -0.5 if scale is None else scale
    _call("fgdm_attention_causal", ("attn causal d=%d n=%d" % (d, n), 2.0 * images * heads * n * n * d,
                                    2.0 * images * heads * d * 4 * n), _p(q), _p(k), _p(vt), _p(out), images, heads, n, d,
          tok_pad, scale, _stream())
    return out


def embed_tokens(ids, tok_emb, pos_emb):
    """ids int64 [B, L]; fp32 tables -> bf16 [B*L, dim] = tok_emb[ids] + pos_emb[position]."""
    b, L = ids.shape
    dim = tok_emb.shape[1]
    out = torch.empty((b * L, dim), device=ids.device, dtype=bf16)
    _call("fgdm_embed_tokens", ("embed_tokens", 0.0, 10.0 * b * L * dim), _p(ids), _p(tok_emb), _p(pos_emb), _p(out), b * L, L,
          dim, tok_emb.shape[0], _stream())
    return out


def _gn_workspace(device, nb, groups):
    # per call (a cached buffer that is later re-allocated would dangle inside captured CUDA graphs)
    return torch.empty(nb * 129 * groups * 2, device=device, dtype=torch.float32)


def groupnorm(x0, gamma, beta, *, x1=None, groups=32, eps=1e-5, silu=False, out=None):
    """x0 (and x1): NHWC bf16 [nb, h, w, c]; returns [nb, h, w, c0+c1] bf16. silu: False, True (one-MUFU tanh form) or
    "exact" (rcp + ex2 form: the VAE decoder)."""
    nb = x0.shape[0]
    c0 = x0.shape[-1]
    c1 = x1.shape[-1] if x1 is not None else 0
    hw = x0.numel() // (nb * c0)
    if out is None:
        out = torch.empty(tuple(x0.shape[:-1]) + (c0 + c1,), device=x0.device, dtype=bf16)
    ws = _gn_workspace(x0.device, nb, groups)
    _call("fgdm_groupnorm", ("gn c=%d hw=%d" % (c0 + c1, hw), 0.0, 4.0 * nb * hw * (c0 + c1)), _p(x0), _p(x1), _p(out), _p(gamma), _p(beta), _p(ws), nb, hw, c0, c1,
                                         groups, eps, (2 if silu == "exact" else 1) if silu else 0, _stream())
    return out


def layernorm(x, gamma, beta, eps=1e-5, out=None):
    c = x.shape[-1]
    rows = x.numel() // c
    if out is None:
        out = torch.empty_like(x)
    _call("fgdm_layernorm", ("ln c=%d" % c, 0.0, 4.0 * rows * c), _p(x), _p(out), _p(gamma), _p(beta), rows, c, eps, _stream())
    return out


def softmax_rows(s, scale, out=None):
    rows, cols = s.shape
    if out is None:
        out = torch.empty((rows, cols), device=s.device, dtype=bf16)
    _call("fgdm_softmax_rows", ("softmax", 0.0, 6.0 * rows * cols), _p(s), _p(out), rows, cols, scale, _stream())
    return out


def cfg_ddim_step(e_uncond, e_cond, x, coef_table, index, scale, *, noise=None, temperature=1.0, index_ptr=None,
                  x_prev=None, pred_x0=None):
    if x_prev is None:
        x_prev = torch.empty_like(x)
    if pred_x0 is None:
        pred_x0 = torch.empty_like(x)
    _call("fgdm_cfg_ddim_step", ("cfg_ddim", 0.0, 20.0 * x.numel()), _p(e_uncond), _p(e_cond), _p(x), _p(noise), _p(coef_table),
                                             _p(index_ptr), index, scale, temperature, _p(x_prev), _p(pred_x0),
                                             x.numel(), _stream())
    return x_prev, pred_x0


def plms_step(e_uncond, e_cond, hist, weights, x, coef_table, index, scale, *, e_out=None, x_prev=None, pred_x0=None):
    """hist: [old_eps[-1], old_eps[-2], old_eps[-3]] (shorter lists allowed); weights: 4 floats (e_t first)."""
    import ctypes as C
    if x_prev is None:
        x_prev = torch.empty_like(x)
    if pred_x0 is None:
        pred_x0 = torch.empty_like(x)
    h = list(hist) + [None] * (3 - len(hist))
    w = (C.c_float * 4)(*[float(v) for v in weights])
    _call("fgdm_plms_step", ("plms_step", 0.0, 4.0 * (4 + len(hist)) * x.numel()), _p(e_uncond), _p(e_cond), _p(h[0]), _p(h[1]),
          _p(h[2]), w, _p(x), _p(coef_table), int(index), float(scale), _p(e_out), _p(x_prev), _p(pred_x0), x.numel(),
          _stream())
    return x_prev, pred_x0


def timestep_embedding(t, freqs, dim):
    """t: int64 (DDIM/PLMS) or float32 (DPM-Solver's fractional timesteps) [rows]."""
    out = torch.empty((t.shape[0], dim), device=t.device, dtype=torch.float32)
    fn = "fgdm_timestep_embedding_f32" if t.dtype == torch.float32 else "fgdm_timestep_embedding"
    assert t.dtype in (torch.float32, torch.int64)
    _call(fn, ("temb", 0.0, 4.0 * t.shape[0] * dim), _p(t), _p(freqs), _p(out), t.shape[0], dim, _stream())
    return out


def lincomb3(a, x, b=0.0, y=None, c=0.0, z=None, out=None):
    """out = a x + b y + c z on fp32 tensors (y, z optional)."""
    if out is None:
        out = torch.empty_like(x)
    _call("fgdm_lincomb3", ("lincomb3", 0.0, 16.0 * x.numel()), float(a), _p(x), float(b), _p(y), float(c), _p(z), _p(out),
          x.numel(), _stream())
    return out


def skinny_linear(x, w, b, in_act=False, out_act=False):
    rows, k = x.shape
    n = w.shape[0]
    y = torch.empty((rows, n), device=x.device, dtype=torch.float32)
    _call("fgdm_skinny_linear", ("skinny k=%d n=%d" % (k, n), 2.0 * rows * k * n, 2.0 * k * n + 4.0 * rows * (k + n)), _p(x), _p(w), _p(b), _p(y), rows, k, n, int(in_act), int(out_act),
                                             _stream())
    return y


def resample(x, mode):
    """mode 0: nearest 2x up; 1: 2x2 avg pool; 2: phase split -> [4, nb, h/2, w/2, c]; 3: pixel shuffle
    [nb, h, w, 4*c'] -> [nb, 2h, 2w, c']."""
    nb, h, w, c = x.shape
    if mode == 0:
        out = torch.empty((nb, 2 * h, 2 * w, c), device=x.device, dtype=bf16)
    elif mode == 3:
        out = torch.empty((nb, 2 * h, 2 * w, c // 4), device=x.device, dtype=bf16)
    elif mode == 1:
        out = torch.empty((nb, h // 2, w // 2, c), device=x.device, dtype=bf16)
    else:
        out = torch.empty((4, nb, h // 2, w // 2, c), device=x.device, dtype=bf16)
    _call("fgdm_resample", ("resample mode=%d" % mode, 0.0, 2.0 * (x.numel() + out.numel())), _p(x), _p(out), nb, h, w, c, mode, _stream())
    return out


def add_bf16(a, b, out=None):
    if out is None:
        out = torch.empty_like(a)
    _call("fgdm_add_bf16", ("add", 0.0, 6.0 * a.numel()), _p(a), _p(b), _p(out), a.numel(), _stream())
    return out


def conv_small_cin(x, w, bias, cout, act=ACT_NONE, in_scale=1.0, nb_out=None):
    """x: NCHW fp32 with 3 or 4 channels -> NHWC bf16 [nb_out, h, w, cout]; output r reads x[r % nb_in]."""
    nb_in, cin, h, wd = x.shape
    nb = nb_out or nb_in
    y = torch.empty((nb, h, wd, cout), device=x.device, dtype=bf16)
    _call("fgdm_conv_small_cin", ("conv_small_cin %d->%d" % (cin, cout), 18.0 * nb * h * wd * cin * cout, 4.0 * nb * cin * h * wd + 2.0 * y.numel()), _p(x), _p(w), _p(bias), _p(y), nb, h, wd, cin, cout, act, in_scale,
                                              nb_in, _stream())
    return y


def im2col_small(x, nb_out=None, in_scale=1.0):
    """x: NCHW fp32 with 3 or 4 channels -> bf16 [nb_out*h*w, 64] im2col rows (k = tap*cin + c)."""
    nb_in, cin, h, wd = x.shape
    nb = nb_out or nb_in
    out = torch.empty((nb * h * wd, 64), device=x.device, dtype=bf16)
    _call("fgdm_im2col_small", ("im2col cin=%d" % cin, 0.0, 4.0 * nb * cin * h * wd + 128.0 * nb * h * wd),
          _p(x), _p(out), nb, h, wd, cin, in_scale, nb_in, _stream())
    return out


def ddim_advance(counter, t_table, idx_table, t_out, index_out):
    _call("fgdm_ddim_advance", ("advance", 0.0, 0.0), _p(counter), _p(t_table), _p(idx_table), _p(t_out), t_out.numel(),
                                            _p(index_out), _stream())


def conv_direct(x, w, bias, cout, stride=1, act=ACT_NONE):
    nb, h, wd, cin = x.shape
    y = torch.empty((nb, h // stride, wd // stride, cout), device=x.device, dtype=bf16)
    _call("fgdm_conv_direct", ("conv_direct %d->%d s%d" % (cin, cout, stride), 18.0 * y.numel() * cin, 2.0 * (x.numel() + y.numel())), _p(x), _p(w), _p(bias), _p(y), nb, h, wd, cin, cout, stride, act,
                                           _stream())
    return y


def conv_small_cout(x, w, bias, cout):
    """x: NHWC bf16 -> NCHW fp32 [nb, cout, h, w] with cout in {3, 4}."""
    nb, h, wd, cin = x.shape
    y = torch.empty((nb, cout, h, wd), device=x.device, dtype=torch.float32)
    _call("fgdm_conv_small_cout", ("conv_small_cout %d->%d" % (cin, cout), 18.0 * nb * h * wd * cin * cout, 2.0 * x.numel() + 4.0 * y.numel()), _p(x), _p(w), _p(bias), _p(y), nb, h, wd, cin, cout, _stream())
    return y


def latent_affine(x, w, b, in_scale):
    nb, c, h, wd = x.shape
    y = torch.empty_like(x)
    _call("fgdm_latent_affine", ("latent_affine", 0.0, 8.0 * x.numel()), _p(x), _p(w), _p(b), _p(y), nb, c, h * wd, in_scale, _stream())
    return y


def image_to_u8(x, mode):
    nb, c, h, wd = x.shape
    y = torch.empty((nb, h, wd, c), device=x.device, dtype=torch.uint8)
    _call("fgdm_image_to_u8", ("image_to_u8", 0.0, 5.0 * x.numel()), _p(x), _p(y), nb, c, h * wd, mode, _stream())
    return y


def resize_u8(src, dh, dw):
    nb, sh, sw, c = src.shape
    dst = torch.empty((nb, dh, dw, c), device=src.device, dtype=torch.uint8)
    _call("fgdm_resize_u8", ("resize_u8", 0.0, 1.0 * (src.numel() + dst.numel())), _p(src), _p(dst), nb, sh, sw, dh, dw, c, _stream())
    return dst


def u8_to_hint(src):
    nb, h, wd, c = src.shape
    dst = torch.empty((nb, c, h, wd), device=src.device, dtype=torch.float32)
    _call("fgdm_u8_to_hint", ("u8_to_hint", 0.0, 5.0 * src.numel()), _p(src), _p(dst), nb, c, h * wd, _stream())
    return dst


# ---------------------------------------------------------------------------------------------
# weight packing (load time, torch plumbing): reference layouts -> the C ABI's packed layouts
# ---------------------------------------------------------------------------------------------
def pack_conv_weight(w: torch.Tensor) -> torch.Tensor:
    """OIHW fp32 (nn.Conv2d) -> [O][kh*kw][I] bf16; Linear [O, I] passes through."""
    if w.dim() == 4:
        o, i, kh, kw = w.shape
        w = w.permute(0, 2, 3, 1).reshape(o, kh * kw * i)
    return w.contiguous().to(bf16)


def pack_geglu(w: torch.Tensor, b: torch.Tensor):
    """GEGLU proj [2*inner, dim] (value rows then gate rows, attention.py:40-44) -> rows interleaved
    per 256-column tile (128 value rows, then the matching 128 gate rows); bias likewise."""
    inner = w.shape[0] // 2
    assert inner % 128 == 0, "GEGLU inner dim must be a multiple of 128"
    wx, wg = w[:inner].reshape(inner // 128, 128, -1), w[inner:].reshape(inner // 128, 128, -1)
    wi = torch.cat([wx, wg], dim=1).reshape(2 * inner, -1)
    bx, bg = b[:inner].reshape(inner // 128, 128), b[inner:].reshape(inner // 128, 128)
    bi = torch.cat([bx, bg], dim=1).reshape(2 * inner)
    return wi.contiguous().to(bf16), bi.contiguous().float()


def timestep_freqs(dim: int, max_period: float = 10000.0) -> torch.Tensor:
    """Frequency table exactly as util.py:171-173 computes it (fp32 on the host)."""
    half = dim // 2
    return torch.exp(-math.log(max_period) * torch.arange(start=0, end=half, dtype=torch.float32) / half)
