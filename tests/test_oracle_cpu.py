"""CPU suite (`-m "not gpu"`): the oracle restatement against the golden vectors produced by the UNMODIFIED
reference (oracle/make_golden.py), the host-side logic (schedules, config plumbing, sharding, gloo
all-gather), and that the C-ABI library loads and exports every symbol include/fgdm_b200.h declares.
No compute entry point is called: there is no GPU here and the product has no CPU fallback."""
import os
import re
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
GOLD = os.path.join(ROOT, "tests", "golden")

import fgdm_oracle as orc  # noqa: E402
from fgdm_b200 import golden_cases as gc  # noqa: E402
from fgdm_b200 import synth  # noqa: E402


def gold(name):
    return np.load(os.path.join(GOLD, name + ".npz"))


def relmax(a, b):
    a, b = torch.as_tensor(a).double(), torch.as_tensor(b).double()
    return float((a - b).abs().max() / b.abs().max())


@pytest.fixture(scope="module")
def sd_fgdm():
    return synth.make_state_dict(gc.fgdm_param_shapes(), gc.FGDM_WEIGHT_SEED)


@pytest.fixture(scope="module")
def sd_control():
    return synth.make_state_dict(gc.control_param_shapes(), gc.CONTROL_WEIGHT_SEED)


def test_oracle_unet_matches_reference_golden(sd_fgdm):
    c = gc.unet_case()
    with torch.no_grad():
        eps = orc.unet_forward(sd_fgdm, c["x"], c["t"], c["ctx"], prefix="model.diffusion_model.")
    assert relmax(eps, gold("unet_fgdm")["eps"]) < 1e-4


def test_oracle_unet_pcond_matches_reference_golden(sd_fgdm):
    """UNetModel.forward(pcond=...) (openaimodel.py:836-844): adapter fed by another factor's latent."""
    c = gc.pcond_case()
    with torch.no_grad():
        eps = orc.unet_forward(sd_fgdm, c["x"], c["t"], c["ctx"], prefix="model.diffusion_model.", pcond=c["pcond"])
        plain = orc.unet_forward(sd_fgdm, c["x"], c["t"], c["ctx"], prefix="model.diffusion_model.")
    assert relmax(eps, gold("unet_pcond")["eps"]) < 1e-4
    assert relmax(plain, gold("unet_pcond")["eps"]) > 1e-2      # the pcond branch really changes the result


def test_oracle_clip_text_matches_transformers_golden():
    """FrozenCLIPEmbedder.forward (modules.py:137-163) = transformers CLIPTextModel(...).last_hidden_state."""
    sd = gc.clip_state_dict()
    with torch.no_grad():
        z = orc.clip_text_forward(sd, gc.clip_case()["ids"])
    g = gold("clip_text")["z"]
    assert z.shape == (2, 77, 768)
    assert relmax(z, g) < 1e-4
    # causal: the hidden state of position i does not depend on later tokens
    ids2 = gc.clip_case()["ids"].clone()
    ids2[:, 5:] = 1234
    with torch.no_grad():
        z2 = orc.clip_text_forward(sd, ids2)
    assert torch.allclose(z2[:, :5], z[:, :5], atol=1e-5) and not torch.allclose(z2[:, 5:], z[:, 5:], atol=1e-3)


def test_oracle_dpm_solver_matches_reference_golden(sd_fgdm):
    """DPMSolverSampler.sample (dpm_solver/sampler.py: DPM-Solver++ 2M, float model timesteps) vs the reference."""
    c, g = gc.dpm_case(), gold("dpm_fgdm")
    rec = []
    with torch.no_grad():
        x0 = orc.dpm_solver_sample(
            lambda x, t, cc: orc.unet_forward(sd_fgdm, x, t, cc, prefix="model.diffusion_model."), c["S"],
            (c["B"],) + tuple(c["shape"]), c["c"], c["uc"], c["scale"], c["x_T"],
            orc.make_schedule_buffers()["alphas_cumprod"], record=rec)
    assert relmax(x0, g["x0"]) < 1e-3
    assert len(rec) == c["S"] and relmax(rec[-1], g["x0"]) < 1e-3


def test_oracle_plms_loop_matches_reference_golden(sd_fgdm):
    """PLMSSampler.sample (plms.py): Euler start + Adams-Bashforth orders 2-4, vs the reference's own sampler."""
    c, g = gc.plms_case(), gold("plms_fgdm")
    rec = []
    with torch.no_grad():
        x0, inter = orc.plms_sample(
            lambda x, t, cc: orc.unet_forward(sd_fgdm, x, t, cc, prefix="model.diffusion_model."), c["S"],
            (c["B"],) + tuple(c["shape"]), c["c"], c["uc"], c["scale"], c["x_T"],
            orc.make_schedule_buffers()["alphas_cumprod"], record=rec)
    assert relmax(x0, g["x0"]) < 1e-3
    assert len(rec) == g["steps"].shape[0]
    for i, (_, xp) in enumerate(rec):
        assert relmax(xp, g["steps"][i]) < 1e-3
    assert len(inter["x_inter"]) == int(g["n_inter"])


def test_oracle_ddim_loop_matches_reference_golden(sd_fgdm):
    c, g = gc.ddim_case(), gold("ddim_fgdm")
    rec = []
    with torch.no_grad():
        x0, inter = orc.ddim_sample(
            lambda x, t, cc: orc.unet_forward(sd_fgdm, x, t, cc, prefix="model.diffusion_model."), c["S"],
            (c["B"],) + tuple(c["shape"]), c["c"], c["uc"], c["scale"], c["x_T"],
            orc.make_schedule_buffers()["alphas_cumprod"], record=rec)
    assert relmax(x0, g["x0"]) < 1e-3
    for i, (_, xp) in enumerate(rec):
        assert relmax(xp, g["steps"][i]) < 1e-3
    assert len(inter["x_inter"]) == int(g["n_inter"])
    assert relmax(inter["pred_x0"][-1], g["pred_x0_last"]) < 1e-3


def test_oracle_vae_decode_matches_reference_golden(sd_fgdm):
    with torch.no_grad():
        img = orc.vae_decode(sd_fgdm, gc.vae_case()["z"])
    assert relmax(img, gold("vae_decode")["img"]) < 1e-4


def test_oracle_controlldm_matches_reference_golden(sd_control):
    c, g = gc.control_case(), gold("control_apply")
    with torch.no_grad():
        eps = orc.control_ldm_apply(sd_control, c["x"], c["t"], c["hint"], c["ctx"])
        assert relmax(eps, g["eps"]) < 1e-4
        ctrl = orc.controlnet_forward(sd_control, c["x"], c["hint"], c["t"], c["ctx"], prefix="control_model.")
        assert len(ctrl) == 13
        assert relmax(ctrl[0], g["ctrl0"]) < 1e-4 and relmax(ctrl[12], g["ctrl12"]) < 1e-4
        eps_mid = orc.control_ldm_apply(sd_control, c["x"], c["t"], c["hint"], c["ctx"], only_mid_control=True)
        assert relmax(eps_mid, g["eps_only_mid"]) < 1e-4


def test_oracle_handoff_bit_exact_vs_reference_golden():
    c, g = gc.handoff_case(), gold("handoff")
    u = orc.seg_image_to_u8(c["img"])
    assert np.array_equal(u, g["seg_u8"])
    assert np.array_equal(np.stack([orc.resize_u8_bilinear(i, c["dh"], c["dw"]) for i in u]), g["resized"])
    assert np.array_equal(np.stack([orc.resize_u8_bilinear(i, 80, 96) for i in u]), g["resized_odd"])
    assert np.array_equal(orc.final_image_to_u8(c["img"]), g["final_u8"])
    same = orc.resize_u8_bilinear(u[0], 64, 64)          # 512 -> 512 in the C2 chain is the identity
    assert np.array_equal(same, u[0])


def test_schedules_match_reference_golden():
    """Host logic of fgdm_b200.ddim.DDIMSampler.make_schedule and LatentDiffusion.register_schedule against the
    reference's own tables (ddim.py:26-55, util.py:46-74, ddpm.py:175-203), bit for bit."""
    from fgdm_b200.ddim import DDIMSampler
    from fgdm_b200.ldm import instantiate_from_config
    g = gold("schedules")
    model = instantiate_from_config(gc.fgdm_config())
    assert np.array_equal(model.betas.numpy(), g["betas"])
    assert np.array_equal(model.alphas_cumprod.numpy(), g["alphas_cumprod"])
    assert np.array_equal(model.alphas_cumprod_prev.numpy(), g["alphas_cumprod_prev"])
    s = DDIMSampler(model)
    for S in (10, 20, 50):
        s.make_schedule(S, ddim_eta=0.0, verbose=False)
        assert (s.ddim_timesteps == g[f"t{S}"]).all()
        assert np.array_equal(s.ddim_alphas.numpy(), g[f"a{S}"])
        assert np.array_equal(s.ddim_alphas_prev, g[f"ap{S}"])
        assert np.array_equal(s.ddim_sqrt_one_minus_alphas.numpy(), g[f"s1m{S}"])
        ct = s.coef_table
        assert ct.shape == (S, 5) and float(ct[:, 4].abs().max()) == 0.0
        a, ap = torch.tensor(g[f"a{S}"]), torch.tensor(g[f"ap{S}"], dtype=torch.float32)
        assert torch.equal(ct[:, 1], a.sqrt()) and torch.equal(ct[:, 2], ap.sqrt())
        assert torch.equal(ct[:, 3], (1.0 - ap).sqrt())
        o = orc.make_ddim_schedule(model.alphas_cumprod, S)
        assert (o["timesteps"] == g[f"t{S}"]).all()
    assert list(g["t50"][:3]) == [1, 21, 41] and list(g["t10"][:2]) == [1, 101]


def test_unsupported_sampler_options_raise():
    from fgdm_b200.ddim import DDIMSampler
    from fgdm_b200.ldm import instantiate_from_config
    s = DDIMSampler(instantiate_from_config(gc.fgdm_config()))
    c = torch.zeros(1, 77, 768)
    for kw in (dict(mask=torch.ones(1)), dict(quantize_x0=True), dict(score_corrector=object()),
               dict(noise_dropout=0.1), dict(inference_loss=True)):
        with pytest.raises(NotImplementedError):
            s.sample(S=2, batch_size=1, shape=[4, 8, 8], conditioning=c, verbose=False, **kw)


def test_reference_target_strings_resolve_to_our_classes():
    from fgdm_b200 import ldm, nets
    assert ldm.get_obj_from_str("ldm.modules.diffusionmodules.openaimodel.UNetModel") is nets.UNetModel
    assert ldm.get_obj_from_str("controlnet.cldm.cldm.ControlNet") is nets.ControlNet
    assert ldm.get_obj_from_str("controlnet.cldm.cldm.ControlLDM") is ldm.ControlLDM
    assert ldm.get_obj_from_str("ldm.models.autoencoder.AutoencoderKL") is nets.AutoencoderKL
    with pytest.raises(NotImplementedError):
        nets.UNetModel(use_spatial_transformer=True, context_dim=768, num_heads=8, num_classes=10)


def test_param_counts_and_synth_determinism():
    sh = gc.fgdm_param_shapes()
    unet = sum(int(np.prod(s)) for k, s in sh.items() if k.startswith("model.diffusion_model."))
    assert unet == 936_348_804                       # 936.35 M (UNet + Adapter)
    shc = gc.control_param_shapes()
    cn = sum(int(np.prod(s)) for k, s in shc.items() if k.startswith("control_model."))
    un = sum(int(np.prod(s)) for k, s in shc.items() if k.startswith("model.diffusion_model."))
    assert round(cn / 1e6, 1) == 361.3 and round(un / 1e6, 1) == 859.5
    small = {k: sh[k] for k in list(sh)[:40]}
    a, b = synth.make_state_dict(small, 5), synth.make_state_dict(small, 5)
    assert all(torch.equal(a[k], b[k]) for k in small)
    assert not torch.equal(a[list(small)[0]], synth.make_state_dict(small, 6)[list(small)[0]])


def test_c_abi_library_loads_and_exports_every_declared_symbol():
    from fgdm_b200 import _lib
    lib = _lib.load()
    hdr = open(os.path.join(ROOT, "include", "fgdm_b200.h")).read()
    declared = set(re.findall(r"\b(fgdm_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/fgdm_b200.h but not exported"
    assert declared == set(_lib.SIGNATURES) | {"fgdm_last_error", "fgdm_launch_count"}
    assert lib.fgdm_abi_version() == _lib.ABI_VERSION


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_product_fails_loudly_without_gpu():
    from fgdm_b200 import _lib
    from fgdm_b200.ldm import instantiate_from_config
    with pytest.raises(RuntimeError, match="no CUDA device|fgdm_device_check"):
        _lib.require_device()
    model = instantiate_from_config(gc.fgdm_config())
    with pytest.raises(RuntimeError):
        model.load_state_dict({})


def _gather_worker(rank, world, port, q):
    import torch.distributed as dist
    from fgdm_b200.pipeline import all_gather_images, shard_indices
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    idx = shard_indices(5, rank, world)
    img = torch.stack([torch.full((4, 4, 3), i, dtype=torch.uint8) for i in idx])
    if len(idx) < 3:                                  # all_gather needs equal shards: pad the short rank
        img = torch.cat([img, torch.full((3 - len(idx), 4, 4, 3), 255, dtype=torch.uint8)])
    out = all_gather_images(img)
    q.put((rank, idx, out[:, 0, 0, 0].tolist()))
    dist.destroy_process_group()


def test_prompt_sharding_and_all_gather_gloo_world2():
    """N>1 path on CPU: contiguous prompt shards per rank, one all-gather of the uint8 images (rank-major)."""
    import torch.multiprocessing as mp
    from fgdm_b200.pipeline import shard_indices
    assert shard_indices(256, 3, 8) == list(range(96, 128))
    assert sum((shard_indices(10, r, 4) for r in range(4)), []) == list(range(10))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_gather_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0][1] == [0, 1, 2] and res[1][1] == [3, 4]
    assert res[0][2] == res[1][2] == [0, 1, 2, 3, 4, 255]


def test_upsample_subpixel_weights_match_interpolate_then_conv():
    """Host logic of the sub-pixel Upsample (layers.Upsample.load, include/fgdm_b200.h FGDM_CONV_UP2_P*): the four phase
    weight sets, applied as 4-tap convolutions at the offsets api.cu programs (dh = a - 1 + r, dw = b - 1 + s), reproduce
    F.conv2d(F.interpolate(x, scale_factor=2, mode="nearest"), w, b, padding=1) (openaimodel.py:118-130) up to the one
    bf16 rounding of the summed weights."""
    import torch.nn.functional as F
    from fgdm_b200 import layers
    torch.manual_seed(3)
    c, nb, h, w = 8, 2, 5, 7
    up = layers.Upsample(c)
    sd = {"conv.weight": torch.randn(c, c, 3, 3), "conv.bias": torch.randn(c)}
    up.load(sd, "", "cpu")
    assert len(up.w_phase) == 4 and all(tuple(p.shape) == (c, 4 * c) and p.dtype == torch.bfloat16 for p in up.w_phase)
    x = torch.randn(nb, c, h, w)
    ref = F.conv2d(F.interpolate(x, scale_factor=2, mode="nearest"), sd["conv.weight"], sd["conv.bias"], padding=1)
    xp = F.pad(x, (1, 1, 1, 1))
    out = torch.empty_like(ref)
    for ph in range(4):
        a, b = ph >> 1, ph & 1
        wp = up.w_phase[ph].float().reshape(c, 4, c)                       # [O][(r, s)][I]
        acc = torch.zeros(nb, c, h, w)
        for t in range(4):
            dh, dw = a - 1 + (t >> 1), b - 1 + (t & 1)
            acc += torch.einsum("oi,nihw->nohw", wp[:, t], xp[:, :, 1 + dh:1 + dh + h, 1 + dw:1 + dw + w])
        out[:, :, a::2, b::2] = acc + sd["conv.bias"][None, :, None, None]
    err = float((out - ref).abs().max() / ref.abs().max())
    assert err < 1e-2, err                                               # bf16 weights: ~2^-9 relative
    # and exactly, with the un-rounded sums
    w32 = sd["conv.weight"]
    sets = (((0,), (1, 2)), ((0, 1), (2,)))
    for ph in range(4):
        a, b = ph >> 1, ph & 1
        exact = torch.stack([sum(w32[:, :, ky, kx] for ky in sets[a][r] for kx in sets[b][s]) for r in range(2) for s in range(2)], 1)
        assert torch.allclose(up.w_phase[ph].float().reshape(c, 4, c), exact, rtol=1e-2, atol=1e-2)


def test_upsample_path_selection_and_call_arguments(monkeypatch):
    """Host logic of layers.Upsample.forward with the kernels stubbed: the sub-pixel form issues the four phase GEMMs
    (conv = FGDM_CONV_UP2_P0 + phase, shared packed output with ldo = 4c and col_off = phase * c) and one pixel shuffle; small
    launches, untileable low-resolution geometries and FGDM_UP_PHASE=0 take the nearest-2x + conv3x3 form."""
    from fgdm_b200 import _lib, layers, ops
    c = 64
    up = layers.Upsample(c)
    up.load({"conv.weight": torch.zeros(c, c, 3, 3), "conv.bias": torch.zeros(c)}, "", "cpu")
    calls = []

    def fake_gemm(a0, weight, n, **kw):
        calls.append(("gemm", kw.get("conv", 0), tuple(a0.shape), tuple(weight.shape), kw.get("ldo"), kw.get("col_off", 0)))
        out = kw.get("out")
        return out if out is not None else torch.zeros(a0.shape[0] * a0.shape[1] * a0.shape[2], n, dtype=torch.bfloat16)

    def fake_resample(x, mode):
        calls.append(("resample", mode, tuple(x.shape)))
        nb, h, w, ch = x.shape
        return torch.zeros(nb, 2 * h, 2 * w, ch // 4 if mode == 3 else ch, dtype=torch.bfloat16)

    monkeypatch.setattr(ops, "conv_gemm", fake_gemm)
    monkeypatch.setattr(ops, "resample", fake_resample)
    monkeypatch.setattr(layers, "UP_PHASE", True)
    monkeypatch.setattr(layers, "UP_PHASE_MIN_ROWS", 4096)
    x = torch.zeros(16, 16, 16, c, dtype=torch.bfloat16)                  # 4096 input pixels: sub-pixel form
    y = up.forward(x)
    assert tuple(y.shape) == (16, 32, 32, c)
    assert [k[0] for k in calls] == ["gemm"] * 4 + ["resample"]
    for ph in range(4):
        kind, conv, ashape, wshape, ldo, col_off = calls[ph]
        assert conv == _lib.CONV_UP2_P0 + ph and ashape == (16, 16, 16, c) and wshape == (c, 4 * c)
        assert ldo == 4 * c and col_off == ph * c
    assert calls[4] == ("resample", 3, (16, 16, 16, 4 * c))
    for xs, flag in (((2, 16, 16, c), True), ((16, 24, 40, c), True), ((16, 16, 16, c), False)):   # small / untileable / off
        calls.clear()
        monkeypatch.setattr(layers, "UP_PHASE", flag)
        y = up.forward(torch.zeros(*xs, dtype=torch.bfloat16))
        assert [k[:2] for k in calls] == [("resample", 0), ("gemm", _lib.CONV_3X3)], calls
        assert tuple(y.shape) == (xs[0], 2 * xs[1], 2 * xs[2], c)
    assert layers._tile_geometry_ok(16, 32, 32) and layers._tile_geometry_ok(8, 256, 256) and not layers._tile_geometry_ok(1, 96, 192)
