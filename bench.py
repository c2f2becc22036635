"""bench.py — FG-DM seg->image chained sampling throughput on B200 (BASELINE.json configs[1]).

  python bench.py --gpus N --steps K --warmup W            # our arm (one process per GPU under torchrun)
  python bench.py --impl reference --steps K --warmup W    # the reference's fp32 PyTorch CPU path (oracle port)

One "step" = one pass of the hot path over one batch: B=8 prompts per GPU through the whole factor chain
(seg factor 64x64 latent x 50 DDIM steps, CFG 7.5 -> VAE decode 512^2 -> uint8 hint -> ControlNet image
factor x 50 steps, CFG 7.5 -> VAE decode -> uint8 images; N>1: + NCCL all-gather of the images).
Prints ONE JSON line (rank 0). Weights are random-init of the SD-v1.4 / FG-DM / ControlNet architecture,
inputs synthetic (fgdm_b200.synth) — there is no network for checkpoints or datasets.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

# algorithmic FLOPs (2*MAC of conv + linear + attention matmuls, FlopCounterMode on the
# reference modules) per image of the C2 chain: 50 * 1717.4 G (seg CFG step) + 2514.5 G (decode) +
# 50 * 2173.2 G (image CFG step) + 2514.5 G (decode).
FLOP_PER_IMAGE = 50 * 1717.4e9 + 2514.5e9 + 50 * 2173.2e9 + 2514.5e9
METRIC = "images/sec (512^2, 50-step DDIM, CFG, FG-DM seg->image chain)"
WORKLOAD = "FG-DM seg->image chained sampling 512x512, batch 8 per GPU, 50+50-step DDIM, CFG 7.5, bf16"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sust=d["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx = max(mx, float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        # median over the samples under load (upper half: the first samples may precede the first kernel)
        med = sm[len(sm) // 2] if sm else None
        return {"sm_mhz": med, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------
# CPU arm: the reference's fp32 PyTorch path (the reference is pure Python/torch and cannot travel to the
# GPU box; oracle/fgdm_oracle.py restates it and is pinned to the reference by tests/golden). This is the
# only place bench.py executes oracle/ — as the timed CPU baseline, never as the product.
# ---------------------------------------------------------------------------------------------------
class CpuChain:
    """Bounded sample of the C2 workload at B=1 on the host cores: one CFG DDIM step of the seg factor and
    one of the image factor at 64x64 latents (each is 1/50 of its loop), plus one 512^2 VAE decode.
    images/s = 1 / (50 t_seg + 50 t_img + 2 t_dec)."""

    def __init__(self, sd_seg=None, sd_img=None):
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import fgdm_oracle as orc
        from fgdm_b200 import golden_cases as gc, synth
        self.orc = orc
        self.cores = os.cpu_count() or 1
        torch.set_num_threads(self.cores)
        if sd_seg is None:
            sd_seg = synth.make_state_dict(gc.fgdm_param_shapes(), gc.FGDM_WEIGHT_SEED)
        if sd_img is None:
            sd_img = synth.make_state_dict(gc.control_param_shapes(), gc.CONTROL_WEIGHT_SEED)
        self.sd_seg, self.sd_img = sd_seg, sd_img
        inp = synth.prompt_inputs([0], (64, 64))
        self.inp = inp
        self.hint = torch.rand((1, 3, 512, 512), generator=torch.Generator().manual_seed(5000))
        self.ac = orc.make_schedule_buffers()["alphas_cumprod"]
        self.t_dec = None

    @torch.no_grad()
    def decode_once(self):
        t0 = time.perf_counter()
        self.orc.vae_decode(self.sd_seg, self.inp["xT_seg"] * 0.18215)
        self.t_dec = time.perf_counter() - t0
        return self.t_dec

    @torch.no_grad()
    def step(self):
        """-> (t_seg, t_img) seconds for one CFG DDIM step of each factor (S=50 schedule, first step)."""
        orc, inp = self.orc, self.inp
        t0 = time.perf_counter()
        orc.ddim_sample(lambda x, t, c: orc.unet_forward(self.sd_seg, x, t, c, prefix="model.diffusion_model."), 50,
                        (1, 4, 64, 64), inp["c"], inp["uc"], 7.5, inp["xT_seg"], self.ac, max_steps=1)
        t1 = time.perf_counter()
        # ddim_hacked.py:190-191: two sequential apply_model calls (cond, then uncond), both run here
        orc.ddim_sample(lambda x, t, c: orc.control_ldm_apply(self.sd_img, x, t, self.hint, c), 50, (1, 4, 64, 64),
                        inp["c2"], inp["uc2"], 7.5, inp["xT_img"], self.ac, sequential_cfg=True, max_steps=1)
        t2 = time.perf_counter()
        return t1 - t0, t2 - t1

    @torch.no_grad()
    def c1(self, steps=2):
        """BASELINE configs[0] (C1): seg-factor DDIM on CPU, 1 prompt, 256x256 (32x32 latent), 10 steps, CFG 7.5.
        Times the first `steps` of the 10-step loop and scales to 10 (every step costs the same)."""
        orc = self.orc
        from fgdm_b200 import synth
        inp = synth.prompt_inputs([0], (32, 32))
        t0 = time.perf_counter()
        orc.ddim_sample(lambda x, t, c: orc.unet_forward(self.sd_seg, x, t, c, prefix="model.diffusion_model."), 10,
                        (1, 4, 32, 32), inp["c"], inp["uc"], 7.5, inp["xT_seg"], self.ac, max_steps=steps)
        dt = time.perf_counter() - t0
        return {"workload": "C1: FG-DM seg factor, 1 prompt, 32x32 latent, 10-step DDIM, CFG 7.5, fp32 CPU (oracle port)",
                "seconds_per_10_steps": 10.0 * dt / steps, "unet_cfg_step_s": dt / steps, "timed_steps": steps,
                "cores": self.cores, "kind": "port"}

    def images_per_s(self, t_seg, t_img):
        return 1.0 / (50.0 * t_seg + 50.0 * t_img + 2.0 * self.t_dec)

    SAMPLE = ("B=1: one CFG DDIM step of the seg factor (one 2-row UNet call) + one CFG step of the image factor (two "
              "sequential ControlLDM.apply_model calls) at 64x64 latents (each 1/50 of its loop) + one 512^2 VAE decode; "
              "images/s = 1/(50 t_seg + 50 t_img + 2 t_dec)")


def run_reference_arm(args, rank):
    if rank != 0:
        return
    cpu = CpuChain()
    cpu.decode_once()
    for _ in range(args.warmup):
        cpu.step()
    ts, ti = 0.0, 0.0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        a, b = cpu.step()
        ts += a
        ti += b
    wall = time.perf_counter() - t0
    v = cpu.images_per_s(ts / args.steps, ti / args.steps)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "images/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * wall / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "timing": "host wall clock; step = bounded sample", "sample": cpu.SAMPLE},
            "cpu_baseline": {"value": v, "unit": "images/s", "cores": cpu.cores, "kind": "port", "sample": cpu.SAMPLE},
            "e2e": {"value": v, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "cpu_c1": cpu.c1()}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------
NOMINAL_BF16_TFLOPS = 2250.0      # B200 dense bf16 (north_star's "B200 dense bf16 peak"); reported next to the measured peaks
NOMINAL_HBM_GBS = 7700.0


def kernel_rooflines(pipe, dev_in, pk, S):
    """Per-kernel pass: ONE eager (non-graph) CFG DDIM step of each factor + one VAE decode, each recorded on its own,
    every launch bracketed by CUDA events on the launching stream; algorithmic FLOPs/bytes come from the launch shapes.
    Families are then weighted as the CHAIN runs them: S x (seg step) + S x (image step) + 2 x (decode)."""
    from fgdm_b200 import ops
    B = dev_in["xT_seg"].shape[0]
    seg, img = pipe.seg_model, pipe.img_model
    t = torch.full((2 * B,), 981, device=pipe.device, dtype=torch.int64)
    ctx = torch.cat([dev_in["uc"], dev_in["c"]])
    hint = torch.rand((B, 3, 512, 512), device=pipe.device)
    unet, cn = img.model.diffusion_model, img.control_model
    kv_s = seg.model.diffusion_model.project_context(ctx)
    kv_u, kv_c, gh = unet.project_context(ctx), cn.project_context(ctx), cn.encode_hint(hint)
    # the same calls the captured sampler step makes (ddim._StepGraph._eps): shared CFG head, time-embedding vectors
    # from the per-schedule table
    from fgdm_b200.ddim import EMB_TABLES, SHARE_CFG_HEAD as share
    es = eu = ec = None
    if EMB_TABLES:
        es = seg.model.diffusion_model.emb_table(t[:1]).expand(2 * B, -1)
        eu, ec = unet.emb_table(t[:1]).expand(2 * B, -1), cn.emb_table(t[:1]).expand(2 * B, -1)
    parts = {"seg_step": lambda: seg.model.diffusion_model.forward(dev_in["xT_seg"], t, kv=kv_s, rows=2 * B,
                                                                   shared_halves=share, emb_all=es),
             "img_step": lambda: img.control_step(dev_in["xT_img"], t, kv_u, kv_c, gh, 2 * B, shared_halves=share,
                                                  emb_unet=eu, emb_ctrl=ec),
             "decode": lambda: seg.decode_first_stage(dev_in["xT_seg"] * 0.18215)}
    weight = {"seg_step": S, "img_step": S, "decode": 2}
    profs = {}
    import fgdm_b200.ldm as ldm_mod
    overlap = ldm_mod.OVERLAP_CONTROL
    ldm_mod.OVERLAP_CONTROL = False      # per-kernel times need ONE stream: the captured step forks the ControlNet aside
    try:
        for name, fn in parts.items():
            fn()                                               # warm
            # let the host run ahead of the device: with every launch already queued, the event intervals are
            # kernel execution time, not Python launch latency
            torch.cuda._sleep(int(0.12 * 1.9e9))
            ops.profile_start()
            fn()
            profs[name] = ops.profile_stop()
    finally:
        ldm_mod.OVERLAP_CONTROL = overlap

    def families(select, weighted):
        fam = {}
        for part in select:
            wgt = weight[part] if weighted else 1
            for name, shapes in profs[part].items():
                f = fam.setdefault(name, dict(launches=0, ms=0.0, flops=0.0, bytes=0.0))
                for v in shapes.values():
                    f["launches"] += wgt * v[0]
                    f["ms"] += wgt * v[1]
                    f["flops"] += wgt * v[2]
                    f["bytes"] += wgt * v[3]
        return fam

    fam = families(parts, True)                            # the whole chain, per batch of B prompts
    total_ms = sum(f["ms"] for f in fam.values())
    top = max(fam, key=lambda k: fam[k]["ms"])
    f = fam[top]
    tensor_bound = f["flops"] / max(f["bytes"], 1.0) > 100.0
    if tensor_bound:
        ach = f["flops"] / (f["ms"] * 1e-3) / 1e12
        roof = {"kernel": top, "bound": "tensor", "achieved": ach, "peak": pk["tf_sust"], "unit": "TFLOP/s",
                "frac": ach / pk["tf_sust"], "traffic": None, "frac_of_nominal_dense_bf16": ach / NOMINAL_BF16_TFLOPS}
    else:
        ach = f["bytes"] / (f["ms"] * 1e-3) / 1e9
        roof = {"kernel": top, "bound": "hbm", "achieved": ach, "peak": pk["hbm"], "unit": "GB/s",
                "frac": ach / pk["hbm"], "traffic": None, "frac_of_nominal_hbm": ach / NOMINAL_HBM_GBS}
    try:      # DRAM traffic of the family's dominant shape from the committed ncu --set full capture (per launch of
        # THAT shape; `achieved` is the family average) - bench.py itself never runs under a profiler
        tr = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r02_traffic.json")))
        if tr.get("kernel") == top:
            roof["traffic"] = tr["dram_bytes_per_launch"]
            roof["traffic_of"] = {k: tr[k] for k in ("shape", "algorithmic_bytes_per_launch", "source")}
    except (OSError, ValueError, KeyError):
        pass
    roof.update(peak_source=pk["src"] + " (sustained)" if tensor_bound else pk["src"], launches=f["launches"],
                avg_launch_us=1e3 * f["ms"] / f["launches"], share_of_chain=f["ms"] / total_ms,
                weighting=f"chain: {S} x seg CFG step + {S} x image CFG step + 2 x VAE decode (per batch of {B} prompts)",
                how="1 eager CFG step of each factor + 1 decode, CUDA events around every launch")

    def table_of(fm):
        tot = sum(v["ms"] for v in fm.values())
        return {k: {"launches": v["launches"], "ms": round(v["ms"], 3), "share": round(v["ms"] / tot, 4),
                    "tflops": round(v["flops"] / (v["ms"] * 1e-3) / 1e12, 1) if v["flops"] else None,
                    "frac_of_sustained_bf16": round(v["flops"] / (v["ms"] * 1e-3) / 1e12 / pk["tf_sust"], 3) if v["flops"] else None,
                    "gbs": round(v["bytes"] / (v["ms"] * 1e-3) / 1e9, 1),
                    "frac_of_hbm": round(v["bytes"] / (v["ms"] * 1e-3) / 1e9 / pk["hbm"], 3)}
                for k, v in sorted(fm.items(), key=lambda kv: -kv[1]["ms"])}

    # the UNet step (what the chain is made of) and the VAE decode, separately and unweighted
    tables = {"chain_weighted": table_of(fam), "unet_steps": table_of(families(("seg_step", "img_step"), False)),
              "vae_decode": table_of(families(("decode",), False))}
    flat = {}
    for part, prof in profs.items():
        for k, sh in prof.items():
            for sk, v in sh.items():
                flat[(part, k, sk)] = v
    executed = sum(v["flops"] for v in fam.values())      # chain-weighted flops of the kernels actually launched
    return roof, tables, flat, executed


def run_ours(args, rank, world, local_rank):
    import torch.distributed as dist
    from fgdm_b200 import _lib, ops, synth
    from fgdm_b200 import golden_cases as gc
    from fgdm_b200.ldm import instantiate_from_config
    from fgdm_b200.pipeline import FGDMPipeline, all_gather_images, shard_indices

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    _lib.require_device()                       # fails loudly without the sm_100a library/device
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    pk = peaks()
    B, S = args.batch, args.ddim_steps

    seg = instantiate_from_config(gc.fgdm_config())
    seg.device = dev
    sd_seg = synth.make_state_dict(seg.param_shapes(), gc.FGDM_WEIGHT_SEED)
    seg.load_state_dict(sd_seg)
    img = instantiate_from_config(gc.control_config())
    img.device = dev
    sd_img = synth.make_state_dict(img.param_shapes(), gc.CONTROL_WEIGHT_SEED)
    img.load_state_dict(sd_img)
    keep_sd = rank == 0 and world == 1 and not args.no_cpu_baseline
    if not keep_sd:
        sd_seg = sd_img = None
    pipe = FGDMPipeline(seg, img, seg_steps=S, img_steps=S, seg_scale=7.5, img_scale=7.5)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    def checksums(local_u8):
        """(sum over the WHOLE gathered uint8 tensor, per-rank sums) - every rank's images count, not rank 0's slice."""
        mine = local_u8.to(torch.int64).sum().reshape(1)
        if world == 1:
            return int(mine), [int(mine)]
        allsum = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allsum, mine)
        per = [int(x) for x in allsum]
        return sum(per), per

    # ------------------------------------------------------------------------------------------------
    # C3 (BASELINE configs[2]): a FIXED set of prompts sharded over the ranks (strong scaling), micro-batched
    # ------------------------------------------------------------------------------------------------
    def make_strong(P):
        idx = shard_indices(P, rank, world)
        assert P % world == 0 and len(idx) % B == 0, "prompts must split evenly into micro-batches on every rank"
        host = {k: v.pin_memory() for k, v in synth.prompt_inputs(idx, (64, 64)).items()}
        local = torch.empty((len(idx), 512, 512, 3), device=dev, dtype=torch.uint8)
        gathered = torch.empty((P, 512, 512, 3), device=dev, dtype=torch.uint8) if world > 1 else local
        host_out = torch.empty((len(idx), 512, 512, 3), dtype=torch.uint8).pin_memory()

        def step():
            for m0 in range(0, len(idx), B):                      # micro-batches of B prompts through the captured graphs
                d = {k: host[k][m0:m0 + B].to(dev, non_blocking=True) for k in host}
                local[m0:m0 + B].copy_(pipe(d["c"], d["uc"], d["xT_seg"], d["c2"], d["uc2"], d["xT_img"]))
            if world > 1:                                         # ONE in-place all-gather of every rank's shard
                dist.all_gather_into_tensor(gathered.view(-1), local.view(-1))
            host_out.copy_(local, non_blocking=True)
        h2d = sum(v.numel() * v.element_size() for v in host.values())
        return step, local, h2d, host_out.numel(), len(idx)

    strong_main = args.prompts > 0
    if strong_main:
        step_fn, local_out, h2d, d2h, n_local = make_strong(args.prompts)
        for _ in range(args.warmup):
            step_fn()
        clk = ClockSampler(local_rank)
        clk.start()
        n0 = ops.launch_count()
        ms = timed(step_fn, args.steps)
        launches = ops.launch_count() - n0
        clocks = clk.stop()
        value = args.prompts * args.steps / (ms * 1e-3)
        total, per_rank = checksums(local_out)
        if rank != 0:
            return
        line = {"metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": f"C3: FG-DM seg->image 512x512, {args.prompts} prompts sharded over {world} B200 "
                                       f"({n_local} per GPU, micro-batches of {B}), 50+50-step DDIM, CFG 7.5, images all-gathered",
                           "prompts": args.prompts, "prompts_per_gpu": n_local, "micro_batch": B,
                           "l2": "working set >> 126 MB L2; no flush needed"},
                "clocks": clocks,
                "e2e": {"value": value, "unit": "images/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "note": "the step itself is end to end: pinned-host inputs in, uint8 images out to pinned host"},
                "gpu_launches": launches, "image_checksum_all_ranks": total, "image_checksum_per_rank": per_rank}
        print(json.dumps(line), flush=True)
        return

    # ------------------------------------------------------------------------------------------------
    # C2 (BASELINE configs[1]) - the headline: B prompts per GPU (weak scaling)
    # ------------------------------------------------------------------------------------------------
    idx = [rank * B + i for i in range(B)]
    host = {k: v.pin_memory() for k, v in synth.prompt_inputs(idx, (64, 64)).items()}
    dev_in = {k: v.to(dev) for k, v in host.items()}
    host_out = torch.empty((B, 512, 512, 3), dtype=torch.uint8).pin_memory()
    last = {}

    def step_resident():
        out = pipe(dev_in["c"], dev_in["uc"], dev_in["xT_seg"], dev_in["c2"], dev_in["uc2"], dev_in["xT_img"])
        return all_gather_images(out)

    def step_e2e():
        d = {k: host[k].to(dev, non_blocking=True) for k in host}
        out = pipe(d["c"], d["uc"], d["xT_seg"], d["c2"], d["uc2"], d["xT_img"])
        last["local"] = out
        out = all_gather_images(out)
        last["gathered"] = out
        host_out.copy_(out[rank * B:(rank + 1) * B] if world > 1 else out, non_blocking=True)

    for _ in range(args.warmup):
        step_resident()
    clk = ClockSampler(local_rank)
    clk.start()
    n0 = ops.launch_count()
    ms = timed(step_resident, args.steps)
    launches = ops.launch_count() - n0
    clocks = clk.stop()
    for _ in range(min(args.warmup, 1)):
        step_e2e()
    ms_e2e = timed(step_e2e, args.steps)
    images = B * world * args.steps
    value = images / (ms * 1e-3)
    e2e_value = images / (ms_e2e * 1e-3)
    h2d = sum(v.numel() * v.element_size() for v in host.values())
    d2h = host_out.numel()
    checksum = int(host_out.to(torch.int64).sum())
    total, per_rank = checksums(last["local"])
    gathered_sum = int(last["gathered"].to(torch.int64).sum())      # what THIS rank received: must equal the total
    # C3 next to it: one pass over 256 prompts sharded across the ranks (same captured graphs, micro-batch B)
    c3 = None
    if args.c3_prompts and args.c3_prompts % (world * B) == 0:
        step_fn, local_out, c3_h2d, c3_d2h, n_local = make_strong(args.c3_prompts)
        ms_c3 = timed(step_fn, 1)
        c3_total, _ = checksums(local_out)
        c3 = {"workload": f"C3: {args.c3_prompts} prompts sharded over {world} GPU(s), {n_local} per GPU, micro-batches of {B}, "
                          "one all-gather; pinned-host in / out inside the timed pass",
              "scaling": "strong", "value": args.c3_prompts / (ms_c3 * 1e-3), "unit": "images/s", "ms": ms_c3, "passes": 1,
              "image_checksum_all_ranks": c3_total}

    if rank != 0:
        return
    # phase timings (CUDA events around graph replays / decodes): the "UNet step ms" half of the metric
    def ms_of(fn, n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n
    gs = next(iter(pipe.seg_sampler._graphs.values()))
    gi = next(iter(pipe.img_sampler._graphs.values()))
    phases = {"seg_unet_cfg_step_ms": ms_of(gs.graph.replay, 10), "img_controlnet_unet_cfg_step_ms": ms_of(gi.graph.replay, 10),
              "vae_decode_ms": ms_of(lambda: seg.decode_first_stage(dev_in["xT_seg"] * 0.18215), 3),
              "rows_per_step": 2 * B, "kernels_per_seg_step": gs.launches_per_step,
              "kernels_per_img_step": gi.launches_per_step}
    roof, tables, flat, executed_flops = kernel_rooflines(pipe, dev_in, pk, S)
    if args.dump_shapes:
        rows = [{"part": part, "kernel": k, "shape": sk, "launches": v[0], "ms": v[1], "flops": v[2], "bytes": v[3]}
                for (part, k, sk), v in flat.items()]
        json.dump(sorted(rows, key=lambda r: -r["ms"]), open(args.dump_shapes, "w"), indent=0)
    flop_img = S * 1717.4e9 + S * 2173.2e9 + 2 * 2514.5e9
    step_tf = flop_img * B / (ms / args.steps * 1e-3) / 1e12
    line = {"metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": WORKLOAD, "batch_per_gpu": B, "ddim_steps": [S, S], "cfg_scale": 7.5,
                       "latent": [4, 64, 64], "parallelism": f"dp{world} (independent samplers, uint8 all-gather)",
                       "l2": "working set (4.5 GB bf16 weights + activations) >> 126 MB L2; no flush needed",
                       "weights": "random-init SD-v1.4/FG-DM/ControlNet architecture (seeded)"},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "images/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps, "image_checksum": checksum,
                    "image_checksum_all_ranks": total, "image_checksum_per_rank": per_rank,
                    "gathered_checksum_rank0": gathered_sum},
            "gpu_launches": launches,
            "roofline": roof,
            # algorithmic = the reference's work (every CFG row through the whole network); executed = what the launched
            # kernels did (the context-free encoder head runs once per latent, ddim.SHARE_CFG_HEAD)
            "roofline_step": {"algorithmic_tflop_per_image": flop_img / 1e12, "achieved_tflops_per_gpu": step_tf,
                              "frac_of_bf16_sustained": step_tf / pk["tf_sust"],
                              "frac_of_nominal_dense_bf16": step_tf / NOMINAL_BF16_TFLOPS,
                              "executed_tflop_per_image": executed_flops / B / 1e12,
                              "executed_tflops_per_gpu": executed_flops / (ms / args.steps * 1e-3) / 1e12},
            "phases": phases, "kernels": tables["chain_weighted"], "kernels_unet_steps": tables["unet_steps"],
            "kernels_vae_decode": tables["vae_decode"], "c3_strong": c3}
    if keep_sd:
        cpu = CpuChain(sd_seg, sd_img)
        cpu.decode_once()
        cpu.step()
        ts, ti = cpu.step()
        line["cpu_baseline"] = {"value": cpu.images_per_s(ts, ti), "unit": "images/s", "cores": cpu.cores,
                                "kind": "port", "sample": cpu.SAMPLE,
                                "t_seg_step_s": ts, "t_img_step_s": ti, "t_decode_s": cpu.t_dec}
        line["cpu_c1"] = cpu.c1()
    else:
        line["cpu_baseline"] = None
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=8, help="prompts per GPU (BASELINE configs[1]: 8)")
    ap.add_argument("--ddim-steps", type=int, default=50)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--prompts", type=int, default=0,
                    help="C3 mode: this many prompts in TOTAL, sharded over the ranks (strong scaling), micro-batches of --batch")
    ap.add_argument("--c3-prompts", type=int, default=256,
                    help="default mode: also time ONE pass over this many prompts sharded over the ranks (0 = skip)")
    ap.add_argument("--dump-shapes", default="", help="write the per-launch-shape table of the roofline pass here")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference_arm(args, rank)
        return
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit("launch with: python -m torch.distributed.run --nnodes=1 --nproc-per-node N "
                         "--master-addr 127.0.0.1 --master-port P bench.py --gpus N ...")
    run_ours(args, rank, world, local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
