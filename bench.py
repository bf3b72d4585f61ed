#!/usr/bin/env python3
"""Headline benchmark: ZeroscopeXL denoise steps/s (BASELINE.json `metric`) + Merkle commit GB/s.

    python bench.py --gpus N --steps K --warmup W            (N>1: launched per rank by torch.distributed.run)
    python bench.py --impl reference ...                      (CPU arm: the oracle port of the reference's path)

A "step" = ONE scheduler timestep advanced for the WHOLE T-frame video: every frame window of the plan goes through
one UNet3D forward (CFG batch 2) and one fused CFG+DDIM(+context) kernel on the rank that owns it
(SURVEY.md §8d). Workload: BASELINE.json configs[2] — ZeroscopeXL 1024x576, 24 frames, fsdp_chunked_coherent
(hybrid_ctx: context injection + ramp smoothing), random-init weights, synthetic prompt embeddings and latents.
Window plan = the reference's own planner for the world size (fsdp_chunked_coherent.py:149-177); at N=8 its defaults do
not terminate, so N=8 runs `--chunk_size 4 --overlap 1` and says so in `config.plan`.
Total work is fixed as N grows -> "scaling": "strong".
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

HEIGHT, WIDTH, FRAMES, TOTAL_STEPS = 576, 1024, 24, 50
GUIDANCE, CTX_W = 7.5, 0.35
METRIC = "ZeroscopeXL denoise steps/sec (1024x576, 24 frames, fsdp_chunked_coherent hybrid_ctx)"


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_burst": d["bf16_tflops"], "bf16_sustained": d["bf16_tflops_sustained"],
                "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm_gbs": 6650.0, "bf16_burst": 1590.0, "bf16_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}


def plan_for(world):
    """(chunk_size, overlap, note) handed to the coherent planner."""
    if world >= 8:
        return 4, 1, "explicit --chunk_size 4 --overlap 1 (reference defaults loop forever at W=8)"
    if world == 4:
        return 8, 2, ("explicit --chunk_size 8 --overlap 2 = one 8-frame window per rank (what chunk_only.py plans at "
                      "W=4); the coherent script's own default at W=4 is 12 windows of 6 frames, 3 per rank, 3x redundant: "
                      "timed too and reported under reference_default_plan")
    return 0, 4, "reference defaults (--chunk_size 0 --overlap 4)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown," \
        "clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            f = [x.strip() for x in line.split(",")]
            if len(f) >= 8:
                self.rows.append(f)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if r[2].replace(".", "").isdigit()]
        pw = [float(r[3]) for r in self.rows if r[3].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[4 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(self.rows), "reasons": reasons}


# ---------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference's path on the host cores (cpu_baseline leg and --impl reference)
# ---------------------------------------------------------------------------------------------------------------
def unet_flops(F, h, w, B=2):
    """Algorithmic FLOPs of one UNet3D forward (same formula as b2v_unet3d_flops), host-side."""
    ch, xd, TL, hd = (320, 640, 1280, 1280), 1024, 77, 64
    N = B * F
    fl = 0.0

    def conv(M, ci, co, k=9):
        return 2.0 * M * ci * co * k

    def block(M, dim, cross, seq, nbatch):
        f = 2.0 * M * dim * dim * 3 + 2.0 * M * dim * dim            # qkv1 + o1
        f += 4.0 * nbatch * (dim // hd) * seq * seq * hd
        if cross:
            f += 2.0 * M * dim * dim + 2.0 * B * TL * xd * 2 * dim + 2.0 * M * dim * dim
            f += 4.0 * nbatch * (dim // hd) * seq * TL * hd
        else:
            f += 2.0 * M * dim * dim * 3 + 2.0 * M * dim * dim
            f += 4.0 * nbatch * (dim // hd) * seq * seq * hd
        f += 2.0 * M * dim * 8 * dim + 2.0 * M * 4 * dim * dim        # GEGLU FF
        return f

    def spatial(M, C, S):
        return 2 * 2.0 * M * C * C + block(M, C, True, S, N)

    def temporal(M, C, inner, S):
        return 2 * 2.0 * M * C * inner + block(M, inner, False, F, B * S)

    def resnet(M, ci, co):
        return conv(M, ci, co) + conv(M, co, co) + (conv(M, ci, co, 1) if ci != co else 0.0)

    def tconv(M, C):
        return 4 * 2.0 * M * C * 3 * C

    hh, ww = h, w
    M = N * hh * ww
    fl += conv(M, 4, ch[0]) + temporal(M, ch[0], 512, hh * ww)
    prev = ch[0]
    dims = []
    for i, c in enumerate(ch):
        M = N * hh * ww
        for j in range(2):
            fl += resnet(M, prev if j == 0 else c, c) + tconv(M, c)
            if i < 3:
                fl += spatial(M, c, hh * ww) + temporal(M, c, c, hh * ww)
        dims.append((hh, ww))
        if i < 3:
            hh, ww = (hh - 1) // 2 + 1, (ww - 1) // 2 + 1
            fl += conv(N * hh * ww, c, c)
        prev = c
    M = N * hh * ww
    fl += 2 * resnet(M, ch[3], ch[3]) + 2 * tconv(M, ch[3]) + spatial(M, ch[3], hh * ww) + temporal(M, ch[3], ch[3], hh * ww)
    rev = ch[::-1]
    prev_out = rev[0]
    for i, c in enumerate(rev):
        hh, ww = dims[3 - i]
        M = N * hh * ww
        inc = rev[min(i + 1, 3)]
        for j in range(3):
            skip = inc if j == 2 else c
            fl += resnet(M, (prev_out if j == 0 else c) + skip, c) + tconv(M, c)
            if i > 0:
                fl += spatial(M, c, hh * ww) + temporal(M, c, c, hh * ww)
        if i < 3:
            fl += conv(N * 4 * hh * ww, c, c)
        prev_out = c
    fl += conv(N * h * w, ch[0], 4)
    return fl


def video_step_flops(ranges, h, w):
    return sum(unet_flops(e - s, h, w) for s, e in ranges)


class CpuOracleArm:
    """The reference's denoise step restated on the CPU (oracle/unet3d_ref.py + oracle/ddim_ref.py), all host threads:
    UNet forward of one frame window (CFG batch 2) + CFG combine + DDIM update, at the REAL window shapes of the plan.
    Attention goes through F.scaled_dot_product_attention (diffusers' default processor; same values as the explicit
    softmax the tests use, which would need 54 GB for the 16-frame window's score tensor)."""

    def __init__(self):
        import torch
        from oracle import unet3d_ref as R
        from oracle.ddim_ref import pipeline_scheduler
        self.torch, self.R = torch, R
        R.Attention.use_sdpa = True
        self.cores = os.cpu_count() or 1
        torch.set_num_threads(self.cores)
        self.net = R.random_init_(R.UNet3DConditionModel(R.ZEROSCOPE), seed=1234).eval()
        self.sched = pipeline_scheduler()
        self.sched.set_timesteps(TOTAL_STEPS)
        g = torch.Generator().manual_seed(1)
        self.emb = torch.randn(2, 77, 1024, generator=g)

    def window_step(self, F, h, w, t):
        """Seconds for one denoise step of one F-frame window."""
        torch = self.torch
        g = torch.Generator().manual_seed(0)
        lat = torch.randn(1, 4, F, h, w, generator=g)
        t0 = time.perf_counter()
        ctx = lat.mean(dim=2, keepdim=True)              # fsdp_chunked_coherent.py:112-118 (stand-in: this window's mean)
        with torch.no_grad():
            x = torch.cat([lat + CTX_W * ctx] * 2)      # fsdp_chunked_coherent.py:134-137 hybrid_ctx injection
            noise = self.net(x, t, self.emb)
            u, c = noise.chunk(2)
            out = self.sched.step(u + GUIDANCE * (c - u), t, lat).prev_sample
        assert bool(torch.isfinite(out).all())
        return time.perf_counter() - t0


def cpu_line(args, ranges, h, w, as_reference):
    """CPU arm. as_reference (``--impl reference``): the whole-video step of the plan — every window at its real frame
    count, one after the other on the host — is MEASURED as often as a ~4 minute budget allows (at least once); nothing
    is extrapolated unless even one whole step does not fit, in which case the largest window is measured and the others
    follow by their FLOP ratio (flagged). Otherwise (the cpu_baseline leg of the GPU line, ~45 s budget): ONE real
    12-frame 72x128 window step (a window of the N = 1 plan) is measured and scaled to the whole-video step by the FLOP
    ratio (flagged; small windows run the host less efficiently, so the largest affordable one is used)."""
    arm = CpuOracleArm()
    ts = [int(t) for t in arm.sched.timesteps]
    step_flops = video_step_flops(ranges, h, w)
    t_cal = arm.window_step(1, 16, 16, ts[0])            # warm-up of the thread pool / primitive caches
    if not as_reference:
        F = min(12, max(e - s for s, e in ranges))       # round 2: 4 frames x 7.00 was 40 % off the measured arm
        dt = arm.window_step(F, h, w, ts[0])
        ratio = step_flops / unet_flops(F, h, w)
        value = 1.0 / (dt * ratio)
        return value, dt * ratio, {
            "value": value, "unit": "steps/s", "cores": arm.cores, "kind": "port", "extrapolated": True,
            "measured_steps": 0, "measured_s": dt,
            "sample": "oracle port (torch fp32 CPU restatement of diffusers UNet3D + DDIM; diffusers not installable "
                      "here): ONE denoise step of a real %d-frame %dx%d window (CFG batch 2, %.1f TFLOP) measured in "
                      "%.1f s, scaled to the whole-video step (%.1f TFLOP) by the algorithmic FLOP ratio %.2f"
                      % (F, h, w, unet_flops(F, h, w) / 1e12, dt, step_flops / 1e12, ratio)}
    budget = float(os.getenv("B2V_REF_BUDGET_S", "240"))
    lens = sorted({e - s for s, e in ranges}, reverse=True)
    # probe: one frame at the real latent size predicts the rate (only used to decide what fits the budget)
    t_probe = arm.window_step(1, h, w, ts[0])
    pred_step = t_probe * step_flops / unet_flops(1, h, w)
    times, measured_steps, extrapolated = {}, 0, False
    if pred_step <= budget * 1.25:
        n_steps = max(1, min(args.steps, int(budget // max(pred_step, 1e-9))))
        per = []
        for i in range(n_steps):
            t = ts[i % len(ts)]
            tot = 0.0
            for s_, e_ in ranges:
                dt = arm.window_step(e_ - s_, h, w, t)
                times.setdefault(e_ - s_, []).append(dt)
                tot += dt
            per.append(tot)
        measured_steps = n_steps
        step_s = sum(per) / len(per)
    else:
        extrapolated = True
        per_tflop = t_probe / unet_flops(1, h, w)
        cands = [f for f in lens + [8, 4, 2, 1] if f <= lens[0] and per_tflop * unet_flops(f, h, w) <= budget]
        F = cands[0] if cands else 1
        dt = arm.window_step(F, h, w, ts[0])
        times[F] = [dt]
        step_s = dt * step_flops / unet_flops(F, h, w)
    value = 1.0 / step_s
    measured_s = sum(sum(v) for v in times.values())
    sample = ("oracle port (torch fp32 CPU restatement; diffusers not installable here). Whole-video step = windows %s "
              "run one after the other on the host at their real shapes (%dx%d latents, CFG batch 2): %s; %s"
              % (ranges, h, w,
                 ", ".join("%d-frame window %.1f s" % (k, sum(v) / len(v)) for k, v in sorted(times.items(), reverse=True)),
                 ("%d whole step(s) measured, nothing extrapolated" % measured_steps) if not extrapolated else
                 "a whole step does not fit the budget: measured window scaled by the FLOP ratio"))
    return value, step_s, {"value": value, "unit": "steps/s", "cores": arm.cores, "kind": "port",
                           "extrapolated": extrapolated, "measured_steps": measured_steps, "measured_s": measured_s,
                           "probe_1_frame_s": t_probe, "calibration_s": t_cal, "sample": sample}


def workload_config(plan, note, h, w):
    """The `config` both arms print, key for key (the driver compares them): the workload, not how an arm ran it."""
    return {"workload": "ZeroscopeXL 1024x576x24f fsdp_chunked_coherent hybrid_ctx, CFG 7.5, DDIM 50-step schedule",
            "plan": {"note": note, "chunk_size": plan.chunk_size, "overlap": plan.overlap, "windows": plan.ranges},
            "latent": [1, 4, FRAMES, h, w],
            "l2": "no flush: each step streams 2.8 GB of weights + >2 GB of activations (>> 126 MB L2)"}


# ---------------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------------
def gpu_main(args):
    import torch
    import torch.distributed as dist
    import b2v  # noqa: F401
    from b2v import ops, planner, proof
    from b2v.pipeline import VideoPipeline
    from b2v.strategies.common import CudaDenoiser, base_noise

    rank, world = int(os.getenv("RANK", 0)), int(os.getenv("WORLD_SIZE", 1))
    local = int(os.getenv("LOCAL_RANK", 0))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus %d must be launched with torch.distributed.run (one rank per GPU)" % args.gpus)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    peaks = measured_peaks()
    h, w = HEIGHT // 8, WIDTH // 8
    cs, ov, plan_note = plan_for(world)
    plan = planner.plan(FRAMES, world, cs, ov, variant="coherent")
    mine = plan.of_rank(rank)

    dtype = {"bf16": torch.bfloat16, "fp16": torch.float16}[args.dtype]
    pipe = VideoPipeline.from_random("cerspense/zeroscope_v2_XL", device=dev, seed=1234, dtype=dtype)
    pipe.scheduler.set_timesteps(TOTAL_STEPS, device="cuda")
    ts = pipe.scheduler.host_timesteps()
    toks = pipe.tokenizer(["a rocket in space, 4k", ""])
    emb = pipe.text_encoder(toks.input_ids)[0]
    cond, uncond = emb[:1], emb[1:]
    base = base_noise(4, FRAMES, h, w, dtype, dev)
    ctx = ops.context_mean(base)
    den = CudaDenoiser(pipe, uncond, cond, GUIDANCE, ctx=ctx, context_weight=CTX_W)
    unet, sched = pipe.unet, pipe.scheduler

    # per-window resident state: latent, UNet input (CFG-duplicated, context injected), eps buffer
    state = []
    for s, e in mine:
        lat = base[:, :, s:e].clone()
        x = ops.make_unet_input(lat, 2, ctx=ctx, ctx_w=CTX_W)
        state.append([lat, x, torch.empty_like(x)])

    def video_step(i):
        t = ts[i % len(ts)]
        for st in state:
            unet.forward_into(st[1], t, den.emb, st[2])
            st[0], st[1] = sched.fused_cfg_step(st[0], st[2], t, GUIDANCE, ctx=ctx, ctx_w=CTX_W, n_dup=2, xnext=st[1])

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    # the dominant kernel once BEFORE the load (clocks not yet pulled down by the power cap): reported next to the
    # in-load measurement the roofline uses
    cold = dominant_kernel_roofline(torch, ops, dev, peaks, mine, dtype=dtype) if rank == 0 else None
    for i in range(args.warmup):
        video_step(i)
    sync_all()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = ops.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        video_step(args.warmup + i)
    e1.record()
    sync_all()
    launches = ops.launch_count() - l0
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_ms = float(ms.item())
    clocks = sampler.stop() if rank == 0 else None
    value = args.steps / (total_ms / 1e3)

    # ---- W = 4 only: the coherent script's own default plan (3 windows of 6 frames per rank), same timing rules ----
    ref_default = None
    if world == 4:
        dplan = planner.plan(FRAMES, world, 0, 4, variant="coherent")
        main_state = state[:]
        del state[:]
        for s, e in dplan.of_rank(rank):
            lat = base[:, :, s:e].clone()
            x = ops.make_unet_input(lat, 2, ctx=ctx, ctx_w=CTX_W)
            state.append([lat, x, torch.empty_like(x)])
        for i in range(args.warmup):
            video_step(i)
        sync_all()
        d0, d1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        d0.record()
        for i in range(args.steps):
            video_step(args.warmup + i)
        d1.record()
        sync_all()
        dms = torch.tensor([d0.elapsed_time(d1)], device=dev, dtype=torch.float64)
        dist.all_reduce(dms, op=dist.ReduceOp.MAX)
        ref_default = {"value": args.steps / (float(dms.item()) / 1e3), "unit": "steps/s", "chunk_size": dplan.chunk_size,
                       "overlap": dplan.overlap, "windows": dplan.ranges, "windows_per_rank": len(dplan.of_rank(rank))}
        del state[:]
        state.extend(main_state)

    # ---- e2e: the same step through the public per-step API with HOST buffers (pinned), copies inside the timing ----
    host_lat = [st[0].to("cpu").pin_memory() for st in state]
    host_emb = den.emb.to("cpu").pin_memory()
    host_out = [torch.empty_like(hl).pin_memory() for hl in host_lat]

    def e2e_step(i):
        t = ts[i % len(ts)]
        embd = host_emb.to(dev, non_blocking=True)
        for k in range(len(state)):
            lat = host_lat[k].to(dev, non_blocking=True)
            x = ops.make_unet_input(lat, 2, ctx=ctx, ctx_w=CTX_W)
            eps2 = unet(x, t, encoder_hidden_states=embd).sample
            out, _ = sched.fused_cfg_step(lat, eps2, t, GUIDANCE)
            host_out[k].copy_(out, non_blocking=True)
        torch.cuda.synchronize(dev)  # the caller holds the step's result on the host

    e2e_step(0)
    sync_all()
    t0 = time.perf_counter()
    for i in range(args.steps):
        e2e_step(i)
    sync_all()
    e2e_s = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_value = args.steps / float(e2e_s.item())
    bi = sum(hl.numel() * hl.element_size() for hl in host_lat) + host_emb.numel() * host_emb.element_size()
    bo = sum(ho.numel() * ho.element_size() for ho in host_out)
    if world > 1:
        tb = torch.tensor([bi, bo], device=dev, dtype=torch.float64)
        dist.all_reduce(tb)
        bi, bo = int(tb[0].item()), int(tb[1].item())

    # ---- N > 1: the end-of-denoise exchange + overlap smoothing on the REAL windows (chunk_only.py:123-150), both
    # transports, bit-compared: the fused NVLink-peer blend kernel (north-star item 3) and NCCL all_gather + blend ----
    exch = None
    if world > 1:
        exch = exchange_leg(torch, dist, dev, plan, rank, world, [(s, e, st[0]) for (s, e), st in zip(mine, state)], h, w)

    # ---- the same timed steps in the OTHER 16-bit arithmetic (fp16 = the reference's torch_dtype, chunk_only.py:44-47,
    # and the strategy CLI's default; bf16 = the wider-range mode): same plan, same timing rules, every rank ----
    other = None
    if not args.timed_only:
        try:
            other = other_dtype_leg(torch, dist, dev, world, mine, h, w, args,
                                    "bf16" if args.dtype == "fp16" else "fp16", sync_all)
        except Exception as ex:  # a report, never a reason to lose the line
            other = {"error": repr(ex)}

    # ---- N = 1: BASELINE configs[1] — Zeroscope v2 576w, 24 frames of 576x320 (latent 40x72), chunk_only windows, bf16
    # as that config names it, and fp16; same timing rules (the schedule there is 25 steps: a step is a step) ----
    config_576w = None
    if world == 1 and not args.timed_only:
        try:
            config_576w = {"workload": "Zeroscope v2 576w, 24 frames 576x320 (latent 40x72), chunk_only windows "
                                       "[(0,16),(12,24)], CFG 7.5, 25-step DDIM schedule (BASELINE configs[1])"}
            for nm in ("bf16", "fp16"):
                leg = other_dtype_leg(torch, dist, dev, world, [(0, 16), (12, 24)], 40, 72, args, nm, sync_all,
                                      use_ctx=False)
                leg["roofline_step"] = {"bound": "tensor", "achieved": leg["tflops"], "peak": peaks["bf16_sustained"],
                                        "unit": "TFLOP/s", "frac": leg["tflops"] / peaks["bf16_sustained"]}
                config_576w[nm] = leg
        except Exception as ex:  # a report, never a reason to lose the line
            config_576w = {"error": repr(ex)}

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- rank 0 only: rooflines, Merkle commit, CPU baseline ------------------------------------------------------
    flops_rank = [sum(unet.flops(2, e - s, h, w) for s, e in plan.of_rank(r)) for r in range(world)]
    step_flops = max(flops_rank)  # the bottleneck rank paces the step
    per_step_ms = total_ms / args.steps
    achieved_step = step_flops / (per_step_ms / 1e3) / 1e12
    roofline_step = {"bound": "tensor", "achieved": achieved_step, "peak": peaks["bf16_sustained"], "unit": "TFLOP/s",
                     "frac": achieved_step / peaks["bf16_sustained"], "traffic": None,
                     "what": "whole UNet3D forward(s) of the bottleneck rank per video-step: algorithmic FLOPs "
                             "(b2v_unet3d_flops, MAC=2) / CUDA-event time; peak = sustained cuBLAS bf16, " + peaks["source"]}
    roofline = dominant_kernel_roofline(torch, ops, dev, peaks, mine, in_load=True, dtype=dtype)
    roofline["when"] = ("right after the timed steps, GPU under the power cap of the long step (clocks as in `clocks`): "
                        "sustained peak as denominator")
    roofline["cold"] = {"achieved": cold["achieved"], "peak": cold["peak"], "frac": cold["frac"],
                        "ms_per_launch": cold["ms_per_launch"],
                        "when": "before the first step of the process (boost clocks, kernel alone): burst peak"}
    if args.timed_only:
        emit({"metric": METRIC, "value": value, "unit": "steps/s", "n_gpus": world, "steps": args.steps,
              "warmup": args.warmup, "ms_per_step": per_step_ms, "gpu_launches": int(launches), "clocks": clocks,
              "roofline": roofline, "roofline_step": roofline_step, "e2e": {"value": e2e_value, "unit": "steps/s"},
              "exchange_blend_ms": exch["ms"] if exch else None, "exchange_parity": exch["parity"] if exch else None,
              "exchange": exch, "note": "--timed-only: auxiliary legs skipped"})
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    merkle = merkle_bench(torch, ops, proof, dev, peaks, h, w, sm_mhz=(clocks or {}).get("sm_max_mhz"),
                          cpu=(world == 1 and not args.no_cpu_baseline))
    try:
        merkle["in_loop"] = in_loop_commit(torch, pipe, emb[:1], dev, h, w)
    except Exception as ex:  # a report, never a reason to lose the line
        merkle["in_loop"] = {"error": repr(ex)}
    decode = None
    if world == 1:
        try:
            decode = frame_decode(torch, dev, peaks, h, w, cpu=not args.no_cpu_baseline)
        except Exception as ex:  # a report, never a reason to lose the line
            decode = {"error": repr(ex)}
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        try:
            _, _, cpu = cpu_line(args, plan.ranges, h, w, as_reference=False)
        except Exception as ex:  # the baseline is a report, never a reason to lose the GPU line
            cpu = {"value": None, "unit": "steps/s", "cores": os.cpu_count(), "kind": "port", "sample": "failed: %r" % (ex,)}
    eager = None
    if world == 1 and not args.no_cpu_baseline:
        try:
            eager = gpu_torch_eager(torch, dev, plan.ranges, h, w)
        except Exception as ex:  # a report, never a reason to lose the line
            eager = {"error": repr(ex)}
    line = {
        "metric": METRIC, "value": value, "unit": "steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": per_step_ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": args.dtype, "other_dtype": other, "config_576w": config_576w,
        "data": "synthetic (random-init weights seed 1234, seeded prompt embeddings, seed-0 base noise)",
        "config": workload_config(plan, plan_note, h, w),
        "run": {"windows_per_rank": len(mine), "unet_flops_per_step_bottleneck_rank": step_flops,
                "reference_default_plan": ref_default},
        "e2e": {"value": e2e_value, "unit": "steps/s", "h2d_bytes_per_step": bi, "d2h_bytes_per_step": bo},
        "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "roofline_step": roofline_step,
        "merkle": merkle, "frame_decode": decode, "cpu_baseline": cpu, "gpu_torch_eager": eager,
    }
    if exch is not None:
        line["exchange_blend_ms"] = exch["ms"]
        line["exchange_parity"] = exch["parity"]
        line["exchange"] = exch
    emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def other_dtype_leg(torch, dist, dev, world, mine, h, w, args, name, sync_all, use_ctx=True):
    """The headline's timed region again with latents and UNet arithmetic in `name`: its own pipeline (same seed, same
    weights rounded to that type), W warm-up steps, K timed steps between CUDA events, max over ranks."""
    from b2v import ops
    from b2v.pipeline import VideoPipeline
    from b2v.strategies.common import CudaDenoiser, base_noise
    dtype = {"bf16": torch.bfloat16, "fp16": torch.float16}[name]
    pipe = VideoPipeline.from_random("cerspense/zeroscope_v2_XL", device=dev, seed=1234, dtype=dtype)
    pipe.scheduler.set_timesteps(TOTAL_STEPS, device="cuda")
    ts = pipe.scheduler.host_timesteps()
    emb = pipe.text_encoder(pipe.tokenizer(["a rocket in space, 4k", ""]).input_ids)[0]
    base = base_noise(4, FRAMES, h, w, dtype, dev)
    ctx = ops.context_mean(base) if use_ctx else None
    ctx_w = CTX_W if use_ctx else 0.0
    den = CudaDenoiser(pipe, emb[1:], emb[:1], GUIDANCE, ctx=ctx, context_weight=ctx_w)
    state = []
    for s, e in mine:
        lat = base[:, :, s:e].clone()
        x = ops.make_unet_input(lat, 2, ctx=ctx, ctx_w=ctx_w)
        state.append([lat, x, torch.empty_like(x)])

    def video_step(i):
        t = ts[i % len(ts)]
        for st in state:
            pipe.unet.forward_into(st[1], t, den.emb, st[2])
            st[0], st[1] = pipe.scheduler.fused_cfg_step(st[0], st[2], t, GUIDANCE, ctx=ctx, ctx_w=ctx_w, n_dup=2,
                                                         xnext=st[1])

    for i in range(args.warmup):
        video_step(i)
    sync_all()
    l0 = ops.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        video_step(args.warmup + i)
    e1.record()
    sync_all()
    launches = ops.launch_count() - l0
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    finite = all(bool(torch.isfinite(st[0].float()).all()) for st in state)
    flops = sum(pipe.unet.flops(2, e - s, h, w) for s, e in mine)
    del state, den, pipe
    torch.cuda.empty_cache()
    per = float(ms.item()) / args.steps
    return {"dtype": name, "value": 1e3 / per, "unit": "steps/s", "ms_per_step": per, "gpu_launches": int(launches),
            "latents_finite": finite, "unet_flops_per_step": flops, "tflops": flops / (per / 1e3) / 1e12,
            "what": "the headline's timed region (same plan, W warm-up + K timed steps, CUDA events, max over ranks) with "
                    "latents and UNet arithmetic in this type"}


def gpu_torch_eager(torch, dev, ranges, h, w):
    """CONTEXT, never the target: the same whole-video step by STOCK torch on the same B200 — the oracle's plain
    torch.nn restatement of diffusers' UNet3DConditionModel moved to the GPU in fp16 (cuDNN convs, cuBLAS linears,
    F.scaled_dot_product_attention = flash attention: what `pipe.unet(...)` of the reference runs in eager mode), CFG
    batch 2, every window of the plan at its real shape. SURVEY.md's stated bar was to beat this."""
    from oracle import unet3d_ref as R
    R.Attention.use_sdpa = True
    net = R.random_init_(R.UNet3DConditionModel(R.ZEROSCOPE), seed=1234).half().to(dev).eval()
    g = torch.Generator().manual_seed(1)
    emb = torch.randn(2, 77, 1024, generator=g).half().to(dev)
    per_len = {}
    with torch.no_grad():
        for F in sorted({e - s for s, e in ranges}, reverse=True):
            x = torch.randn(2, 4, F, h, w, generator=g).half().to(dev)
            net(x, 981, emb)  # warm-up: cuDNN / cuBLAS heuristics
            torch.cuda.synchronize(dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(2):
                out = net(x, 981, emb)
            e1.record()
            torch.cuda.synchronize(dev)
            assert bool(torch.isfinite(out).all())
            per_len[F] = e0.elapsed_time(e1) / 2
    step_ms = sum(per_len[e - s] for s, e in ranges)
    del net
    torch.cuda.empty_cache()
    return {"value": 1e3 / step_ms, "unit": "steps/s", "ms_per_step": step_ms,
            "ms_per_window_forward": {str(k): v for k, v in per_len.items()},
            "tflops": video_step_flops(ranges, h, w) / step_ms / 1e9,
            "what": "stock torch %s fp16 eager on this GPU (oracle/unet3d_ref.py module, SDPA attention), UNet forwards of "
                    "every window only (no CFG / DDIM kernels): context for the B200-native number, not a target"
                    % torch.__version__}


def exchange_leg(torch, dist, dev, plan, rank, world, my_windows, h, w, iters=5):
    """All ranks. Times (CUDA events, max over ranks) the exchange + blend of the denoised windows over both transports
    and compares the blended videos bit for bit across transports and across ranks."""
    from b2v import exchange
    from b2v.strategies.common import cuda_blend
    T, ov = FRAMES, plan.overlap
    C_ = my_windows[0][2].shape[1]
    dtype = my_windows[0][2].dtype
    pw = exchange.PeerWindows(plan, world, rank, C_, h, w, dtype, dev)

    def p2p():
        return pw.gather_blend(my_windows, T, ov)[0]

    def nccl():
        wins, _ = exchange.nccl_all_gather(my_windows, plan, rank, world, dev)
        return cuda_blend(wins, T, ov)

    res, ms = {}, {}
    for name, fn in (("p2p_fused_blend", p2p), ("nccl_all_gather_then_blend", nccl)):
        res[name] = fn()  # warm-up (peer rendezvous / NCCL channels exist afterwards)
        torch.cuda.synchronize(dev)
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            res[name] = fn()
        e1.record()
        torch.cuda.synchronize(dev)
        t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms[name] = float(t.item())
    a, b = res["p2p_fused_blend"], res["nccl_all_gather_then_blend"]
    same = torch.tensor([1 if torch.equal(a, b) else 0], device=dev)
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    # every rank must hold the SAME video: compare a checksum of the bytes across ranks
    chk = a.view(torch.int32).sum(dtype=torch.int64).reshape(1)
    lo, hi = chk.clone(), chk.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    ok = bool(same.item()) and int(lo.item()) == int(hi.item())
    payload = sum(t.numel() * t.element_size() for _, _, t in my_windows)
    # per-step halo exchange (--halo_sync): the neighbours' overlap frames read out of their HBM over NVLink
    # and crossfaded, against the same synchronisation done by ONE process holding every window (gathered over NCCL)
    halo = None
    try:
        hs = exchange.HaloSync(plan, world, rank, C_, h, w, dtype, dev)
        lats = [t for _, _, t in my_windows]
        got = hs.sync(lats)
        torch.cuda.synchronize(dev)
        dist.barrier()
        h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        h0.record()
        for _ in range(iters):
            got = hs.sync(lats)
        h1.record()
        torch.cuda.synchronize(dev)
        ht = torch.tensor([h0.elapsed_time(h1) / iters], device=dev, dtype=torch.float64)
        dist.all_reduce(ht, op=dist.ReduceOp.MAX)
        every, _ = exchange.nccl_all_gather(my_windows, plan, rank, world, dev)

        class OneRank:
            ranges, overlap, chunk_size = [(s, e) for s, e, _ in every], plan.overlap, plan.chunk_size

            @staticmethod
            def of_rank(r):
                return [(s, e) for s, e, _ in every]
        one = exchange.HaloSync(OneRank, 1, 0, C_, h, w, dtype, dev)
        ref = {(s, e): t for (s, e, _), t in zip(every, one.sync([t for _, _, t in every]))}
        eq = torch.tensor([1 if all(torch.equal(g_, ref[(s, e)]) for (s, e, _), g_ in zip(my_windows, got)) else 0],
                          device=dev)
        dist.all_reduce(eq, op=dist.ReduceOp.MIN)
        halo = {"ms": float(ht.item()), "parity": "bit-exact" if bool(eq.item()) else "MISMATCH",
                "what": "HaloSync.sync of the denoised windows over NVLink peer mappings vs one process holding every window"}
    except ValueError as ex:  # plans the halo exchange does not take (a frame under three windows)
        halo = {"ms": None, "parity": "n/a", "what": str(ex)}
    except Exception as ex:  # a report, never a reason to lose the line
        halo = {"ms": None, "parity": "error", "what": repr(ex)}
    return {"ms": ms, "halo_sync": halo, "parity": "bit-exact" if ok else "MISMATCH", "transports_equal": bool(same.item()),
            "ranks_equal": int(lo.item()) == int(hi.item()), "window_bytes_this_rank": payload,
            "what": "stage -> device barrier -> ONE blend kernel reading every window from its owner's HBM over NVLink "
                    "peer mappings -> barrier (p2p_fused_blend) vs all_gather_into_tensor + the same blend kernel on "
                    "local memory; windows = this run's denoised latents; CUDA events, max over ranks, %d iterations" % iters}


def frame_decode(torch, dev, peaks, h, w, cpu=True):
    """SURVEY.md §8f N1 (chunk_only.py:152-158): the FRAMES blended latents -> uint8 frames, SD VAE decoder with
    random-init weights. Device time with CUDA events; e2e adds the device->pinned-host copy of the frame bytes."""
    from b2v.vae import AutoencoderKL
    net = AutoencoderKL.from_random(device=dev)
    lat = (torch.randn(1, 4, FRAMES, h, w, generator=torch.Generator().manual_seed(3)) * 0.3).to(dev)
    host = torch.empty((FRAMES, 8 * h, 8 * w, 3), dtype=torch.uint8).pin_memory()
    l0 = ops_launches()
    ms = _time_cuda(torch, lambda: net.decode_frames(lat), iters=2, warm=1)
    launches = (ops_launches() - l0) // 3
    ms_e2e = _time_cuda(torch, lambda: host.copy_(net.decode_frames(lat), non_blocking=True), iters=2, warm=1)
    fl = net.flops(FRAMES, h, w)
    out = {"workload": "%d frames %dx%d, SD VAE decoder (49.5 M params, random init), bf16" % (FRAMES, 8 * w, 8 * h),
           "ms": ms, "frames_per_s": FRAMES / ms * 1e3, "e2e_ms": ms_e2e, "e2e_frames_per_s": FRAMES / ms_e2e * 1e3,
           "d2h_bytes": host.numel(), "gpu_launches": int(launches),
           "roofline": {"bound": "tensor", "achieved": fl / ms / 1e9, "peak": peaks["bf16_sustained"], "unit": "TFLOP/s",
                        "frac": fl / ms / 1e9 / peaks["bf16_sustained"],
                        "what": "algorithmic FLOPs of convs + projections + attention (MAC=2) / CUDA-event time"}}
    if cpu:
        from oracle import vae_ref as V  # CPU baseline leg: the oracle as the reported baseline, never the product
        from oracle.unet3d_ref import random_init_
        ref = random_init_(V.AutoencoderKLDecoder(V.SD_VAE), seed=4321).eval()
        ch, cw = 40, 72  # one 576x320 frame: a bounded sample (1.8 TFLOP fp32)
        z = torch.randn(1, 4, 1, ch, cw, generator=torch.Generator().manual_seed(3)) * 0.3
        t0 = time.perf_counter()
        V.frames_u8(ref, z)
        dt = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": 1.0 / dt * (ch * cw) / (h * w), "unit": "frames/s at %dx%d" % (8 * w, 8 * h),
                               "cores": torch.get_num_threads(), "kind": "port",
                               "sample": "1 frame of 576x320 decoded by oracle/vae_ref.py (fp32 torch CPU) in %.2f s, "
                                         "scaled by the pixel ratio" % dt}
    return out


def ops_launches():
    from b2v import ops
    return ops.launch_count()


def _time_cuda(torch, fn, iters=3, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def dominant_kernel_roofline(torch, ops, dev, peaks, mine, in_load=False, dtype=None):
    """The kernel with the largest share of a step (profiles/: attn_kernel at UNet level 0, spatial self-attention,
    seq 9216, 5 heads, batch 2*F) timed live with CUDA events. in_load: right after the timed steps, the GPU still under
    the power cap of the long step -> the SUSTAINED measured peak is the denominator; otherwise (cold, boost clocks, the
    kernel alone) the BURST peak (MEASURED_PEAKS.json gives both for exactly these two situations)."""
    F = max(e - s for s, e in mine)
    nb, heads, S = 2 * F, 5, (HEIGHT // 8) * (WIDTH // 8)
    g = torch.Generator(device=dev).manual_seed(0)
    q, k, v = (torch.randn(nb, S, heads * 64, device=dev, generator=g).to(dtype or torch.bfloat16) for _ in range(3))
    ms = _time_cuda(torch, lambda: ops.attention(q, k, v, heads))
    flops = 4.0 * nb * heads * S * S * 64
    ach = flops / (ms / 1e3) / 1e12
    # dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of this shape from the committed `ncu --set full` capture
    # (profiles/rd2b_ncu_full_summary.md: 566.6 MB read + 171.6 MB written; algorithmic q+k+v+out = 755 MB)
    traffic = 738.2e6 if (nb, heads, S) == (32, 5, 9216) else None
    peak = peaks["bf16_sustained"] if in_load else peaks["bf16_burst"]
    return {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s",
            "frac": ach / peak, "traffic": traffic,
            "traffic_source": "ncu --set full, profiles/rd2b_ncu_full_summary.md (bytes per launch)",
            "kernel": "attn_kernel (spatial self-attention L0)",
            "shape": {"batch": nb, "heads": heads, "seq": S, "head_dim": 64}, "ms_per_launch": ms,
            "operand_dtype": str(q.dtype).replace("torch.", ""),
            "flops_per_launch": flops,
            "peak_source": ("sustained" if in_load else "burst") + " cuBLAS bf16, " + peaks["source"]}


def in_loop_commit(torch, pipe, ehs, dev, h, w, steps=TOTAL_STEPS, loop_steps=10):
    """Cost of the commitment on the miner's recording loop (b2v.miner.record_trajectory: batch 1, no CFG, 24 frames) for
    the trajectory BASELINE configs[4] names (50 DDIM steps): no commitment / all leaves hashed side by side after the loop
    (the default, commit='end'). Also, on a shorter run, commit='loop': leaf i hashed by one shared-memory-free warp on a
    side stream under the UNet of step i+1."""
    from b2v import miner, ops
    z0 = torch.randn(1, 4, FRAMES, h, w, device=dev, generator=torch.Generator(device=dev).manual_seed(3)).half()
    ehs = ehs.to(torch.float16)
    out = {"steps": steps, "loop_steps": loop_steps}

    def run(commit, n):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        miner.record_trajectory(pipe, z0, ehs, n, commit=commit)
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) * 1e3

    for commit in (False, "end", "loop"):
        run(commit, 2)  # warm-up (graph capture for this shape)
    # alternate plain / committed runs and keep the fastest of each: one 5 s run alone is at the mercy of a power-cap
    # excursion (a single pair once read +16 %; alternating pairs read +1.8 .. +2.1 %, profiles/rd2b_commit_end_overhead.log)
    plain, end = [], []
    for _ in range(2):
        plain.append(run(False, steps))
        end.append(run("end", steps))
    out["ms_trajectory_plain"], out["ms_trajectory_commit_end"] = min(plain), min(end)
    out["ms_trajectory_runs"] = {"plain": plain, "commit_end": end}
    out["ms_per_step_plain"] = out["ms_trajectory_plain"] / steps
    out["ms_per_step_commit_end"] = out["ms_trajectory_commit_end"] / steps
    out["overhead_frac_end"] = out["ms_trajectory_commit_end"] / out["ms_trajectory_plain"] - 1.0
    p10, l10 = run(False, loop_steps), run("loop", loop_steps)
    out["ms_per_step_plain_short"] = p10 / loop_steps
    out["ms_per_step_commit_loop"] = l10 / loop_steps
    out["overhead_frac_loop"] = l10 / p10 - 1.0
    e = torch.randn_like(z0)
    for mode, key in ((3, "leaf_ms_warp_kernel"), (1, "leaf_ms_chain_kernel")):
        out[key] = _time_cuda(torch, lambda: ops.sha256_leaves([z0], [e], prefixes=[981], mode=mode), iters=2, warm=1)
    out["note"] = ("miner path B=1 x %d frames, one 3.5 MB leaf per step. 'end' (default) hashes the %d leaves side by side "
                   "after the loop: a fixed ~55 ms per trajectory whatever the step count. 'loop' (%d-step run) hides each "
                   "leaf under the next UNet but its warp shares an SM with a persistent GEMM CTA and stretches that "
                   "GEMM's tail" % (FRAMES, steps, loop_steps))
    return out


def merkle_bench(torch, ops, proof, dev, peaks, h, w, sm_mhz=None, cpu=True):
    """Merkle commit GB/s (BASELINE.json metric, second half): leaves of a 50-step XL trajectory (miner path: batch 1,
    24 frames; leaf = be16(t) ‖ z fp16 ‖ eps fp16 = 3 538 946 B) hashed in place + tree, device resident; next to it the
    reference's own construction (miner.py:207-233 leaf hashing + proof.py:250-288 tree, hashlib) on the SAME bytes on the
    host cores: single thread as written, and a thread pool over leaves."""
    import hashlib
    out = {}
    elems = 4 * FRAMES * h * w
    msg = 2 + 4 * elems
    nblocks = (msg + 9 + 63) // 64
    for n in (50, 400):
        g = torch.Generator(device=dev).manual_seed(n)
        zs = torch.randn(n, elems, device=dev, generator=g).half()
        es = torch.randn(n, elems, device=dev, generator=g).half()
        zl, el = [zs[i] for i in range(n)], [es[i] for i in range(n)]
        ts = [(981 - 20 * i) % 1000 for i in range(n)]
        ms = _time_cuda(torch, lambda: proof.commit_trajectory_device(zl, el, ts), iters=2, warm=1)
        alg = n * nblocks * 64 + (n - 1) * 128
        out["leaves_%d" % n] = {"gb_per_s": alg / (ms / 1e3) / 1e9, "ms": ms, "bytes": alg}
        if n == 50:
            proof.build_merkle_tree(zl, el, ts)  # warm-up: pins the host staging buffer once, as a long-lived miner does
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            root, leaf = proof.build_merkle_tree(zl, el, ts)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            out["e2e_build_merkle_tree_50"] = {"gb_per_s": alg / dt / 1e9, "ms": dt * 1e3,
                                               "note": "drop-in call incl. D2H of every leaf's bytes (return value): one "
                                                       "pinned DMA and the host slicing into bytes both under the hashing"}
            if cpu:
                # CPU baseline leg (the oracle / hashlib as the reported baseline, never the product): identical bytes
                from oracle import merkle_ref as M
                torch.cuda.synchronize()
                t0 = time.perf_counter()   # what the reference does before it can hash (miner.py:203-204), per leaf
                zb = [z.cpu().numpy().tobytes() for z in zl]
                eb = [e.cpu().numpy().tobytes() for e in el]
                t_fetch = time.perf_counter() - t0
                t0 = time.perf_counter()
                hs = [M.leaf_hash(t, z, e) for t, z, e in zip(ts, zb, eb)]
                lv = M.tree_levels(hs)
                t1 = time.perf_counter() - t0
                cores = os.cpu_count() or 1
                from concurrent.futures import ThreadPoolExecutor
                with ThreadPoolExecutor(cores) as ex:
                    t0 = time.perf_counter()
                    hs2 = list(ex.map(lambda a: M.leaf_hash(*a), zip(ts, zb, eb)))
                    lv2 = M.tree_levels(hs2)
                    tp = time.perf_counter() - t0
                assert lv[-1][0] == lv2[-1][0] == root, "device root != hashlib root"
                out["cpu_baseline"] = {
                    "value": alg / tp / 1e9, "unit": "GB/s", "cores": cores, "kind": "reference",
                    "single_thread_gb_per_s": alg / t1 / 1e9, "single_thread_ms": t1 * 1e3, "pool_ms": tp * 1e3,
                    "roots_equal": True, "fetch_ms": t_fetch * 1e3,
                    "like_for_like": {
                        "what": "miner.py:203-233 as the reference runs it from device latents: z.cpu().numpy().tobytes() "
                                "per leaf + hashlib leaves + tree; compares with e2e_build_merkle_tree_50 (device latents "
                                "in, root + bytes + paths out)",
                        "single_thread_ms": (t_fetch + t1) * 1e3, "pool_ms": (t_fetch + tp) * 1e3,
                        "single_thread_gb_per_s": alg / (t_fetch + t1) / 1e9, "pool_gb_per_s": alg / (t_fetch + tp) / 1e9},
                    "sample": "the reference's construction (miner.py:207-233 leaf = sha256(be16(t) + z + eps) via hashlib, "
                              "proof.py:260-286 tree) on the same 50 leaves (%.1f MB): single thread as written, and "
                              "ThreadPoolExecutor(%d) over leaves; hashing only (bytes already on the host)"
                              % (alg / 1e6, cores)}
                del zb, eb
        del zs, es, zl, el
    # The bound of this kernel is not HBM: one SHA-256 chain is serial. Per 64-byte block the compression is 64 rounds of
    # ~12 dependent alu-pipe instructions at 2 cycles each on ONE sub-partition (csrc/sha256.cu) => >= 1536 cycles per
    # block per chain; `chains` leaves run side by side. Peak = chains x 64 B x f_clk / 1536.
    clk = (sm_mhz or 1965.0) * 1e6
    per_chain = 64.0 * clk / 1536.0 / 1e9
    best = out["leaves_50"]["gb_per_s"]
    out["roofline"] = {"bound": "integer (alu pipe, serial chain per leaf)", "achieved": best, "peak": 50 * per_chain,
                       "unit": "GB/s", "frac": best / (50 * per_chain), "traffic": None, "chains": 50,
                       "per_chain_peak_gb_per_s": per_chain, "sm_mhz": clk / 1e6,
                       "hbm_frac": best / peaks["hbm_gbs"],
                       "note": "SHA-256 is one Merkle-Damgard chain per leaf (the construction is the reference's and "
                               "cannot change without changing the root): the device parallelism of ONE trajectory is "
                               "its 50 leaves. 400 leaves (8 trajectories): %.1f GB/s. ncu: profiles/rd2b_ncu_sha256.md"
                               % out["leaves_400"]["gb_per_s"]}
    out["metric"] = "Merkle commit GB/s"
    return out


_REAL_STDOUT = None


def emit(line):
    """The ONE JSON line goes to the process's original stdout; everything else written to fd 1 while the bench runs
    (NCCL prints its version banner there) is diverted to stderr by main()."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b2v", choices=["b2v", "reference"])
    ap.add_argument("--no-cpu-baseline", dest="no_cpu_baseline", action="store_true")
    ap.add_argument("--dtype", default="fp16", choices=["fp16", "bf16"],
                    help="16-bit arithmetic of the headline steps (latents + UNet). fp16 is the reference's torch_dtype "
                         "(chunk_only.py:44-47), the strategy CLI's default and the mode that meets Appendix D's 1e-2 "
                         "parity bar; the other type is timed too and reported under other_dtype")
    ap.add_argument("--timed-only", dest="timed_only", action="store_true",
                    help="only the timed denoise steps + e2e leg (no Merkle / decode / baseline legs): for ncu launch lists")
    args = ap.parse_args()
    if args.impl == "reference":
        if int(os.getenv("RANK", 0)) != 0:
            return
        import b2v  # noqa: F401  (host-side planner only; no GPU work on this arm)
        from b2v import planner
        world = int(os.getenv("WORLD_SIZE", args.gpus))
        cs, ov, note = plan_for(world)
        plan = planner.plan(FRAMES, world, cs, ov, variant="coherent")
        h, w = HEIGHT // 8, WIDTH // 8
        # the CPU arm has one host: it executes every rank's windows, i.e. the whole-video step
        value, per, cpu = cpu_line(args, plan.ranges, h, w, as_reference=True)
        emit({
            "impl": "reference", "metric": METRIC, "value": value, "unit": "steps/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": per * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "extrapolated": cpu["extrapolated"], "measured_steps": cpu["measured_steps"], "measured_s": cpu["measured_s"],
            "config": workload_config(plan, note, h, w),
            "note": "CPU oracle port of the reference's step on the host cores; each whole-video step takes minutes on a "
                    "CPU, so `steps` whole steps are measured only as far as a ~4 min budget allows (measured_steps); "
                    "ms_per_step is the measured mean, not steps x anything",
            "cpu_baseline": cpu, "e2e": {"value": value, "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0})
        return
    gpu_main(args)


if __name__ == "__main__":
    main()
