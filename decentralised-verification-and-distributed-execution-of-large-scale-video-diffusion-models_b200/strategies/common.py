"""Shared engine of the strategy entry points (Distribution/strategies/{chunk_only,fsdp_chunked_coherent,fsdp_chunked,
fsdp}.py): plan windows -> denoise this rank's windows -> exchange -> blend -> (decode) -> metrics row.

What changes relative to the reference: weights are replicated per rank instead of FSDP-sharded; the
denoised windows stay on the device and are exchanged with ONE NCCL all_gather of device buffers instead of
``all_gather_object`` of pickled CPU tensors; the ramp blend runs as a CUDA kernel on the gathered buffers
(b2v_blend_overlap) reproducing the reference's fp16-accumulator arithmetic; per step there is one UNet forward and one
fused CFG+DDIM(+context) kernel. There is no per-step communication, exactly like the reference.
"""
import csv
import datetime
import logging
import os
import random
import socket
import time

import torch
import torch.distributed as dist

from .. import exchange, ops, planner

LOG = logging.getLogger("b2v.strategy")

CSV_HEADER = ["timestamp", "host", "mode", "world_size", "num_frames", "chunk_size", "overlap", "latency_s",
              "throughput_fps", "peak_vram_mb", "end_vram_mb", "network_bytes", "net_gather_s", "net_reduce_s",
              "temp_instab", "flow_err"]


def add_common_args(p, num_frames=32, out_csv="benchmarks.csv", chunking=True):
    p.add_argument("--model_id", default="cerspense/zeroscope_v2_XL")
    p.add_argument("--prompt", default="a rocket in space, 4k")
    p.add_argument("--num_frames", type=int, default=num_frames)
    p.add_argument("--steps", type=int, default=50)
    if chunking:
        p.add_argument("--guidance_scale", type=float, default=7.5)
        p.add_argument("--chunk_size", type=int, default=0)
        p.add_argument("--overlap", type=int, default=4)
        p.add_argument("--fps", type=int, default=8)
    p.add_argument("--height", type=int, default=576)
    p.add_argument("--width", type=int, default=1024)
    p.add_argument("--device", default="cuda")
    p.add_argument("--out_csv", default=out_csv)
    p.add_argument("--emu_bw_mbps", type=float, default=0, help="throttle bandwidth in Mbps (0 = no throttle)")
    p.add_argument("--emu_rtt_ms", type=float, default=0, help="one-way latency in ms (0 = none)")
    p.add_argument("--emu_jitter_ms", type=float, default=0, help="jitter stddev in ms")
    # additions that do not exist in the reference. Defaults keep its behaviour: fp16 latents and fp16 UNet arithmetic
    # (torch_dtype=torch.float16, chunk_only.py:44-47), so the fp16-accumulator blend of chunk_only.py:136-150 is
    # reproduced bit for bit without flags. Deviations from the reference at default flags, on purpose: (1) frames are
    # decoded (out.mp4, temp_instab / flow_err CSV columns) only when a VAE with real weights is there — a local
    # checkpoint folder as --model_id — or when --decode is given; (2) weights are seeded random values unless
    # --model_id is such a folder (there is no network to fetch the hub model).
    p.add_argument("--seed", type=int, default=1234, help="seed of the random-init weights (no checkpoint download)")
    p.add_argument("--dtype", default="fp16", choices=["bf16", "fp16", "fp32"],
                   help="dtype of the latents (base noise, windows, blend accumulator); fp16 = the reference")
    p.add_argument("--unet_dtype", default=None, choices=["bf16", "fp16"],
                   help="arithmetic dtype inside the UNet (default: fp16 when --dtype fp16, else bf16)")
    p.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"],
                   help="end-of-denoise window exchange: p2p = blend kernel reads the peers' windows over NVLink "
                        "(fused exchange + blend), nccl = all_gather then blend")
    p.add_argument("--synthetic_text", action="store_true",
                   help="with a checkpoint folder that has no tokenizer/ + text_encoder/: run on seeded random prompt "
                        "embeddings instead of failing (the prompt is then ignored)")
    p.add_argument("--halo_sync", action="store_true",
                   help="per-step halo exchange (not in the reference): after every DDIM step adjacent windows swap "
                        "their overlapping frames over NVLink and adopt the same crossfade, so windows evolve as one "
                        "video; needs a plan where at most two windows cover a frame")
    p.add_argument("--pre_steps", type=int, default=0,
                   help="legacy coordinator's full-latent pre-denoise (legacy/Latent Chunking/precond_overlap/"
                        "run_worker.py:161-176, its default there is 3; 0 = the strategies' behaviour): the whole video "
                        "goes through the first N timesteps on every rank (replicated, bit-identical), then the windows "
                        "are cut from that latent and run the complete schedule, as the legacy workers do")
    p.add_argument("--decode", dest="decode", action="store_true", default=None,
                   help="decode the blended latents to frames on rank 0 (b2v.vae) and write the mp4 + boundary metrics "
                        "like the reference (chunk_only.py:152-208). Default: on when --model_id is a checkpoint folder "
                        "with a VAE, off for random-init weights (their frames are noise and the steps/s metric "
                        "excludes the decode; the CSV columns temp_instab / flow_err then stay empty)")
    p.add_argument("--no_decode", dest="decode", action="store_false", help="never decode")
    return p


def torch_dtype(name):
    return {"bf16": torch.bfloat16, "fp16": torch.float16, "fp32": torch.float32}[name]


def load_text_tower(path, device, dtype):
    """The CLIP tokenizer / text encoder of a local diffusers-layout checkpoint through ``transformers`` (library code,
    outside the accelerated path: SURVEY.md §8a U11). Returns (None, None) when the folder has no text tower."""
    tdir, edir = os.path.join(path, "tokenizer"), os.path.join(path, "text_encoder")
    if not (os.path.isdir(tdir) and os.path.isdir(edir)):
        return None, None
    from transformers import CLIPTextModel, CLIPTokenizer
    tok = CLIPTokenizer.from_pretrained(tdir)
    enc = CLIPTextModel.from_pretrained(edir, torch_dtype=dtype if dtype != torch.float32 else None).to(device).eval()
    return tok, enc


def init_distributed(device="cuda"):
    """One process per GPU (torchrun env). Returns (rank, world, local_rank)."""
    local_rank = int(os.getenv("LOCAL_RANK", 0))
    if device.startswith("cuda"):
        torch.cuda.set_device(local_rank)
    if not dist.is_initialized():
        if "RANK" in os.environ and "WORLD_SIZE" in os.environ:
            backend = "nccl" if device.startswith("cuda") else "gloo"
            kw = {}
            if backend == "nccl":
                kw["device_id"] = torch.device("cuda", local_rank)
            dist.init_process_group(backend, **kw)
        else:
            return 0, 1, local_rank
    return dist.get_rank(), dist.get_world_size(), local_rank


def base_noise(in_channels, T, h, w, dtype, device, init_noise_sigma=1.0, seed=0):
    """Shared seed-0 noise (chunk_only.py:110-113). Drawn on the CPU generator in fp32 and cast, so every rank — and
    the CPU oracle — sees identical values (the reference's CUDA Philox draw is not reproducible off-device,
    SURVEY.md §8d)."""
    g = torch.Generator().manual_seed(seed)
    base = torch.randn(1, in_channels, T, h, w, generator=g, dtype=torch.float32).to(dtype)
    base = base * init_noise_sigma
    return base.to(device)


class CudaDenoiser:
    """The per-window hot loop (chunk_only.py:65-74 / fsdp_chunked_coherent.py:129-143) on libb2v."""

    def __init__(self, pipe, uncond_emb, cond_emb, guidance_scale, ctx=None, context_weight=0.0, on_step=None):
        self.pipe, self.gs, self.ctx, self.ctx_w = pipe, guidance_scale, ctx, float(context_weight)
        self.emb = torch.cat([uncond_emb, cond_emb], 0).contiguous() if guidance_scale is not None else cond_emb
        self.on_step = on_step  # callback(step_index, t, lat_in, eps2) before the update (commit-in-the-loop)
        self.unet_forwards = 0
        self.launches0 = ops.launch_count()

    def __call__(self, lat, first=0, last=None, use_ctx=True):
        """``first`` / ``last`` restrict the loop to ``timesteps[first:last]`` (the full-latent pre-denoise)."""
        sched, unet = self.pipe.scheduler, self.pipe.unet
        n_dup = 2 if self.gs is not None else 1
        ctx, ctx_w = (self.ctx, self.ctx_w) if use_ctx else (None, 0.0)
        x = ops.make_unet_input(lat, n_dup, ctx=ctx, ctx_w=ctx_w)
        eps2 = torch.empty_like(x)
        for i, t in list(enumerate(sched.host_timesteps()))[first:last]:
            unet.forward_into(x, t, self.emb, eps2)
            self.unet_forwards += 1
            if self.on_step is not None:
                self.on_step(i, t, lat, eps2)
            lat, x = sched.fused_cfg_step(lat, eps2, t, self.gs, ctx=ctx, ctx_w=ctx_w, n_dup=n_dup, xnext=x)
        return lat


    def lockstep(self, lats, halo, on_sync=None):
        """all of this rank's windows advance one DDIM step at a time; after every step the windows
        (here and on the peer GPUs) exchange their overlapping frames and adopt the common crossfade
        (exchange.HaloSync), then the next UNet inputs are rebuilt from the synchronised latents."""
        sched, unet = self.pipe.scheduler, self.pipe.unet
        n_dup = 2 if self.gs is not None else 1
        xs = [ops.make_unet_input(lat, n_dup, ctx=self.ctx, ctx_w=self.ctx_w) for lat in lats]
        eps = [torch.empty_like(x) for x in xs]
        for i, t in enumerate(sched.host_timesteps()):
            for k in range(len(lats)):
                unet.forward_into(xs[k], t, self.emb, eps[k])
                self.unet_forwards += 1
                lats[k], _ = sched.fused_cfg_step(lats[k], eps[k], t, self.gs, ctx=self.ctx, ctx_w=self.ctx_w, n_dup=n_dup,
                                                  xnext=xs[k])
            lats = halo.sync(lats)
            if on_sync is not None:
                on_sync(i, t, lats)
            xs = [ops.make_unet_input(lat, n_dup, ctx=self.ctx, ctx_w=self.ctx_w) for lat in lats]
        return lats


def exchange_windows(my_windows, plan, rank, world, device):
    """NCCL / gloo transport: every rank receives every window (exchange.nccl_all_gather)."""
    return exchange.nccl_all_gather(my_windows, plan, rank, world, device)


def cuda_blend(windows, T, ov):
    """chunk_only.py:136-150 on the device; windows in gather order. Returns fp32 [1,C,T,h,w]."""
    weights = [planner.window_weights(e - s, ov) for s, e, _ in windows]
    return ops.blend_overlap([t for _, _, t in windows], [(s, e) for s, e, _ in windows], weights, T)


def emulate_network(payload_bytes, cfg):
    """Sleep-based bandwidth / latency emulation of the reference (chunk_only.py:127-131); never touches data."""
    if cfg.emu_bw_mbps > 0:
        time.sleep(payload_bytes / (cfg.emu_bw_mbps * 1e6 / 8))
    if cfg.emu_rtt_ms > 0:
        time.sleep(max(0.0, random.gauss(cfg.emu_rtt_ms, cfg.emu_jitter_ms) / 1000.0))


def append_csv(path, row):
    new = not os.path.exists(path)
    with open(path, "a", newline="") as f:
        w = csv.DictWriter(f, fieldnames=CSV_HEADER)
        if new:
            w.writeheader()
        w.writerow({k: row.get(k, "") for k in CSV_HEADER})


def finish_row(res, mode, num_frames, t0):
    elapsed = time.time() - t0
    res.update({"timestamp": datetime.datetime.now(datetime.timezone.utc).replace(tzinfo=None).isoformat(timespec="seconds"), "host": socket.gethostname(),
                "mode": mode, "latency_s": elapsed, "throughput_fps": round(num_frames / elapsed, 3)})
    return res


def vram_mb(device):
    if device.type != "cuda":
        return 0
    try:
        free, total = torch.cuda.mem_get_info(device)
        return int((total - free) // 1024 ** 2)
    except Exception:
        return 0


MAX_FRAMES_PER_FORWARD = 32   # b2v_unet3d_forward / b2v_op_temporal_attention: one window's frames stay in registers


def check_pre_steps(pre_steps, num_frames):
    """--pre_steps sends the WHOLE video through the UNet in one forward; one forward takes at most 32 frames. The
    legacy coordinator (run_worker.py:161-176) pre-denoises at any length — this port refuses up front, before any model
    is loaded, instead of failing with B2V_ERR_ARG inside the first pre-denoise step."""
    if int(pre_steps or 0) > 0 and int(num_frames) > MAX_FRAMES_PER_FORWARD:
        raise ValueError("--pre_steps %d: the pre-denoise runs all %d frames through the UNet in one forward, and one "
                         "forward takes at most %d frames; use --pre_steps 0 for longer videos"
                         % (int(pre_steps), int(num_frames), MAX_FRAMES_PER_FORWARD))


class WindowedDiffuser:
    """Engine behind ChunkedVideoDiffuser (chunk_only.py:36) and DistributedVideoDiffuser
    (fsdp_chunked_coherent.py:47, fsdp_chunked.py:45). ``cfg`` is the argparse namespace of the script.

    Injection points (used by the CPU gloo tests, which have no GPU): ``pipe`` (any pipe-shaped object),
    ``denoise_fn(lat) -> lat``, ``blend_fn(windows, T, ov) -> lat`` and ``base`` (the shared noise)."""

    variant = "chunk_only"   # planner variant
    mode_name = "chunk"

    def __init__(self, cfg, pipe=None, denoise_fn=None, blend_fn=None, base=None):
        self.cfg = cfg
        if denoise_fn is None:
            check_pre_steps(getattr(cfg, "pre_steps", 0), cfg.num_frames)
        self.rank, self.world, self.local_rank = init_distributed(cfg.device)
        self.device = torch.device("cuda", self.local_rank) if cfg.device.startswith("cuda") else torch.device("cpu")
        self.dtype = torch_dtype(getattr(cfg, "dtype", "bf16"))
        if pipe is None:
            from ..pipeline import VideoPipeline
            udt = torch_dtype(cfg.unet_dtype) if getattr(cfg, "unet_dtype", None) else None
            if os.path.isdir(cfg.model_id):  # a downloaded diffusers-layout checkpoint (chunk_only.py:44-56)
                LOG.info("rank %d: loading %s on %s", self.rank, cfg.model_id, self.device)
                tok, enc = load_text_tower(cfg.model_id, self.device, self.dtype)
                pipe = VideoPipeline.from_pretrained(cfg.model_id, device=self.device, dtype=self.dtype,
                                                     unet_dtype=udt, tokenizer=tok, text_encoder=enc,
                                                     allow_synthetic_text=bool(getattr(cfg, "synthetic_text", False)))
                if getattr(cfg, "decode", None) is None:
                    cfg.decode = pipe.vae is not None
            else:                            # no network here: the named architecture with seeded random weights
                LOG.info("rank %d: building random-init %s on %s", self.rank, cfg.model_id, self.device)
                pipe = VideoPipeline.from_random(cfg.model_id, device=self.device, seed=getattr(cfg, "seed", 1234),
                                                 dtype=self.dtype, unet_dtype=udt)
        if getattr(cfg, "decode", False) and getattr(pipe, "vae", None) is None and self.rank == 0:
            from ..vae import AutoencoderKL
            pipe.vae = AutoencoderKL.from_random(device=self.device, seed=getattr(cfg, "seed", 1234) + 1)
        self.pipe = pipe
        self.unet = pipe.unet
        pipe.scheduler.set_timesteps(cfg.steps, device=cfg.device)
        toks = pipe.tokenizer([cfg.prompt, ""], padding="max_length", max_length=pipe.tokenizer.model_max_length,
                              truncation=True, return_tensors="pt")
        with torch.no_grad():
            emb = pipe.text_encoder(toks.input_ids.to(self.device))[0]
        self.cond_emb, self.uncond_emb = emb[:1], emb[1:]
        self.use_ctx = bool(getattr(cfg, "use_ctx", False))
        self.no_chunking = bool(getattr(cfg, "no_chunking", False))
        self.guidance_scale = getattr(cfg, "guidance_scale", 7.5)
        self.context_weight = float(getattr(cfg, "context_weight", 0.35))
        self._denoise_fn, self._blend_fn, self._base = denoise_fn, blend_fn, base
        self.ctx = None
        self.timings = {}

    # ---- pieces -------------------------------------------------------------------------------------------
    def make_plan(self):
        c = self.cfg
        return planner.plan(c.num_frames, self.world, getattr(c, "chunk_size", 0), getattr(c, "overlap", 4),
                            variant=self.variant, no_chunking=self.no_chunking)

    def make_base(self):
        if self._base is not None:
            return self._base.to(self.device)
        c = self.cfg
        return base_noise(self.unet.config.in_channels, c.num_frames, c.height // 8, c.width // 8, self.dtype,
                          self.device, self.pipe.scheduler.init_noise_sigma)

    def make_context(self, base):
        """fsdp_chunked_coherent.py:105-127: ctx = mean over frames of the base noise. Every rank holds the same base
        (same seed), so it is recomputed locally instead of broadcast — identical values, no collective."""
        if not self.use_ctx:
            return None
        if base.is_cuda:
            return ops.context_mean(base)
        return base.mean(dim=2, keepdim=True)

    def _denoise(self, lat):
        if self._denoise_fn is not None:
            return self._denoise_fn(lat)
        if getattr(self, "_cuda_denoiser", None) is None:
            self._cuda_denoiser = CudaDenoiser(self.pipe, self.uncond_emb, self.cond_emb, self.guidance_scale,
                                               ctx=self.ctx, context_weight=self.context_weight if self.use_ctx else 0.0)
        return self._cuda_denoiser(lat)

    def _blend(self, windows, T, ov):
        if self._blend_fn is not None:
            return self._blend_fn(windows, T, ov)
        return cuda_blend(windows, T, ov)

    # ---- the run ------------------------------------------------------------------------------------------
    def pre_denoise(self, base):
        """--pre_steps N (N4, second half): the full latent through the first N timesteps before it is cut into windows.
        Every rank does the same work on the same values (same seed, same kernels: bit-identical, no broadcast)."""
        n = int(getattr(self.cfg, "pre_steps", 0) or 0)
        if n <= 0 or self._denoise_fn is not None:
            return base
        check_pre_steps(n, base.shape[2])
        if getattr(self, "_cuda_denoiser", None) is None:
            self._cuda_denoiser = CudaDenoiser(self.pipe, self.uncond_emb, self.cond_emb, self.guidance_scale,
                                               ctx=self.ctx, context_weight=self.context_weight if self.use_ctx else 0.0)
        with torch.no_grad():
            return self._cuda_denoiser(base.clone(), 0, n, use_ctx=False)

    def denoise_my_windows(self, plan, base):
        base = self.pre_denoise(base)
        if getattr(self.cfg, "halo_sync", False) and self._denoise_fn is None:
            return self.denoise_my_windows_lockstep(plan, base)
        out = []
        for s, e in plan.of_rank(self.rank):
            chunk = base[:, :, s:e].clone()
            with torch.no_grad():
                out.append((s, e, self._denoise(chunk)))
        return out

    def _ensure_halo(self, plan, base):
        """One-off: peer-map the halo slots (symmetric-memory rendezvous, ~1 s) — setup, not part of the denoise loop."""
        if getattr(self, "_halo", None) is None:
            self._halo = exchange.HaloSync(plan, self.world, self.rank, base.shape[1], base.shape[3], base.shape[4],
                                           base.dtype, self.device)

    def denoise_my_windows_lockstep(self, plan, base, on_sync=None):
        """--halo_sync (N4): per-step halo exchange between adjacent windows instead of independent windows."""
        ranges = plan.of_rank(self.rank)
        lats = [base[:, :, s:e].clone() for s, e in ranges]
        self._ensure_halo(plan, base)
        self._halo.step = 0
        if getattr(self, "_cuda_denoiser", None) is None:
            self._cuda_denoiser = CudaDenoiser(self.pipe, self.uncond_emb, self.cond_emb, self.guidance_scale,
                                               ctx=self.ctx, context_weight=self.context_weight if self.use_ctx else 0.0)
        with torch.no_grad():
            lats = self._cuda_denoiser.lockstep(lats, self._halo, on_sync=on_sync)
        return [(s, e, lat) for (s, e), lat in zip(ranges, lats)]

    def __call__(self):
        cfg = self.cfg
        T = cfg.num_frames
        plan = self.make_plan()
        LOG.info("chunk_size=%d, overlap=%d, world=%d", plan.chunk_size, plan.overlap, self.world)
        base = self.make_base()
        self.ctx = self.make_context(base)
        cuda = self.device.type == "cuda"
        if getattr(cfg, "halo_sync", False) and self._denoise_fn is None:
            self._ensure_halo(plan, base)
        if cuda:
            torch.cuda.synchronize(self.device)
        t0 = time.time()
        mine = self.denoise_my_windows(plan, base)
        if cuda:
            torch.cuda.synchronize(self.device)
        self.timings["denoise_s"] = time.time() - t0
        if self.world > 1:
            dist.barrier()
        payload = planner.payload_bytes([(s, e) for s, e, _ in mine], self.unet.config.in_channels)
        emulate_network(payload, cfg)
        if cuda and self.world > 1 and self._blend_fn is None and getattr(cfg, "exchange", "p2p") == "p2p":
            # fused exchange + blend: ONE kernel pulls every window out of its owner's HBM over NVLink and blends
            if getattr(self, "_peer_windows", None) is None:
                t = mine[0][2]
                self._peer_windows = exchange.PeerWindows(plan, self.world, self.rank, t.shape[1], t.shape[3],
                                                          t.shape[4], t.dtype, self.device)
            lat, gather_s = self._peer_windows.gather_blend(mine, T, plan.overlap)
            self.timings["exchange"] = "p2p"
        else:
            windows, gather_s = exchange_windows(mine, plan, self.rank, self.world, self.device)
            lat = self._blend(windows, T, plan.overlap)
            self.timings["exchange"] = "nccl" if self.world > 1 else "local"
        self.latents = lat
        t0d = time.time()
        frames = self.decode(lat)
        self.timings["decode_s"] = time.time() - t0d
        if self.rank == 0 and frames:
            self.write_video(frames)
        peak_local = torch.cuda.max_memory_allocated(self.device) // 1024 ** 2 if cuda else 0
        peak = torch.tensor(int(peak_local), device=self.device)
        end = torch.tensor(int(vram_mb(self.device)), device=self.device)
        if cfg.emu_rtt_ms > 0:
            time.sleep(cfg.emu_rtt_ms / 1000.0)
        t0r = time.time()
        if self.world > 1:
            dist.barrier()
            dist.all_reduce(peak, op=dist.ReduceOp.MAX)
            dist.all_reduce(end, op=dist.ReduceOp.MAX)
        reduce_s = time.time() - t0r
        temp_instab, flow_err = (None, None)
        if self.rank == 0 and len(frames) > 1 and not self.no_chunking:
            temp_instab, flow_err = boundary_metrics(frames, plan.ranges)
        return {"world_size": self.world, "chunk_size": plan.chunk_size, "overlap": plan.overlap, "num_frames": T,
                "peak_vram_mb": int(peak.item()), "end_vram_mb": int(end.item()), "network_bytes": int(payload),
                "net_gather_s": gather_s, "net_reduce_s": reduce_s, "temp_instab": temp_instab, "flow_err": flow_err,
                "denoise_s": self.timings["denoise_s"],
                "steps_per_s": cfg.steps / max(self.timings["denoise_s"], 1e-9)}

    # ---- decode / video (outside the denoise loop; runs only when a VAE is attached) ----------
    def decode(self, lat):
        vae = getattr(self.pipe, "vae", None)
        if vae is None:
            return []
        if hasattr(vae, "decode_frames"):
            # b2v.vae.AutoencoderKL: the whole loop below on the device, frames leave as bytes in one copy
            return list(vae.decode_frames(lat).cpu().numpy())
        frames = []
        for i in range(lat.shape[2]):
            z = lat[:, :, i].to(self.device, dtype=self.dtype)
            with torch.no_grad():
                img = vae.decode(z / 0.18215).sample
            img = (img[0].permute(1, 2, 0) * 0.5 + 0.5).clamp(0, 1)
            frames.append((img * 255).byte().cpu().numpy())
        return frames

    def write_video(self, frames, path="out.mp4"):
        import cv2
        h, w = frames[0].shape[:2]
        vw = cv2.VideoWriter(path, cv2.VideoWriter_fourcc(*"mp4v"), getattr(self.cfg, "fps", 8), (w, h))
        for f in frames:
            vw.write(cv2.cvtColor(f, cv2.COLOR_RGB2BGR))
        vw.release()


def boundary_metrics(frames, ranges):
    """Mean L1 and Farneback flow-warp error across window boundaries (chunk_only.py:183-208)."""
    import cv2
    import numpy as np
    ends = [e for (_, e) in sorted(ranges, key=lambda r: r[0])[:-1]]
    l1, fl = [], []
    for idx in ends:
        if not 0 < idx < len(frames):
            continue
        a, b = frames[idx - 1], frames[idx]
        l1.append(np.mean(np.abs(b.astype(np.float32) - a.astype(np.float32))))
        ga, gb = cv2.cvtColor(a, cv2.COLOR_BGR2GRAY), cv2.cvtColor(b, cv2.COLOR_BGR2GRAY)
        flow = cv2.calcOpticalFlowFarneback(ga, gb, None, 0.5, 3, 15, 3, 5, 1.2, 0)
        hh, ww = ga.shape
        mx = (np.arange(ww)[None, :] + flow[:, :, 0]).astype(np.float32)
        my = (np.arange(hh)[:, None] + flow[:, :, 1]).astype(np.float32)
        fl.append(np.mean(np.abs(cv2.remap(a, mx, my, cv2.INTER_LINEAR).astype(np.float32) - b.astype(np.float32))))
    return (float(np.mean(l1)) if l1 else None, float(np.mean(fl)) if fl else None)


def run_cli(diffuser_cls, cfg, mode):
    """main() body shared by the four scripts (chunk_only.py:246-281)."""
    if cfg.device.startswith("cuda"):
        if not torch.cuda.is_available():
            raise ops._lib.B2VError("no CUDA device: the strategies have no CPU path")
        torch.cuda.empty_cache()
        torch.cuda.reset_peak_memory_stats()
    t0 = time.time()
    bench = diffuser_cls(cfg)
    res = bench()
    if bench.rank == 0:
        finish_row(res, mode, cfg.num_frames, t0)
        append_csv(cfg.out_csv, res)
        LOG.warning("Metrics appended ->  %s (denoise %.3fs, %.3f steps/s)", cfg.out_csv, res["denoise_s"],
                    res["steps_per_s"])
    if dist.is_initialized():
        dist.destroy_process_group()
    return res
