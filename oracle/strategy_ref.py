"""ORACLE — test infrastructure only (tests/, __graft_entry__.smoke(), bench.py's CPU baseline / reference arm).

CPU restatement of the strategy layer of the reference's hot path (Distribution/strategies), single process, with the
rank loop made explicit and the base noise passed in (the reference draws it from the CUDA Philox stream, which cannot
be reproduced on CPU — SURVEY.md §8a S4):

  plan_chunk_only      chunk_only.py:80-105  (== fsdp_chunked.py:136-171):  ov = min(overlap, cs // 3)
  plan_coherent        fsdp_chunked_coherent.py:149-177: ov = overlap if overlap > 0 else max(4, cs // 3); does not
                       terminate when cs <= ov in the reference — here that case raises
  ramp_weights         chunk_only.py:138-147
  blend                chunk_only.py:136-150 executed with the same torch CPU ops (fp16 accumulator, fp32 weights)
  context_mean         fsdp_chunked_coherent.py:112-118
  denoise              chunk_only.py:65-74 / fsdp_chunked_coherent.py:129-143
  run_strategy         chunk_only.py:107-150 for all ranks in rank order (the all_gather_object order)

Pinned by tests/golden/strategy_golden.json, which oracle/gen_golden.py produced by executing the reference's own
planner / blend source lines (sliced out of the files under /root/reference at generation time).
"""
import torch


def _auto_chunk(T, world):
    lo = max(4, T // (world * 2))
    hi = min(16, T // world)
    return min(hi, max(lo, T // world))


def _ranges(T, size, ov):
    if size - ov <= 0:
        raise ValueError("chunk_size %d <= overlap %d: the reference's planner loops forever here" % (size, ov))
    out, i = [], 0
    while i < T:
        out.append((i, min(i + size, T)))
        i += size - ov
    return out


def _fit_world(T, cs, ov, world):
    ranges = _ranges(T, cs, ov)
    if len(ranges) % world:
        for d in range(1, cs):
            trial = _ranges(T, cs + d, ov)
            if len(trial) % world == 0:
                cs, ranges = cs + d, trial
                break
    if len(ranges) % world:
        ranges = ranges + [ranges[-1]] * (world - len(ranges) % world)
    return cs, ranges


def plan_chunk_only(T, world, chunk_size=0, overlap=4):
    cs = chunk_size if chunk_size > 0 else _auto_chunk(T, world)
    ov = min(overlap, cs // 3)
    cs, ranges = _fit_world(T, cs, ov, world)
    return cs, ov, ranges


def plan_coherent(T, world, chunk_size=0, overlap=4, no_chunking=False):
    if no_chunking:
        cs, ov = T, 0
    else:
        cs = chunk_size if chunk_size > 0 else _auto_chunk(T, world)
        ov = overlap if overlap > 0 else max(4, cs // 3)
    cs, ranges = _fit_world(T, cs, ov, world)
    return cs, ov, ranges


def ranges_of_rank(ranges, world, rank):
    return [r for i, r in enumerate(ranges) if i % world == rank]


def ramp_weights(length, ov):
    """Per-frame blend weight of one chunk as a [length] fp32 tensor."""
    w = torch.ones(length)
    if ov > 0:
        ramp = torch.linspace(0, 1, ov)
        k = min(ov, length)
        if k > 0:
            w[:k] = ramp[:k]
            w[-k:] = torch.flip(ramp[:k], [0])
    return w


def blend(gathered, T, ov, like):
    """gathered: list over ranks of lists of (s, e, cpu tensor [1,C,e-s,h,w]); `like`: the base noise (dtype of the
    accumulator). Returns (lat fp32 [1,C,T,h,w], full accumulator, weight)."""
    full = torch.zeros_like(like.cpu())
    weight = torch.zeros((1, 1, T, 1, 1))
    for per_rank in gathered:
        for s, e, latc in per_rank:
            w = ramp_weights(e - s, ov).view(1, 1, e - s, 1, 1)
            full[:, :, s:e] += latc * w
            weight[:, :, s:e] += w
    return full / weight.clamp(min=1e-6), full, weight


def context_mean(base):
    return base.mean(dim=2, keepdim=True)


def denoise(unet, sched, lat, uncond_emb, cond_emb, guidance_scale, ctx=None, context_weight=0.35, record=None,
            first=0, last=None):
    """One chunk through every timestep (or timesteps[first:last]) with classifier-free guidance.
    `record(step, t, lat_in, eps, lat_out)`."""
    emb = torch.cat([uncond_emb, cond_emb], 0)
    for i, t in list(enumerate(sched.timesteps))[first:last]:
        x = sched.scale_model_input(torch.cat([lat] * 2), t)
        if ctx is not None:
            x = x + context_weight * ctx.repeat(1, 1, lat.shape[2], 1, 1)
        with torch.no_grad():
            noise = unet(x, t, emb)
        u, c = noise.chunk(2)
        eps = u + guidance_scale * (c - u)
        nxt = sched.step(eps, t, lat).prev_sample
        if record is not None:
            record(i, int(t), lat, eps, nxt)
        lat = nxt
    return lat


def denoise_miner(unet, sched, z, ehs):
    """InferNet/neurons/miner.py:571-589: batch 1, no CFG; returns (z_final, latents, noise_preds, timesteps)."""
    latents, noise_preds, timesteps = [], [], []
    for t in sched.timesteps:
        latents.append(z)
        timesteps.append(int(t))
        with torch.no_grad():
            eps = unet(z, t, ehs)
        noise_preds.append(eps)
        z = sched.step(eps, t, z).prev_sample
    return z, latents, noise_preds, timesteps


def run_strategy(unet, sched, base, uncond_emb, cond_emb, world, guidance_scale=7.5, chunk_size=0, overlap=4,
                 coherent=False, context_weight=0.35, record=None, pre_steps=0):
    """All ranks of chunk_only.py (coherent=False) or fsdp_chunked_coherent.py --mode hybrid_ctx (coherent=True),
    sequentially. Returns dict(lat, chunks, ranges, cs, ov).
    pre_steps > 0: the legacy coordinator's full-latent pre-denoise (legacy/Latent Chunking/precond_overlap/
    run_worker.py:161-176): the WHOLE video goes through the first pre_steps timesteps (guidance as given, no context),
    then the windows are cut from that latent and — as the legacy workers do, run_worker.py:184 'pre_steps': 0, :44 —
    run the complete schedule again from its first timestep."""
    T = base.shape[2]
    if pre_steps > 0:
        base = denoise(unet, sched, base.clone(), uncond_emb, cond_emb, guidance_scale, None, 0.0, None, 0, pre_steps)
    cs, ov, ranges = (plan_coherent if coherent else plan_chunk_only)(T, world, chunk_size, overlap)
    ctx = context_mean(base) if coherent else None
    gathered = []
    for rank in range(world):
        mine = []
        for (s, e) in ranges_of_rank(ranges, world, rank):
            rec = (lambda i, t, a, b, c, s=s, e=e: record(rank, s, e, i, t, a, b, c)) if record else None
            den = denoise(unet, sched, base[:, :, s:e].clone(), uncond_emb, cond_emb, guidance_scale, ctx,
                          context_weight, rec)
            mine.append((s, e, den))
        gathered.append(mine)
    lat, full, weight = blend(gathered, T, ov, base)
    return dict(lat=lat, full=full, weight=weight, chunks=gathered, ranges=ranges, cs=cs, ov=ov, ctx=ctx)


# ---------------------------------------------------------------------------------------------------------------
# Per-step halo exchange variant. NOT in the reference (its windows never communicate during
# the loop; the idea is sketched in Distribution/legacy/Latent Chunking/precond_overlap/run_worker.py:165-173): this is
# the single-process statement of the algorithm that b2v's --halo_sync runs across GPUs, so the two can be compared.
# ---------------------------------------------------------------------------------------------------------------
def halo_sync(lats, ranges):
    """Adjacent windows (sorted by start, at most two per frame) replace their overlapping frames by the same
    crossfade: later window's weight of overlap frame k = (k + 0.5) / ov, earlier window's = 1 - that."""
    order = sorted(range(len(ranges)), key=lambda i: ranges[i])
    out = [l.clone() for l in lats]
    for a, b in zip(order[:-1], order[1:]):
        ov = ranges[a][1] - ranges[b][0]
        if ov <= 0:
            continue
        later = torch.tensor([(k + 0.5) / ov for k in range(ov)], dtype=torch.float32).view(1, 1, ov, 1, 1)
        la = ranges[a][1] - ranges[a][0]
        mix = (1.0 - later) * lats[a][:, :, la - ov:].float() + later * lats[b][:, :, :ov].float()
        out[a][:, :, la - ov:] = mix.to(lats[a].dtype)
        out[b][:, :, :ov] = mix.to(lats[b].dtype)
    return out


def denoise_halo_sync(unet, sched, base, ranges, uncond_emb, cond_emb, guidance_scale, ctx=None, context_weight=0.35,
                      record=None):
    """All windows in lockstep: one DDIM step each (as `denoise`), then `halo_sync`. Returns the windows' latents.
    `record(step, t, [latents after the sync])`."""
    emb = torch.cat([uncond_emb, cond_emb], 0)
    lats = [base[:, :, s:e].clone() for s, e in ranges]
    for i, t in enumerate(sched.timesteps):
        nxt = []
        for lat in lats:
            x = sched.scale_model_input(torch.cat([lat] * 2), t)
            if ctx is not None:
                x = x + context_weight * ctx.repeat(1, 1, lat.shape[2], 1, 1)
            with torch.no_grad():
                noise = unet(x, t, emb)
            u, c = noise.chunk(2)
            nxt.append(sched.step(u + guidance_scale * (c - u), t, lat).prev_sample)
        lats = halo_sync(nxt, ranges)
        if record is not None:
            record(i, int(t), lats)
    return lats
