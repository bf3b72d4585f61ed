"""not-gpu: the N>1 host path of the strategies with world_size 2 on the gloo backend (two spawned processes):
planning, window->rank assignment, the all_gather exchange order and the blend weights must reproduce the single-process
oracle (oracle/strategy_ref.run_strategy). The denoiser and blend are injected CPU stand-ins (test doubles — there is no
GPU here and the product has no CPU path)."""
import argparse
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _fake_denoise(lat):
    # deterministic, position dependent, cheap: stands in for the UNet loop
    return (lat * 0.5 + torch.arange(lat.shape[2], dtype=lat.dtype).view(1, 1, -1, 1, 1) * 0.25).to(lat.dtype)


class _FakePipe:
    class _U:
        config = argparse.Namespace(in_channels=4)
        device = torch.device("cpu")
    unet = _U()

    class _S:
        init_noise_sigma = 1.0

        def set_timesteps(self, n, device=None):
            self.n = n
    scheduler = _S()

    class _Tok:
        model_max_length = 77

        def __call__(self, texts, **kw):
            return argparse.Namespace(input_ids=torch.zeros(len(texts), 77, dtype=torch.int64))
    tokenizer = _Tok()

    @staticmethod
    def text_encoder(ids):
        return (torch.zeros(ids.shape[0], 77, 8),)


def _worker(rank, world, port, variant, tmp, chunk_size, overlap):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import b2v  # noqa: F401
    from b2v.strategies.common import WindowedDiffuser
    from oracle import strategy_ref as S

    class D(WindowedDiffuser):
        pass
    D.variant = variant
    cfg = argparse.Namespace(model_id="fake", prompt="p", num_frames=12, steps=3, guidance_scale=7.5,
                             chunk_size=chunk_size, overlap=overlap, fps=8, height=32, width=48, device="cpu",
                             out_csv=os.path.join(tmp, "r.csv"), emu_bw_mbps=0, emu_rtt_ms=0, emu_jitter_ms=0,
                             dtype="fp16", use_ctx=False)

    def blend(windows, T, ov):
        # windows arrive in gather order; regroup per rank for the oracle's nested-list signature
        return S.blend([windows], T, ov, torch.zeros(1, 4, T, 4, 6, dtype=torch.float16))[0]

    d = D(cfg, pipe=_FakePipe(), denoise_fn=_fake_denoise, blend_fn=blend)
    res = d()
    torch.save({"lat": d.latents, "res": res}, os.path.join(tmp, "out%d.pt" % rank))
    import torch.distributed as dist
    dist.destroy_process_group()


@pytest.mark.parametrize("variant,chunk_size,overlap", [("chunk_only", 0, 4), ("coherent", 6, 2), ("chunk_only", 5, 1)])
def test_world2_gloo_matches_single_process_oracle(tmp_path, variant, chunk_size, overlap):
    from oracle import strategy_ref as S
    import b2v  # noqa: F401
    from b2v.strategies.common import base_noise
    port = _free_port()
    mp.spawn(_worker, args=(2, port, variant, str(tmp_path), chunk_size, overlap), nprocs=2, join=True)
    outs = [torch.load(os.path.join(str(tmp_path), "out%d.pt" % r)) for r in range(2)]
    assert torch.equal(outs[0]["lat"], outs[1]["lat"])  # every rank ends with the same blended video
    T = 12
    base = base_noise(4, T, 4, 6, torch.float16, torch.device("cpu"))
    plan = (S.plan_coherent if variant == "coherent" else S.plan_chunk_only)(T, 2, chunk_size, overlap)
    cs, ov, ranges = plan
    gathered = [[(s, e, _fake_denoise(base[:, :, s:e].clone())) for s, e in S.ranges_of_rank(ranges, 2, r)]
                for r in range(2)]
    want = S.blend(gathered, T, ov, base)[0]
    assert torch.equal(outs[0]["lat"], want)
    r = outs[0]["res"]
    assert (r["world_size"], r["chunk_size"], r["overlap"], r["num_frames"]) == (2, cs, ov, T)
    assert r["network_bytes"] == sum((e - s) * 4 * 2 for s, e in S.ranges_of_rank(ranges, 2, 0))
