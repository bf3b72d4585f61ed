"""-m gpu: the miner's recording loop + in-loop commitment (InferNet/neurons/miner.py:571-589, 180-264) on libb2v against
the oracle's restatement (oracle/strategy_ref.denoise_miner, oracle/merkle_ref): per-step latents within the stated
bf16-compute tolerance, Merkle root / sibling paths bit-exact for the device's own leaf bytes."""
import hashlib

import pytest
import torch

pytestmark = pytest.mark.gpu

STEP_REL_L2 = 5e-2  # per-step latent rel-l2 (bf16 tensor-core compute vs the fp32 CPU oracle), as in test_gpu_strategy


def _setup(cuda, frames=4, hw=16, steps=3, io_dtype=torch.float16):
    from b2v.pipeline import VideoPipeline
    from oracle import unet3d_ref as R
    from oracle.ddim_ref import pipeline_scheduler
    ref = R.random_init_(R.UNet3DConditionModel(R.TINY), seed=1234, round_bf16=True).eval()
    pipe = VideoPipeline.from_state_dict(ref.state_dict(), config=R.TINY.to_dict(), device=cuda, dtype=io_dtype)
    g = torch.Generator().manual_seed(7)
    ehs = torch.randn(1, 77, R.TINY.cross_attention_dim, generator=g)
    z0 = torch.randn(1, 4, frames, hw, hw, generator=g).to(io_dtype)
    sr = pipeline_scheduler()
    sr.set_timesteps(steps)
    return ref, pipe, sr, ehs, z0


@pytest.mark.parametrize("commit", ["end", "loop"])
def test_record_trajectory_matches_oracle_and_commits(cuda, commit):
    from b2v import miner, proof
    from oracle import merkle_ref as M
    from oracle import strategy_ref as S
    steps = 3
    ref, pipe, sr, ehs, z0 = _setup(cuda, steps=steps)
    z_ref, lat_ref, eps_ref, ts_ref = S.denoise_miner(ref, sr, z0.float(), ehs)
    traj = miner.record_trajectory(pipe, z0.to(cuda), ehs.to(cuda, z0.dtype), steps, commit=commit)
    torch.cuda.synchronize()
    assert traj.timesteps == ts_ref
    assert len(traj.alphas) == steps and all(0.0 < a < 1.0 for a in traj.alphas)
    assert torch.equal(traj.latents[0].cpu(), z0)  # leaf 0 holds the initial noise bytes untouched
    for i in range(steps):
        for got, want in ((traj.latents[i], lat_ref[i]), (traj.noise_preds[i], eps_ref[i])):
            assert got.dtype == torch.float16
            rel = float((got.float().cpu() - want).norm() / want.norm())
            assert rel < STEP_REL_L2, "step %d: rel-l2 %g" % (i, rel)
    rel = float((traj.final.float().cpu() - z_ref).norm() / z_ref.norm())
    assert rel < STEP_REL_L2
    # commitment: bit-exact against hashlib / the oracle tree over the DEVICE's own leaf bytes
    root, leaf = traj.merkle()
    zb = [t.cpu().numpy().tobytes() for t in traj.latents]
    eb = [t.cpu().numpy().tobytes() for t in traj.noise_preds]
    want_hashes = [hashlib.sha256(int(t).to_bytes(2, "big") + z + e).digest() for t, z, e in zip(traj.timesteps, zb, eb)]
    assert [bytes(r) for r in traj.digests.cpu().numpy()] == want_hashes  # hashed on the device (in the loop on a side stream, or all leaves together at the end)
    o_root, o_leaf = M.commit_trajectory([t.cpu() for t in traj.latents], [t.cpu() for t in traj.noise_preds], traj.timesteps)
    assert root == o_root
    for t in traj.timesteps:
        assert leaf[t][0] == o_leaf[t][0] and leaf[t][1] == o_leaf[t][1] and leaf[t][2] == o_leaf[t][2]
    assert all(proof.verify_merkle_leaf(h, leaf[t][2], root) for h, t in zip(want_hashes, traj.timesteps))
    # the drop-in call over the same tensors gives the same commitment
    root2, _ = proof.build_merkle_tree(traj.latents, traj.noise_preds, traj.timesteps, with_bytes=False)
    assert root2 == root


def test_record_trajectory_without_commit(cuda):
    from b2v import miner
    ref, pipe, sr, ehs, z0 = _setup(cuda, steps=2, io_dtype=torch.bfloat16)
    a = miner.record_trajectory(pipe, z0.to(cuda), ehs.to(cuda, z0.dtype), 2, commit=False)
    b = miner.record_trajectory(pipe, z0.to(cuda), ehs.to(cuda, z0.dtype), 2, commit="loop")
    torch.cuda.synchronize()
    assert a.digests is None and torch.equal(a.final, b.final)  # the side-stream commitment never perturbs the trajectory
    assert a.merkle(with_bytes=False)[0] == b.merkle(with_bytes=False)[0]


def test_wire_format_matches_reference_encoding(cuda):
    """Trajectory.wire() == the base64 fields the reference builds on the host (miner.py:643-666)."""
    import base64
    from b2v import miner
    ref, pipe, sr, ehs, z0 = _setup(cuda, steps=3)
    traj = miner.record_trajectory(pipe, z0.to(cuda), ehs.to(cuda, z0.dtype), 3)
    w = traj.wire()
    root, leaf = traj.merkle()
    assert w["merkle_root"] == base64.b64encode(root).decode("ascii")
    assert w["timesteps"] == traj.timesteps and w["alphas"] == traj.alphas
    assert w["latents"] == [base64.b64encode(z.cpu().numpy().tobytes()).decode("ascii") for z in traj.latents]
    assert w["noise_preds"] == [base64.b64encode(e.cpu().numpy().tobytes()).decode("ascii") for e in traj.noise_preds]
    for t, (zb, eb, path) in leaf.items():
        assert w["leaf_data"][t] == (base64.b64encode(zb).decode("ascii"), base64.b64encode(eb).decode("ascii"),
                                     [base64.b64encode(p).decode("ascii") for p in path])
    # separately allocated tensors (not views of one buffer) take the per-tensor path
    loose = [z.clone() for z in traj.latents]
    assert miner.leaves_b64(loose) == w["latents"]


def test_reexecution_spot_check(cuda):
    """Validator-side re-execution of revealed leaves: an honest trajectory passes (bit-exact re-execution), a forged
    eps that was committed honestly passes membership but fails re-execution, a flipped byte fails membership."""
    from b2v import miner, proof
    steps = 4
    ref, pipe, sr, ehs, z0 = _setup(cuda, steps=steps)
    e_dev = ehs.to(cuda, z0.dtype)
    traj = miner.record_trajectory(pipe, z0.to(cuda), e_dev, steps)
    root, leaf = traj.merkle()
    idx = proof.commit_then_reveal_merkle_spotcheck(root, steps, 2, 7)
    revealed = {traj.timesteps[i]: leaf[traj.timesteps[i]] for i in idx}
    shape = tuple(z0.shape)
    rep = miner.reexecute_spot_check(pipe, e_dev, revealed, root, traj.timesteps, shape, steps)
    assert rep["ok"], rep
    assert all(v["eps_exact"] and v["member"] for v in rep["leaves"].values()), rep
    assert any(v["chain"] is True for v in rep["leaves"].values()), rep  # consecutive pairs were revealed
    # sharded over two ranks: every leaf is checked exactly once
    a = miner.reexecute_spot_check(pipe, e_dev, revealed, root, traj.timesteps, shape, steps, rank=0, world=2)
    b = miner.reexecute_spot_check(pipe, e_dev, revealed, root, traj.timesteps, shape, steps, rank=1, world=2)
    assert sorted(list(a["leaves"]) + list(b["leaves"])) == sorted(revealed) and a["ok"] and b["ok"]
    # forged noise prediction, committed honestly: the tree is consistent, the UNet is not fooled
    forged = [e.clone() for e in traj.noise_preds]
    forged[idx[0]] = torch.randn_like(forged[idx[0]])
    root_f, leaf_f = proof.build_merkle_tree(traj.latents, forged, traj.timesteps)
    rev_f = {traj.timesteps[i]: leaf_f[traj.timesteps[i]] for i in idx}
    rep_f = miner.reexecute_spot_check(pipe, e_dev, rev_f, root_f, traj.timesteps, shape, steps)
    bad_t = traj.timesteps[idx[0]]
    assert not rep_f["ok"] and rep_f["leaves"][bad_t]["member"] and not rep_f["leaves"][bad_t]["reexec"], rep_f
    # a flipped byte in a revealed leaf breaks membership
    t0 = traj.timesteps[idx[0]]
    zb = bytearray(revealed[t0][0]); zb[5] ^= 0x40
    rev_b = dict(revealed); rev_b[t0] = (bytes(zb), revealed[t0][1], revealed[t0][2])
    rep_b = miner.reexecute_spot_check(pipe, e_dev, rev_b, root, traj.timesteps, shape, steps)
    assert not rep_b["ok"] and not rep_b["leaves"][t0]["member"]


def test_decode_final_frames_reference_view(cuda):
    """miner.py:593-600: final latents -> frames through pipe.vae, with the reference's `z.view(B*T, C, H, W)` memory
    reinterpretation (default) and with true frames; against the fp32 oracle decoder, rel-l2 <= 2.5e-2."""
    from types import SimpleNamespace
    from b2v import miner
    from b2v.vae import AutoencoderKL
    from oracle import vae_ref as V
    from oracle.unet3d_ref import random_init_
    ref = random_init_(V.AutoencoderKLDecoder(V.TINY), seed=9, round_bf16=True).eval()
    vae = AutoencoderKL(V.TINY.to_dict(), device=cuda).load_state_dict(ref.state_dict())
    pipe = SimpleNamespace(vae=vae)
    z = torch.randn(1, 4, 3, 8, 8, generator=torch.Generator().manual_seed(1)) * 0.4
    for ref_view in (True, False):
        got = miner.decode_final_frames(pipe, z.to(cuda), reference_view=ref_view)
        flat = z.view(3, 4, 8, 8) if ref_view else z.permute(0, 2, 1, 3, 4).reshape(3, 4, 8, 8)
        with torch.no_grad():
            want = ref.decode(flat / 0.18215).permute(0, 2, 3, 1)
        assert got.shape == (3, 64, 64, 3)
        rel = float((torch.from_numpy(got) - want).norm() / want.norm())
        assert rel <= 2.5e-2, rel
    a = miner.decode_final_frames(pipe, z.to(cuda))
    b = miner.decode_final_frames(pipe, z.to(cuda), reference_view=False)
    assert not (a == b).all()  # the quirk is real: the two readings are different videos
