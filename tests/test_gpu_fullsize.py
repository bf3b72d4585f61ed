"""-m gpu: the hot path at BASELINE.json's FULL sizes.

* the full-WIDTH Zeroscope UNet3D (1.41 B parameters, channels 320/640/1280/1280 — the tile shapes the benchmark runs)
  against the CPU fp32 oracle on a latent small enough for the oracle to finish in seconds;
* the ZeroscopeXL latent size (72 x 128) through size-independent properties: run-to-run bit reproducibility, batch
  independence of the CFG pair, DDIM/CFG step linearity;
* the Proof-of-Inference commitment of a full 50-step ZeroscopeXL trajectory (50 leaves x 3 538 946 bytes) against
  hashlib on the same bytes, bit-exact, plus the spot-check reveal round trip.
"""
import hashlib

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def full_pair(cuda):
    from oracle import unet3d_ref as R
    from b2v.unet import UNet3DConditionModel
    torch.set_num_threads(max(1, torch.get_num_threads()))
    ref = R.random_init_(R.UNet3DConditionModel(R.ZEROSCOPE), seed=1234, round_bf16=True).eval()
    net = UNet3DConditionModel(R.ZEROSCOPE.to_dict(), device=cuda).load_state_dict(ref.state_dict())
    return ref, net


def test_full_width_unet_vs_oracle(cuda, full_pair):
    """bf16 tensor-core arithmetic against the fp32 oracle through the full 1.41 B-parameter network. Stated tolerance
    (confirmed by measurement: cos 0.99987, rel-l2 1.6e-2, max-abs 2.4e-2 max|ref| on B200):
    cosine >= 0.9995, rel-l2 <= 3e-2, max-abs <= 4e-2 max|ref|."""
    ref, net = full_pair
    n_params = sum(p.numel() for p in ref.parameters())
    assert n_params == 1411233860  # the published 1.41 B (oracle anchor)
    g = torch.Generator().manual_seed(11)
    B, F, H, W = 2, 3, 16, 16
    x = torch.randn(B, 4, F, H, W, generator=g)
    ehs = torch.randn(B, 77, 1024, generator=g)
    with torch.no_grad():
        want = ref(x, 501, ehs)
    got = net(x.to(cuda, torch.bfloat16), torch.tensor(501), encoder_hidden_states=ehs.to(cuda, torch.bfloat16)).sample
    out, r = got.float().cpu(), want.float()
    cos = float(torch.nn.functional.cosine_similarity(out.flatten(), r.flatten(), dim=0))
    err = float((out - r).abs().max())
    rel = float((out - r).norm() / r.norm())
    print("full-width unet", (B, F, H, W), "max-abs", err, "cos", cos, "rel-l2", rel, "max|ref|", float(r.abs().max()))
    assert cos >= 0.9995, (err, cos, rel)
    assert rel <= 3e-2, (err, cos, rel)
    assert err <= 4e-2 * float(r.abs().max()), (err, cos, rel)


def test_xl_latent_properties(cuda, full_pair):
    """ZeroscopeXL latent 72 x 128, 4 frames, CFG batch 2, the full-width model."""
    _, net = full_pair
    g = torch.Generator().manual_seed(5)
    x = torch.randn(2, 4, 4, 72, 128, generator=g).to(cuda, torch.bfloat16)
    ehs = torch.randn(2, 77, 1024, generator=g).to(cuda, torch.bfloat16)
    a = net(x, torch.tensor(981), encoder_hidden_states=ehs).sample
    b = net(x, torch.tensor(981), encoder_hidden_states=ehs).sample
    assert torch.isfinite(a).all()
    assert torch.equal(a, b), "the forward must be bit-reproducible (the commitment re-executes steps)"
    # the two CFG batch entries never mix: entry 1 alone gives the same prediction up to the bf16 rounding floor of the
    # network (GroupNorm partial sums are folded in another split order and the GEMM tile widths change with M, which
    # flips bf16 roundings that then propagate through ~400 layers: measured 1.9e-2 rel-l2, the same size as the
    # bf16-vs-fp32 deviation of the full-width test above). A mixed batch would be O(1) off.
    c = net(x[1:], torch.tensor(981), encoder_hidden_states=ehs[1:]).sample
    rel = float((a[1:].float() - c.float()).norm() / c.float().norm())
    assert rel < 4e-2, rel
    swapped = net(x.flip(0), torch.tensor(981), encoder_hidden_states=ehs.flip(0)).sample.flip(0)
    rel_swap = float((a.float() - swapped.float()).norm() / a.float().norm())
    assert rel_swap < 4e-2, rel_swap  # batch order is irrelevant
    cross = float((a[:1].float() - a[1:].float()).norm() / a[1:].float().norm())
    assert cross > 0.3, cross          # ... while the two entries themselves are far apart
    # another timestep / other text really change the output (no stale graph inputs)
    d = net(x, torch.tensor(21), encoder_hidden_states=ehs).sample
    e = net(x, torch.tensor(981), encoder_hidden_states=ehs.flip(0)).sample
    assert not torch.equal(a, d) and not torch.equal(a, e)


def test_xl_step_linearity(cuda):
    """The fused CFG + DDIM step is affine in (lat, eps): step(2 lat, 2 eps) == 2 step(lat, eps) (no clipping in the
    pipeline scheduler), at the XL latent size, fp32 I/O so the check is tight."""
    from b2v.scheduler import DDIMScheduler
    sm = DDIMScheduler.from_pipeline_config()
    sm.set_timesteps(50)
    t = sm.host_timesteps()[7]
    g = torch.Generator().manual_seed(9)
    lat = torch.randn(1, 4, 24, 72, 128, generator=g).to(cuda)
    eps2 = torch.randn(2, 4, 24, 72, 128, generator=g).to(cuda)
    y1, _ = sm.fused_cfg_step(lat, eps2, t, 7.5)
    y2, _ = sm.fused_cfg_step(2 * lat, 2 * eps2, t, 7.5)
    assert float((y2 - 2 * y1).abs().max()) <= 1e-4 * float(y1.abs().max())


def test_xl_trajectory_commitment_bit_exact(cuda):
    """BASELINE config 5: 50 leaves of [1,4,24,72,128] fp16 x 2 (z, eps): 176.9 MB hashed on the device == hashlib."""
    from b2v import proof
    from oracle import merkle_ref as M
    n, shape = 50, (1, 4, 24, 72, 128)
    g = torch.Generator(device=cuda).manual_seed(50)
    zs = [torch.randn(shape, device=cuda, generator=g).half() for _ in range(n)]
    es = [torch.randn(shape, device=cuda, generator=g).half() for _ in range(n)]
    ts = [981 - 20 * i for i in range(n)]
    root, leaf = proof.build_merkle_tree(zs, es, ts)
    hashes = []
    for i, t in enumerate(ts):
        zb, eb = zs[i].cpu().numpy().tobytes(), es[i].cpu().numpy().tobytes()
        assert leaf[t][0] == zb and leaf[t][1] == eb and len(zb) == 1769472
        hashes.append(hashlib.sha256(int(t).to_bytes(2, "big") + zb + eb).digest())
    levels = M.tree_levels(hashes)
    assert root == levels[-1][0]
    paths = M.sibling_paths(levels)
    assert all(leaf[t][2] == paths[i] for i, t in enumerate(ts))
    # reveal round trip (proof.py:397-416 + 291-304): every revealed leaf verifies, a flipped byte does not
    idx = proof.commit_then_reveal_merkle_spotcheck(root, n, 5, 1234)
    assert idx == M.spotcheck_indices(n, 5, 1234)
    for i in idx:
        assert proof.verify_merkle_leaf(hashes[i], leaf[ts[i]][2], root)
    bad = bytearray(hashes[idx[0]])
    bad[0] ^= 1
    assert not proof.verify_merkle_leaf(bytes(bad), leaf[ts[idx[0]]][2], root)
