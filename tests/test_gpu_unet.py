"""-m gpu: UNet3DConditionModel forward (libb2v, bf16 tensor-core arithmetic) vs the CPU fp32 oracle restatement
(oracle/unet3d_ref.py) on the same seeded weights and inputs. Tolerances follow SURVEY.md Appendix D:
max-abs <= 2e-2 * max|ref| and cosine >= 0.9995 for a single forward."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _compare(out, ref):
    out, ref = out.float().cpu(), ref.float()
    err = float((out - ref).abs().max())
    cos = float(torch.nn.functional.cosine_similarity(out.flatten(), ref.flatten(), dim=0))
    rel = float((out - ref).norm() / ref.norm())
    return err, cos, rel


@pytest.fixture(scope="module")
def tiny_pair(cuda):
    from oracle import unet3d_ref as R
    from b2v.unet import UNet3DConditionModel
    torch.manual_seed(0)
    ref = R.random_init_(R.UNet3DConditionModel(R.TINY), seed=1234, round_bf16=True).eval()
    cfg = R.TINY.to_dict()
    net = UNet3DConditionModel(cfg, device=cuda).load_state_dict(ref.state_dict())
    return ref, net


@pytest.mark.parametrize("B,F,H,W,io", [(1, 2, 8, 8, torch.float32), (2, 4, 16, 16, torch.float32),
                                        (2, 5, 16, 24, torch.float16), (1, 8, 32, 32, torch.bfloat16)])
def test_unet_tiny_vs_oracle(cuda, tiny_pair, B, F, H, W, io):
    ref, net = tiny_pair
    g = torch.Generator().manual_seed(B * 1000 + F)
    x = torch.randn(B, 4, F, H, W, generator=g).to(io).float()
    ehs = torch.randn(B, 77, ref.cfg.cross_attention_dim, generator=g).to(io).float()
    t = 501
    with torch.no_grad():
        want = ref(x, t, ehs)
    got = net(x.to(cuda, io), torch.tensor(t), encoder_hidden_states=ehs.to(cuda, io)).sample
    assert got.shape == want.shape and got.dtype == io
    err, cos, rel = _compare(got, want)
    print("unet tiny", (B, F, H, W), "max-abs", err, "cos", cos, "rel-l2", rel, "max|ref|", float(want.abs().max()))
    assert cos >= 0.9995, (err, cos, rel)
    assert err <= 2e-2 * float(want.abs().max()) + (4e-3 if io != torch.float32 else 0.0), (err, cos, rel)


def test_unet_param_count_and_flops(cuda, tiny_pair):
    ref, net = tiny_pair
    assert net.param_count() == sum(p.numel() for p in ref.parameters())
    assert net.flops(2, 4, 16, 16) > 0
    assert net.workspace_bytes(2, 4, 16, 16) > 0


def test_unet_rejects_bad_shapes(cuda, tiny_pair):
    from b2v import B2VError
    _, net = tiny_pair
    x = torch.zeros(1, 4, 2, 12, 12, device=cuda)  # not a multiple of 8
    ehs = torch.zeros(1, 77, 128, device=cuda)
    with pytest.raises(B2VError):
        net(x, 1, encoder_hidden_states=ehs)


def test_unet_deterministic(cuda, tiny_pair):
    _, net = tiny_pair
    g = torch.Generator().manual_seed(5)
    x = torch.randn(2, 4, 4, 16, 16, generator=g).to(cuda)
    ehs = torch.randn(2, 77, 128, generator=g).to(cuda)
    a = net(x, 251, encoder_hidden_states=ehs).sample
    b = net(x, 251, encoder_hidden_states=ehs).sample
    assert torch.equal(a, b)
