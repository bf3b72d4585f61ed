"""not-gpu: structural anchors of the UNet3D oracle restatement (parity with diffusers is UNPINNED — SURVEY.md §8c —
so these are the anchors that exist): parameter total, state_dict key set vs the product's shape table, shape round
trip, frame-permutation equivariance of the temporal transformer (no positional encoding), zero-init conv4 identity."""
import torch

import b2v  # noqa: F401
from b2v.pipeline import parameter_shapes, ZEROSCOPE_CONFIG
from oracle import unet3d_ref as R


def test_param_count_matches_zeroscope():
    assert R.param_count(R.ZEROSCOPE) == 1411233860


def test_state_dict_keys_equal_product_table():
    with torch.device("meta"):
        m = R.UNet3DConditionModel(R.ZEROSCOPE)
    assert {k: tuple(v.shape) for k, v in m.state_dict().items()} == parameter_shapes(ZEROSCOPE_CONFIG)


def test_forward_shape_and_temporal_properties():
    torch.manual_seed(0)
    m = R.random_init_(R.UNet3DConditionModel(R.TINY), seed=3).eval()
    x = torch.randn(2, 4, 3, 8, 8)
    ehs = torch.randn(2, 77, 128)
    with torch.no_grad():
        y = m(x, 501, ehs)
        assert y.shape == x.shape and bool(torch.isfinite(y).all())
        # TransformerTemporalModel alone is permutation-equivariant over frames
        tt = m.transformer_in
        h = torch.randn(2 * 3, 64, 4, 4)
        perm = torch.tensor([2, 0, 1])
        a = tt(h, 3).reshape(2, 3, 64, 4, 4)[:, perm]
        b = tt(h.reshape(2, 3, 64, 4, 4)[:, perm].reshape(6, 64, 4, 4), 3).reshape(2, 3, 64, 4, 4)
        assert torch.allclose(a, b, atol=1e-5)
    # upstream zero-init: a freshly constructed TemporalConvLayer is the identity
    tc = R.TemporalConvLayer(64, 32).eval()
    with torch.no_grad():
        assert torch.equal(tc(h, 3), h)
