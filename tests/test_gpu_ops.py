"""-m gpu: every building-block kernel of libb2v, called through the C-ABI, against a plain torch fp32 reference of the
same op (floating-point kernels; tolerance written at each assert). Every test runs in both activation dtypes of the
library: bf16 (default) and fp16 (the reference's torch_dtype, chunk_only.py:44-47). The 2-byte inputs are shared by
both sides, so the error budget is fp32-accumulate + one output rounding (bf16: 2^-8 relative; fp16: 2^-11, gated 4x
tighter) unless stated. The largest cases are the ZeroscopeXL shapes the benchmark runs (72 x 128 latent: 9216 tokens)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def rel_l2(a, b):
    a, b = a.float(), b.float()
    return float((a - b).norm() / b.norm().clamp_min(1e-12))


@pytest.fixture(params=[torch.bfloat16, torch.float16], ids=["bf16", "fp16"])
def act(request):
    return request.param


class _Cast:
    """A(x): cast to the activation dtype of the running test (set by the autouse fixture below)."""
    dtype = torch.bfloat16

    def __call__(self, x):
        return x.to(self.dtype)


A = _Cast()


@pytest.fixture(autouse=True)
def _set_cast(act):
    A.dtype = act
    yield


def tol(act, bf16_tol):
    """Gate of a check whose error is dominated by the output rounding: fp16 has 3 more mantissa bits than bf16."""
    return bf16_tol if act == torch.bfloat16 else bf16_tol / 4


def pack_geglu(w, b):
    """[2H,K] (value rows then gate rows) -> kernel layout: per 256-row tile, 128 value rows then 128 gate rows."""
    H = w.shape[0] // 2
    idx = []
    for j in range(H // 128):
        idx += list(range(j * 128, (j + 1) * 128)) + list(range(H + j * 128, H + (j + 1) * 128))
    idx = torch.tensor(idx, device=w.device)
    return w[idx].contiguous(), b[idx].contiguous()


@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (256, 128, 128), (1000, 320, 192), (4096, 256, 320),
                                   (300, 640, 1280), (77 * 2, 2560, 1024), (20000, 1280, 640)])
def test_gemm_plain(cuda, act, M, N, K):
    from b2v import ops
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    a = A(torch.randn(M, K, device=cuda, generator=g))
    w = A(torch.randn(N, K, device=cuda, generator=g) / math.sqrt(K))
    bias = torch.randn(N, device=cuda, generator=g)
    out = ops.gemm(a, w, bias=bias)
    ref = a.float() @ w.float().t() + bias
    assert out.shape == (M, N)
    assert rel_l2(out, ref) < tol(act, 4e-3), rel_l2(out, ref)
    assert float((out.float() - ref).abs().max()) < 0.05 * float(ref.abs().max())


def test_gemm_residual_rowbias(cuda, act):
    from b2v import ops
    g = torch.Generator(device="cuda").manual_seed(7)
    B, rows_per_group, N, K = 3, 200, 320, 256
    M = B * rows_per_group
    a = A(torch.randn(M, K, device=cuda, generator=g))
    w = A(torch.randn(N, K, device=cuda, generator=g) / math.sqrt(K))
    bias = torch.randn(N, device=cuda, generator=g)
    rowbias = torch.randn(B, N, device=cuda, generator=g)
    res = A(torch.randn(M, N, device=cuda, generator=g))
    out = ops.gemm(a, w, bias=bias, rowbias=rowbias, rows_per_group=rows_per_group, residual=res,
                   rows_per_batch=rows_per_group)
    ref = a.float() @ w.float().t() + bias + rowbias.repeat_interleave(rows_per_group, 0) + res.float()
    assert rel_l2(out, ref) < tol(act, 4e-3), rel_l2(out, ref)


@pytest.mark.parametrize("M,dim", [(256, 64), (1000, 320), (640, 1280)])
def test_gemm_geglu(cuda, act, M, dim):
    from b2v import ops, _lib
    g = torch.Generator(device="cuda").manual_seed(dim)
    a = A(torch.randn(M, dim, device=cuda, generator=g))
    w = A(torch.randn(8 * dim, dim, device=cuda, generator=g) / math.sqrt(dim))
    b = torch.randn(8 * dim, device=cuda, generator=g)
    wp, bp = pack_geglu(w, b)
    out = ops.gemm(a, wp, bias=bp, flags=_lib.EPI_GEGLU)
    proj = a.float() @ w.float().t() + b
    h, gate = proj.chunk(2, dim=-1)
    ref = h * torch.nn.functional.gelu(gate)
    assert out.shape == (M, 4 * dim)
    assert rel_l2(out, ref) < tol(act, 5e-3), rel_l2(out, ref)


@pytest.mark.parametrize("B,F,HW,C", [(1, 4, 128, 64), (2, 5, 100, 128), (2, 8, 256, 320), (1, 16, 45, 1280)])
def test_gemm_temporal_taps(cuda, act, B, F, HW, C):
    """TemporalConvLayer Conv3d (3,1,1), zero padded in time, as a 3-tap GEMM over row strides."""
    from b2v import ops
    g = torch.Generator(device="cuda").manual_seed(B * 100 + F)
    x = A(torch.randn(B, F, HW, C, device=cuda, generator=g))
    w3 = A(torch.randn(C, C, 3, device=cuda, generator=g) / math.sqrt(3 * C))  # [Cout][Cin][tap]
    bias = torch.randn(C, device=cuda, generator=g)
    wk = w3.permute(0, 2, 1).contiguous().reshape(C, 3 * C)  # [Cout][tap][Cin]
    out = ops.gemm(x.reshape(-1, C), wk, bias=bias, taps=3, tap_row_stride=HW, rows_per_batch=F * HW)
    xr = x.float().permute(0, 3, 1, 2).reshape(B, C, F, HW, 1)
    ref = torch.nn.functional.conv3d(xr, w3.float().reshape(C, C, 3, 1, 1), bias, padding=(1, 0, 0))
    ref = ref.reshape(B, C, F, HW).permute(0, 2, 3, 1).reshape(-1, C)
    assert rel_l2(out, ref) < tol(act, 4e-3), rel_l2(out, ref)


@pytest.mark.parametrize("mode", [1, 0], ids=["explicit", "tmaim2col"])
@pytest.mark.parametrize("N,H,W,Cin,Cout,stride", [(2, 16, 16, 64, 64, 1), (3, 20, 36, 128, 320, 1),
                                                   (2, 16, 24, 64, 128, 2), (1, 8, 8, 640, 1280, 1),
                                                   (2, 40, 72, 320, 320, 1), (2, 10, 18, 320, 320, 2),
                                                   (2, 72, 128, 320, 320, 1), (1, 72, 128, 960, 320, 1)])
def test_conv3x3(cuda, act, mode, N, H, W, Cin, Cout, stride):
    from b2v import ops
    g = torch.Generator(device="cuda").manual_seed(H * W + Cin)
    x = A(torch.randn(N, H, W, Cin, device=cuda, generator=g))
    w = A(torch.randn(Cout, 3, 3, Cin, device=cuda, generator=g) / math.sqrt(9 * Cin))
    bias = torch.randn(Cout, device=cuda, generator=g)
    out = ops.conv3x3(x, w, bias=bias, stride=stride, mode=mode)
    ref = torch.nn.functional.conv2d(x.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), bias,
                                     stride=stride, padding=1).permute(0, 2, 3, 1)
    assert out.shape == ref.shape
    assert rel_l2(out, ref) < tol(act, 4e-3), (mode, rel_l2(out, ref))


def test_conv3x3_rowbias_residual(cuda, act):
    from b2v import ops
    g = torch.Generator(device="cuda").manual_seed(3)
    B, F, H, W, Cin, Cout = 2, 3, 16, 16, 64, 128
    x = A(torch.randn(B * F, H, W, Cin, device=cuda, generator=g))
    w = A(torch.randn(Cout, 3, 3, Cin, device=cuda, generator=g) / math.sqrt(9 * Cin))
    bias = torch.randn(Cout, device=cuda, generator=g)
    rowbias = torch.randn(B, Cout, device=cuda, generator=g)
    res = A(torch.randn(B * F, H, W, Cout, device=cuda, generator=g))
    out = ops.conv3x3(x, w, bias=bias, rowbias=rowbias, rows_per_group=F * H * W, residual=res)
    ref = torch.nn.functional.conv2d(x.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), bias,
                                     padding=1).permute(0, 2, 3, 1)
    ref = ref + rowbias.repeat_interleave(F, 0)[:, None, None, :] + res.float()
    assert rel_l2(out, ref) < tol(act, 4e-3), rel_l2(out, ref)


@pytest.mark.parametrize("items,rows,C,silu,eps", [(4, 256, 64, 1, 1e-5), (6, 1024, 320, 1, 1e-5),
                                                   (2, 2880, 640, 0, 1e-6), (3, 45, 1280, 1, 1e-5)])
def test_group_norm(cuda, act, items, rows, C, silu, eps):
    from b2v import ops
    g = torch.Generator(device="cuda").manual_seed(C)
    x = A(torch.randn(items * rows, C, device=cuda, generator=g) * 2 + 0.5)
    gamma = 1 + 0.1 * torch.randn(C, device=cuda, generator=g)
    beta = 0.1 * torch.randn(C, device=cuda, generator=g)
    y = ops.group_norm(x, gamma, beta, items, 32, eps, silu)
    xr = x.float().reshape(items, rows, C).permute(0, 2, 1)
    ref = torch.nn.functional.group_norm(xr, 32, gamma, beta, eps)
    if silu:
        ref = torch.nn.functional.silu(ref)
    ref = ref.permute(0, 2, 1).reshape(-1, C)
    # tolerance: one bf16 rounding of an O(1) value
    # one output rounding of an O(1..8) value + tanh.approx (2^-11 relative) inside the fused SiLU
    assert float((y.float() - ref).abs().max()) < (0.03 if act == torch.bfloat16 else 0.012)
    assert rel_l2(y, ref) < tol(act, 4e-3)


def test_group_norm_concat(cuda, act):
    from b2v import ops
    g = torch.Generator(device="cuda").manual_seed(11)
    items, rows, C1, C2 = 4, 180, 1280, 640
    x1 = A(torch.randn(items * rows, C1, device=cuda, generator=g))
    x2 = A(torch.randn(items * rows, C2, device=cuda, generator=g) * 3)
    C = C1 + C2
    gamma = 1 + 0.1 * torch.randn(C, device=cuda, generator=g)
    beta = 0.1 * torch.randn(C, device=cuda, generator=g)
    y = ops.group_norm(x1, gamma, beta, items, 32, 1e-5, 1, x2=x2)
    xr = torch.cat([x1, x2], 1).float().reshape(items, rows, C).permute(0, 2, 1)
    ref = torch.nn.functional.silu(torch.nn.functional.group_norm(xr, 32, gamma, beta, 1e-5))
    ref = ref.permute(0, 2, 1).reshape(-1, C)
    assert rel_l2(y, ref) < tol(act, 4e-3)


@pytest.mark.parametrize("rows,C", [(1000, 64), (4096, 320), (333, 512), (512, 1280)])
def test_layer_norm(cuda, act, rows, C):
    from b2v import ops
    g = torch.Generator(device="cuda").manual_seed(C)
    x = A(torch.randn(rows, C, device=cuda, generator=g) * 3 + 1)
    gamma = 1 + 0.1 * torch.randn(C, device=cuda, generator=g)
    beta = 0.1 * torch.randn(C, device=cuda, generator=g)
    y = ops.layer_norm(x, gamma, beta)
    ref = torch.nn.functional.layer_norm(x.float(), (C,), gamma, beta, 1e-5)
    assert rel_l2(y, ref) < tol(act, 4e-3)


@pytest.mark.parametrize("nb,heads,Sq,Skv,kv_group", [(2, 1, 128, 128, 1), (3, 2, 256, 256, 1), (4, 5, 1000, 1000, 1),
                                                      (6, 2, 300, 77, 3), (2, 20, 180, 180, 1), (1, 5, 2880, 2880, 1),
                                                      (2, 5, 9216, 9216, 1), (2, 5, 9216, 77, 2)])
def test_attention(cuda, act, nb, heads, Sq, Skv, kv_group):
    from b2v import ops
    g = torch.Generator(device="cuda").manual_seed(Sq + Skv)
    D = heads * 64
    q = A(torch.randn(nb, Sq, D, device=cuda, generator=g))
    k = A(torch.randn(nb // kv_group, Skv, D, device=cuda, generator=g))
    v = A(torch.randn(nb // kv_group, Skv, D, device=cuda, generator=g))
    out = ops.attention(q, k, v, heads, kv_group=kv_group)
    qh = q.float().reshape(nb, Sq, heads, 64).transpose(1, 2)
    kh = k.float().reshape(-1, Skv, heads, 64).transpose(1, 2).repeat_interleave(kv_group, 0)
    vh = v.float().reshape(-1, Skv, heads, 64).transpose(1, 2).repeat_interleave(kv_group, 0)
    ref = torch.softmax(qh @ kh.transpose(-1, -2) * 0.125, -1) @ vh
    ref = ref.transpose(1, 2).reshape(nb, Sq, D)
    # P is rounded to the activation dtype before the PV product: 2^-8 (bf16) / 2^-11 (fp16) relative per probability
    err = rel_l2(out, ref)
    print("attention", (nb, heads, Sq, Skv), act, "rel-l2 %.3e" % err)
    assert err < tol(act, 1e-2), err


@pytest.mark.parametrize("amp,drift", [(4.0, 0.0), (10.0, 0.0), (2.0, 40.0), (1.0, 400.0)])
def test_attention_sharp_rows_and_moving_maximum(cuda, act, amp, drift):
    """The softmax paths the unit-variance test does not reach: near one-hot rows (amp), a row maximum that keeps growing
    along the kv axis (drift: every kv tile beats the reference of the one before it, so O is rescaled in TMEM tile after
    tile) and score ranges beyond the exponent field (drift 400: the FMA-pipe exponentials clamp at 2^-125, the MUFU ones
    flush to zero, both vanish against the row's maximum)."""
    from b2v import ops
    g = torch.Generator(device="cuda").manual_seed(int(amp * 10 + drift))
    nb, heads, Sq, Skv = 2, 2, 384, 1024
    D = heads * 64
    q = torch.randn(nb, Sq, D, device=cuda, generator=g) * amp
    k = torch.randn(nb, Skv, D, device=cuda, generator=g)
    if drift:
        # a component of k that grows with the kv index, aligned with a constant component of q: scores rise by `drift`
        # (in logit units) from the first to the last kv position
        q[:, :, 0::64] = 8.0
        k[:, :, 0::64] = torch.linspace(0, drift, Skv, device=cuda)[None, :, None] / 8.0 * 8.0
    v = torch.randn(nb, Skv, D, device=cuda, generator=g)
    q, k, v = A(q), A(k), A(v)
    out = ops.attention(q, k, v, heads)
    qh, kh, vh = (t.float().reshape(nb, -1, heads, 64).transpose(1, 2) for t in (q, k, v))
    ref = (torch.softmax(qh @ kh.transpose(-1, -2) * 0.125, -1) @ vh).transpose(1, 2).reshape(nb, Sq, D)
    assert bool(torch.isfinite(out).all())
    err = rel_l2(out, ref)
    print("attention sharp", act, (amp, drift), "rel-l2 %.3e" % err)
    assert err < tol(act, 1.2e-2), err


@pytest.mark.parametrize("B,F,HW,heads", [(1, 4, 64, 1), (2, 8, 100, 5), (2, 16, 45, 20), (1, 24, 33, 8), (1, 1, 16, 2)])
def test_temporal_attention(cuda, act, B, F, HW, heads):
    from b2v import ops
    g = torch.Generator(device="cuda").manual_seed(F * HW)
    D = heads * 64
    qkv = A(torch.randn(B * F * HW, 3 * D, device=cuda, generator=g))
    out = ops.temporal_attention(qkv, B, F, HW, heads)
    t = qkv.float().reshape(B, F, HW, 3, heads, 64).permute(3, 0, 2, 4, 1, 5)  # [3][B][HW][heads][F][64]
    q, k, v = t[0], t[1], t[2]
    ref = torch.softmax(q @ k.transpose(-1, -2) * 0.125, -1) @ v  # [B][HW][heads][F][64]
    ref = ref.permute(0, 3, 1, 2, 4).reshape(B * F * HW, D)
    assert rel_l2(out, ref) < tol(act, 4e-3), rel_l2(out, ref)


# ---- GroupNorm statistics out of the producing GEMM's epilogue -------------------------------------
def _ref_block_stats(out_f32, groups):
    M, N = out_f32.shape
    v = out_f32.reshape(M // 32, 32, groups, N // groups).double()
    return torch.stack([v.sum(dim=(1, 3)), (v * v).sum(dim=(1, 3))], dim=-1)


@pytest.mark.parametrize("adt", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("M,N,K,taps,res,items", [(1024, 320, 320, 1, False, 4), (2048, 640, 320, 1, True, 2),
                                                   (1536, 320, 64, 3, True, 2), (36864, 320, 128, 1, False, 2),
                                                   (2304, 1280, 128, 1, False, 1),
                                                   # SD VAE decoder widths: 256-wide tiles with 16 / 8 groups per epilogue
                                                   # warp, 128-wide tiles with 16
                                                   (37888, 256, 64, 1, True, 2), (37888, 512, 64, 1, False, 4),
                                                   (2048, 128, 64, 1, True, 1)])
def test_gemm_emits_group_norm_stats(cuda, adt, M, N, K, taps, res, items):
    """gemm_gnstats: same output bits as gemm(); statistics = {sum, sumsq} per 32-row block and group of the fp32 results
    (before rounding); group_norm_pre on them == the two-pass group_norm within rounding of the statistics, and both
    within the op gate of torch fp32. (36864 rows / 2 items = 576 blocks per item: the pre-fold kernel path.)"""
    from b2v import ops
    g = torch.Generator(device=cuda).manual_seed(M + N + taps)
    rpb = M // items if taps == 3 else 0
    if not ops.gn_stats_ok(M, N, rpb, 32):
        pytest.skip("the cost model picks a tile width without a statistics epilogue for this shape")
    a = torch.randn(M, K, device=cuda, generator=g).to(adt)
    w = (torch.randn(N, taps * K, device=cuda, generator=g) / math.sqrt(taps * K)).to(adt)
    bias = torch.randn(N, device=cuda, generator=g)
    r = torch.randn(M, N, device=cuda, generator=g).to(adt) if res else None
    kw = dict(bias=bias, residual=r, taps=taps, tap_row_stride=(M // items) // 4 if taps == 3 else 0, rows_per_batch=rpb)
    plain = ops.gemm(a, w, **kw)
    out, stats = ops.gemm_gnstats(a, w, groups=32, **kw)
    assert torch.equal(out, plain)
    want = _ref_block_stats(out.float(), 32)          # from the ROUNDED output: differs by the rounding noise only
    err = (stats.double() - want).abs().max() / want.abs().max()
    print("gn stats", adt, (M, N, K, taps, res), "max rel err vs stats of the rounded output %.3e" % float(err))
    assert float(err) < (2e-3 if adt == torch.bfloat16 else 3e-4)
    gam = 1 + 0.1 * torch.randn(N, device=cuda, generator=g)
    bet = 0.1 * torch.randn(N, device=cuda, generator=g)
    for silu in (0, 1):
        two_pass = ops.group_norm(out, gam, bet, items, 32, 1e-5, silu)
        fused = ops.group_norm_pre(out, gam, bet, items, 32, 1e-5, silu, stats)
        ref = torch.nn.functional.group_norm(out.float().reshape(items, M // items, N).permute(0, 2, 1), 32, gam, bet, 1e-5)
        ref = (torch.nn.functional.silu(ref) if silu else ref).permute(0, 2, 1).reshape(M, N)
        gate = 1e-2 if adt == torch.bfloat16 else 2.5e-3
        assert rel_l2(fused, ref) < gate and rel_l2(two_pass, ref) < gate
        assert rel_l2(fused, two_pass) < gate / 2
        assert torch.equal(fused, ops.group_norm_pre(out, gam, bet, items, 32, 1e-5, silu, stats))   # bit-reproducible


@pytest.mark.parametrize("adt", [torch.bfloat16, torch.float16])
def test_conv3x3_emits_group_norm_stats(cuda, adt):
    from b2v import ops
    g = torch.Generator(device=cuda).manual_seed(9)
    Nimg, H, W, Cin, Cout = 2, 16, 32, 64, 320
    assert ops.gn_stats_ok(Nimg * H * W, Cout, 0, 32)
    x = torch.randn(Nimg, H, W, Cin, device=cuda, generator=g).to(adt)
    w = (torch.randn(Cout, 3, 3, Cin, device=cuda, generator=g) / 24.0).to(adt)
    bias = torch.randn(Cout, device=cuda, generator=g)
    out, stats = ops.conv3x3_gnstats(x, w, bias=bias)
    assert torch.equal(out, ops.conv3x3(x, w, bias=bias))
    want = _ref_block_stats(out.float().reshape(-1, Cout), 32)
    err = float((stats.double() - want).abs().max() / want.abs().max())
    assert err < (2e-3 if adt == torch.bfloat16 else 3e-4), err
    gam, bet = torch.ones(Cout, device=cuda), torch.zeros(Cout, device=cuda)
    rows = out.reshape(-1, Cout)
    a = ops.group_norm_pre(rows, gam, bet, Nimg, 32, 1e-5, 1, stats)
    b = ops.group_norm(rows, gam, bet, Nimg, 32, 1e-5, 1)
    assert rel_l2(a, b) < 5e-3


# ---- LayerNorm fused across two GEMMs (producer emits row partials, consumer finishes the normalisation) ---------------
@pytest.mark.parametrize("adt", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("M,C,Nout,res,geglu", [(512, 320, 960, False, False), (1000, 320, 320, True, False),
                                                (2048, 640, 1920, True, False), (768, 1280, 1280, False, False),
                                                (1024, 512, 1536, True, False), (512, 320, 2560, True, True),
                                                (40000, 320, 960, True, False)])
def test_layer_norm_fused_across_gemms(cuda, adt, M, C, Nout, res, geglu):
    """h = a W1^T + b1 (+ h0)  [row partials out];  y = LayerNorm(h) W2^T + b2  [normalisation in the epilogue]
    vs the unfused chain gemm -> layer_norm -> gemm and vs torch fp32 on the same rounded h."""
    from b2v import ops, _lib
    g = torch.Generator(device=cuda).manual_seed(M + C + Nout)
    a = torch.randn(M, 128, device=cuda, generator=g).to(adt)
    w1 = (torch.randn(C, 128, device=cuda, generator=g) / math.sqrt(128)).to(adt)
    b1 = torch.randn(C, device=cuda, generator=g) * 0.5 + 0.3          # a row mean that is not zero
    h0 = torch.randn(M, C, device=cuda, generator=g).to(adt) if res else None
    h, stats = ops.gemm_rowstats(a, w1, bias=b1, residual=h0)
    assert torch.equal(h, ops.gemm(a, w1, bias=b1, residual=h0))
    hf = h.float()
    tot = stats.sum(dim=1)                                              # [M, 2]
    assert rel_l2(tot[:, 0], hf.sum(1)) < 2e-3 and rel_l2(tot[:, 1], (hf * hf).sum(1)) < 2e-3
    gamma = 1 + 0.2 * torch.randn(C, device=cuda, generator=g)
    beta = 0.2 * torch.randn(C, device=cuda, generator=g)
    w2 = torch.randn(Nout, C, device=cuda, generator=g) / math.sqrt(C)
    b2 = torch.randn(Nout, device=cuda, generator=g)
    w2g = (w2 * gamma[None, :]).to(adt)                                 # W' = W diag(gamma), rounded once
    csum = w2g.float().sum(1)
    b2f = b2 + w2 @ beta
    flags = 0
    if geglu:
        _, csum = pack_geglu(w2, csum)           # vectors are permuted like the rows of W
        w2g, b2f = pack_geglu(w2g, b2f)
        flags = _lib.EPI_GEGLU
    y = ops.gemm_lnin(h, w2g, b2f, csum, stats, eps=1e-5, flags=flags)
    # reference: torch fp32 on the same (rounded) hidden states
    ln = torch.nn.functional.layer_norm(hf, (C,), gamma, beta, 1e-5)
    proj = ln @ w2.t() + b2
    if geglu:
        v, gt = proj.chunk(2, dim=-1)
        ref = v * torch.nn.functional.gelu(gt)
    else:
        ref = proj
    # unfused chain of the same library
    l = ops.layer_norm(h, gamma, beta)
    if geglu:
        wp, bp = pack_geglu(w2.to(adt), b2)
        chain = ops.gemm(l, wp, bias=bp, flags=_lib.EPI_GEGLU)
    else:
        chain = ops.gemm(l, w2.to(adt), bias=b2)
    e_f, e_c = rel_l2(y, ref), rel_l2(chain, ref)
    print("ln fused", adt, (M, C, Nout, res, geglu), "fused %.3e  unfused %.3e" % (e_f, e_c))
    assert e_f < tol(adt, 8e-3), e_f
    assert e_f < 1.5 * e_c + 1e-4          # no worse than the chain it replaces
    assert torch.equal(y, ops.gemm_lnin(h, w2g, b2f, csum, stats, eps=1e-5, flags=flags))
