"""GPU box: launch each hot kernel ONCE on its dominant shape (for `ncu --set full`); prints nothing but timings.
usage: python tools/ncu_targets.py [gemm|attn|norm|all]"""
import math
import sys

import torch

sys.path.insert(0, ".")
import b2v  # noqa: E402
from b2v import ops, _lib  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "all"
dev = torch.device("cuda:0")
M = 294912


def t(fn, name):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    print("%-40s %.3f ms" % (name, e0.elapsed_time(e1)), flush=True)


if what in ("gemm", "gemm2", "all"):
    shapes = [(M, 960, 320, 0, 0), (M, 320, 320, 0, 1), (M, 2560, 320, 1, 0), (18432, 3840, 1280, 0, 0)]
    if what == "gemm2":
        shapes = [(73728, 640, 640, 0, 1), (73728, 1920, 640, 0, 0), (18432, 1280, 1280, 0, 1), (73728, 5120, 640, 1, 0)]
    for (m, n, k, flags, r) in shapes:
        a = torch.randn(m, k, device=dev).bfloat16()
        w = (torch.randn(n, k, device=dev) / math.sqrt(k)).bfloat16()
        bias = torch.zeros(n, device=dev)
        out = torch.randn(m, n // 2 if flags & 1 else n, device=dev).bfloat16()
        t(lambda: ops.gemm(a, w, bias=bias, residual=out if r else None, flags=flags, out=out), "gemm %d %d %d f%d r%d" % (m, n, k, flags, r))
if what in ("attn", "all"):
    nb, heads, S = 32, 5, 9216  # the shape bench.py's roofline times (2 x 16 frames)
    q, k, v = (torch.randn(nb, S, heads * 64, device=dev).bfloat16() for _ in range(3))
    t(lambda: ops.attention(q, k, v, heads), "attention nb=32 h=5 S=9216")
if what in ("norm", "all"):
    x = torch.randn(M, 320, device=dev).bfloat16()
    g, b = torch.ones(320, device=dev), torch.zeros(320, device=dev)
    t(lambda: ops.group_norm(x, g, b, 32, 32, 1e-5, True), "group_norm+silu L0 (items=32)")
    t(lambda: ops.layer_norm(x, g, b), "layer_norm L0")
if what in ("sha", "all"):
    # the commitment of one 50-step ZeroscopeXL trajectory: 50 leaves of be16(t) || z fp16 || eps fp16 (3 538 946 B each)
    elems = 4 * 24 * 72 * 128
    zs = torch.randn(50, elems, device=dev).half()
    es = torch.randn(50, elems, device=dev).half()
    zl, el = [zs[i] for i in range(50)], [es[i] for i in range(50)]
    t(lambda: ops.sha256_leaves(zl, el, prefixes=list(range(50)), mode=1), "sha256 chain kernel, 50 XL leaves")
    t(lambda: ops.sha256_leaves(zl[:1], el[:1], prefixes=[981], mode=3), "sha256 warp kernel, 1 XL leaf")
if what in ("fuse", "all"):
    # round 2 fusions at the level-0 shape (2 x 16 frames of 72 x 128, C = 320):
    # (a) proj_out-like GEMM emitting GroupNorm statistics -> group_norm_pre (no statistics pass): traffic of the norm
    # (b) to_out-like GEMM emitting row partials -> QKV GEMM that finishes the LayerNorm in its epilogue
    a = torch.randn(M, 320, device=dev).bfloat16()
    w = (torch.randn(320, 320, device=dev) / math.sqrt(320)).bfloat16()
    bias = torch.zeros(320, device=dev)
    res = torch.randn(M, 320, device=dev).bfloat16()
    g, b = torch.ones(320, device=dev), torch.zeros(320, device=dev)
    out, st = ops.gemm_gnstats(a, w, bias=bias, residual=res, groups=32)
    t(lambda: ops.gemm_gnstats(a, w, bias=bias, residual=res, groups=32), "gemm N=320 K=320 res + GN stats")
    t(lambda: ops.group_norm_pre(out, g, b, 32, 32, 1e-5, True, st), "group_norm_pre+silu L0 (items=32, 288 blocks)")
    t(lambda: ops.group_norm_pre(out, g, b, 2, 32, 1e-5, True, st), "group_norm_pre+silu L0 temporal (items=2: prefold)")
    h, rs = ops.gemm_rowstats(a, w, bias=bias, residual=res)
    t(lambda: ops.gemm_rowstats(a, w, bias=bias, residual=res), "gemm N=320 K=320 res + row stats")
    w3 = (torch.randn(960, 320, device=dev) / math.sqrt(320)).bfloat16()
    cs = w3.float().sum(1)
    b3 = torch.zeros(960, device=dev)
    t(lambda: ops.gemm_lnin(h, w3, b3, cs, rs), "gemm N=960 K=320 LayerNorm in the epilogue")
    t(lambda: ops.gemm(h, w3), "gemm N=960 K=320 plain (for comparison)")
