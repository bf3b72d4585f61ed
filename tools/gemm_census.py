"""GPU box: run one UNet forward with B2V_TRACE=1 in a child process to collect the GEMM / conv shape census, then time
every distinct shape standalone (CUDA events) and rank by total time per forward.
usage: python tools/gemm_census.py [B F H W]"""
import collections
import math
import os
import re
import subprocess
import sys

import torch

sys.path.insert(0, ".")
args = [a for a in sys.argv[1:] if not a.startswith("--")]
B, F, H, W = (int(a) for a in args) if len(args) == 4 else (2, 16, 72, 128)

if os.environ.get("B2V_CENSUS_CHILD"):
    import b2v  # noqa
    from b2v.pipeline import VideoPipeline
    pipe = VideoPipeline.from_random(device="cuda:0")
    x = torch.randn(B, 4, F, H, W, device="cuda:0", dtype=torch.bfloat16)
    e = torch.randn(B, 77, 1024, device="cuda:0", dtype=torch.bfloat16)
    sys.stderr.write("B2VTRACE begin\n")
    pipe.unet(x, 501, encoder_hidden_states=e)
    torch.cuda.synchronize()
    sys.exit(0)

env = dict(os.environ, B2V_TRACE="1", B2V_CENSUS_CHILD="1")
r = subprocess.run([sys.executable, __file__] + args, env=env, capture_output=True, text=True)
lines = r.stderr.split("\n")
lines = lines[lines.index("B2VTRACE begin") + 1:]
cnt = collections.Counter(l for l in lines if l.startswith("B2VTRACE "))
import b2v  # noqa
from b2v import ops, _lib
dev = torch.device("cuda:0")


def timeit(fn, n=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


rows = []
for line, c in cnt.items():
    kv = dict(re.findall(r"(\w+)=(-?\d+)", line))
    kv = {k: int(v) for k, v in kv.items()}
    if " gemm " in line:
        M, N, K, taps, flags = kv["M"], kv["N"], kv["K"], kv["taps"], kv["flags"]
        a = torch.randn(M, K, device=dev).bfloat16()
        w = (torch.randn(N, taps * K, device=dev) / math.sqrt(taps * K)).bfloat16()
        bias = torch.zeros(N, device=dev)
        res = torch.randn(M, N // 2 if flags & 1 else N, device=dev).bfloat16() if kv["res"] else None
        rpb = kv["rpb"] if taps == 3 else 0
        ms = timeit(lambda: ops.gemm(a, w, bias=bias, residual=res, taps=taps, tap_row_stride=kv["trs"], rows_per_batch=rpb, flags=flags))
        fl = 2.0 * M * N * K * taps
        by = 2.0 * (M * K + N * K * taps + M * (N // 2 if flags & 1 else N) * (2 if kv["res"] else 1))
        name = "gemm M=%d N=%d K=%d taps=%d f=%d r=%d" % (M, N, K, taps, flags, kv["res"])
    else:
        n, h, w_, ci, co, st = kv["N"], kv["H"], kv["W"], kv["Cin"], kv["Cout"], kv["stride"]
        x = torch.randn(n, h, w_, ci, device=dev).bfloat16()
        wt = (torch.randn(co, 3, 3, ci, device=dev) / math.sqrt(9 * ci)).bfloat16()
        bias = torch.zeros(co, device=dev)
        ms = timeit(lambda: ops.conv3x3(x, wt, bias=bias, stride=st))
        P, Q = (h - 1) // st + 1, (w_ - 1) // st + 1
        fl = 2.0 * n * P * Q * co * 9 * ci
        by = 2.0 * (n * h * w_ * ci + co * 9 * ci + n * P * Q * co)
        name = "conv N=%d %dx%d %d->%d s%d" % (n, h, w_, ci, co, st)
    rows.append((c * ms, c, ms, fl / ms / 1e9, by / ms / 1e6, name))
rows.sort(reverse=True)
tot = sum(r[0] for r in rows)
print("total %.1f ms over %d distinct shapes" % (tot, len(rows)))
print("%8s %4s %9s %8s %8s  %s" % ("tot ms", "cnt", "ms each", "TFLOP/s", "GB/s", "shape"))
for r in rows:
    print("%8.2f %4d %9.3f %8.0f %8.0f  %s" % r)
