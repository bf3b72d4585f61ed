"""GPU box: A/B of the attn_kernel variants (B2V_ATTN_POLY, csrc/attention.cu): time (CUDA events, 5 launches after 2
warm-ups) and error vs torch fp32 on the benchmark's attention shapes.
usage: for v in 0 1 2 3; do B2V_ATTN_POLY=$v python tools/attn_ab.py; done"""
import os
import sys

import torch

sys.path.insert(0, ".")
import b2v  # noqa: E402,F401
from b2v import ops  # noqa: E402

dev = torch.device("cuda:0")
v = os.environ.get("B2V_ATTN_POLY", "default")
names = {"0": "all ex2 on the MUFU", "1": "1/8 of the pairs on the FMA pipe", "2": "2/8 on the FMA pipe", "3": "3/8 on the FMA pipe"}


def ref_attn(q, k, v, heads):
    nb, Sq, C = q.shape
    qh, kh, vh = (t.float().reshape(nb, -1, heads, 64).transpose(1, 2) for t in (q, k, v))
    o = torch.nn.functional.scaled_dot_product_attention(qh, kh, vh)
    return o.transpose(1, 2).reshape(nb, Sq, C)


for dt in (torch.bfloat16, torch.float16):
    for (nb, heads, Sq, Skv, amp) in ((32, 5, 9216, 9216, 1.0), (32, 10, 2304, 2304, 1.0), (32, 5, 9216, 77, 1.0),
                                      (4, 5, 9216, 9216, 4.0)):
        g = torch.Generator(device=dev).manual_seed(1)
        q = (torch.randn(nb, Sq, heads * 64, device=dev, generator=g) * amp).to(dt)
        k = (torch.randn(nb, Skv, heads * 64, device=dev, generator=g) * amp).to(dt)
        vv = torch.randn(nb, Skv, heads * 64, device=dev, generator=g).to(dt)
        for _ in range(2):
            out = ops.attention(q, k, vv, heads)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            out = ops.attention(q, k, vv, heads)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        n_chk = min(nb, 2)
        want = ref_attn(q[:n_chk], k[:n_chk], vv[:n_chk], heads)
        rel = float((out[:n_chk].float() - want).norm() / want.norm())
        print("variant %s (%s) %s nb=%d heads=%d Sq=%d Skv=%d amp=%g: %.3f ms  %.1f TFLOP/s  rel-l2 %.3e" % (
            v, names.get(v, "?"), str(dt)[6:], nb, heads, Sq, Skv, amp, ms, 4.0 * nb * heads * Sq * Skv * 64 / ms / 1e9, rel),
            flush=True)
