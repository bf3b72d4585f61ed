"""GPU box: time one full-size Zeroscope UNet3D forward and print a per-kernel breakdown (torch.profiler / CUPTI).
usage: python tools/time_unet.py [B F H W] [--profile]"""
import sys
import time

import torch

sys.path.insert(0, ".")
import b2v  # noqa: E402
from b2v.unet import UNet3DConditionModel  # noqa: E402
from b2v.pipeline import random_state_dict  # noqa: E402
from b2v.unet import ZEROSCOPE_CONFIG  # noqa: E402


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    B, F, H, W = (int(a) for a in args) if len(args) == 4 else (2, 16, 72, 128)
    dev = torch.device("cuda:0")
    t0 = time.time()
    net = UNet3DConditionModel(ZEROSCOPE_CONFIG, device=dev)
    net.load_state_dict(random_state_dict(net.config.__dict__, dev))
    print("weights loaded in %.1fs, params %d" % (time.time() - t0, net.param_count()), flush=True)
    x = torch.randn(B, 4, F, H, W, device=dev, dtype=torch.bfloat16)
    ehs = torch.randn(B, 77, 1024, device=dev, dtype=torch.bfloat16)
    print("workspace %.2f GB, flops %.2f TF" % (net.workspace_bytes(B, F, H, W) / 1e9, net.flops(B, F, H, W) / 1e12),
          flush=True)
    out = net(x, 501, encoder_hidden_states=ehs).sample
    torch.cuda.synchronize()
    print("finite:", bool(torch.isfinite(out.float()).all()), "absmax", float(out.float().abs().max()), flush=True)
    for _ in range(2):
        net(x, 501, encoder_hidden_states=ehs)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 3
    e0.record()
    for _ in range(n):
        net(x, 501, encoder_hidden_states=ehs)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    print("forward %.2f ms  -> %.1f TFLOP/s" % (ms, net.flops(B, F, H, W) / ms / 1e9), flush=True)
    if "--profile" in sys.argv:
        from torch.profiler import profile, ProfilerActivity
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            net(x, 501, encoder_hidden_states=ehs)
            torch.cuda.synchronize()
        print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=30, max_name_column_width=90))


if __name__ == "__main__":
    main()
