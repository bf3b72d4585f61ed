"""GPU box: A/B two builds of libb2v on a fixed list of GEMM shapes under in-forward conditions: L2 flushed before every
timed launch (a 512 MB write), residual either a separate tensor or IN PLACE (out == residual, as the UNet does).
usage: python tools/gemm_ab.py [libA.so libB.so ...]     (default: the in-tree build only)
Each library runs in its own child process (B2V_LIB)."""
import math
import os
import subprocess
import sys

SHAPES = [  # M, N, K, taps, flags, residual
    (294912, 2560, 320, 1, 1, 0), (294912, 320, 320, 1, 0, 1), (294912, 960, 320, 1, 0, 0),
    (73728, 5120, 640, 1, 1, 0), (73728, 640, 640, 1, 0, 1), (73728, 1920, 640, 1, 0, 0),
    (294912, 320, 1280, 1, 0, 1), (73728, 640, 2560, 1, 0, 1), (18432, 1280, 1280, 1, 0, 1),
    (18432, 3840, 1280, 1, 0, 0), (18432, 1280, 5120, 1, 0, 1), (18432, 10240, 1280, 1, 1, 0),
    (294912, 320, 320, 3, 0, 0), (73728, 640, 640, 3, 0, 0), (18432, 1280, 1280, 3, 0, 0), (4608, 1280, 1280, 3, 0, 0),
    (4608, 1280, 1280, 1, 0, 1),
]

if os.environ.get("B2V_AB_CHILD"):
    import torch
    sys.path.insert(0, ".")
    import b2v  # noqa
    from b2v import ops
    dev = torch.device("cuda:0")
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)

    def timeit(fn, n=8):
        fn()
        tot = 0.0
        for _ in range(n):
            flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            tot += e0.elapsed_time(e1)
        return tot / n

    for M, N, K, taps, flags, r in SHAPES:
        a = torch.randn(M, K, device=dev).bfloat16()
        w = (torch.randn(N, taps * K, device=dev) / math.sqrt(taps * K)).bfloat16()
        bias = torch.zeros(N, device=dev)
        n_out = N // 2 if flags & 1 else N
        out = torch.randn(M, n_out, device=dev).bfloat16()
        res = torch.randn(M, n_out, device=dev).bfloat16() if r else None
        kw = dict(bias=bias, taps=taps, flags=flags)
        if taps == 3:
            F = 16
            kw.update(tap_row_stride=M // 2 // F, rows_per_batch=M // 2)
        ms_sep = timeit(lambda: ops.gemm(a, w, residual=res, out=out, **kw))
        ms_inp = timeit(lambda: ops.gemm(a, w, residual=out, out=out, **kw)) if r else float("nan")
        fl = 2.0 * M * N * K * taps
        by = 2.0 * (M * K + N * K * taps + M * n_out * (2 if r else 1))
        print("RES %d %d %d %d %d %d %.4f %.4f %.0f %.0f" % (M, N, K, taps, flags, r, ms_sep, ms_inp, fl / ms_sep / 1e9,
                                                             by / ms_sep / 1e6), flush=True)
    sys.exit(0)

libs = sys.argv[1:] or [""]
results = []
for lib in libs:
    env = dict(os.environ, B2V_AB_CHILD="1")
    if lib:
        env["B2V_LIB"] = os.path.abspath(lib)
    r = subprocess.run([sys.executable, __file__], env=env, capture_output=True, text=True)
    rows = [l.split()[1:] for l in r.stdout.split("\n") if l.startswith("RES ")]
    if not rows:
        print("no results for", lib, r.stderr[-2000:])
    results.append(rows)
print("libs:", libs)
print("%-44s" % "shape (M N K taps flags res)" + "".join("  | %8s %8s %7s %6s" % ("ms", "ms inpl", "TFLOP/s", "GB/s") for _ in libs))
for i in range(len(SHAPES)):
    line = "%-44s" % " ".join(results[0][i][:6]) if results[0] else ""
    for rows in results:
        if i < len(rows):
            line += "  | %8s %8s %7s %6s" % tuple(rows[i][6:10])
    print(line)
