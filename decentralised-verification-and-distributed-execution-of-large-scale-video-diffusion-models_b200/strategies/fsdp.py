#!/usr/bin/env python3
"""Drop-in for Distribution/strategies/fsdp.py (FSDP-only baseline): the full latent on every rank, no chunking,
guidance hard-coded to 7.5 (fsdp.py:151), no --guidance_scale/--chunk_size/--overlap/--fps flags, legacy ``--out``.
CLI parity only: the parameter sharding it benchmarks does not exist on replicated weights."""
import argparse
import logging

if __package__ in (None, ""):  # run by path like the reference: torchrun --nproc_per_node=W <pkg>/strategies/fsdp.py ...
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
    import b2v  # noqa: F401
    from b2v.strategies.common import WindowedDiffuser, add_common_args, run_cli
else:
    from .common import WindowedDiffuser, add_common_args, run_cli


class FSDPBenchmark(WindowedDiffuser):
    variant = "coherent"
    mode_name = "fsdp"

    def __init__(self, cfg, **kw):
        cfg.no_chunking, cfg.use_ctx, cfg.guidance_scale = True, False, 7.5
        cfg.chunk_size, cfg.overlap, cfg.fps = 0, 0, 8
        super().__init__(cfg, **kw)

    run = WindowedDiffuser.__call__


def parse_args(argv=None):
    p = add_common_args(argparse.ArgumentParser(), num_frames=25, chunking=False)
    # fsdp.py:278: a dead legacy flag in the reference (default fsdp_results.npy); the CSV always goes to --out_csv
    p.add_argument("--out", default="fsdp_results.npy", help="legacy flag, parsed and ignored like the reference does")
    p.set_defaults(prompt="a rocket flying off into space, hd, detailed")  # fsdp.py's own default prompt
    return p.parse_args(argv)


def main(argv=None):
    logging.basicConfig(level=logging.INFO, format="%(asctime)s %(levelname)s: %(message)s")
    cfg = parse_args(argv)
    return run_cli(FSDPBenchmark, cfg, "fsdp")


if __name__ == "__main__":
    main()
