#!/usr/bin/env python3
"""Drop-in for Distribution/strategies/fsdp_chunked.py: the chunk_only planner (ov = min(overlap, cs // 3)) without
context injection; ``--mode`` / ``--context_weight`` are accepted and ignored like the reference (:340-341)."""
import argparse
import logging

from .common import WindowedDiffuser, add_common_args, run_cli


class DistributedVideoDiffuser(WindowedDiffuser):
    variant = "chunk_only"
    mode_name = "split"


def parse_args(argv=None):
    p = add_common_args(argparse.ArgumentParser(), num_frames=25)
    p.add_argument("--mode", default="split")
    p.add_argument("--context_weight", type=float, default=0.35)
    return p.parse_args(argv)


def main(argv=None):
    logging.basicConfig(level=logging.INFO, format="%(asctime)s %(levelname)s: %(message)s")
    cfg = parse_args(argv)
    return run_cli(DistributedVideoDiffuser, cfg, cfg.mode)


if __name__ == "__main__":
    main()
