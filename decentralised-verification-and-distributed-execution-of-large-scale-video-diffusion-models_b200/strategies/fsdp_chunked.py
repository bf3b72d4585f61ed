#!/usr/bin/env python3
"""Drop-in for Distribution/strategies/fsdp_chunked.py: the chunk_only planner (ov = min(overlap, cs // 3)) without
context injection; ``--mode`` / ``--context_weight`` are accepted and ignored like the reference (:340-341)."""
import argparse
import logging

if __package__ in (None, ""):  # run by path like the reference: torchrun --nproc_per_node=W <pkg>/strategies/fsdp_chunked.py ...
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
    import b2v  # noqa: F401
    from b2v.strategies.common import WindowedDiffuser, add_common_args, run_cli
else:
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
