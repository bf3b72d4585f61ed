#!/usr/bin/env python3
"""Drop-in for Distribution/strategies/chunk_only.py: pure frame-window chunking, every rank holds the full model,
round-robin windows from a shared seed-0 noise tensor, end-of-run exchange + ramp blend. Same flags, same CSV row."""
import argparse
import logging

if __package__ in (None, ""):  # run by path like the reference: torchrun --nproc_per_node=W <pkg>/strategies/chunk_only.py ...
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
    import b2v  # noqa: F401
    from b2v.strategies.common import WindowedDiffuser, add_common_args, run_cli
else:
    from .common import WindowedDiffuser, add_common_args, run_cli


class ChunkedVideoDiffuser(WindowedDiffuser):
    variant = "chunk_only"
    mode_name = "chunk"


def parse_args(argv=None):
    return add_common_args(argparse.ArgumentParser()).parse_args(argv)


def main(argv=None):
    logging.basicConfig(level=logging.INFO, format="%(asctime)s %(levelname)s: %(message)s")
    cfg = parse_args(argv)
    return run_cli(ChunkedVideoDiffuser, cfg, "chunk")


if __name__ == "__main__":
    main()
