#!/usr/bin/env python3
"""Drop-in for Distribution/strategies/chunk_only.py: pure frame-window chunking, every rank holds the full model,
round-robin windows from a shared seed-0 noise tensor, end-of-run exchange + ramp blend. Same flags, same CSV row."""
import argparse
import logging

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
