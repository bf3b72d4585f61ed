"""Strategy entry points with the reference's CLI (Distribution/full_experiments_ZeroscopeXL.sh:25-87):
``torchrun --nproc_per_node=W strategies/chunk_only.py ...`` (repo-root shims) or ``-m b2v.strategies.chunk_only``."""
