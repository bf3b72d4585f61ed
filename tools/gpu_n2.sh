#!/bin/bash
# GPU box with N >= 2 GPUs: multi-GPU parity test, the strategy CLI under torchrun, bench.py at N. usage: tools/gpu_n2.sh <tag> [N]
TAG=${1:-rXX}; N=${2:-2}; O=gpurun_out/$TAG; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
PKG=$(ls -d decentralised-*_b200)
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q -s > $O/pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -5 $O/pytest_multi.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 \
  "$PKG/strategies/fsdp_chunked_coherent.py" --num_frames 24 --steps 2 --out_csv $O/strategy_n$N.csv > $O/strategy_n$N.log 2>&1; echo "strategy rc=$?"; tail -3 $O/strategy_n$N.log; cat $O/strategy_n$N.csv
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 \
  bench.py --gpus $N --steps 3 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench rc=$?"; tail -c 1500 $O/bench_n$N.json
