#!/bin/bash
# Runs the -m gpu tests in separate processes (a trapped kernel poisons its CUDA context; the next group still runs).
# usage: tools/gpu_groups.sh [outdir]   (from the repo root, on the GPU box)
OUT=${1:-gpurun_out/groups}
mkdir -p "$OUT"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > "$OUT/nvsmi.txt" 2>&1
run() { # name, file, -k expr
  local name=$1 file=$2 k=$3
  timeout 600 python -m pytest "$file" -m gpu -q --no-header -p no:cacheprovider -k "$k" -s > "$OUT/$name.log" 2>&1
  echo "$name exit=$? $(tail -1 "$OUT/$name.log")"
}
run proof      tests/test_gpu_proof_kernels.py ""
run gemm_plain tests/test_gpu_ops.py "gemm_plain or gemm_residual"
run gemm_geglu tests/test_gpu_ops.py "gemm_geglu"
run gemm_taps  tests/test_gpu_ops.py "temporal_taps"
run conv_m1    tests/test_gpu_ops.py "conv3x3 and explicit"
run conv_m0    tests/test_gpu_ops.py "conv3x3 and tmaim2col"
run conv_epi   tests/test_gpu_ops.py "conv3x3_rowbias"
run norms      tests/test_gpu_ops.py "group_norm or layer_norm"
run attn       tests/test_gpu_ops.py "test_attention"
run tattn      tests/test_gpu_ops.py "temporal_attention"
run unet       tests/test_gpu_unet.py ""
run proof_api  tests/test_gpu_proof_api.py ""
run strategy   tests/test_gpu_strategy.py ""
