#!/bin/bash
# GPU box: the round's measurement pass. usage: tools/gpu_round.sh <tag>   (from the repo root)
#   1. -m gpu tests  2. smoke()  3. bench.py (N=1, default K/W)  4. per-kernel timing of one UNet forward
#   5. ncu launch list of the bench command  6. ncu --set full of the dominant kernels (one launch each)
TAG=${1:-rXX}
O=gpurun_out/$TAG
mkdir -p "$O"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > "$O/nvsmi.txt" 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > "$O/pytest_gpu.log" 2>&1; echo "pytest rc=$?" | tee -a "$O/pytest_gpu.log"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > "$O/smoke.log" 2>&1; echo "smoke rc=$?" | tee -a "$O/smoke.log"
timeout 900 python bench.py > "$O/bench_n1.json" 2> "$O/bench_n1.err"; echo "bench rc=$?" | tee -a "$O/bench_n1.err"
timeout 300 python tools/time_unet.py 2 16 72 128 --profile > "$O/time_unet_16f.log" 2>&1; echo "time_unet rc=$?"
if [ -z "$SKIP_NCU" ]; then
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file "$O/launches.csv" \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline > "$O/ncu_bench.log" 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_kernel -c 1 -s 1 -f -o "$O/attn_full" \
  python tools/ncu_targets.py attn > "$O/ncu_attn.log" 2>&1; echo "ncu attn rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -f -o "$O/gemm_full" \
  python tools/ncu_targets.py gemm > "$O/ncu_gemm.log" 2>&1; echo "ncu gemm rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"gn_|layer_norm" -f -o "$O/norm_full" \
  python tools/ncu_targets.py norm > "$O/ncu_norm.log" 2>&1; echo "ncu norm rc=$?"
fi
ls -la "$O"
