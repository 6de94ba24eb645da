#!/bin/bash
# Run each diagnostic stage in its own process under a timeout; logs into gpurun_out/.
# usage: tools/gpu_debug.sh [stage ...]   (default: all)
mkdir -p gpurun_out
stages=("$@")
[ ${#stages[@]} -eq 0 ] && stages=(frontend fp32 gemm attention bf16 perf)
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/debug_gpu.txt 2>&1
for s in "${stages[@]}"; do
  echo "===== stage $s ====="
  timeout 240 python tools/gpu_debug.py "$s" 2>&1 | tee "gpurun_out/debug_${s}.log" | tail -60
  echo "(exit ${PIPESTATUS[0]})"
done
