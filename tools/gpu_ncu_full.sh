#!/bin/bash
# ncu --set full captures of the three hot kernels on the short prof_case workload.
# usage: tools/gpu_ncu_full.sh [tag]    reports -> gpurun_out/<tag>_{gemm,attn,logmel}.ncu-rep
tag=${1:-r01}
mkdir -p gpurun_out
CMD="python tools/prof_case.py"
$CMD > gpurun_out/${tag}_prof_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/${tag}_prof_plain.log; exit 1; }
NCU="ncu --set full --clock-control none --import-source on"
# second iteration's launches (skip the first iteration = warm-up)
timeout 600 $NCU -k regex:gemm_bf16 -s 12 -c 6 -f -o gpurun_out/${tag}_gemm $CMD > gpurun_out/${tag}_ncu_gemm.log 2>&1; echo "gemm ncu exit $?"
timeout 600 $NCU -k regex:attention -s 2 -c 1 -f -o gpurun_out/${tag}_attn $CMD > gpurun_out/${tag}_ncu_attn.log 2>&1; echo "attn ncu exit $?"
timeout 600 $NCU -k regex:logmel_kernel -s 1 -c 1 -f -o gpurun_out/${tag}_logmel $CMD > gpurun_out/${tag}_ncu_logmel.log 2>&1; echo "logmel ncu exit $?"
timeout 600 $NCU -k regex:layernorm -s 5 -c 1 -f -o gpurun_out/${tag}_ln $CMD > gpurun_out/${tag}_ncu_ln.log 2>&1; echo "ln ncu exit $?"
ls -la gpurun_out/*.ncu-rep
