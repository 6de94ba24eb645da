#!/bin/bash
# ncu --set full of chosen GEMM launches of a short encoder forward.  Launch order of gemm_bf16_kernel in prof_case:
# 0 conv0, 1 conv1 [LN producer], 2 qkv [LN consumer], 3 out_proj [producer, residual], 4 fc1 [consumer], 5 fc2 [producer]
tag=${1:-r02}; skip=${2:-3}; count=${3:-1}
mkdir -p gpurun_out
CMD="python tools/prof_case.py encoder --iters 1"
$CMD > gpurun_out/${tag}_g_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/${tag}_g_plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_bf16_kernel -s $skip -c $count -f -o gpurun_out/${tag}_gemm $CMD > gpurun_out/${tag}_ncu_gemm.log 2>&1; echo "gemm ncu exit $?"
ls -la gpurun_out/${tag}_gemm.ncu-rep
