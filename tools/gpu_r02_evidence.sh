#!/bin/bash
# Round-2 ncu evidence at the bench's batch (256 clips): `ncu --set full` of the first layer's GEMMs (conv0, conv1, qkv,
# out_proj, fc1, fc2), one attention launch, one log-mel launch and the final LayerNorm.  Reports -> gpurun_out/r02_*.ncu-rep
tag=${1:-r02}; clips=${2:-256}
mkdir -p gpurun_out
CMD="python tools/prof_case.py --clips $clips --iters 1"
$CMD > gpurun_out/${tag}_prof_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/${tag}_prof_plain.log; exit 1; }
tail -1 gpurun_out/${tag}_prof_plain.log
NCU="ncu --set full --clock-control none --import-source on"
timeout 900 $NCU -k regex:gemm_bf16 -c 6 -f -o gpurun_out/${tag}_gemm $CMD > gpurun_out/${tag}_ncu_gemm.log 2>&1; echo "gemm ncu exit $?"
timeout 600 $NCU -k regex:attention -c 1 -f -o gpurun_out/${tag}_attn $CMD > gpurun_out/${tag}_ncu_attn.log 2>&1; echo "attn ncu exit $?"
timeout 600 $NCU -k regex:logmel_kernel -s 1 -c 1 -f -o gpurun_out/${tag}_logmel $CMD > gpurun_out/${tag}_ncu_logmel.log 2>&1; echo "logmel ncu exit $?"
timeout 600 $NCU -k regex:layernorm -c 1 -f -o gpurun_out/${tag}_ln $CMD > gpurun_out/${tag}_ncu_ln.log 2>&1; echo "ln ncu exit $?"
ls -la gpurun_out/${tag}_*.ncu-rep
