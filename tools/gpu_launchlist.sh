#!/bin/bash
# per-launch device times of one short encoder forward (prof_case: whisper-small, 32 clips, 2 layers), LayerNorm fused and not;
# cold-cache, serialised: compare shares
tag=${1:-r02}
mkdir -p gpurun_out
for f in 1 0; do
  CMD="python tools/prof_case.py encoder --iters 2"
  YOHO_LN_FUSE=$f $CMD > gpurun_out/${tag}_ll_plain_f$f.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/${tag}_ll_plain_f$f.log; exit 1; }
  YOHO_LN_FUSE=$f timeout 600 ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_f$f.csv $CMD > gpurun_out/${tag}_ll_ncu_f$f.log 2>&1
  echo "ncu exit $?"
done
