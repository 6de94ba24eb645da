#!/bin/bash
# ncu --set full capture of one logmel_kernel launch (prof_case frontend workload: 128 clips in one launch; the second launch is captured)
tag=${1:-r02}
mkdir -p gpurun_out
CMD="python tools/prof_case.py frontend"
$CMD > gpurun_out/${tag}_prof_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/${tag}_prof_plain.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:logmel_kernel -s 1 -c 1 -f -o gpurun_out/${tag}_logmel $CMD > gpurun_out/${tag}_ncu_logmel.log 2>&1; echo "logmel ncu exit $?"
tail -3 gpurun_out/${tag}_ncu_logmel.log
ls -la gpurun_out/*.ncu-rep
